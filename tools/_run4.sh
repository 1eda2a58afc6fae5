set -x
timeout 600 python -m pytest tests/test_gpu_pair.py -q -rA -p no:cacheprovider -x > gpurun_out/pytest4.log 2>&1; echo "pytest pair rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed\|^E " gpurun_out/pytest4.log | head -30
timeout 600 python tools/layer_bench.py > gpurun_out/layers4.txt 2>&1; echo "layer_bench rc=$?"; head -30 gpurun_out/layers4.txt
timeout 900 python bench.py --single --no-library-baseline --no-cpu-baseline > gpurun_out/bench4.json 2> gpurun_out/bench4.err; echo "bench rc=$?"; cat gpurun_out/bench4.json | cut -c1-1500; tail -5 gpurun_out/bench4.err
