timeout 600 python -m pytest tests/test_gpu_pair.py -q -p no:cacheprovider -x > gpurun_out/pytest5.log 2>&1; echo "pytest pair rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed\|^E " gpurun_out/pytest5.log | head -30
timeout 600 python tools/layer_bench.py --only pair > gpurun_out/layers5.txt 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers5.txt
for th in 2 8; do RTSEG_PAIR_TH=$th timeout 600 python tools/layer_bench.py --only "c64 pair" 2>&1 | sed "s/^/TH=$th /"; done
RTSEG_PAIR_TH=16 timeout 600 python tools/layer_bench.py --only "c16 pair" 2>&1 | sed "s/^/TH=16 /"
RTSEG_PAIR_TH=4 timeout 600 python tools/layer_bench.py --only "c16 pair" 2>&1 | sed "s/^/TH=4 /"
timeout 900 python bench.py --single --no-library-baseline --no-cpu-baseline > gpurun_out/bench5.json 2> gpurun_out/bench5.err; echo "bench rc=$?"; cat gpurun_out/bench5.json | cut -c1-400; tail -5 gpurun_out/bench5.err
