timeout 600 python -m pytest tests/test_gpu_pair.py -q -p no:cacheprovider -x > gpurun_out/pytest7.log 2>&1; echo "pytest pair rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed\|^E " gpurun_out/pytest7.log | head -30
timeout 600 python tools/layer_bench.py --only pair > gpurun_out/layers7.txt 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers7.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_pair -s 4 -c 1 -o gpurun_out/pair_c64_v4 -f python tools/layer_bench.py --only "c64 pair 3x1->1x3 (one" > gpurun_out/ncu7.log 2>&1; echo "ncu rc=$?"
timeout 900 python bench.py --single --no-library-baseline --no-cpu-baseline > gpurun_out/bench7.json 2> gpurun_out/bench7.err; echo "bench rc=$?"; cat gpurun_out/bench7.json | cut -c1-300; tail -5 gpurun_out/bench7.err
