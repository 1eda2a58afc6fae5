timeout 600 python -m pytest tests/test_gpu_pair.py -q -p no:cacheprovider -x > gpurun_out/pytest6.log 2>&1; echo "pytest pair rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed\|^E " gpurun_out/pytest6.log | head -30
timeout 600 python tools/layer_bench.py --only pair > gpurun_out/layers6.txt 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers6.txt
RTSEG_PAIR_TH=16 timeout 600 python tools/layer_bench.py --only "c16 pair" 2>&1 | sed "s/^/TH=16 /"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_pair -s 4 -c 1 -o gpurun_out/pair_c64 -f python tools/layer_bench.py --only "c64 pair 3x1->1x3 (one" > gpurun_out/ncu6.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu6.log
