RTSEG_TC_SPLIT=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 4 -c 1 -o gpurun_out/c128_nosplit -f python tools/layer_bench.py --only "c128 3x1 d1" > gpurun_out/ncu9a.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 4 -c 1 -o gpurun_out/c128_split -f python tools/layer_bench.py --only "c128 3x1 d1" > gpurun_out/ncu9b.log 2>&1; echo "ncu rc=$?"
RTSEG_TC_SPLIT=0 timeout 900 python bench.py --single --no-library-baseline --no-cpu-baseline 2>/dev/null | cut -c1-200
