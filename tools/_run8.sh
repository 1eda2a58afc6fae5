timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_steady_state.py -q -p no:cacheprovider -x -k "tensor_core or steady or residual or deconv" > gpurun_out/pytest8.log 2>&1; echo "pytest rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed\|^E " gpurun_out/pytest8.log | head -30
timeout 600 python tools/layer_bench.py --only c128 > gpurun_out/layers8.txt 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers8.txt
RTSEG_TC_SPLIT=0 timeout 600 python tools/layer_bench.py --only c128 2>&1 | sed "s/^/nosplit /" | head -8
timeout 900 python bench.py --single --no-library-baseline --no-cpu-baseline > gpurun_out/bench8.json 2> gpurun_out/bench8.err; echo "bench rc=$?"; cat gpurun_out/bench8.json | cut -c1-300; tail -5 gpurun_out/bench8.err
