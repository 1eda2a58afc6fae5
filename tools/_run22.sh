timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_steady_state.py tests/test_gpu_nets.py -q -p no:cacheprovider -x > gpurun_out/pytest22.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest22.log; grep -n "^E " gpurun_out/pytest22.log | head
timeout 600 python tools/layer_bench.py --only "c128 3x1" 2>&1 | grep -v "d1 "
timeout 900 python bench.py --no-library-baseline --no-cpu-baseline --single 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('erfnet', d['value'], d['ms_per_step'], d['roofline']['families'])"
