timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_cfg4.py -q -p no:cacheprovider -x -k "depthwise or gather or nets" 2>&1 | tail -3
python tools/layer_bench.py --set net:bisenetv2 --reps 5 --only left_branch.1 2>&1 | cut -c1-150
timeout 900 python bench.py --workload bisenetv2 --no-library-baseline --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('bisenetv2', d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel'][-120:], d['roofline']['frac'])"
