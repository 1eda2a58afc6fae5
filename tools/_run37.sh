timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_nets.py tests/test_zz_no_alloc.py -q -p no:cacheprovider -x -k "argmax or head or upsample or confmat or metric or miou" 2>&1 | tail -2
python tools/layer_bench.py --set head --batch 32 --reps 5 --only upsample 2>&1 | cut -c1-160 | head -2
timeout 600 python bench.py --workload head_miou --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('head_miou', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['parity_check'])"
