timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_nets.py -q -p no:cacheprovider -x -k "deconv or labels or argmax or erfnet" 2>&1 | tail -2
python tools/layer_bench.py --set erfnet --reps 10 --only "deconv 16" 2>&1 | cut -c1-130
timeout 900 python bench.py --no-library-baseline --no-cpu-baseline --single 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('erfnet', d['value'], d['ms_per_step'], d['parity_check']['argmax_agree'])"
