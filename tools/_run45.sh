python tools/_dwpw_bench.py 2>&1 | cut -c1-150 | grep "fused=True"
python tools/layer_bench.py --set erfnet --reps 10 --only "deconv 16" 2>&1 | cut -c1-130
timeout 900 python bench.py --no-library-baseline --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('erfnet', d['value'], d['ms_per_step']); f=d['workloads']['fastscnn']; print('fastscnn', f['value'], f['ms_per_step'])"
