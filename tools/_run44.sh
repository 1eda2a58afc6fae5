timeout 600 python -m pytest tests/test_gpu_dwpw.py -q -p no:cacheprovider -x 2>&1 | tail -2
python tools/_dwpw_bench.py 2>&1 | cut -c1-150
timeout 900 python bench.py --workload fastscnn --no-library-baseline --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('fastscnn', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
