for k in 1 2; do
RTSEG_LIB_PATH=/root/repo/gpurun_scratch/ab/libhead.so timeout 600 python tools/layer_bench.py --only "c128 1x3 d1" 2>&1 | head -3 | sed "s/^/head     /"
RTSEG_TC_DIRECT=0 timeout 600 python tools/layer_bench.py --only "c128 1x3 d1" 2>&1 | head -3 | sed "s/^/nodirect /"
done
RTSEG_TC_DIRECT=0 timeout 900 python bench.py --no-library-baseline --no-cpu-baseline --single 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('nodirect erfnet', d['value'], d['ms_per_step'])"
RTSEG_LIB_PATH=/root/repo/gpurun_scratch/ab/libhead.so timeout 900 python bench.py --no-library-baseline --no-cpu-baseline --single 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('head erfnet', d['value'], d['ms_per_step'])"
