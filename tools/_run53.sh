for k in 1 2; do
RTSEG_LIB_PATH=/root/repo/gpurun_scratch/ab/libhead.so python tools/layer_bench.py --set erfnet --reps 20 --only "deconv 16" 2>&1 | cut -c1-100 | sed "s/^/head /"
python tools/layer_bench.py --set erfnet --reps 20 --only "deconv 16" 2>&1 | cut -c1-100 | sed "s/^/new  /"
done
RTSEG_LIB_PATH=/root/repo/gpurun_scratch/ab/libhead.so timeout 900 python bench.py --no-library-baseline --no-cpu-baseline --single 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('head erfnet', d['value'], d['ms_per_step'])"
timeout 900 python bench.py --no-library-baseline --no-cpu-baseline --single 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('new  erfnet', d['value'], d['ms_per_step'])"
