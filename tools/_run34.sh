timeout 900 python -m pytest tests/test_gpu_uint8_input.py tests/test_gpu_kernels.py -q -p no:cacheprovider -x -k "u8 or uint8 or stem" 2>&1 | tail -2
for k in 0 1; do
RTSEG_STEM_U8_WORDS=$k timeout 900 python bench.py --workload fastscnn --no-library-baseline --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('fastscnn words=$k', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['parity_check']['argmax_agree'])"
done
