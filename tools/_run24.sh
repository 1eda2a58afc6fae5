for w in ddrnet bisenetv2; do
timeout 900 python bench.py --workload $w --no-library-baseline --steps 10 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; tail -3 gpurun_out/bench_$w.err | cut -c1-300
python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_$w.json').read().splitlines() if l.startswith('{')][-1])
    print('$w', d['value'], d['ms_per_step'], 'e2e', d['e2e'], 'launches', d['gpu_launches']); print(d['parity_check']); print(d['roofline']['kernel'][:300], d['roofline']['frac']); print(d['cpu_baseline'])
except Exception as e: print('no json', e)
PY
done
