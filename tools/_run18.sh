timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_steady_state.py -q -p no:cacheprovider -x -k "tensor_core or steady or residual or deconv or nan" > gpurun_out/pytest18.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest18.log
timeout 600 python tools/layer_bench.py --only c128 > gpurun_out/layers18.txt 2>&1; cat gpurun_out/layers18.txt
timeout 900 python bench.py --no-library-baseline --no-cpu-baseline > gpurun_out/bench18.json 2> gpurun_out/bench18.err; python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench18.json').read().splitlines() if l.startswith('{')][-1])
print('erfnet', d['value'], d['ms_per_step'], d['roofline']['families']); f=d['workloads']['fastscnn']; print('fastscnn', f['value'], f['ms_per_step'])
PY
