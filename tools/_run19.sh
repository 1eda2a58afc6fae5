timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_steady_state.py -q -p no:cacheprovider -x -k "tensor_core or steady or residual or deconv or nan or dab or ss_nbt" > gpurun_out/pytest19.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest19.log; grep -n "^E " gpurun_out/pytest19.log | head
timeout 600 python tools/layer_bench.py --only c128 > gpurun_out/layers19.txt 2>&1; cat gpurun_out/layers19.txt
RTSEG_TC_DIRECT=0 timeout 600 python tools/layer_bench.py --only "c128 1x3 d1" 2>&1 | sed "s/^/nodirect /"
timeout 900 python bench.py --no-library-baseline --no-cpu-baseline > gpurun_out/bench19.json 2> gpurun_out/bench19.err; python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench19.json').read().splitlines() if l.startswith('{')][-1])
print('erfnet', d['value'], d['ms_per_step'], d['roofline']['families']); f=d['workloads']['fastscnn']; print('fastscnn', f['value'], f['ms_per_step'])
PY
tail -3 gpurun_out/bench19.err
