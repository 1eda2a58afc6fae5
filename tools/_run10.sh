timeout 600 python -m pytest tests/test_gpu_pair.py -q -p no:cacheprovider -x > gpurun_out/pytest10.log 2>&1; echo "pytest pair rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed\|^E " gpurun_out/pytest10.log | head -30
timeout 600 python tools/layer_bench.py --only pair > gpurun_out/layers10.txt 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers10.txt
timeout 900 python bench.py --single --no-library-baseline --no-cpu-baseline > gpurun_out/bench10.json 2> gpurun_out/bench10.err; echo "bench rc=$?"; cat gpurun_out/bench10.json | cut -c1-300; tail -5 gpurun_out/bench10.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench10.json')); print(d["parity_check"])
PY
