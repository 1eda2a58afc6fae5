timeout 600 python -m pytest tests/test_gpu_dwpw.py -q -p no:cacheprovider -x 2>&1 | tail -2
timeout 900 python bench.py --workload fastscnn --no-library-baseline --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][-1]); print('fastscnn', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
cat > /tmp/one.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch, rtseg_b200
m = rtseg_b200.FastSCNN(19).eval().cuda().set_act_dtype(torch.bfloat16)
x = torch.randint(0, 256, (32, 1024, 2048, 3), dtype=torch.uint8, device="cuda")
with torch.no_grad():
    for _ in range(3): m.learning_to_downsample(x)
torch.cuda.synchronize()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stem_tc --launch-skip 2 --launch-count 1 -o gpurun_out/stem_u8_r02 -f python /tmp/one.py > gpurun_out/ncu33.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu33.log
