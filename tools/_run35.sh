cat > /tmp/one.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import ops
x = torch.randn((32, 32, 128, 256), device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
for _ in range(3): ops.upsample_argmax(x, (1024, 2048), True, num_classes=19)
torch.cuda.synchronize()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:upsample_argmax --launch-skip 2 --launch-count 1 -o gpurun_out/head_r02 -f python /tmp/one.py > gpurun_out/ncu35.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu35.log
