cat > /tmp/one.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand
mod = modules.DeConvBNAct(16, 19).eval().cuda()
x = nhwc_rand((16, 16, 512, 1024))
labels = torch.empty((16, 1024, 2048), dtype=torch.uint8, device="cuda")
with torch.no_grad():
    for _ in range(3): mod(x, labels=labels, want_logits=False)
torch.cuda.synchronize()
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc --launch-skip 2 --launch-count 1 -o gpurun_out/labels_head_r02 -f python /tmp/one.py > gpurun_out/ncu41.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu41.log
