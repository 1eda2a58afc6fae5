cap() { # name regex script
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $4 --launch-count 1 -o gpurun_out/$1 -f python $3 > gpurun_out/ncu_$1.log 2>&1; echo "$1 rc=$?"; }
cat > /tmp/a.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand
m = modules.DSConvBNAct(128, 128, 3, 1).eval().cuda(); x = nhwc_rand((32, 128, 128, 256))
with torch.no_grad():
    for _ in range(3): m(x)
torch.cuda.synchronize()
PY
cat > /tmp/b.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import ops
x = torch.randn((32, 32, 128, 256), device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
for _ in range(3): ops.upsample_argmax(x, (1024, 2048), True, num_classes=19)
torch.cuda.synchronize()
PY
cat > /tmp/c.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch, rtseg_b200
m = rtseg_b200.FastSCNN(19).eval().cuda().set_act_dtype(torch.bfloat16)
x = torch.randint(0, 256, (32, 1024, 2048, 3), dtype=torch.uint8, device="cuda")
with torch.no_grad():
    for _ in range(3): m.learning_to_downsample(x)
torch.cuda.synchronize()
PY
cat > /tmp/d.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand
mod = modules.DeConvBNAct(16, 19).eval().cuda()
x = nhwc_rand((16, 16, 512, 1024)); labels = torch.empty((16, 1024, 2048), dtype=torch.uint8, device="cuda")
with torch.no_grad():
    for _ in range(3): mod(x, labels=labels, want_logits=False)
torch.cuda.synchronize()
PY
cat > /tmp/e.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand
m = modules.ConvBNAct(128, 128, (3, 1), 1, 16).eval().cuda(); x = nhwc_rand((16, 128, 128, 256))
with torch.no_grad():
    for _ in range(3): m(x)
torch.cuda.synchronize()
PY
cap dwpw_ds128_r02 dwpw /tmp/a.py 2
cap head_argmax_r02 upsample_argmax /tmp/b.py 2
cap stem_u8_r02 stem_tc /tmp/c.py 2
cap labels_head_r02 conv_tc /tmp/d.py 2
cap conv_tc_c128_3x1_d16_r02 conv_tc /tmp/e.py 2
