import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, rtseg_b200
from rtseg_b200 import nets
torch.manual_seed(0)
m = rtseg_b200.FastSCNN(num_class=19).eval().cuda().set_act_dtype(torch.bfloat16)
x = torch.randint(0, 256, (32, 1024, 2048, 3), dtype=torch.uint8, device="cuda")
ltd = m.learning_to_downsample if hasattr(m, "learning_to_downsample") else None
first = None
for name, mod in m.named_modules():
    if name.endswith("conv") or "downsample" in name:
        pass
import time
from rtseg_b200 import ops
# time the whole first conv via the model's own stem call path: run the model, take kernel time by events around repeated forward of only the first layer
conv0 = [mod for n, mod in m.named_modules() if hasattr(mod, "forward") and n.count(".") == 1][0]
print(type(conv0).__name__)
xv = ops.image_view(x)
print(type(xv), getattr(xv, "shape", None))
def run(): return conv0(xv)
for _ in range(3): run()
torch.cuda.synchronize()
fl = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for _ in range(10):
    fl.zero_()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record(); run(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
print("stem us", sorted(ts)[len(ts)//2] * 1000)
