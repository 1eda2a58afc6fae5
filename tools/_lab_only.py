import sys, torch
sys.path.insert(0, '.')
import rtseg_b200
from rtseg_b200 import modules, ops
torch.manual_seed(0)
m = modules.DeConvBNAct(16, 19, act_type="relu").eval().cuda()
x = torch.randn((16, 16, 512, 1024), device='cuda').to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
labels = torch.empty((16, 1024, 2048), dtype=torch.uint8, device='cuda')
with torch.no_grad():
    for _ in range(3):
        m(x, labels=labels, want_logits=False)
torch.cuda.synchronize()
print(labels.float().mean().item())
