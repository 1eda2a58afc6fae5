import sys, torch
sys.path.insert(0, '.')
from rtseg_b200 import ops
x = torch.randn((8, 19, 128, 256), device='cuda').to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
for _ in range(4):
    ops.upsample_argmax(x, (1024, 2048), True)
torch.cuda.synchronize()
