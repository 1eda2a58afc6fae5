import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand, time_ms, report
for c, hw, b in [(576, (32, 64), 32), (576, (32, 64), 128), (768, (32, 64), 32), (384, (64, 128), 32), (384, (128, 256), 32)]:
    m = modules.DWConvBNAct(c, c, 3, 1).eval().cuda()
    x = nhwc_rand((b, c) + hw)
    with torch.no_grad():
        ms = time_ms(lambda: m(x), 10)
    report(f"dw {c}@{hw[0]}x{hw[1]} b{b}", ms, x.numel() * 4)
