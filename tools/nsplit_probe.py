import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rtseg_b200 import modules, ops
from tools.layer_bench import nhwc_rand, time_ms
b, c, h, w = 16, 128, 128, 256
x = nhwc_rand((b, c, h, w)); out = ops.empty_nhwc(b, c, h, w, torch.bfloat16, "cuda")
for k, d in (((1, 3), 1), ((3, 1), 1), ((3, 1), 16)):
    full = modules.ConvBNAct(c, c, k, 1, d).eval().cuda()
    half = modules.ConvBNAct(c, c // 2, k, 1, d).eval().cuda()
    with torch.no_grad():
        t_full = time_ms(lambda: full(x, out=out), 20)
        t_half = time_ms(lambda: half(x, out=out, y_coff=0), 20)
        t_two = time_ms(lambda: (half(x, out=out, y_coff=0), half(x, out=out, y_coff=64)), 20)
    print(f"k={k} d={d}: full {t_full:.4f} ms, one half {t_half:.4f} ms, two halves back to back {t_two:.4f} ms", flush=True)
