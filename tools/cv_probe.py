import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rtseg_b200 import modules
from tools.layer_bench import nhwc_rand, time_ms, report
for cin, cout, hw, s in [(128, 128, (128, 256), 1), (64, 128, (256, 512), 1), (128, 128, (256, 512), 2), (64, 64, (512, 1024), 1)]:
    m = modules.ConvBNAct(cin, cout, 3, s).eval().cuda()
    x = nhwc_rand((16, cin) + hw)
    with torch.no_grad():
        ms = time_ms(lambda: m(x), 10)
    fl = 16 * hw[0] * hw[1] // (s * s) * cin * cout * 9 * 2
    report(f"cv {cin}->{cout} s{s} @{hw[0]}x{hw[1]}  {fl / ms / 1e9:.0f} TF/s", ms, x.numel() * 2 + x.numel() // cin * cout * 2 // (s * s))
