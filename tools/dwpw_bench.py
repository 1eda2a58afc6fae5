import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rtseg_b200 import modules, nets, ops
from tools.layer_bench import nhwc_rand, time_ms, report
b = 32
cases = [("DS 128->128 @128x256 (classifier)", modules.DSConvBNAct(128, 128, 3, 1), (b, 128, 128, 256), 128),
         ("IR tail 64->384->64 @64x128 (+x)", nets.InvertedResidual(64, 64, 1), (b, 64, 64, 128), 64)]
for name, m, shp, cout in cases:
    m = m.eval().cuda(); x = nhwc_rand(shp)
    with torch.no_grad():
        for fused in (False, True):
            ops.USE_DWPW_FUSION = fused
            ms = time_ms(lambda: m(x), 10)
            report(f"{name} fused={fused}", ms, (x.numel() + x.numel() // shp[1] * cout) * 2)
ops.USE_DWPW_FUSION = True
