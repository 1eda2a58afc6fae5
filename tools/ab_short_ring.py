#!/usr/bin/env python
"""A/B probe for the residual-ring ordering in conv_tc_kernel: residual convolutions whose mbarrier ring is shorter
than one unit's boxes, launched many times on the same inputs; counts launches whose output differs from the first.

  python tools/ab_short_ring.py [--tree DIR] [--reps 400]     # DIR: a checkout whose rtseg_b200 package to load
"""
import argparse
import json
import os
import sys


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tree", default=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    ap.add_argument("--reps", type=int, default=400)
    args = ap.parse_args()
    sys.path.insert(0, args.tree)
    import torch
    import rtseg_b200
    from rtseg_b200 import modules, ops

    cases = [(576, 96, 1, 1, "post", 19, 64, 128), (768, 128, 1, 1, "post", 10, 64, 128), (128, 128, 3, 2, "pre", 4, 256, 512),
             (384, 64, 1, 1, "post", 5, 128, 256)]
    out = {"tree": args.tree, "package": os.path.dirname(rtseg_b200.__file__), "reps": args.reps, "cases": []}
    torch.manual_seed(0)
    for cin, cout, k, stride, mode, n, h, w in cases:
        m = modules.ConvBNAct(cin, cout, k, stride, act_type="none" if mode == "post" else "relu").eval().cuda()
        x = torch.randn((n, cin, h, w), device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        r = torch.randn((n, cout, h // stride, w // stride), device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        rm = ops.RES_PRE if mode == "pre" else ops.RES_POST
        bad = 0
        err = None
        try:
            with torch.no_grad():
                first = m(x, res=r, res_mode=rm).clone()
                torch.cuda.synchronize()
                for i in range(args.reps):
                    y = m(x, res=r, res_mode=rm)
                    if i % 8 == 7:
                        bad += int(not torch.equal(y, first))
            torch.cuda.synchronize()
            flag = ops.tc_timeout_flag()
        except Exception as e:  # a device fault ends the probe
            err, flag = f"{type(e).__name__}: {str(e)[:200]}", None
        out["cases"].append({"conv": f"{cin}->{cout} k{k} s{stride} +res({mode})", "checked": args.reps // 8,
                             "differing_outputs": bad, "watchdog": flag, "error": err})
        if err:
            break
    print(json.dumps(out))


if __name__ == "__main__":
    main()
