#!/usr/bin/env python
"""Per-op time breakdown of one forward pass of a net (CUDA events around every rtseg op call).

    python tools/net_profile.py --net dabnet [--batch 16] [--reps 3] [--top 40]

Each `rtseg_b200.ops` entry point is wrapped with a pair of events on the current stream; what is left
between the sum of those and the whole pass is torch glue (cat / mul / add).  Algorithmic GB/s for
convolutions = (input slice + output [+ residual]) bytes / time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import rtseg_b200  # noqa: E402,F401
from rtseg_b200 import nets, ops  # noqa: E402

WRAPPED = ["conv", "stem", "maxpool3x3s2", "upsample_bilinear", "upsample_argmax", "avgpool3x3s2", "channel_shuffle",
           "adaptive_avgpool", "to_nhwc", "argmax_channels"]


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--net", default="dabnet", choices=sorted(nets.MODEL_HUB))
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--height", type=int, default=1024)
    ap.add_argument("--width", type=int, default=2048)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--top", type=int, default=60)
    ap.add_argument("--fp32", action="store_true")
    args = ap.parse_args()

    torch.manual_seed(1)
    model = nets.MODEL_HUB[args.net](num_class=19).eval().cuda()
    if not args.fp32:
        model.set_act_dtype(torch.bfloat16)
    x = torch.randn((args.batch, 3, args.height, args.width), device="cuda")
    records = []
    stream = torch.cuda.current_stream()
    originals = {n: getattr(ops, n) for n in WRAPPED}

    def wrap(name, fn):
        def inner(*a, **kw):
            if name == "to_nhwc" and ops.is_nhwc(a[0]) and (len(a) < 2 or a[1] is None or a[1] == a[0].dtype):
                return fn(*a, **kw)  # no-op call
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            out = fn(*a, **kw)
            e1.record(stream)
            e1.synchronize()
            desc, nbytes = name, 0
            if name == "conv":
                f, xin = a[0], a[1]
                n, _, h, w = xin.shape
                ho, wo = ops.out_hw(f, h, w)
                es = xin.element_size()
                nbytes = (n * h * w * f.cin + n * ho * wo * f.cout * (2 if kw.get("res") is not None else 1)) * es
                desc = (f"conv {f.kind} {f.cin}->{f.cout} k{f.kh}x{f.kw} s{f.stride} d{f.dil} g{f.groups} @{h}x{w}"
                        f"{' +res' if kw.get('res') is not None else ''}{' slice' if xin.shape[1] != f.cin else ''}")
            else:
                t = [v for v in a if isinstance(v, torch.Tensor)]
                if t:
                    desc = f"{name} {tuple(t[0].shape)}"
                    nbytes = t[0].numel() * t[0].element_size()
                    if isinstance(out, torch.Tensor):
                        nbytes += out.numel() * out.element_size()
            records.append((desc, e0.elapsed_time(e1), nbytes))
            return out
        return inner

    with torch.no_grad():
        for _ in range(2):
            model.predict(x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.reps):
            model.predict(x)
        e1.record(stream)
        e1.synchronize()
        whole = e0.elapsed_time(e1) / args.reps
        for n in WRAPPED:
            setattr(ops, n, wrap(n, originals[n]))
        try:
            for _ in range(args.reps):
                model.predict(x)
        finally:
            for n in WRAPPED:
                setattr(ops, n, originals[n])
    agg = OrderedDict()
    for desc, ms, nb in records:
        a = agg.setdefault(desc, [0, 0.0, 0])
        a[0] += 1
        a[1] += ms
        a[2] += nb
    pk = peak()
    tot = sum(a[1] for a in agg.values()) / args.reps
    print(f"{args.net} batch {args.batch} {args.height}x{args.width} {'fp32' if args.fp32 else 'bf16'}: "
          f"whole pass {whole:.3f} ms ({args.batch / whole * 1e3:.1f} img/s); rtseg ops {tot:.3f} ms; "
          f"torch glue + launch gaps {whole - tot:.3f} ms")
    print(f"{'op':78s} {'calls':>5s} {'ms/pass':>9s} {'share':>6s} {'GB/s':>8s} {'of peak':>7s}")
    for desc, (cnt, ms, nb) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:args.top]:
        gbs = nb / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        print(f"{desc:78s} {cnt // args.reps:5d} {ms / args.reps:9.3f} {100 * ms / args.reps / whole:5.1f}% {gbs:8.0f} {gbs / pk:7.2f}")


if __name__ == "__main__":
    main()
