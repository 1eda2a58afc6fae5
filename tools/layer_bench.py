#!/usr/bin/env python
"""Per-layer micro-benchmark of the hot kernels at BASELINE sizes (CUDA events, inputs > L2).

    python tools/layer_bench.py [--set erfnet|fastscnn|head] [--only SUBSTR] [--reps 10] [--batch 16]

Prints one line per layer: milliseconds, algorithmic GB/s (SURVEY §8d byte count) and fraction of the
measured HBM peak.  Also the ncu target: `ncu --set full -k regex:conv_tc ... python tools/layer_bench.py --only c64`.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import rtseg_b200  # noqa: E402
from rtseg_b200 import modules, nets, ops  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def nhwc_rand(shape, dtype=torch.bfloat16):
    return torch.randn(shape, device="cuda").to(dtype).contiguous(memory_format=torch.channels_last)


def time_ms(fn, reps):
    for _ in range(3):
        fn()
    s = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.synchronize()
    e0.record(s)
    for _ in range(reps):
        fn()
    e1.record(s)
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, ms, nbytes, flops=0.0):
    gbs = nbytes / ms / 1e6
    print(f"{name:44s} {ms:9.4f} ms  {gbs:8.1f} GB/s  {gbs / peak():6.3f} of HBM peak"
          + (f"  {flops / ms / 1e9:8.1f} TFLOP/s" if flops else ""), flush=True)


def erfnet_layers(batch, only, reps):
    cfgs = [("c64", 64, 256, 512, [1]), ("c128", 128, 128, 256, [1, 2, 4, 8, 16]), ("c16", 16, 512, 1024, [1])]
    for tag, c, h, w, dils in cfgs:
        x, r = nhwc_rand((batch, c, h, w)), nhwc_rand((batch, c, h, w))
        elems = batch * c * h * w
        for d in dils:
            blk = nets.NonBt1DBlock(c, d).eval().cuda()
            cases = [(f"{tag} 3x1 d1", lambda: blk.conv[0](x), 2), (f"{tag} 1x3 d1", lambda: blk.conv[1](x), 2),
                     (f"{tag} 3x1 d{d}", lambda: blk.conv[2](x), 2),
                     (f"{tag} 1x3 d{d} +res+bn+act", lambda: blk.conv[3](x, res=r, res_mode=ops.RES_PRE, bn=blk.bn_act[0],
                                                                         act=blk.bn_act[1].activation), 3)]
            for name, fn, passes in cases:
                if only and only not in name:
                    continue
                with torch.no_grad():
                    report(name, time_ms(fn, reps), passes * elems * 2 + 3 * c * c * 2, 2.0 * 3 * c * c * batch * h * w)
            if d == 1 and c in (16, 64):   # fused pairs: algorithmic bytes = read x (+ res), write y — no intermediate
                fa, fb = blk.conv[0]._folded()[0], blk.conv[1]._folded()[0]
                fc, fd = blk.conv[2]._folded()[0], blk.conv[3].folded(blk.bn_act[0], blk.bn_act[1].activation)[0]
                for name, fn, passes in [(f"{tag} pair 3x1->1x3 (one launch)", lambda: ops.conv_pair(fa, fb, x), 2),
                                         (f"{tag} pair 3x1->1x3 +res+bn+act (one launch)", lambda: ops.conv_pair(fc, fd, x, res=r), 3)]:
                    if only and only not in name:
                        continue
                    with torch.no_grad():
                        report(name, time_ms(fn, reps), passes * elems * 2 + 6 * c * c * 2, 2.0 * 6 * c * c * batch * h * w)
        del x, r
    # downsamplers and deconvs
    for name, mod, shape in [
        ("stem 3->16 (conv+pool) fp32 NCHW in", nets.InitialBlock(3, 16, "relu"), (batch, 3, 1024, 2048)),
        ("down 16->64", nets.InitialBlock(16, 64, "relu"), (batch, 16, 512, 1024)),
        ("down 64->128", nets.InitialBlock(64, 128, "relu"), (batch, 64, 256, 512)),
        ("deconv 128->64", modules.DeConvBNAct(128, 64), (batch, 128, 128, 256)),
        ("deconv 64->16", modules.DeConvBNAct(64, 16), (batch, 64, 256, 512)),
    ]:
        if only and only not in name:
            continue
        mod = mod.eval().cuda()
        if shape[1] == 3:
            x = torch.randn(shape, device="cuda")
            mod.act_dtype = torch.bfloat16
        else:
            x = nhwc_rand(shape)
        with torch.no_grad():
            y = mod(x)
            report(name, time_ms(lambda: mod(x), reps), x.numel() * x.element_size() + y.numel() * y.element_size())
        del x, y
    if not only or "deconv 16->19" in only or "argmax" in only:
        mod = modules.DeConvBNAct(16, 19).eval().cuda()
        x = nhwc_rand((batch, 16, 512, 1024))
        labels = torch.empty((batch, 1024, 2048), dtype=torch.uint8, device="cuda")
        with torch.no_grad():
            report("deconv 16->19 + argmax (labels only)", time_ms(lambda: mod(x, labels=labels, want_logits=False), reps),
                   x.numel() * 2 + labels.numel())


def fastscnn_layers(batch, only, reps):
    model = rtseg_b200.FastSCNN(19).eval().cuda().set_act_dtype(torch.bfloat16)
    shapes = {}
    hooks = []
    for name, m in model.named_modules():
        if isinstance(m, (modules.DWConvBNAct, modules.PWConvBNAct, modules.ConvBNAct, modules.Conv2d)):
            hooks.append(m.register_forward_pre_hook(
                lambda mod, inp, name=name: shapes.__setitem__(name, (mod, tuple(inp[0].shape)))))
    with torch.no_grad():
        model.lowres_logits(torch.randn((batch, 3, 1024, 2048), device="cuda"))
    for h in hooks:
        h.remove()
    for name, (m, shp) in shapes.items():
        if only and only not in name:
            continue
        if shp[1] == 3:
            continue
        x = nhwc_rand(shp)
        with torch.no_grad():
            y = m(x)
            conv = m if isinstance(m, modules.Conv2d) else m[0]
            kind = "dw" if conv.groups > 1 else "pw"
            flops = 2.0 * y.numel() * conv.in_channels // conv.groups * conv.kernel_size[0] * conv.kernel_size[1]
            report(f"{kind} {name[-34:]} {shp[1]}->{y.shape[1]}@{shp[2]}x{shp[3]}", time_ms(lambda: m(x), reps),
                   (x.numel() + y.numel()) * 2, flops)
        del x, y


def net_layers(name, batch, only, reps):
    """Every fused conv layer of a net from nets.MODEL_HUB, with the tensors / keyword arguments the net really passes."""
    from rtseg_b200.modules import Conv2d, _FusedConvMixin
    model = nets.MODEL_HUB[name](19).eval().cuda().set_act_dtype(torch.bfloat16)
    calls, hooks = [], []
    for mname, m in model.named_modules():
        if isinstance(m, (Conv2d, _FusedConvMixin)):
            hooks.append(m.register_forward_pre_hook(
                lambda mod, a, kw, mname=mname: calls.append((mname, mod, a, dict(kw))), with_kwargs=True))
    with torch.no_grad():
        model.lowres_logits(torch.randn((batch, 3, 1024, 2048), device="cuda"))
    for h in hooks:
        h.remove()
    total = 0.0
    with torch.no_grad():
        for mname, mod, a, kw in calls:
            if only and only not in mname:
                continue
            conv = mod if isinstance(mod, Conv2d) else mod[0]
            x = a[0]
            if x.dtype != torch.bfloat16:
                continue
            y = mod(*a, **kw)
            ms = time_ms(lambda: mod(*a, **kw), reps)
            total += ms
            n, _, hh, ww = x.shape
            by = (n * hh * ww * conv.in_channels + y.shape[0] * y.shape[2] * y.shape[3] * conv.out_channels *
                  (2 if kw.get("res") is not None else 1)) * 2
            kind = "dw" if conv.groups > 1 else "pw" if tuple(conv.kernel_size) == (1, 1) else "cv"
            k = "x".join(map(str, conv.kernel_size))
            report(f"{kind} {mname[-30:]} {conv.in_channels}->{conv.out_channels} k{k} s{conv.stride[0]} @{hh}x{ww}", ms, by,
                   2.0 * y.shape[0] * y.shape[2] * y.shape[3] * conv.out_channels * conv.in_channels // conv.groups *
                   conv.kernel_size[0] * conv.kernel_size[1])
    print(f"sum of the conv layers: {total:.3f} ms", flush=True)


def head_layers(batch, only, reps):
    x = nhwc_rand((batch, 19, 128, 256))
    target = torch.randint(0, 19, (batch, 1024, 2048), device="cuda")
    cm = torch.zeros((19, 19), dtype=torch.int64, device="cuda")
    report("upsample x8 + argmax (u8)", time_ms(lambda: ops.upsample_argmax(x, (1024, 2048), True), reps),
           x.numel() * 2 + batch * 1024 * 2048)
    report("upsample x8 + argmax + confmat", time_ms(
        lambda: ops.upsample_argmax(x, (1024, 2048), True, target=target, confmat=cm, want_labels=False), reps),
        x.numel() * 2 + batch * 1024 * 2048 * 8)
    labels = ops.upsample_argmax(x, (1024, 2048), True)
    report("confmat update (u8 pred, i64 target)", time_ms(lambda: ops.confmat_update(cm, labels, target), reps),
           batch * 1024 * 2048 * 9)
    lg = torch.randn((8, 19, 1024, 1024), device="cuda")
    tg = torch.randint(0, 19, (8, 1024, 1024), device="cuda")
    crit = rtseg_b200.OhemCELoss(0.7)
    report("OHEM fwd fp32 NCHW 8x19x1024x1024 (all hard)", time_ms(lambda: crit(lg, tg), reps), lg.numel() * 4 + tg.numel() * 8)
    lg2 = lg.clone()
    lg2.scatter_add_(1, tg.unsqueeze(1), torch.full((8, 1, 1024, 1024), 12.0, device="cuda"))
    report("OHEM fwd fp32 (top-k branch)", time_ms(lambda: crit(lg2, tg), reps), lg.numel() * 4 + tg.numel() * 8)
    lg2.requires_grad_(True)

    def fb():
        lg2.grad = None
        crit(lg2, tg).backward()
    report("OHEM fwd+bwd fp32 (top-k branch)", time_ms(fb, reps), 3 * lg.numel() * 4 + tg.numel() * 16)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--set", default="erfnet")
    ap.add_argument("--only", default="")
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--batch", type=int, default=16)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    for s in a.set.split(","):
        if s.startswith("net:"):
            net_layers(s[4:], a.batch, a.only, a.reps)
            continue
        {"erfnet": erfnet_layers, "fastscnn": fastscnn_layers, "head": head_layers}[s](a.batch, a.only, a.reps)
