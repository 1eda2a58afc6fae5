#!/usr/bin/env python
"""bench.py — images/s at 1024x2048 on N B200s (BASELINE.json metric), one JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload erfnet|fastscnn|dabnet|lednet|enet|bisenetv2|ddrnet|head_miou|erfnet_train] [--batch B]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      # the reference's CPU path (oracle port) on the host cores

A step = one forward pass of the workload over one batch of synthetic images per GPU
(default: ERFNet, batch 16, 1024x2048 = BASELINE.json configs[1]); ranks shard by batch with no
data-path collective (weak scaling).  `value` is device-timed with inputs resident in HBM;
`e2e` goes through the public host API (pinned host images in, class map out) with both copies
inside the timed region.  See DESIGN.md §Measurement for the roofline arithmetic.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, NUM_CLASS = 1024, 2048, 19
# ENet is BASELINE configs[0]: the reference's own speed protocol (tools/test_speed.py: ratio 0.5 -> 512 x 1024)
WORKLOAD_HW = {"enet": (512, 1024)}
DEFAULT_BATCH = {"erfnet": 16, "fastscnn": 32, "dabnet": 16, "lednet": 16, "enet": 16, "erfnet_train": 8, "head_miou": 16,
                 "bisenetv2": 16, "ddrnet": 16}
# BASELINE configs[3] (mIoU evaluation): model -> fused head + confusion matrix -> int64 all-reduce -> IoU
EVAL_WORKLOADS = ("bisenetv2", "ddrnet")


def hw_of(workload):
    return WORKLOAD_HW.get(workload, (H, W))
WORKLOAD_NAME = {
    "erfnet": "ERFNet inference batch {b} at 1024x2048 (BASELINE configs[1])",
    "fastscnn": "Fast-SCNN inference batch {b} at 1024x2048 (BASELINE configs[2])",
    "dabnet": "DABNet inference batch {b} at 1024x2048 (SURVEY §8 a10)",
    "lednet": "LEDNet inference batch {b} at 1024x2048 (SURVEY §8 a11)",
    "enet": "ENet inference batch {b} at 512x1024 (BASELINE configs[0]: tools/test_speed.py size; CPU arm = that protocol)",
    "bisenetv2": "BiSeNetv2 mIoU evaluation, batch {b} per GPU at 1024x2048 (BASELINE configs[3]): model -> fused x8 bilinear + argmax "
                 "+ confusion matrix against int64 masks -> one int64 all-reduce -> IoU",
    "ddrnet": "DDRNet-23-slim mIoU evaluation, batch {b} per GPU at 1024x2048 (BASELINE configs[3]): model -> fused x8 bilinear + "
              "argmax + confusion matrix against int64 masks -> one int64 all-reduce -> IoU",
    "head_miou": "segmentation head + mIoU of BASELINE configs[3] (BiSeNetv2 / DDRNet-23-slim output geometry: 19 x 128 x 256 "
                 "logits per 1024x2048 image), batch {b} per GPU: fused x8 bilinear + argmax + confusion matrix, int64 all-reduce, IoU",
    "erfnet_train": "ERFNet training step (fwd + OHEM + bwd + SGD), batch {b} per GPU at 1024x1024 (SURVEY §8d config 5)",
}
NET_LABEL = {"erfnet": "ERFNet", "fastscnn": "Fast-SCNN", "dabnet": "DABNet", "lednet": "LEDNet", "enet": "ENet", "head_miou": "head + mIoU",
             "bisenetv2": "BiSeNetv2 + mIoU", "ddrnet": "DDRNet-23-slim + mIoU",
             "erfnet_train": "ERFNet train step"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="rtseg", choices=["rtseg", "reference"])
    ap.add_argument("--workload", default="erfnet", choices=sorted(DEFAULT_BATCH))
    ap.add_argument("--batch", type=int, default=0, help="images per GPU per step (default 16 ERFNet / 32 Fast-SCNN)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-library-baseline", action="store_true")
    ap.add_argument("--single", action="store_true", help="default run: skip the Fast-SCNN sub-record")
    ap.add_argument("--masks", default="i64", choices=["i64", "u8"],
                    help="evaluation workloads: host masks as the reference's dataset yields them (int64) or as stored (uint8); "
                         "the engine widens uint8 masks on the device")
    ap.add_argument("--input", default="u8", choices=["u8", "f32"],
                    help="u8: decoded uint8 HWC images, normalised inside the stem kernel; f32: normalised fp32 NCHW")
    return ap.parse_args()


# ------------------------------------------------------------------------- CPU baseline ----
def cpu_reference_run(workload: str, steps: int, warmup: int):
    """The reference's CPU implementation of the path: oracle port (functional fp32 PyTorch on the
    host cores, same op sequence as the reference modules — the Python reference itself cannot
    travel to the GPU box).  One step = ONE 1024x2048 image (bounded sample of the batch)."""
    import torch
    from oracle import ref_nets
    from tests.helpers import load_specs

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = load_specs()[workload]["spec"]
    sd = ref_nets.make_state(spec, 1, True)
    H, W = hw_of(workload)
    if workload == "enet":      # config 1: 10 warm-ups, >= 20 timed iterations (reference tools/test_speed.py:29-60 minus the cuda syncs)
        steps, warmup = max(steps, 20), max(warmup, 10)
    x = ref_nets.make_input(0, 1, H, W)
    fwd = ref_nets.FORWARD[workload]
    with torch.no_grad():
        for _ in range(warmup):
            fwd(sd, x)
        t0 = time.perf_counter()
        for _ in range(steps):
            fwd(sd, x)
        dt = time.perf_counter() - t0
    return {"value": steps / dt, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{steps} forward passes of 1 image 1x3x{H}x{W} fp32 after {warmup} warm-up "
                      f"(oracle/ref_nets.py, torch CPU)", "ms_per_image": 1e3 * dt / steps}


def cpu_head_run(steps: int, warmup: int):
    """CPU arm of the head + mIoU workload: F.interpolate(bilinear, align_corners=True) -> max(1)[1]
    (reference core/seg_trainer.py:172) -> restated torchmetrics confusion matrix (utils/metrics.py:4-7) -> IoU, fp32,
    one 1024x2048 image per step."""
    import torch
    from oracle import ref_head

    torch.set_num_threads(os.cpu_count() or 1)
    g = torch.Generator().manual_seed(0)
    x = torch.randn((1, NUM_CLASS, H // 8, W // 8), generator=g)
    t = torch.randint(0, NUM_CLASS, (1, H, W), generator=g)
    t[torch.rand((1, H, W), generator=g) < 0.1] = 255

    def one():
        cm = ref_head.confusion_matrix(ref_head.upsample_argmax(x, (H, W), True).numpy(), t.numpy(), NUM_CLASS, 255)
        return ref_head.iou_from_confmat(cm)
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    return {"value": steps / dt, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{steps} passes over 1 image: interpolate x8 + argmax + confusion matrix + IoU (oracle/ref_head.py, torch / numpy CPU)",
            "ms_per_image": 1e3 * dt / steps}


def cpu_reference_train_run(steps: int, warmup: int, size: int = 1024):
    """The reference's CPU training step restated (oracle, torch autograd on the host cores): ERFNet forward with
    batch statistics + OHEM loss + backward on ONE size x size crop per step (bounded sample of the batch of 8)."""
    import torch
    from oracle import ref_blocks, ref_nets, ref_ohem
    from tests.helpers import load_specs

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = load_specs()["erfnet"]["spec"]
    sd0 = ref_nets.make_state(spec, 1, True)
    x = ref_nets.make_input(0, 1, size, size)
    g = torch.Generator().manual_seed(1)
    labels = torch.randint(0, 19, (1, size, size), generator=g)

    def one():
        sd = {k: v.clone() for k, v in sd0.items()}
        for k, v in sd.items():
            if v.dtype.is_floating_point and "running" not in k:
                v.requires_grad_(True)
        ref_blocks.TRAINING = True
        try:
            logits = ref_nets.erfnet_forward(sd, x)
        finally:
            ref_blocks.TRAINING = False
        ref_ohem.ohem_ce_loss(logits, labels, 0.7, 255)[0].backward()

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = time.perf_counter() - t0
    return {"value": steps / dt, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{steps} training steps (fwd + OHEM + bwd) of 1 crop 1x3x{size}x{size} fp32 after {warmup} warm-up "
                      f"(oracle/ref_nets.py with batch-statistics BatchNorm, torch autograd, CPU)",
            "ms_per_image": 1e3 * dt / steps}


# --------------------------------------------------------------------------- clocks ----
class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(2)
        except Exception:
            self.proc.kill()
        sm, mx, power, reasons = [], [], [], set()
        for r in self.rows:
            f = [t.strip() for t in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                power.append(float(f[2]))
            except ValueError:
                continue
            for name, flag in zip(self.NAMES, f[3:7]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------- dominant kernel ----
def dominant_kernel_roofline(model, workload: str, batch: int, peak_gbs: float, peak_src: str):
    """Times the dominant kernel family alone, launch by launch, with CUDA events on the launch stream.

    ERFNet: the tcgen05 implicit-GEMM kernel behind the 68 dense 3x1/1x3 dilated convs of the
    NonBt1D blocks (93 % of the net's FLOPs, 79 % of its bytes).  Fast-SCNN: the NHWC depthwise kernel.
    achieved = sum(algorithmic bytes per launch) / sum(event-timed duration per launch) over one
    forward's launches of that family; algorithmic bytes = (in + out [+ residual]) elements x 2 B
    + weights + 8 B/channel of folded BN.
    """
    import torch
    from rtseg_b200 import nets, ops
    from rtseg_b200.modules import DWConvBNAct

    dev = torch.device("cuda", torch.cuda.current_device())
    reps, total_bytes, total_ms, launches, per_cfg = 5, 0.0, 0.0, 0, {}
    stream = torch.cuda.current_stream()

    def time_call(fn):
        fn()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1) / reps

    if workload == "erfnet":
        # Every NonBt1D block is timed the way the net runs it: undilated 16- / 64-channel blocks as two fused pair
        # launches (conv_pair_tc_kernel: algorithmic bytes = read x [+ residual], write y — the intermediate never
        # reaches HBM), the dilated 128-channel blocks as four conv_tc_kernel launches.  The family with the larger
        # share of the step is reported; the other one is listed next to it.
        res_hw = {64: (256, 512), 128: (128, 256), 16: (512, 1024)}
        fam = {"pair": [0.0, 0.0, 0], "conv_tc": [0.0, 0.0, 0]}
        for blk in [m for m in model.modules() if isinstance(m, nets.NonBt1DBlock)]:
            c = blk.conv[0][0].in_channels
            h, w = res_hw[c]
            key = (c, blk.conv[2][0].dilation[0])
            if key not in per_cfg:
                x = torch.randn((batch, c, h, w), device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
                r = torch.randn_like(x)
                elems = batch * c * h * w
                wbytes = 3 * c * c * 2 + 8 * c
                with torch.no_grad():
                    fused = blk._pairs(x) is not None
                if fused:
                    fa, fb = blk.conv[0]._folded()[0], blk.conv[1]._folded()[0]
                    fc, fd = blk.conv[2]._folded()[0], blk.conv[3].folded(blk.bn_act[0], blk.bn_act[1].activation)[0]
                    ms = [time_call(lambda: ops.conv_pair(fa, fb, x)), time_call(lambda: ops.conv_pair(fc, fd, x, res=r))]
                    by = [2 * elems * 2 + 2 * wbytes, 3 * elems * 2 + 2 * wbytes]
                    per_cfg[key] = ("pair", ms, by)
                else:
                    ms = [time_call(lambda i=i: blk.conv[i](x)) for i in range(3)]
                    ms.append(time_call(lambda: blk.conv[3](x, res=r, res_mode=ops.RES_PRE, bn=blk.bn_act[0],
                                                            act=blk.bn_act[1].activation)))
                    by = [2 * elems * 2 + wbytes] * 3 + [3 * elems * 2 + wbytes]
                    per_cfg[key] = ("conv_tc", ms, by)
                del x, r
            name, ms, by = per_cfg[key]
            fam[name][0] += sum(ms)
            fam[name][1] += sum(by)
            fam[name][2] += len(ms)
        top = max(fam, key=lambda k: fam[k][0])
        total_ms, total_bytes, launches = fam[top]
        label = {"pair": "conv_pair_tc_kernel (tcgen05: 3x1 -> 1x3 pair of a NonBt1D block in one launch, intermediate tile in "
                         "shared memory, BatchNorm in the MMA, fused residual)",
                 "conv_tc": "conv_tc_kernel (tcgen05 implicit GEMM, 3x1/1x3 dilated, fused BN/act/residual)"}
        kernel = label[top]
        families = {label[k].split(" ")[0]: {"launches_per_step": v[2], "ms_per_step": round(v[0], 4), "bytes_per_step": int(v[1]),
                                             "achieved_gbs": round(v[1] / (v[0] * 1e-3) / 1e9, 1) if v[0] else None,
                                             "frac": round(v[1] / (v[0] * 1e-3) / 1e9 / peak_gbs, 4) if v[0] else None}
                    for k, v in fam.items()}
    elif workload in ("dabnet", "lednet", "enet", "bisenetv2", "ddrnet"):
        # every fused conv layer is captured with the tensors / keyword arguments the net really passes, timed
        # alone, and binned by kernel family; the family with the largest share of the time is reported
        from rtseg_b200.modules import Conv2d, _FusedConvMixin
        calls, hooks = [], []
        for m in model.modules():
            if isinstance(m, (Conv2d, _FusedConvMixin)):
                hooks.append(m.register_forward_pre_hook(
                    lambda mod, a, kw: calls.append((mod, a, dict(kw))), with_kwargs=True))
        x_in = torch.randn((batch, 3, *hw_of(workload)), device=dev)
        with torch.no_grad():
            model.lowres_logits(x_in)
        for hk in hooks:
            hk.remove()
        del x_in
        fam = {}
        with torch.no_grad():
            for mod, a, kw in calls:
                conv = mod if isinstance(mod, Conv2d) else mod[0]
                x = a[0]
                if x.dtype != torch.bfloat16:
                    continue
                y = mod(*a, **kw)
                ms = time_call(lambda: mod(*a, **kw))
                n, _, hh, ww = x.shape
                by = (n * hh * ww * conv.in_channels + y.shape[0] * y.shape[2] * y.shape[3] * conv.out_channels *
                      (2 if kw.get("res") is not None else 1)) * 2 + conv.weight.numel() * 2 + 8 * conv.out_channels
                name = ("dw" if conv.groups > 1 else "tc" if conv.in_channels % 16 == 0 else "dense")
                f = fam.setdefault(name, [0.0, 0.0, 0])
                f[0] += ms
                f[1] += by
                f[2] += 1
        top = max(fam, key=lambda k: fam[k][0])
        total_ms, total_bytes, launches = fam[top]
        kernel = {"dw": "dw_win_kernel / dw_vec_kernel (NHWC depthwise 3x1 / 1x3 dilated + BN + act + residual)",
                  "tc": "conv_tc_kernel (tcgen05 implicit GEMM, fused bias/BN/act/residual, channel-slice I/O)",
                  "dense": "conv_dense_kernel (CUDA-core implicit GEMM for channel counts that are not multiples of 16)"}[top]
        kernel += "; family shares of conv time: " + ", ".join(
            f"{k} {100 * v[0] / sum(t[0] for t in fam.values()):.0f}% ({v[2]} launches)" for k, v in sorted(fam.items()))
    else:
        x_in = torch.randn((batch, 3, H, W), device=dev)
        shapes = {}
        hooks = []
        for m in model.modules():
            if isinstance(m, DWConvBNAct):
                hooks.append(m.register_forward_pre_hook(lambda mod, inp: shapes.__setitem__(mod, tuple(inp[0].shape))))
        with torch.no_grad():
            model.lowres_logits(x_in)
        for hk in hooks:
            hk.remove()
        del x_in
        for m, shp in shapes.items():
            x = torch.randn(shp, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            with torch.no_grad():
                y = m(x)
                ms = time_call(lambda: m(x))
            total_ms += ms
            total_bytes += (x.numel() + y.numel()) * 2 + m[0].weight.numel() * 4 + 8 * y.shape[1]
            launches += 1
            del x, y
        kernel = "dw_tma_kernel / dw_win_kernel (NHWC depthwise 3x3 + BN + act; TMA-staged halo tiles for the large maps)"
    achieved = total_bytes / (total_ms * 1e-3) / 1e9
    # DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the same kernel family over one step, from the
    # committed ncu launch list of this command (profiles/launches_r01.md); null for workloads not captured.
    traffic, traffic_note = None, None
    try:   # ncu dram__bytes_read.sum + dram__bytes_write.sum of the same family over one step (tools/parse_ncu_launches.py)
        with open(os.path.join(ROOT, "profiles", "traffic_r02.json")) as f:
            tr = json.load(f).get(workload, {})
        if tr.get("batch") == batch:
            fam_key = kernel.split(" ")[0].split("_kernel")[0]
            for k, v in tr.get("families", {}).items():
                if k.startswith(fam_key):
                    traffic, traffic_note = int(v["dram_bytes_per_step"]), v.get("note")
    except Exception:
        traffic = None
    out_families = locals().get("families")
    return {"bound": "hbm", "kernel": kernel, "achieved": round(achieved, 1), "peak": peak_gbs, "unit": "GB/s",
            "frac": round(achieved / peak_gbs, 4), "traffic": traffic, "traffic_note": traffic_note,
            "families": out_families,
            "peak_source": peak_src,
            "launches_per_step": launches, "bytes_per_step": int(total_bytes), "ms_per_step": round(total_ms, 4),
            "frac_of_8TBps_nominal": round(achieved / 8000.0, 4)}


PARITY_MIN_AGREE = 0.985  # bf16 class map vs the fp32 oracle on the benchmark's own weights (default init)
# ENet (no normalisation behind its residual adds): bf16 storage alone moves ~4 % of a random-init class map; the
# parity tests hold its bf16 path to the storage-emulating oracle instead (tests/test_gpu_enet.py)
PARITY_MIN_BY_WORKLOAD = {"enet": 0.93, "bisenetv2": 0.93}   # (BiSeNetv2: a sigmoid gate on un-normalised features)


def parity_check(model, eng, workload: str):
    """Checks the output of the timed configuration itself: the class maps the engine produced for the first and
    the last image of slot 0's resident batch (bf16 path, CUDA graph) against the CPU oracle's fp32 argmax on the
    same images and the same weights.  The run fails below PARITY_MIN_AGREE."""
    import torch
    from oracle import ref_nets

    labels = eng.run_device(0)
    torch.cuda.synchronize()
    idx = sorted({0, eng.batch - 1})
    x = eng.slots[0].x[idx]
    if x.dtype == torch.uint8:   # decoded images: the oracle normalises them as the reference's dataset does
        from oracle import ref_input
        x = ref_input.normalize_u8(x.permute(0, 2, 3, 1).contiguous().cpu())
    else:
        x = x.float().cpu()
    got = labels[idx].cpu().long()
    sd = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.no_grad():
        ref = ref_nets.FORWARD[workload](sd, x)
    want = ref.max(1)[1]
    per_image = (got == want).float().mean(dim=(1, 2)).tolist()
    agree = float((got == want).float().mean().item())
    out = {"argmax_agree": round(agree, 6), "per_image": [round(v, 6) for v in per_image], "n_images": len(idx),
           "images": idx, "classes_present": int(want.unique().numel()),
           "min_required": PARITY_MIN_BY_WORKLOAD.get(workload, PARITY_MIN_AGREE),
           "oracle": "oracle/ref_nets.py fp32 on the host, same weights and inputs; labels from the timed bf16 graph"}
    if min(per_image) < out["min_required"]:
        raise SystemExit(f"bench.py: parity check failed: {json.dumps(out)}")
    return out


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


# ------------------------------------------------------------------- training workload ----
def wgrad_roofline(model, batch, size, peak_gbs, peak_src):
    """One eager forward + backward with the launches of the step's two big kernel families — the BatchNorm backward
    passes (reduce + apply) and the weight gradients — bracketed by CUDA events on their stream; the family with the
    larger share of the time is reported: achieved = sum of its algorithmic bytes / sum of its durations."""
    import torch
    from rtseg_b200 import _lib
    from rtseg_b200 import loss as rloss

    lib = _lib.load()
    stream = torch.cuda.current_stream()
    rec = []

    def conv_bytes(a):
        d = a[0]._obj
        gh = (d.h + 2 * d.pad_h - d.dil * (d.kh - 1) - 1) // d.stride + 1
        gw = (d.w + 2 * d.pad_w - d.dil * (d.kw - 1) - 1) // d.stride + 1
        es = 2 if d.dtype == 1 else 4
        return (d.n * d.h * d.w * d.cin + d.n * gh * gw * d.cout) * es + d.kh * d.kw * d.cin * d.cout * 4

    def bn_bytes(extra):   # args: npix, c, dtype, dz, .., .., y, .., .., res, ...: dz + y (+ res) read, `extra` tensors written
        return lambda a: a[0] * a[1] * (2 if a[2] == 1 else 4) * (2 + (1 if a[9] else 0) + extra)

    fams = {"rtseg_conv_wgrad": ("wgrad", conv_bytes), "rtseg_conv_wgrad_tc": ("wgrad", conv_bytes),
            "rtseg_bn_act_bwd_reduce": ("bn_bwd", bn_bytes(0)), "rtseg_bn_act_bwd_apply": ("bn_bwd", bn_bytes(1))}
    origs = {name: getattr(lib, name) for name in fams}

    def make_timed(name):
        orig, (fam, nbytes) = origs[name], fams[name]

        def timed(*a):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            rc = orig(*a)
            e1.record(stream)
            e1.synchronize()
            rec.append((fam, name, e0.elapsed_time(e1), nbytes(a)))
            return rc
        return timed

    m = model.module if hasattr(model, "module") else model
    x = torch.randn((batch, 3, size, size), device="cuda")
    y = torch.randint(0, 19, (batch, size, size), device="cuda")
    for name in origs:
        setattr(lib, name, make_timed(name))
    try:
        rloss.OhemCELoss(0.7, 255)(m(x), y).backward()
        torch.cuda.synchronize()
    finally:
        for name, fn in origs.items():
            setattr(lib, name, fn)
    m.zero_grad(set_to_none=True)
    tot = {}
    for fam, name, ms, by in rec:
        t = tot.setdefault(fam, [0.0, 0.0, 0])
        t[0] += ms
        t[1] += by
        t[2] += 1
    top = max(tot, key=lambda k: tot[k][0])
    ms, by, cnt = tot[top]
    achieved = by / (ms * 1e-3) / 1e9
    label = {"bn_bwd": "bn_act_bwd_reduce_vec_kernel + bn_act_bwd_apply_vec_kernel (BatchNorm + activation backward: one "
                       "reduction and one apply pass over dz, y [, res])",
             "wgrad": "conv_wgrad_tc_kernel (tcgen05 weight gradient, MN-major operands, TMEM-resident accumulators)"}[top]
    others = "; ".join(f"{k}: {v[2]} launches, {v[0]:.2f} ms, {v[1] / max(v[0], 1e-9) / 1e6:.0f} GB/s" for k, v in sorted(tot.items()))
    return {"bound": "hbm", "kernel": label + " — the largest kernel family of the step",
            "achieved": round(achieved, 1), "peak": peak_gbs, "unit": "GB/s", "frac": round(achieved / peak_gbs, 4),
            "traffic": None, "peak_source": peak_src, "launches_per_step": cnt, "bytes_per_step": int(by),
            "ms_per_step": round(ms, 4), "families": others,
            "note": "algorithmic bytes: BatchNorm backward = dz + y (+ res) read per pass, du written by the apply pass; "
                    "weight gradient = x + dy read once + fp32 dW"}


def train_main(args, rank, local_rank, world, batch, workload_name, metric):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import train_bench

    sampler = ClockSampler(local_rank) if rank == 0 else None
    r = train_bench.run(batch=batch, size=1024, steps=args.steps, warmup=max(3, args.warmup), fp32=False, graph=True)
    clocks = sampler.stop() if sampler else None
    line = None
    if rank == 0:
        peak, peak_src = measured_peak()
        # a fresh single-GPU copy of the net (plain BatchNorm, no DDP hooks): the instrumented pass must not issue
        # collectives that the other ranks do not join
        del r["model"]
        torch.cuda.empty_cache()
        import rtseg_b200
        probe = rtseg_b200.nets.ERFNet(num_class=19).cuda().train().set_act_dtype(torch.bfloat16)
        roof = wgrad_roofline(probe, batch, 1024, peak, peak_src)
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            b = cpu_reference_train_run(2, 1)
            cpu = {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")}
            cpu["value"] = round(cpu["value"], 4)
        launches = r["launches_per_step"] if r["launches_per_step"] is not None else r["capture_launches"] // 5
        line = {
            "metric": metric, "value": round(r["images_per_s"], 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": round(r["ms_per_step"], 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name, "per_gpu_batch": batch, "global_batch": batch * world,
                       "weights": "random init, torch.manual_seed(1); fp32 masters, bf16 activations and gradients",
                       "loss": "OHEM cross entropy thresh 0.7, ignore 255", "optimizer": "SGD lr 0.01*world, momentum 0.9, wd 1e-4",
                       "batchnorm": "SyncBatchNorm over ranks" if world > 1 else "BatchNorm, batch statistics",
                       "l2": "every activation and gradient tensor (>= 268 MB at batch 8) exceeds the 126 MB L2",
                       "parallelism": f"dp{world}: DistributedDataParallel, NCCL gradient all-reduce" if world > 1 else "dp1",
                       "cuda_graph": r["graph"]},
            "clocks": clocks,
            "e2e": {"value": round(r["e2e_images_per_s"], 2), "unit": "images/s", "h2d_bytes_per_step": r["h2d_bytes"],
                    "d2h_bytes_per_step": r["d2h_bytes"]},
            "gpu_launches": int(launches * args.steps),
            "roofline": roof, "cpu_baseline": cpu, "loss": round(r["loss"], 5),
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------- config 4: head + mIoU ----
def head_main(args, rank, local_rank, world, batch, workload_name):
    """BASELINE configs[3] on the part of it that lies on the path: low-resolution logits as BiSeNetv2 / DDRNet-23-slim
    emit them (reference models/bisenetv2.py:35, models/ddrnet.py:53: 19 classes at 1/8 resolution) -> fused x8
    bilinear upsample + argmax + confusion-matrix update against int64 labels with 10 % ignore pixels -> ONE int64
    all-reduce of the 19 x 19 matrix over the ranks (NCCL) -> IoU.  The two backbones themselves are not built here
    (their blocks — GatherExpansionLayer, RB / DAPPM — run on the fused leaf modules through inject.install).
    A step = one batch through the head kernel; the all-reduce + IoU of the evaluation sit inside the timed region once,
    as `metrics.compute()` does in the reference's validate loop (core/seg_trainer.py:136)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import rtseg_b200
    from oracle import ref_head
    from rtseg_b200 import _lib, ops

    C, lh, lw = NUM_CLASS, H // 8, W // 8
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    logits = [torch.randn((batch, C, lh, lw), device="cuda", generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
              for _ in range(2)]
    target = torch.randint(0, C, (batch, H, W), device="cuda", generator=g)
    target[torch.rand((batch, H, W), device="cuda", generator=g) < 0.1] = 255
    met = rtseg_b200.SegMetrics(C, 255).to("cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i):
        return ops.upsample_argmax(logits[i % 2], (H, W), True, target=target, confmat=met.confmat, ignore_index=255)

    for i in range(max(3, args.warmup)):
        step(i)
    met.compute()
    met.reset()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    stream = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = _lib.launch_count()
    e0.record(stream)
    for i in range(args.steps):
        labels = step(i)
    iou = met.compute()                      # all-reduce of the int64 matrix + IoU: once per evaluation
    e1.record(stream)
    barrier()
    launches = _lib.launch_count() - l0
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dev_ms = float(ms.item())
    clocks = sampler.stop() if sampler else None
    value = world * batch * args.steps / (dev_ms * 1e-3)
    total = met.synced_confmat()
    # head kernel alone, event-timed: the roofline line
    reps = 10
    barrier()
    e0.record(stream)
    for i in range(reps):
        step(i)
    e1.record(stream)
    e1.synchronize()
    k_ms = e0.elapsed_time(e1) / reps
    # end to end: pinned host logits + labels in, class map + IoU out
    _PHASE[0] = "head_miou: end to end"
    h_logits = logits[0].cpu().pin_memory()
    h_target = target.cpu().pin_memory()
    h_labels = torch.empty((batch, H, W), dtype=torch.uint8).pin_memory()
    met.reset()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        logits[0].copy_(h_logits, non_blocking=True)
        target.copy_(h_target, non_blocking=True)
        h_labels.copy_(step(0), non_blocking=True)
    iou_host = met.compute().cpu()
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    line = None
    if rank == 0:
        peak, peak_src = measured_peak()
        # self-check: confusion matrix of the first and last image against the oracle's argmax of the same logits
        _PHASE[0] = "head_miou: parity check"
        idx = sorted({0, batch - 1})
        cm = torch.zeros((C, C), dtype=torch.int64, device="cuda")
        lab = ops.upsample_argmax(logits[0][idx].contiguous(memory_format=torch.channels_last), (H, W), True,
                                  target=target[idx].contiguous(), confmat=cm, ignore_index=255)
        want_lab = ref_head.upsample_argmax(logits[0][idx].float().cpu(), (H, W), True)
        agree = float((lab.cpu().long() == want_lab).float().mean().item())
        want_cm = ref_head.confusion_matrix(lab.cpu().numpy(), target[idx].cpu().numpy(), C, 255)
        cm_exact = bool(np.array_equal(cm.cpu().numpy(), want_cm))
        oracle_cm = ref_head.confusion_matrix(want_lab.numpy(), target[idx].cpu().numpy(), C, 255)
        cm_l1 = int(np.abs(cm.cpu().numpy() - oracle_cm).sum())
        if not cm_exact or agree < 0.999:
            raise SystemExit(f"bench.py: head parity check failed: labels agree {agree}, confmat from the same labels exact: {cm_exact}")
        nbytes = batch * C * lh * lw * 2 + batch * H * W * (1 + 8) + C * C * 8
        achieved = nbytes / (k_ms * 1e-3) / 1e9
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            torch.set_num_threads(os.cpu_count() or 1)
            xl, tg = logits[0][:4].float().cpu(), target[:4].cpu()
            t0 = time.perf_counter()
            for _ in range(3):
                ref_head.confusion_matrix(ref_head.upsample_argmax(xl, (H, W), True).numpy(), tg.numpy(), C, 255)
            dt = (time.perf_counter() - t0) / 3
            cpu = {"value": round(4 / dt, 3), "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
                   "sample": "3 passes over a 4-image batch: F.interpolate(bilinear, align_corners=True) -> max(1)[1] -> restated "
                             "torchmetrics confusion matrix (oracle/ref_head.py), fp32 on the host"}
        line = {
            "metric": "images/sec @1024x2048 (fused head + mIoU, BASELINE configs[3] geometry)", "value": round(value, 2),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": round(dev_ms / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name, "per_gpu_batch": batch, "global_batch": batch * world,
                       "labels": "randint(0, 19) with 10 % set to 255 (ignored)", "collective": "one all-reduce of the 19 x 19 int64 "
                       "confusion matrix per evaluation (NCCL)" if world > 1 else "none (one rank)",
                       "l2": "labels in + class map out are 151 MB per step: larger than the 126 MB L2",
                       "parallelism": f"dp{world}: images sharded across ranks"},
            "clocks": clocks,
            "e2e": {"value": round(world * batch * args.steps / float(e2e_s.item()), 2), "unit": "images/s",
                    "h2d_bytes_per_step": h_logits.numel() * 2 + h_target.numel() * 8, "d2h_bytes_per_step": h_labels.numel()},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "upsample_argmax_kernel (x8 bilinear + argmax + confusion-matrix histogram, one pass)",
                         "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                         "traffic": None, "peak_source": peak_src, "bytes_per_step": int(nbytes), "ms_per_step": round(k_ms, 4),
                         "note": "int64 labels in (8 B/pixel) dominate the algorithmic bytes; the kernel is ALU-bound on the "
                                 "19-class interpolation, not bandwidth-bound"},
            "cpu_baseline": cpu,
            "parity_check": {"argmax_agree": round(agree, 6), "confmat_exact_from_kernel_labels": cm_exact,
                             "confmat_l1_vs_oracle_labels": cm_l1, "n_images": len(idx), "pixels_counted": int(total.sum().item()),
                             "miou": round(float(iou.mean().item()), 6), "miou_e2e": round(float(iou_host.mean().item()), 6)},
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------- main ----
_PHASE = ["set-up, warm-up and device-timed steps"]


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    batch = args.batch or DEFAULT_BATCH[args.workload]
    workload_name = WORKLOAD_NAME[args.workload].format(b=batch)
    metric = "images/sec @%dx%d (%s)" % (*hw_of(args.workload), NET_LABEL[args.workload])

    train = args.workload == "erfnet_train"
    if train:
        metric = "images/sec ERFNet training step @1024x1024 (fwd + OHEM + bwd + SGD)"
    if args.impl == "reference":
        if rank != 0:
            return 0
        base = (cpu_reference_train_run(max(1, args.steps), max(0, args.warmup)) if train else
                cpu_head_run(max(1, args.steps), max(0, args.warmup)) if args.workload == "head_miou" else
                cpu_reference_run(args.workload, max(1, args.steps), max(0, args.warmup)))
        line = {
            "impl": "reference", "metric": metric, "value": round(base["value"], 4), "unit": "images/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(base["ms_per_image"], 2), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name, "sample": "1 image per step on the host cores",
                       "weights": "seeded random init (oracle.ref_nets.make_state)"},
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": round(base["value"], 4), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        print(json.dumps(line), flush=True)
        return 0

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU path in rtseg_b200)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if args.workload == "erfnet_train":
            os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")   # DDP step captured in a CUDA graph (tools/train_bench.py)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import rtseg_b200
    from rtseg_b200 import _lib

    _lib.check(_lib.load().rtseg_device_check(), "rtseg_device_check")
    if train:
        return train_main(args, rank, local_rank, world, batch, workload_name, metric)
    if args.workload == "head_miou":
        return head_main(args, rank, local_rank, world, batch, workload_name)
    rec = run_inference(args.workload, batch, args, rank, local_rank, world)
    extra = {}
    if args.workload == "erfnet" and not args.batch and not args.single:
        # BASELINE.json's metric names both nets: time Fast-SCNN (configs[2]) in the same run, same clocks record style
        _PHASE[0] = "Fast-SCNN sub-record"
        torch.cuda.empty_cache()
        sub = run_inference("fastscnn", DEFAULT_BATCH["fastscnn"], args, rank, local_rank, world)
        if sub is not None:
            extra["workloads"] = {"fastscnn": sub}
    if rank == 0 and world == 1 and not args.no_library_baseline:
        _PHASE[0] = "library baseline (reference modules' op sequence through stock PyTorch / cuDNN on this GPU)"
        torch.cuda.empty_cache()
        lib = {args.workload: library_baseline(args.workload, batch)}
        if "workloads" in extra:
            torch.cuda.empty_cache()
            lib["fastscnn"] = library_baseline("fastscnn", DEFAULT_BATCH["fastscnn"])
        for name, r in [(args.workload, rec)] + [(k, v) for k, v in extra.get("workloads", {}).items()]:
            best = max(lib[name][k]["value"] for k in ("fp32_nchw", "bf16_channels_last") if lib[name].get(k))
            r["vs_library"] = round(r["value"] / best, 3)
        extra["library_baseline"] = lib
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rec is not None:
        rec.update(extra)
        print(json.dumps(rec), flush=True)
    return 0


def library_baseline(workload: str, batch: int, steps: int = 5):
    """The bar SURVEY §2.1 names: the reference modules' own op sequence (oracle/ref_nets.py: F.conv2d -> F.batch_norm
    -> activation, unfused, exactly what the reference's nn.Modules dispatch) run by stock PyTorch on this GPU
    (cuDNN / ATen), (i) fp32 NCHW eager as reference tools/test_speed.py:22-60 runs it and (ii) bf16 channels_last.
    Device-timed with CUDA events, inputs resident, output = full-resolution logits + argmax like the reference's
    predict (core/seg_trainer.py:172).  None of this repo's kernels is on this path."""
    import torch
    from oracle import ref_nets
    from tests.helpers import load_specs

    torch.backends.cudnn.benchmark = True
    spec = load_specs()[workload]["spec"]
    sd0 = ref_nets.make_state(spec, 1, True)
    out = {}
    for key, dt, cl in (("fp32_nchw", torch.float32, False), ("bf16_channels_last", torch.bfloat16, True)):
        try:
            sd = {k: (v.cuda().to(dt) if v.dtype.is_floating_point else v.cuda()) for k, v in sd0.items()}
            if cl:
                sd = {k: (v.contiguous(memory_format=torch.channels_last) if v.dim() == 4 else v) for k, v in sd.items()}
            b = batch
            while True:
                try:
                    x = torch.randn((b, 3, *hw_of(workload)), device="cuda").to(dt)
                    if cl:
                        x = x.contiguous(memory_format=torch.channels_last)
                    fwd = ref_nets.FORWARD[workload]
                    with torch.no_grad():
                        for _ in range(3):
                            fwd(sd, x).max(1)[1]
                        torch.cuda.synchronize()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        for _ in range(steps):
                            fwd(sd, x).max(1)[1]
                        e1.record()
                        e1.synchronize()
                    break
                except torch.OutOfMemoryError:
                    torch.cuda.empty_cache()
                    b //= 2
                    if b < 1:
                        raise
            ms = e0.elapsed_time(e1) / steps
            out[key] = {"value": round(b * 1e3 / ms, 2), "unit": "images/s", "ms_per_step": round(ms, 3), "batch": b}
            del sd, x
            torch.cuda.empty_cache()
        except Exception as e:   # a library failure must not void our own line
            out[key] = None
            out[key + "_error"] = f"{type(e).__name__}: {str(e)[:160]}"
    out["what"] = ("reference op sequence (oracle/ref_nets.py) on cuda through stock PyTorch: cuDNN / ATen convolutions, "
                   "separate batch_norm / activation / interpolate / argmax kernels; cudnn.benchmark on; eager")
    return out


def run_inference(workload, batch, args, rank, local_rank, world):
    """One inference workload: device-timed value, end-to-end e2e, self-check, roofline, CPU baseline.
    Returns the JSON record on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist
    import rtseg_b200
    from rtseg_b200 import _lib

    workload_name = WORKLOAD_NAME[workload].format(b=batch)
    metric = "images/sec @%dx%d (%s)" % (*hw_of(workload), NET_LABEL[workload])
    _PHASE[0] = f"{workload}: set-up, warm-up and device-timed steps"
    torch.manual_seed(1)  # the reference's seed (configs/base_config.py:70); default random init
    model = rtseg_b200.nets.MODEL_HUB[workload](num_class=NUM_CLASS).eval().cuda()
    model.set_act_dtype(torch.bfloat16)
    u8 = args.input == "u8"
    H, W = hw_of(workload)
    evaluate = workload in EVAL_WORKLOADS
    met = rtseg_b200.SegMetrics(NUM_CLASS, 255).to("cuda") if evaluate else None
    n_slots = int(os.environ.get("RTSEG_BENCH_SLOTS", "3"))   # copy / compute slots of the end-to-end pipeline
    eng = rtseg_b200.InferenceEngine(model, batch=batch, height=H, width=W, use_graph=True, slots=n_slots,
                                     in_dtype=torch.uint8 if u8 else torch.float32,
                                     confmat=met.confmat if evaluate else None)

    # synthetic inputs: per-rank seeds, resident on the device for `value`, pinned on the host for `e2e`
    g = torch.Generator(device="cuda").manual_seed(rank)
    for s in eng.slots:
        if u8:
            s.x.copy_(torch.randint(0, 256, s.x.shape, generator=g, device="cuda", dtype=torch.uint8))
        else:
            s.x.copy_(torch.randn(s.x.shape, generator=g, device="cuda"))
    if u8:
        host_in = [torch.randint(0, 256, (batch, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
    else:
        host_in = [torch.randn(eng.slots[0].x.shape).pin_memory() for _ in range(2)]
    host_masks = [None, None]
    if evaluate:    # int64 masks as the reference's dataset yields them, 10 % ignored pixels
        for s in eng.slots:
            s.target.copy_(torch.randint(0, NUM_CLASS, s.target.shape, generator=g, device="cuda"))
            s.target[torch.rand(s.target.shape, generator=g, device="cuda") < 0.1] = 255
        host_masks = [eng.slots[i].target.cpu().to(torch.uint8 if args.masks == "u8" else torch.int64).pin_memory() for i in range(2)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value) -------------------------------------------------------
    stream = eng.slots[0].stream
    for _ in range(max(3, args.warmup)):
        eng.run_device(0)
    if evaluate:
        with torch.cuda.stream(stream):
            met.compute()      # warms the all-reduce up
            met.reset()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = _lib.launch_count()
    e0.record(stream)
    for _ in range(args.steps):
        eng.run_device(0)
    if evaluate:               # once per evaluation, inside the timed region: all-reduce of the int64 matrix + IoU
        with torch.cuda.stream(stream):
            iou = met.compute()
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    eager_launches = _lib.launch_count() - launches0  # 0 under graph replay; the graph replays launches_per_step
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([dev_ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    value = world * batch * args.steps / (dev_ms * 1e-3)

    # ---- end to end through the public API (pinned host in, host class map out) --------------------
    _PHASE[0] = f"{workload}: end to end (pinned host in, host class map out)"
    eval_check = None
    if evaluate:
        # the matrix the timed steps accumulated: K passes over slot 0's batch = K x the matrix of one pass, and that
        # one is bit-exact against the restated torchmetrics matrix of the engine's own class map (oracle/ref_head.py)
        import numpy as np
        from oracle import ref_head
        total = met.confmat.clone()
        met.reset()
        lab = eng.run_device(0)
        torch.cuda.synchronize()
        one = met.confmat.cpu().numpy()
        want = ref_head.confusion_matrix(lab.cpu().numpy(), eng.slots[0].target.cpu().numpy(), NUM_CLASS, 255)
        eval_check = {"confmat_exact_from_kernel_labels": bool(np.array_equal(one, want)),
                      "timed_matrix_is_steps_x_one_pass": bool(torch.equal(total.cpu(), torch.from_numpy(one) * args.steps)),
                      "pixels_counted_per_pass": int(one.sum()), "miou": round(float(iou.mean().item()), 6)}
        if not (eval_check["confmat_exact_from_kernel_labels"] and eval_check["timed_matrix_is_steps_x_one_pass"]):
            raise SystemExit(f"bench.py: confusion-matrix check failed: {json.dumps(eval_check)}")
        met.reset()
    for i in range(n_slots):
        eng(host_in[i % 2], host_masks[i % 2])
    if evaluate:
        met.reset()
    barrier()
    t0 = time.perf_counter()
    pending = []
    for i in range(args.steps):
        pending.append(eng.submit(host_in[i % 2], host_masks[i % 2]))
        if len(pending) == n_slots:
            eng.collect(pending.pop(0))
    while pending:
        eng.collect(pending.pop(0))
    if evaluate:
        eval_check["miou_e2e"] = round(float(met.compute().mean().item()), 6)      # all-reduce + IoU + read-back
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * batch * args.steps / float(t.item())

    # ---- self-check: watchdog flag on every rank, class maps of the first and last image against the oracle ----
    _PHASE[0] = f"{workload}: parity check (class maps of image 0 and image B-1 against the CPU oracle)"
    from rtseg_b200 import ops as _ops
    flag = torch.tensor([_ops.tc_timeout_flag()], device="cuda", dtype=torch.int32)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MAX)
    if int(flag.item()) != 0:
        raise SystemExit("bench.py: a tensor-core kernel hit its barrier watchdog during the timed region "
                         "(rtseg_tc_timeout_flag set): the numbers of this run are void")
    parity = parity_check(model, eng, workload) if rank == 0 else None

    line = None
    if rank == 0:
        peak, peak_src = measured_peak()
        _PHASE[0] = f"{workload}: roofline pass (dominant kernel family timed with CUDA events)"
        roof = dominant_kernel_roofline(model, workload, batch, peak, peak_src)
        cpu = None
        _PHASE[0] = f"{workload}: cpu baseline"
        if world == 1 and not args.no_cpu_baseline:
            b = cpu_reference_run(workload, 3, 1)
            cpu = {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")}
            cpu["value"] = round(cpu["value"], 4)
        line = {
            "metric": metric, "value": round(value, 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": round(dev_ms / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name, "per_gpu_batch": batch, "global_batch": batch * world,
                       "weights": "random init, torch.manual_seed(1)", "activations": "NHWC bf16, fp32 accumulate",
                       "input": ("decoded uint8 HWC images, normalised (datasets/cityscapes.py:122-123) inside the stem kernel"
                                 if u8 else "normalised fp32 NCHW images"),
                       "output": "uint8 class map (fused argmax head)",
                       "l2": f"input {eng.h2d_bytes() / 1e6:.0f} MB and every activation tensor exceed the 126 MB L2",
                       "parallelism": f"dp{world}: batch sharded across ranks, no data-path collective",
                       "cuda_graph": True},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 2), "unit": "images/s",
                    "h2d_bytes_per_step": eng.slots[0].x.numel() * eng.slots[0].x.element_size() +
                    (host_masks[0].numel() * host_masks[0].element_size() if evaluate else 0),
                    "d2h_bytes_per_step": eng.d2h_bytes(), "pipelined_slots": n_slots},
            "gpu_launches": int(eng.launches_per_step * args.steps + eager_launches),
            "roofline": roof,
            "cpu_baseline": cpu,
            "parity_check": parity,
        }
        if evaluate:
            line["parity_check"].update(eval_check)
            line["config"]["masks"] = (f"{'int64' if args.masks == 'i64' else 'uint8'} [N, H, W] on the host, randint(0, 19) with 10 % set "
                                       "to 255 (ignored), copied in with the images (the device copy is int64)")
            line["config"]["collective"] = ("one all-reduce of the 19 x 19 int64 confusion matrix per evaluation (NCCL), inside the "
                                            "timed region" if world > 1 else "none (one rank)")
            line["config"]["output"] = "uint8 class map + int64 confusion matrix -> IoU"
    del eng, model
    torch.cuda.empty_cache()
    return line


if __name__ == "__main__":
    try:
        sys.exit(main())
    except Exception:
        # say which phase of the run raised (a CUDA error surfaces at the next synchronising call)
        print(f"bench.py failed in phase: {_PHASE[0]}", file=sys.stderr, flush=True)
        raise
