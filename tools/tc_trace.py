#!/usr/bin/env python
"""Development aid: per-tile timeline of conv_tc_kernel (CTA 0) from a trace build of librtseg.

  python tools/tc_trace.py --build          # realtime-.../librtseg_tctrace.so with -DRTSEG_TC_TRACE
  python tools/tc_trace.py [--c 128] [--kernel 3x1] [--dil 1] [--res]
"""
import argparse
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = os.path.join(ROOT, "realtime-semantic-segmentation-pytorch_b200")
TRACE_LIB = os.path.join(PKG, "librtseg_tctrace.so")
EV = ["tma_kb0", "tma_kb1", "acc_free", "kb0_landed", "kb1_landed", "mma_issued", "epi_start", "epi_end"]


def build():
    from rtseg_b200 import build as b
    objs = []
    for src in b.SOURCES:
        obj = os.path.join(PKG, "build", src.replace(".cu", ".o"))
        if src == "conv_tc.cu":
            obj = os.path.join(PKG, "build", "conv_tc_trace.o")
            subprocess.check_call([b._nvcc(), *b.NVCC_FLAGS, "-DRTSEG_TC_TRACE", "-c", os.path.join(b.CSRC, src), "-o", obj])
        objs.append(obj)
    subprocess.check_call([b._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", TRACE_LIB, *objs])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build", action="store_true")
    ap.add_argument("--c", type=int, default=128)
    ap.add_argument("--kernel", default="3x1")
    ap.add_argument("--dil", type=int, default=1)
    ap.add_argument("--res", action="store_true")
    a = ap.parse_args()
    if a.build:
        build()
        return
    os.environ["RTSEG_LIB_PATH"] = TRACE_LIB
    import torch
    from rtseg_b200 import _lib, modules, ops
    hw = {128: (128, 256), 64: (256, 512), 16: (512, 1024)}[a.c]
    k = tuple(int(v) for v in a.kernel.split("x"))
    m = modules.ConvBNAct(a.c, a.c, k, dilation=a.dil).eval().cuda()
    x = torch.randn((16, a.c, *hw), device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    kw = dict(res=torch.randn_like(x), res_mode=ops.RES_PRE) if a.res else {}
    with torch.no_grad():
        for _ in range(3):
            m(x, **kw)
    torch.cuda.synchronize()
    lib = _lib.load()
    lib.rtseg_tc_trace_read.restype = C.c_int
    buf = (C.c_longlong * (64 * 16))()
    assert lib.rtseg_tc_trace_read(buf) == 0
    t0 = buf[0]
    print("tile " + " ".join(f"{e:>11s}" for e in EV))
    for t in range(0, 30):
        print(f"{t:4d} " + " ".join(f"{buf[t * 16 + e] - t0:11d}" for e in range(len(EV))))


if __name__ == "__main__":
    main()
