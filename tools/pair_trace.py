#!/usr/bin/env python
"""Development aid: per-tile timeline of the fused pair kernel (CTA 0), from a trace build of librtseg.

  python tools/pair_trace.py            # builds realtime-.../librtseg_trace.so with -DRTSEG_PAIR_TRACE, runs one C=64 pair
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = os.path.join(ROOT, "realtime-semantic-segmentation-pytorch_b200")
TRACE_LIB = os.path.join(PKG, "librtseg_trace.so")

EV = ["tma_issued", "x_landed", "s1acc_free", "s2_issue", "A_start", "A_end", "B_s2ready", "B_start", "B_written", "B_synced", "store_done"]


def build():
    from rtseg_b200 import build as b
    objs = []
    for src in b.SOURCES:
        obj = os.path.join(PKG, "build", src.replace(".cu", ".o"))
        if src == "pair_tc.cu":
            obj = os.path.join(PKG, "build", "pair_tc_trace.o")
            subprocess.check_call([b._nvcc(), *b.NVCC_FLAGS, "-DRTSEG_PAIR_TRACE", "-c", os.path.join(b.CSRC, src), "-o", obj])
        objs.append(obj)
    subprocess.check_call([b._nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", TRACE_LIB, *objs])


def main():
    if "--build" in sys.argv:
        build()
        return
    os.environ["RTSEG_LIB_PATH"] = TRACE_LIB
    import torch
    from rtseg_b200 import _lib, nets, ops
    c = int(os.environ.get("C", "64"))
    h, w = (256, 512) if c == 64 else (512, 1024)
    blk = nets.NonBt1DBlock(c, 1).eval().cuda()
    x = torch.randn((16, c, h, w), device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = "--res" in sys.argv
    fa = blk.conv[2 if res else 0]._folded()[0]
    fb = blk.conv[3].folded(blk.bn_act[0], blk.bn_act[1].activation)[0] if res else blk.conv[1]._folded()[0]
    with torch.no_grad():
        for _ in range(3):
            ops.conv_pair(fa, fb, x, res=x if res else None)
    torch.cuda.synchronize()
    lib = _lib.load()
    lib.rtseg_pair_trace_read.restype = C.c_int
    buf = (C.c_longlong * (64 * 16))()
    assert lib.rtseg_pair_trace_read(buf) == 0
    t0 = buf[0]
    print("tile " + " ".join(f"{e:>11s}" for e in EV))
    for t in range(0, 40):
        print(f"{t:4d} " + " ".join(f"{buf[t * 16 + e] - t0:11d}" for e in range(len(EV))))


if __name__ == "__main__":
    main()
