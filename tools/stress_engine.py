#!/usr/bin/env python
"""Stress run for the inference engine at the benchmark shape: two CUDA-graph slots replayed concurrently on
their own streams (what bench.py's end-to-end phase does), many times, checking after every block of replays that
  * no launch faulted (the synchronise raises),
  * the barrier watchdog flag stayed clear (rtseg_tc_timeout_flag),
  * every class map equals the slot's first result (bit-exact run to run).

  python tools/stress_engine.py --workload dabnet --iters 300 [--batch 16] [--host]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="dabnet")
    ap.add_argument("--iters", type=int, default=300)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--host", action="store_true", help="also drive the pinned-host submit/collect path")
    args = ap.parse_args()
    import torch
    import rtseg_b200
    from rtseg_b200 import ops

    torch.manual_seed(1)
    model = rtseg_b200.nets.MODEL_HUB[args.workload](num_class=19).eval().cuda().set_act_dtype(torch.bfloat16)
    eng = rtseg_b200.InferenceEngine(model, batch=args.batch, height=1024, width=2048, use_graph=True, slots=2)
    g = torch.Generator(device="cuda").manual_seed(0)
    for s in eng.slots:
        s.x.copy_(torch.randn(s.x.shape, generator=g, device="cuda"))
    torch.cuda.synchronize()
    first = []
    for i in range(2):
        out = eng.run_device(i)           # enqueued on the slot's own stream
        torch.cuda.synchronize()
        first.append(out.clone())
    torch.cuda.synchronize()
    bad = flags = 0
    t0 = time.time()
    for it in range(args.iters):
        a, b = eng.run_device(0), eng.run_device(1)
        if it % 20 == 19:
            torch.cuda.synchronize()
            bad += int(not torch.equal(a, first[0])) + int(not torch.equal(b, first[1]))
            flags += ops.tc_timeout_flag()
    torch.cuda.synchronize()
    host_bad = 0
    if args.host:
        host_in = [s.x.cpu().pin_memory() for s in eng.slots]
        pending = []
        for it in range(max(20, args.iters // 10)):
            pending.append((it % 2, eng.submit(host_in[it % 2])))
            if len(pending) == 2:
                k, slot = pending.pop(0)
                host_bad += int(not torch.equal(eng.collect(slot), first[k].cpu()))
        while pending:
            k, slot = pending.pop(0)
            host_bad += int(not torch.equal(eng.collect(slot), first[k].cpu()))
        flags += ops.tc_timeout_flag()
    print(json.dumps({"workload": args.workload, "batch": args.batch, "concurrent_replays": 2 * args.iters,
                      "mismatching_outputs": bad, "host_path_mismatches": host_bad, "watchdog_flags": flags,
                      "seconds": round(time.time() - t0, 1)}))
    return 0 if bad == 0 and flags == 0 and host_bad == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
