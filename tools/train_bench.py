#!/usr/bin/env python
"""ERFNet training step: fwd + OHEM loss + bwd + (DDP all-reduce) + SGD step.

    python tools/train_bench.py [--batch 8] [--size 1024] [--steps 5] [--fp32] [--graph] [--profile]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_bench.py ...

Per rank: `batch` random 3 x size x size crops, labels randint(0,19) with 10 % ignore; bf16 activations with fp32
master weights (the reference uses fp16 autocast + GradScaler), OHEM thresh 0.7, SGD lr 0.01*world,
momentum 0.9, weight decay 1e-4 (reference utils/optimizer.py:8-12), SyncBatchNorm when world > 1
(utils/parallel.py:29-32).  `bench.py --workload erfnet_train` prints the contract's JSON line from `run()`.
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def build(batch, size, fp32, rank, local, world, model_name="erfnet"):
    import rtseg_b200  # noqa: F401
    from rtseg_b200 import loss as rloss
    from rtseg_b200 import nets
    torch.manual_seed(1)
    model = nets.MODEL_HUB[model_name](num_class=19).cuda().train()   # erfnet | fastscnn | dabnet | enet train on this path
    if not fp32:
        model.set_act_dtype(torch.bfloat16)
    if world > 1:
        model = torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
        side = torch.cuda.Stream()      # DDP built on a side stream: a requirement for capturing its step in a CUDA graph
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
        torch.cuda.current_stream().wait_stream(side)
    crit = rloss.OhemCELoss(0.7, 255)
    if os.environ.get("RTSEG_TORCH_SGD") == "1":     # A/B: the stock optimizer (~5 ATen kernels per parameter tensor)
        opt = torch.optim.SGD(model.parameters(), lr=0.01 * world, momentum=0.9, weight_decay=1e-4)
    else:                                             # one rtseg launch per step
        opt = rtseg_b200.optim.FusedSGD(model.parameters(), lr=0.01 * world, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator(device="cuda").manual_seed(rank)
    x = torch.randn((batch, 3, size, size), device="cuda", generator=g)
    y = torch.randint(0, 19, (batch, size, size), device="cuda", generator=g)
    y[torch.rand(y.shape, device="cuda", generator=g) < 0.1] = 255
    return model, crit, opt, x, y


def make_step(model, crit, opt, x, y, graph, world):
    def eager():
        opt.zero_grad(set_to_none=True)
        loss = crit(model(x), y)
        loss.backward()
        opt.step()
        return loss

    ddp_graph = os.environ.get("RTSEG_DDP_GRAPH", "1") == "1"
    if not graph or (world > 1 and not ddp_graph):
        return eager
    # One CUDA graph for the whole step.  Under DDP the graph also holds the NCCL collectives (the bucketed gradient
    # all-reduce and the 148 small SyncBatchNorm all-reduces): ~3,000 kernel launches and the collectives' launch
    # latency leave the host's critical path.  DDP rebuilds its buckets after the first iterations, hence 11 warm-ups.
    for _ in range(2):
        eager()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):   # warm-up on a side stream with the gradients kept allocated
        for _ in range(2 if world == 1 else 11):
            opt.zero_grad(set_to_none=False)
            crit(model(x), y).backward()
            opt.step()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):       # the whole step — ~3,000 launches — replays as one graph
        opt.zero_grad(set_to_none=False)
        static_loss = crit(model(x), y)
        static_loss.backward()
        opt.step()

    def replay():
        g.replay()
        return static_loss
    return replay


def run(batch=8, size=1024, steps=5, warmup=2, fp32=False, graph=True, model_name="erfnet"):
    """Times the step; returns a dict (rank 0) with device-timed and end-to-end images/s.  Call under torchrun for N > 1."""
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    own_pg = False
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")   # the watchdog thread must not touch a capturing stream
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        own_pg = True
    from rtseg_b200 import _lib
    model, crit, opt, x, y = build(batch, size, fp32, rank, local, world, model_name)
    l0 = _lib.launch_count()
    step = make_step(model, crit, opt, x, y, graph, world)
    for _ in range(max(1, warmup)):
        loss = step()
    torch.cuda.synchronize()
    launches_before = _lib.launch_count()
    graphed = getattr(step, "__name__", "") == "replay"
    if not graphed:   # count one eager step's librtseg launches
        step()
        torch.cuda.synchronize()
    per_step = (_lib.launch_count() - launches_before) if not graphed else None
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda", dtype=torch.float64)
    # end to end: pinned host images + labels in, loss value out, every step
    hx, hy = x.cpu().pin_memory(), y.cpu().pin_memory()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        x.copy_(hx, non_blocking=True)
        y.copy_(hy, non_blocking=True)
        lv = step().item()
    torch.cuda.synchronize()
    e2e = torch.tensor([(time.perf_counter() - t0) / steps], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
    out = {"ms_per_step": ms.item(), "images_per_s": world * batch / ms.item() * 1e3,
           "e2e_images_per_s": world * batch / e2e.item(), "h2d_bytes": hx.numel() * 4 + hy.numel() * 8, "d2h_bytes": 4,
           "loss": lv, "world": world, "batch": batch, "size": size, "graph": bool(graphed),
           "launches_per_step": per_step, "capture_launches": _lib.launch_count() - l0,
           "peak_mem_gib": torch.cuda.max_memory_allocated() / 2 ** 30, "model": model}
    if own_pg:
        dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--fp32", action="store_true")
    ap.add_argument("--model", default="erfnet", choices=["erfnet", "fastscnn", "dabnet", "enet"])
    ap.add_argument("--graph", action="store_true", help="capture the whole step in one CUDA graph (1 GPU)")
    ap.add_argument("--profile", action="store_true")
    args = ap.parse_args()
    r = run(args.batch, args.size, args.steps, args.warmup, args.fp32, args.graph, args.model)
    if int(os.environ.get("RANK", 0)) == 0:
        print(f"{args.model} train step: {r['ms_per_step']:.1f} ms, {r['images_per_s']:.1f} images/s device-timed, "
              f"{r['e2e_images_per_s']:.1f} images/s end to end ({r['world']} GPU(s), batch {r['batch']}/GPU, "
              f"{r['size']}x{r['size']}, {'fp32' if args.fp32 else 'bf16'}, graph={r['graph']}), loss {r['loss']:.4f}, "
              f"peak mem {r['peak_mem_gib']:.1f} GiB")
    if args.profile and int(os.environ.get("RANK", 0)) == 0 and not r["graph"]:
        from torch.profiler import ProfilerActivity, profile
        model, crit, opt, x, y = build(args.batch, args.size, args.fp32, 0, 0, 1, args.model)
        step = make_step(model, crit, opt, x, y, False, 1)
        step()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=60))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
