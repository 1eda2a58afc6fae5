#!/usr/bin/env python
"""ERFNet training step: fwd + OHEM loss + bwd + (DDP all-reduce) + SGD step.

    python tools/train_bench.py [--batch 8] [--size 1024] [--steps 5] [--fp32] [--profile]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_bench.py ...

Per rank: `batch` random 3 x size x size crops, labels randint(0,19) with 10 % ignore; bf16 activations with fp32
master weights (the reference uses fp16 autocast + GradScaler), OHEM thresh 0.7, SGD lr 0.01*world,
momentum 0.9, weight decay 1e-4 (reference utils/optimizer.py:8-12), SyncBatchNorm when world > 1
(utils/parallel.py:29-32).  Prints images/s (all ranks) and, with --profile, the kernel time table of one step.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import rtseg_b200  # noqa: E402,F401
from rtseg_b200 import loss as rloss  # noqa: E402
from rtseg_b200 import nets  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--fp32", action="store_true")
    ap.add_argument("--profile", action="store_true")
    args = ap.parse_args()
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.manual_seed(1)
    model = nets.ERFNet(num_class=19).cuda().train()
    if not args.fp32:
        model.set_act_dtype(torch.bfloat16)
    if world > 1:
        model = torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
    crit = rloss.OhemCELoss(0.7, 255)
    opt = torch.optim.SGD(model.parameters(), lr=0.01 * world, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator(device="cuda").manual_seed(rank)
    x = torch.randn((args.batch, 3, args.size, args.size), device="cuda", generator=g)
    y = torch.randint(0, 19, (args.batch, args.size, args.size), device="cuda", generator=g)
    y[torch.rand(y.shape, device="cuda", generator=g) < 0.1] = 255

    def step():
        opt.zero_grad(set_to_none=True)
        loss = crit(model(x), y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(args.warmup):
        l = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        l = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"ERFNet train step: {ms.item():.1f} ms, {world * args.batch / ms.item() * 1e3:.1f} images/s "
              f"({world} GPU(s), batch {args.batch}/GPU, {args.size}x{args.size}, {'fp32' if args.fp32 else 'bf16'}), "
              f"loss {l.item():.4f}, peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
    if args.profile and rank == 0:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step()
            torch.cuda.synchronize()
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=60))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
