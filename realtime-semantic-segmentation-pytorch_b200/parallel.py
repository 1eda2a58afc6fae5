"""One process per GPU — the contract of reference utils/parallel.py:12-53.

Inference shards by batch with no data-path collective; the only cross-rank traffic is the
C*C int64 confusion matrix (metrics.SegMetrics.synced_confmat) and, for training, DDP's
gradient all-reduce over NCCL/NVLink.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn.parallel import DistributedDataParallel


def is_parallel(model) -> bool:
    return type(model) in (nn.parallel.DataParallel, DistributedDataParallel)


def de_parallel(model):
    return model.module if is_parallel(model) else model


def set_device(config, rank):
    """reference utils/parallel.py:17-31; DataParallel (the non-DDP branch) is not carried over."""
    if config.DDP:
        torch.cuda.set_device(rank)
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group(backend="nccl", init_method="env://")
        device = torch.device("cuda", rank)
        config.gpu_num = dist.get_world_size()
    else:
        device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        config.gpu_num = 1
    config.num_workers = config.gpu_num * getattr(config, "base_workers", 0)
    return device


def parallel_model(config, model, rank, device):
    """reference utils/parallel.py:34-43 (SyncBN conversion + DDP wrap)."""
    if config.DDP:
        if getattr(config, "synBN", False):
            model = nn.SyncBatchNorm.convert_sync_batchnorm(model)
        return DistributedDataParallel(model.to(rank), device_ids=[rank], output_device=rank)
    return model.to(device)


def destroy_ddp_process(config):
    if config.DDP and dist.is_initialized():
        dist.destroy_process_group()


def sampler_set_epoch(config, loader, cur_epochs):
    if config.DDP:
        loader.sampler.set_epoch(cur_epochs)


def shard_range(total: int, rank: int, world: int):
    """Images [lo, hi) of a batch owned by `rank` (contiguous, sizes differ by at most one)."""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)
