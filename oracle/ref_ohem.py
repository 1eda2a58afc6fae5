"""OHEM cross entropy — CPU restatement of reference core/loss.py:6-20 (without its .cuda() call)."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def ohem_thresh(thresh: float) -> torch.Tensor:
    # core/loss.py:9 : -log(thresh) evaluated in fp32
    return -torch.log(torch.tensor(thresh, dtype=torch.float))


def ohem_ce_loss(logits: torch.Tensor, labels: torch.Tensor, thresh: float = 0.7, ignore_index: int = 255):
    """Returns (loss, branch) with branch 0 = 'loss > thresh' mean, 1 = top-k mean (core/loss.py:13-20)."""
    th = ohem_thresh(thresh)
    n_min = labels[labels != ignore_index].numel() // 16
    loss = F.cross_entropy(logits.float(), labels, ignore_index=ignore_index, reduction="none").view(-1)
    hard = loss[loss > th]
    branch = 0
    if hard.numel() < n_min:
        hard, _ = loss.topk(n_min)
        branch = 1
    return torch.mean(hard), branch
