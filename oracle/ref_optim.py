"""Optimizer step and weight mirror — CPU restatement.  TEST INFRASTRUCTURE ONLY.

SGD as the reference configures it (utils/optimizer.py:8-12: torch.optim.SGD(lr, momentum, weight_decay), dampening 0,
no Nesterov) written out as its update rule, and ModelEmaV2.update (utils/model_ema.py:28-40).  Pinned in
tests/test_oracle_golden.py against torch.optim.SGD itself (torch is the reference's own dependency for this step).
"""
from __future__ import annotations

import torch


def sgd_step(params, grads, bufs, lr, momentum, weight_decay):
    """In place on plain tensors; `bufs[i]` is None before the first step (torch clones the gradient then)."""
    for i, (p, g) in enumerate(zip(params, grads)):
        g = g + weight_decay * p if weight_decay else g.clone()
        if momentum:
            bufs[i] = g.clone() if bufs[i] is None else bufs[i] * momentum + g
            g = bufs[i]
        p.sub_(lr * g)


def ema_update(ema_values, model_values, cur_itrs, total_itrs, use_ema=True):
    """utils/model_ema.py:28-40: for every state-dict entry e.copy_(decay * e + (1 - decay) * m) or e.copy_(m)."""
    decay = min(max(cur_itrs / total_itrs, 0), 1)
    for e, m in zip(ema_values, model_values):
        e.copy_(decay * e + (1. - decay) * m if use_ema else m)
