"""Confusion-matrix mIoU on the rtseg kernels — drop-in for reference utils/metrics.py:4-7.

The reference returns `torchmetrics.JaccardIndex(task='multiclass', num_classes=C,
ignore_index=255, average='none')`; the trainer uses `.to / .update / .compute / .reset`
(core/seg_trainer.py:22,131,136,151).  State is one int64 [C, C] matrix, rows = target.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import ops


def iou_from_confmat(confmat: torch.Tensor) -> torch.Tensor:
    """Per-class IoU = diag / (rowsum + colsum - diag), 0 where the denominator is 0 (fp32)."""
    tp = confmat.diag()
    denom = confmat.sum(0) + confmat.sum(1) - tp
    tp_f, denom_f = tp.float(), denom.float()
    return torch.where(denom_f != 0, tp_f / denom_f, torch.zeros_like(tp_f))


class SegMetrics:
    def __init__(self, num_classes: int, ignore_index: int = 255, device=None):
        if not 0 < num_classes <= 32:
            raise NotImplementedError("the confusion-matrix kernel supports 1..32 classes")
        self.num_classes = num_classes
        self.ignore_index = ignore_index
        self.confmat = torch.zeros((num_classes, num_classes), dtype=torch.int64, device=device)

    def to(self, device):
        self.confmat = self.confmat.to(device)
        return self

    def reset(self):
        self.confmat.zero_()

    def update(self, preds: torch.Tensor, target: torch.Tensor):
        """preds: float [N, C, H, W] scores (argmax over dim 1 is taken) or integer [N, H, W] labels."""
        if preds.dim() == target.dim() + 1:
            preds = ops.argmax_channels(preds, torch.uint8)
        elif preds.dtype not in (torch.uint8, torch.int64):
            preds = preds.long()
        if target.dtype != torch.int64:
            target = target.long()
        ops.confmat_update(self.confmat, preds, target, self.ignore_index)

    def update_from_lowres_logits(self, logits: torch.Tensor, target: torch.Tensor, align_corners: bool = True):
        """Fused head: x8 bilinear + argmax + confusion matrix in one kernel (no full-res logits)."""
        ops.upsample_argmax(logits, target.shape[-2:], align_corners, target=target.contiguous(), confmat=self.confmat,
                            ignore_index=self.ignore_index, want_labels=False)

    def synced_confmat(self) -> torch.Tensor:
        """Sum over ranks: one NCCL all-reduce of C*C int64 (replaces torchmetrics' all-gather)."""
        cm = self.confmat
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            cm = cm.clone()
            dist.all_reduce(cm, op=dist.ReduceOp.SUM)
        return cm

    def compute(self) -> torch.Tensor:
        return iou_from_confmat(self.synced_confmat())


def get_seg_metrics(config, task="multiclass", reduction="none"):
    if task != "multiclass" or reduction != "none":
        raise NotImplementedError("only task='multiclass', average='none' (what the reference trainer uses)")
    return SegMetrics(config.num_class, config.ignore_index)
