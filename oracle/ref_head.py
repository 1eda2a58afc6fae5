"""Segmentation head, confusion matrix and IoU — CPU restatements."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def upsample_argmax(logits_lowres: torch.Tensor, size, align_corners=True) -> torch.Tensor:
    """F.interpolate(bilinear) then preds.max(dim=1)[1] (reference models/fastscnn.py:30 +
    core/seg_trainer.py:172).  torch.max returns the first maximal index."""
    up = F.interpolate(logits_lowres.float(), tuple(size), mode="bilinear", align_corners=align_corners)
    return up.max(dim=1)[1]


def confusion_matrix(pred: np.ndarray, target: np.ndarray, num_classes: int, ignore_index: int = 255) -> np.ndarray:
    """torchmetrics multiclass confusion matrix as the reference configures it (utils/metrics.py:4-7):
    flatten, drop target == ignore_index, bincount(target * C + pred, C*C).reshape(C, C); rows = target."""
    pred = np.asarray(pred).reshape(-1).astype(np.int64)
    target = np.asarray(target).reshape(-1).astype(np.int64)
    keep = target != ignore_index
    idx = target[keep] * num_classes + pred[keep]
    return np.bincount(idx, minlength=num_classes * num_classes).reshape(num_classes, num_classes).astype(np.int64)


def confusion_matrix_loops(pred, target, num_classes: int, ignore_index: int = 255) -> np.ndarray:
    """Pure-Python double check of `confusion_matrix` (small inputs only)."""
    cm = [[0] * num_classes for _ in range(num_classes)]
    for p, t in zip(np.asarray(pred).reshape(-1).tolist(), np.asarray(target).reshape(-1).tolist()):
        if t != ignore_index:
            cm[t][p] += 1
    return np.array(cm, dtype=np.int64)


def iou_from_confmat(cm: np.ndarray) -> np.ndarray:
    """Per-class IoU: diag / (sum over rows + sum over cols - diag) in fp32, 0 where the denominator is 0
    (torchmetrics `_safe_divide(..., zero_division=0)` with average='none')."""
    cm = np.asarray(cm, dtype=np.int64)
    tp = np.diag(cm).astype(np.float32)
    denom = (cm.sum(0) + cm.sum(1) - np.diag(cm)).astype(np.float32)
    out = np.zeros_like(tp)
    nz = denom != 0
    out[nz] = tp[nz] / denom[nz]
    return out


def miou(cm: np.ndarray) -> float:
    """validate(): iou.mean() over ALL classes, absent ones count as 0 (reference core/seg_trainer.py:136-137)."""
    return float(iou_from_confmat(cm).mean(dtype=np.float32))


def colorize_blend(labels: np.ndarray, colormap, image=None, alpha: float = 0.0) -> np.ndarray:
    """Predict epilogue of reference core/seg_trainer.py:172-191: colormap[labels] as uint8 HWC and, with `image`,
    Pillow's Image.blend(image, pred, alpha) restated: (uint8)(in1 + alpha * (in2 - in1)) in float32, truncated
    (libImaging/Blend.c, interpolation branch 0 <= alpha <= 1).  Pinned against Pillow itself in the CPU tests."""
    lut = np.asarray(colormap, dtype=np.uint8)
    pred = lut[labels]
    if image is None:
        return pred
    in1, in2 = image.astype(np.int32), pred.astype(np.int32)
    t = in1.astype(np.float32) + np.float32(alpha) * (in2 - in1).astype(np.float32)
    return t.astype(np.int32).astype(np.uint8)
