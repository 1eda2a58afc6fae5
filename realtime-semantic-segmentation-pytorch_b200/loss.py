"""OHEM cross-entropy on the rtseg kernels — drop-in for reference core/loss.py:6-20,55-71."""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib, ops


class _OhemFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, thresh: float, ignore_index: int):
        ops._require_cuda(logits, "OhemCELoss")
        if logits.dtype not in (torch.float32, torch.bfloat16):
            logits = logits.float()  # fp16 autocast outputs: CE is an fp32 op under autocast anyway
        n_, c_, h_, w_ = logits.shape
        pitch = logits.stride(3) if logits.dim() == 4 else 0
        if logits.is_contiguous():
            layout = 0
        elif ops.is_nhwc(logits):
            layout = 1
        elif (logits.stride(1) == 1 and pitch > c_ and logits.stride(2) == w_ * pitch and logits.stride(0) == h_ * w_ * pitch
              and pitch < (1 << 20)):
            layout = 1 | (pitch << 8)   # the first c channels of a wider (16-byte aligned) NHWC buffer, read in place
        else:
            logits, layout = logits.contiguous(), 0
        if labels.dtype != torch.int64:
            labels = labels.long()
        labels = labels.contiguous()
        n, c, h, w = logits.shape
        if tuple(labels.shape) != (n, h, w):
            raise ValueError(f"labels {tuple(labels.shape)} do not match logits {tuple(logits.shape)}")
        lib = _lib.load()
        dev = logits.device
        loss_px = torch.empty(n * h * w, dtype=torch.float32, device=dev)
        ws = torch.empty(lib.rtseg_ohem_workspace_bytes(), dtype=torch.uint8, device=dev)
        out = torch.empty(4, dtype=torch.float32, device=dev)
        _lib.check(lib.rtseg_ohem_fwd(n, c, h, w, layout, ops._dtype_code(logits.dtype), logits.data_ptr(),
                                      labels.data_ptr(), int(ignore_index), float(thresh), loss_px.data_ptr(),
                                      ws.data_ptr(), out.data_ptr(), ops._stream()), "rtseg_ohem_fwd")
        ctx.save_for_backward(logits, labels, loss_px, ws)
        ctx.meta = (layout, int(ignore_index))
        ctx.stats = out
        return out[0].clone()

    @staticmethod
    def backward(ctx, grad_out):
        logits, labels, loss_px, ws = ctx.saved_tensors
        layout, ignore_index = ctx.meta
        n, c, h, w = logits.shape
        if layout >> 8:   # same padded NHWC geometry as the logits; the kernel zeroes the padding channels
            pitch = layout >> 8
            grad = torch.empty((n, h, w, pitch), dtype=logits.dtype, device=logits.device).permute(0, 3, 1, 2)[:, :c]
        else:
            grad = torch.empty_like(logits)  # preserves the NCHW / channels_last layout
        g = grad_out.detach().to(device=logits.device, dtype=torch.float32).reshape(1).contiguous()
        _lib.check(_lib.load().rtseg_ohem_bwd(n, c, h, w, layout, ops._dtype_code(logits.dtype), logits.data_ptr(),
                                              labels.data_ptr(), ignore_index, loss_px.data_ptr(), ws.data_ptr(),
                                              g.data_ptr(), grad.data_ptr(), ops._stream()), "rtseg_ohem_bwd")
        return grad, None, None, None


class OhemCELoss(nn.Module):
    """mean of the per-pixel CE losses above -log(thresh), or of the top n_valid//16 if fewer qualify.

    Same constructor and call signature as the reference (core/loss.py:6-20); unlike it, the threshold
    is a Python float, so constructing the module does not need a GPU.
    """

    def __init__(self, thresh, ignore_index=255):
        super().__init__()
        self.thresh = float(-torch.log(torch.tensor(thresh, dtype=torch.float)))  # fp32 value, core/loss.py:9
        self.ignore_index = ignore_index

    def forward(self, logits, labels):
        return _OhemFunction.apply(logits, labels, self.thresh, self.ignore_index)


def loss_with_aux_heads(loss_fn, preds, preds_aux, masks, aux_coef=None):
    """Main loss + coefficient-weighted loss of every auxiliary head against the label map resized ('nearest') to that
    head's resolution — the `config.use_aux` branch of reference core/seg_trainer.py:53-65, with the resize on the
    rtseg kernel (int64 in, int64 out) instead of a float round trip through F.interpolate."""
    loss = loss_fn(preds, masks)
    if aux_coef is None:
        aux_coef = [1.0] * len(preds_aux)
    elif len(preds_aux) != len(aux_coef):
        raise ValueError("Auxiliary loss coefficient length does not match.")
    for coef, aux in zip(aux_coef, preds_aux):
        loss = loss + float(coef) * loss_fn(aux, ops.resize_labels_nearest(masks, aux.shape[2:]))
    return loss


def get_loss_fn(config, device):
    """Factory with the reference's contract (core/loss.py:55-71)."""
    if config.loss_type == "ohem":
        if config.class_weights is not None:
            raise NotImplementedError("class_weights are ignored by the reference's OHEM branch and unsupported here")
        return OhemCELoss(thresh=config.ohem_thrs, ignore_index=config.ignore_index)
    if config.loss_type == "ce":
        weights = None if config.class_weights is None else torch.Tensor(config.class_weights).to(device)
        return nn.CrossEntropyLoss(ignore_index=config.ignore_index, reduction=config.reduction, weight=weights)
    raise NotImplementedError(f"Unsupport loss type: {config.loss_type}")
