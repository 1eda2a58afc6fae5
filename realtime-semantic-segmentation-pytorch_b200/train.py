"""Training-mode forward/backward of the conv + BatchNorm + activation blocks on the rtseg kernels.

In training the reference runs Conv2d -> BatchNorm2d (batch statistics) -> activation as separate autograd
nodes (models/modules.py:73-85).  Here the same two nodes exist as `torch.autograd.Function`s whose forward and
backward are librtseg launches — autograd is only the tape:

* `_ConvFn`: forward = the inference convolution kernel without an epilogue; data gradient = the same kernel
  with transposed / flipped weights (a stride-2 3x3 convolution's data gradient is the 3x3 stride-2 transposed
  convolution and vice versa); weight gradient = `rtseg_conv_wgrad`; bias gradient = a per-channel sum.
* `_BNActFn`: batch statistics (`rtseg_bn_stats`, all-reduced over ranks for `nn.SyncBatchNorm`), running-stat
  update, normalise + activation in one pass (`rtseg_affine_act`), optionally on `y + res` (NonBt1DBlock's
  `+= x` before its last BatchNorm, models/erfnet.py:76-82); backward = one reduction + one apply pass.
* `_MaxPoolFn`: InitialBlock's pool branch.
* `_DWConvFn`: depthwise convolutions (DWConvBNAct, reference models/modules.py:46-59): forward = the inference
  depthwise kernel, data gradient = `rtseg_conv_dw_bwd_data`, weight gradient = `rtseg_conv_dw_wgrad`.
* PReLU (reference models/modules.py:119): the slope rides through `_BNActFn` / `_ActFn`; its gradient is
  `rtseg_prelu_slope_grad`.
* `_BilinearFn`, `_AdaptiveAvgPoolFn`: the resize / pooling steps of Fast-SCNN's pyramid pooling, feature fusion
  and output head (reference models/modules.py:134-158, fastscnn.py:30, 88) with their adjoint kernels.

Nothing here falls back to ATen convolutions or cuDNN; unsupported cases (grouped convolutions other than
depthwise with multiplier 1, glu / softmax blocks) raise.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib, ops
from ._lib import ConvDesc


WGRAD_TC = os.environ.get("RTSEG_WGRAD_TC", "1") != "0"   # tcgen05 weight gradient where it applies (else mma.sync / CUDA cores)


class _Spec:
    """Quacks like an nn.Conv2d / nn.ConvTranspose2d for ops.fold_conv, around a bare weight tensor."""

    def __init__(self, weight, bias, stride, dilation, padding, transposed, groups=1):
        self.weight, self.bias = weight, bias
        self.kernel_size = tuple(weight.shape[2:])
        self.stride, self.dilation, self.padding = (stride, stride), (dilation, dilation), tuple(padding)
        self.transposed = transposed
        self.groups = groups
        self.padding_mode = "zeros"
        self.output_padding = (1, 1)
        if transposed:
            self.in_channels, self.out_channels = weight.shape[0], weight.shape[1]
        else:
            self.out_channels, self.in_channels = weight.shape[0], weight.shape[1] * groups


def _conv_raw(x, weight, bias, stride, dilation, padding, transposed):
    """y = conv(x) (+ bias) through the inference kernels, no BatchNorm, no activation."""
    f, _ = ops.fold_conv(_Spec(weight.detach(), None if bias is None else bias.detach(), stride, dilation, padding,
                               transposed))
    if (not transposed and f.cin <= 3 and f.cout <= 32 and (f.kh, f.kw, f.stride, f.dil) == (3, 3, 2, 1)
            and tuple(padding) == (1, 1) and x.shape[1] == f.cin):
        # the image stem: the (tensor-core) stem kernel without pooling, BatchNorm or activation — same operand
        # rounding as in inference instead of the CUDA-core catch-all for 3 input channels
        return ops.stem(f, x, pool=False, out_dtype=x.dtype)
    return ops.conv(f, x)


def _zeros64(c, device):
    return torch.zeros(c, dtype=torch.float64, device=device)


def _channel_sums(t, c, res=None, want_sq=True):
    """Per-channel sum (and sum of squares) of an NHWC tensor (+ res) over all pixels, fp64."""
    n, ct, h, w = t.shape
    s = _zeros64(c, t.device)
    sq = _zeros64(c, t.device) if want_sq else None
    _lib.check(_lib.load().rtseg_bn_stats(n * h * w, c, ops._dtype_code(t.dtype), t.data_ptr(), ct, 0,
                                          ops._ptr(res), res.shape[1] if res is not None else 0, 0, s.data_ptr(),
                                          ops._ptr(sq), ops._stream()), "rtseg_bn_stats")
    return s, sq


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, dilation, padding, transposed):
        x = ops.to_nhwc(x)
        ctx.save_for_backward(x, weight)
        ctx.meta = (stride, dilation, tuple(padding), transposed, bias is not None)
        return _conv_raw(x, weight, bias, stride, dilation, padding, transposed)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        stride, dilation, padding, transposed, has_bias = ctx.meta
        dy = ops.to_nhwc(dy.to(x.dtype))
        n, _, h, w = x.shape
        kh, kw = weight.shape[2:]
        lib = _lib.load()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if not transposed and stride == 1:
                wt = weight.detach().flip(2, 3).transpose(0, 1)                      # (cin, cout, kh, kw): a conv cout -> cin
                pad = (dilation * (kh - 1) - padding[0], dilation * (kw - 1) - padding[1])
                dx = _conv_raw(dy, wt, None, 1, dilation, pad, False)
            elif not transposed:   # 3x3 stride 2 pad 1 on even sizes: the adjoint is the k3 s2 p1 op1 transposed conv
                if not (stride == 2 and (kh, kw) == (3, 3) and padding == (1, 1) and dilation == 1 and h % 2 == 0 and w % 2 == 0):
                    raise NotImplementedError("data gradient of this strided convolution has no kernel")
                dx = _conv_raw(dy, weight, None, 2, 1, (1, 1), True)
            else:                  # transposed conv: the adjoint is the stride-2 convolution with the same weight tensor
                dx = _conv_raw(dy, weight, None, 2, 1, (1, 1), False)
        if ctx.needs_input_grad[1]:
            # a transposed conv's weight gradient is that of its adjoint convolution (input := dy, output gradient := x)
            a, g = (dy, x) if transposed else (x, dy)
            cin, cout = a.shape[1], g.shape[1]
            d = ConvDesc(n if not transposed else dy.shape[0], a.shape[2], a.shape[3], cin, cout, kh, kw, stride, dilation,
                         padding[0], padding[1], 1, ops._dtype_code(a.dtype), 0, 0.0, 0.0, 0, 0, cin, 0, cout, 0, 0, 0)
            acc = torch.zeros((kh, kw, cin, cout), dtype=torch.float32, device=x.device)
            if WGRAD_TC and lib.rtseg_conv_wgrad_tc_supported(C.byref(d)):
                _lib.check(lib.rtseg_conv_wgrad_tc(C.byref(d), a.data_ptr(), g.data_ptr(), acc.data_ptr(), ops._stream()),
                           "rtseg_conv_wgrad_tc")
            else:
                _lib.check(lib.rtseg_conv_wgrad(C.byref(d), a.data_ptr(), g.data_ptr(), acc.data_ptr(), ops._stream()),
                           "rtseg_conv_wgrad")
            dw = acc.permute(3, 2, 0, 1).contiguous().to(weight.dtype)               # both layouts: (g-ch, a-ch, kh, kw)
        if has_bias and ctx.needs_input_grad[2]:
            db = _channel_sums(dy, dy.shape[1], want_sq=False)[0].to(weight.dtype)
        return dx, dw, db, None, None, None, None


class _DWConvFn(torch.autograd.Function):
    """Depthwise convolution (groups = cin = cout): reference models/modules.py:46-59 under autograd."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, dilation, padding):
        x = ops.to_nhwc(x)
        c = weight.shape[0]
        ctx.save_for_backward(x, weight)
        ctx.meta = (stride, dilation, tuple(padding), bias is not None)
        f, _ = ops.fold_conv(_Spec(weight.detach(), None if bias is None else bias.detach(), stride, dilation, padding, False,
                                   groups=c))
        return ops.conv(f, x)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        stride, dilation, padding, has_bias = ctx.meta
        dy = ops.to_nhwc(dy.to(x.dtype))
        n, c, h, w = x.shape
        kh, kw = weight.shape[2:]
        lib = _lib.load()
        d = ConvDesc(n, h, w, c, c, kh, kw, stride, dilation, padding[0], padding[1], c, ops._dtype_code(x.dtype), 0, 0.0, 0.0,
                     0, 0, c, 0, c, 0, 0, 0)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            w32 = weight.detach()[:, 0].permute(1, 2, 0).float().contiguous()          # [kh][kw][c]
            dx = torch.empty_like(x, memory_format=torch.channels_last)
            _lib.check(lib.rtseg_conv_dw_bwd_data(C.byref(d), dy.data_ptr(), w32.data_ptr(), dx.data_ptr(), ops._stream()),
                       "rtseg_conv_dw_bwd_data")
        if ctx.needs_input_grad[1]:
            acc = torch.zeros((kh, kw, c), dtype=torch.float32, device=x.device)
            _lib.check(lib.rtseg_conv_dw_wgrad(C.byref(d), x.data_ptr(), dy.data_ptr(), acc.data_ptr(), ops._stream()),
                       "rtseg_conv_dw_wgrad")
            dw = acc.permute(2, 0, 1).unsqueeze(1).contiguous().to(weight.dtype)
        if has_bias and ctx.needs_input_grad[2]:
            db = _channel_sums(dy, c, want_sq=False)[0].to(weight.dtype)
        return dx, dw, db, None, None, None


def _slope32(prelu_w, c):
    """The PReLU weight (1 or c parameters) as c contiguous fp32 slopes."""
    s = prelu_w.detach().float().reshape(-1)
    if s.numel() not in (1, c):
        raise ValueError("PReLU num_parameters must be 1 or the number of channels")
    return s.expand(c).contiguous()


def _slope_grad(prelu_w, dz, y, res, scale, shift):
    """d loss / d PReLU weight: per-channel sums of dz * v * [v <= 0] (summed over channels for a single slope)."""
    n, c, h, w = y.shape
    acc = _zeros64(c, y.device)
    _lib.check(_lib.load().rtseg_prelu_slope_grad(n * h * w, c, ops._dtype_code(y.dtype), dz.data_ptr(), c, 0, y.data_ptr(), c, 0,
                                                  ops._ptr(res), c if res is not None else 0, 0, ops._ptr(scale), ops._ptr(shift),
                                                  acc.data_ptr(), ops._stream()), "rtseg_prelu_slope_grad")
    if prelu_w.numel() == 1:
        acc = acc.sum()
    return acc.reshape(prelu_w.shape).to(prelu_w.dtype)


class _BNActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, res, gamma, beta, prelu_w, bn, act_module):
        y = ops.to_nhwc(y)
        n, c, h, w = y.shape
        if res is not None:
            res = ops.to_nhwc(res)
            if res.shape != y.shape or res.dtype != y.dtype:
                raise ValueError("residual must match the convolution output")
        code, p0, p1, _, post = ops.act_spec(act_module)
        if post is not None:
            raise NotImplementedError("training through glu / softmax blocks is not built")
        slope = _slope32(prelu_w, c) if prelu_w is not None else None
        dev = y.device
        lib = _lib.load()
        npix = n * h * w
        sync = (getattr(bn, "rtseg_sync", isinstance(bn, nn.SyncBatchNorm)) and dist.is_available() and dist.is_initialized()
                and dist.get_world_size() > 1)
        # [sum | sumsq | count] in one fp64 buffer: one all-reduce under SyncBatchNorm (reference utils/parallel.py:29-32)
        st64 = torch.zeros(2 * c + 1, dtype=torch.float64, device=dev)
        _lib.check(lib.rtseg_bn_stats(npix, c, ops._dtype_code(y.dtype), y.data_ptr(), c, 0, ops._ptr(res),
                                      c if res is not None else 0, 0, st64.data_ptr(), st64.data_ptr() + 8 * c, ops._stream()),
                   "rtseg_bn_stats")
        count = None
        if sync:
            st64[2 * c:].fill_(float(npix))
            dist.all_reduce(st64)
            count = st64[2 * c:]
        running = bn.track_running_stats and bn.running_mean is not None
        fast = (bn.momentum is not None and all(t is None or (t.dtype == torch.float32 and t.is_cuda and t.is_contiguous())
                                                for t in (gamma, beta, bn.running_mean if running else None,
                                                          bn.running_var if running else None)))
        if fast:    # mean / invstd / scale / shift / running statistics in one launch
            out32 = torch.empty(5 * c, dtype=torch.float32, device=dev)
            _lib.check(lib.rtseg_bn_finalize(c, st64.data_ptr(), ops._ptr(count), float(npix), ops._ptr(gamma), ops._ptr(beta),
                                             float(bn.eps), float(bn.momentum), ops._ptr(bn.running_mean) if running else None,
                                             ops._ptr(bn.running_var) if running else None, out32.data_ptr(), ops._stream()),
                       "rtseg_bn_finalize")
            mean32, invstd32, scale, shift, kk = (out32[i * c:(i + 1) * c] for i in range(5))
            if running:
                with torch.no_grad():
                    bn.num_batches_tracked += 1
        else:
            s_, sq = st64[:c], st64[c:2 * c]
            cnt = count if sync else torch.full((1,), float(npix), dtype=torch.float64, device=dev)
            mean = s_ / cnt
            var = (sq / cnt - mean * mean).clamp_(min=0.0)
            with torch.no_grad():
                if running:
                    m = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked + 1)
                    unbiased = var * (cnt / (cnt - 1).clamp_(min=1.0))
                    bn.running_mean.mul_(1 - m).add_(mean.to(bn.running_mean.dtype), alpha=m)
                    bn.running_var.mul_(1 - m).add_(unbiased.to(bn.running_var.dtype), alpha=m)
                    bn.num_batches_tracked += 1
            invstd = torch.rsqrt(var + bn.eps)
            g64 = gamma.detach().double() if gamma is not None else torch.ones(c, dtype=torch.float64, device=dev)
            b64 = beta.detach().double() if beta is not None else torch.zeros(c, dtype=torch.float64, device=dev)
            scale = (g64 * invstd).float()
            shift = (b64 - mean * g64 * invstd).float()
            mean32, invstd32, kk = mean.float(), invstd.float(), (g64 * invstd).float()
        z = torch.empty_like(y, memory_format=torch.channels_last)
        _lib.check(_lib.load().rtseg_affine_act(n * h * w, c, ops._dtype_code(y.dtype), y.data_ptr(), c, 0, ops._ptr(res),
                                                c if res is not None else 0, 0, scale.data_ptr(), shift.data_ptr(), code,
                                                p0, p1, ops._ptr(slope), z.data_ptr(), c, 0, ops._stream()), "rtseg_affine_act")
        ctx.save_for_backward(y, res, scale, shift, mean32, invstd32, kk, slope, prelu_w)
        ctx.meta = (code, p0, p1, sync, float(npix) if not sync else None, count if sync else None,
                    gamma is not None, beta is not None, gamma.dtype if gamma is not None else torch.float32)
        return z

    @staticmethod
    def backward(ctx, dz):
        y, res, scale, shift, mean, invstd, k, slope, prelu_w = ctx.saved_tensors
        code, p0, p1, sync, count_f, count_t, has_g, has_b, pdtype = ctx.meta
        dz = ops.to_nhwc(dz.to(y.dtype))
        n, c, h, w = y.shape
        lib = _lib.load()
        dt, st = ops._dtype_code(y.dtype), ops._stream()
        rc = c if res is not None else 0
        local = torch.zeros(2 * c, dtype=torch.float64, device=y.device)     # [sum g | sum g * xhat]
        _lib.check(lib.rtseg_bn_act_bwd_reduce(n * h * w, c, dt, dz.data_ptr(), c, 0, y.data_ptr(), c, 0, ops._ptr(res), rc, 0,
                                               scale.data_ptr(), shift.data_ptr(), mean.data_ptr(), invstd.data_ptr(), code,
                                               p0, p1, ops._ptr(slope), local.data_ptr(), local.data_ptr() + 8 * c, st),
                   "rtseg_bn_act_bwd_reduce")
        glob = local
        if sync:     # the normalisation's own gradient needs the sums over every rank's pixels
            glob = local.clone()
            dist.all_reduce(glob)
        out32 = torch.empty(4 * c, dtype=torch.float32, device=y.device)
        _lib.check(lib.rtseg_bn_bwd_finalize(c, local.data_ptr(), glob.data_ptr(), ops._ptr(count_t), count_f or 0.0,
                                             out32.data_ptr(), st), "rtseg_bn_bwd_finalize")
        mg, mgx = out32[:c], out32[c:2 * c]
        # parameter gradients from this rank's sums: DDP averages them itself
        dgamma = out32[2 * c:3 * c].to(pdtype) if has_g else None
        dbeta = out32[3 * c:].to(pdtype) if has_b else None
        du = torch.empty_like(y, memory_format=torch.channels_last)
        _lib.check(lib.rtseg_bn_act_bwd_apply(n * h * w, c, dt, dz.data_ptr(), c, 0, y.data_ptr(), c, 0, ops._ptr(res), rc, 0,
                                              scale.data_ptr(), shift.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                              k.data_ptr(), mg.data_ptr(), mgx.data_ptr(), code, p0, p1, ops._ptr(slope),
                                              du.data_ptr(), c, 0, st), "rtseg_bn_act_bwd_apply")
        dslope = None
        if prelu_w is not None and ctx.needs_input_grad[4]:
            dslope = _slope_grad(prelu_w, dz, y, res, scale, shift)
        return du, (du if res is not None else None), dgamma, dbeta, dslope, None, None


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = ops.to_nhwc(x)
        n, c, h, w = x.shape
        out = ops.empty_nhwc(n, c, (h - 1) // 2 + 1, (w - 1) // 2 + 1, x.dtype, x.device)
        ops.maxpool3x3s2(x, out, y_coff=0)
        ctx.save_for_backward(x)
        return out

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dy = ops.to_nhwc(dy.to(x.dtype))
        n, c, h, w = x.shape
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        _lib.check(_lib.load().rtseg_maxpool3x3s2_bwd(n, h, w, c, ops._dtype_code(x.dtype), x.data_ptr(), c, 0, dy.data_ptr(),
                                                      c, 0, dx.data_ptr(), c, 0, ops._stream()), "rtseg_maxpool3x3s2_bwd")
        return dx


class _ActFn(torch.autograd.Function):
    """act(x + res) without a BatchNorm in front (reference enet.py:165-181 `self.act(x_left + x_right)`)."""

    @staticmethod
    def forward(ctx, x, res, prelu_w, act_module):
        x = ops.to_nhwc(x)
        n, c, h, w = x.shape
        if res is not None:
            res = ops.to_nhwc(res)
            if res.shape != x.shape or res.dtype != x.dtype:
                raise ValueError("residual must match the activation input")
        code, p0, p1, _, post = ops.act_spec(act_module)
        if post is not None:
            raise NotImplementedError("training through glu / softmax is not built")
        slope = _slope32(prelu_w, c) if prelu_w is not None else None
        z = torch.empty_like(x, memory_format=torch.channels_last)
        _lib.check(_lib.load().rtseg_affine_act(n * h * w, c, ops._dtype_code(x.dtype), x.data_ptr(), c, 0, ops._ptr(res),
                                                c if res is not None else 0, 0, None, None, code, p0, p1, ops._ptr(slope),
                                                z.data_ptr(), c, 0, ops._stream()), "rtseg_affine_act")
        ctx.save_for_backward(x, res, slope, prelu_w)
        ctx.meta = (code, p0, p1)
        return z

    @staticmethod
    def backward(ctx, dz):
        x, res, slope, prelu_w = ctx.saved_tensors
        code, p0, p1 = ctx.meta
        dz = ops.to_nhwc(dz.to(x.dtype))
        n, c, h, w = x.shape
        # du = dz * act'(u): the BatchNorm backward pass with scale 1, shift 0 and no statistics terms
        one = torch.ones(c, dtype=torch.float32, device=x.device)
        zero = torch.zeros(c, dtype=torch.float32, device=x.device)
        du = torch.empty_like(x, memory_format=torch.channels_last)
        rc = c if res is not None else 0
        _lib.check(_lib.load().rtseg_bn_act_bwd_apply(n * h * w, c, ops._dtype_code(x.dtype), dz.data_ptr(), c, 0, x.data_ptr(), c,
                                                      0, ops._ptr(res), rc, 0, one.data_ptr(), zero.data_ptr(), zero.data_ptr(),
                                                      zero.data_ptr(), one.data_ptr(), zero.data_ptr(), zero.data_ptr(), code,
                                                      p0, p1, ops._ptr(slope), du.data_ptr(), c, 0, ops._stream()),
                   "rtseg_bn_act_bwd_apply")
        dslope = None
        if prelu_w is not None and ctx.needs_input_grad[2]:
            dslope = _slope_grad(prelu_w, dz, x, res, None, None)
        return du, (du if res is not None else None), dslope, None


class _BilinearFn(torch.autograd.Function):
    """F.interpolate(x, size, mode='bilinear', align_corners=...) (reference fastscnn.py:30, 88; modules.py:153-156)."""

    @staticmethod
    def forward(ctx, x, size, align_corners):
        x = ops.to_nhwc(x)
        ctx.meta = (tuple(x.shape), int(size[0]), int(size[1]), bool(align_corners))
        return ops.upsample_bilinear(x, size, align_corners)

    @staticmethod
    def backward(ctx, dy):
        (n, c, hi, wi), ho, wo, align = ctx.meta
        dy = ops.to_nhwc(dy)
        dx = ops.empty_nhwc(n, c, hi, wi, dy.dtype, dy.device)
        _lib.check(_lib.load().rtseg_upsample_bilinear_bwd(n, hi, wi, c, ho, wo, int(align), ops._dtype_code(dy.dtype),
                                                           dy.data_ptr(), c, 0, dx.data_ptr(), c, 0, ops._stream()),
                   "rtseg_upsample_bilinear_bwd")
        return dx, None, None


class _AdaptiveAvgPoolFn(torch.autograd.Function):
    """nn.AdaptiveAvgPool2d(p) (reference models/modules.py:145-149)."""

    @staticmethod
    def forward(ctx, x, p):
        x = ops.to_nhwc(x)
        ctx.meta = (tuple(x.shape), int(p))
        return ops.adaptive_avgpool(x, int(p))

    @staticmethod
    def backward(ctx, dy):
        (n, c, h, w), p = ctx.meta
        dy = ops.to_nhwc(dy)
        dx = ops.empty_nhwc(n, c, h, w, dy.dtype, dy.device)
        _lib.check(_lib.load().rtseg_adaptive_avgpool_bwd(n, h, w, c, p, ops._dtype_code(dy.dtype), dy.data_ptr(), dx.data_ptr(),
                                                          c, 0, ops._stream()), "rtseg_adaptive_avgpool_bwd")
        return dx, None


class _MaxPool2x2IdxFn(torch.autograd.Function):
    """nn.MaxPool2d(2, 2, return_indices=True) (reference models/enet.py:130-131): the adjoint is the unpooling kernel."""

    @staticmethod
    def forward(ctx, x):
        x = ops.to_nhwc(x)
        pooled, idx = ops.maxpool2x2_with_indices(x)
        ctx.save_for_backward(idx)
        ctx.size = (x.shape[2], x.shape[3])
        ctx.mark_non_differentiable(idx)
        return pooled, idx

    @staticmethod
    def backward(ctx, dy, _didx):
        (idx,) = ctx.saved_tensors
        return ops.maxunpool2x2(ops.to_nhwc(dy), idx, ctx.size)


class _MaxUnpoolFn(torch.autograd.Function):
    """nn.MaxUnpool2d(2, 2)(x, indices) (reference models/enet.py:176-178)."""

    @staticmethod
    def forward(ctx, x, idx, size):
        x = ops.to_nhwc(x)
        ctx.save_for_backward(idx)
        ctx.shape = tuple(x.shape)
        return ops.maxunpool2x2(x, idx, size)

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        n, c, hp, wp = ctx.shape
        dy = ops.to_nhwc(dy)
        dx = ops.empty_nhwc(n, c, hp, wp, dy.dtype, dy.device)
        _lib.check(_lib.load().rtseg_maxunpool2x2_bwd(n, hp, wp, c, dy.shape[2], dy.shape[3], ops._dtype_code(dy.dtype),
                                                      dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), ops._stream()),
                   "rtseg_maxunpool2x2_bwd")
        return dx, None, None


# ------------------------------------------------------------------ module-level entry points ----
def conv(module: nn.Module, x: torch.Tensor) -> torch.Tensor:
    """Training forward of a bare nn.Conv2d / nn.ConvTranspose2d (dense, zero padding)."""
    transposed = bool(getattr(module, "transposed", False))
    if module.groups != 1:
        if transposed or module.groups != module.in_channels or module.out_channels != module.in_channels \
                or module.stride[0] != module.stride[1] or module.dilation[0] != module.dilation[1]:
            raise NotImplementedError("training through grouped convolutions other than depthwise (multiplier 1) is not built")
        return _DWConvFn.apply(x, module.weight, module.bias, module.stride[0], module.dilation[0], tuple(module.padding))
    if transposed and not (tuple(module.kernel_size) == (3, 3) and tuple(module.stride) == (2, 2) and
                           tuple(module.padding) == (1, 1) and tuple(module.output_padding) == (1, 1)):
        raise NotImplementedError("only ConvTranspose2d(k=3, s=2, p=1, output_padding=1) can be trained through")
    if module.stride[0] != module.stride[1] or module.dilation[0] != module.dilation[1]:
        raise NotImplementedError("anisotropic stride/dilation is not supported")
    return _ConvFn.apply(x, module.weight, module.bias, module.stride[0], module.dilation[0], tuple(module.padding), transposed)


def bn_act(bn: nn.Module, act_module: Optional[nn.Module], y: torch.Tensor, res: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Training forward of BatchNorm2d (batch statistics) + activation on y (+ res)."""
    return _BNActFn.apply(y, res, bn.weight, bn.bias, _prelu_weight(act_module), bn, act_module)


def _prelu_weight(act_module):
    return act_module.weight if isinstance(act_module, nn.PReLU) else None


def act(act_module: Optional[nn.Module], x: torch.Tensor, res: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Training forward of a stand-alone activation on x (+ res)."""
    return _ActFn.apply(x, res, _prelu_weight(act_module), act_module)


def conv_bn_act(seq: nn.Module, x: torch.Tensor) -> torch.Tensor:
    """Training forward of a Sequential(conv, bn, Activation) block (ConvBNAct / PWConvBNAct / DeConvBNAct.up_conv)."""
    return bn_act(seq[1], seq[2].activation, conv(seq[0], x))


def maxpool3x3s2(x: torch.Tensor) -> torch.Tensor:
    return _MaxPoolFn.apply(x)


def upsample_bilinear(x: torch.Tensor, size, align_corners: bool) -> torch.Tensor:
    return _BilinearFn.apply(x, tuple(size), align_corners)


def adaptive_avgpool(x: torch.Tensor, p: int) -> torch.Tensor:
    return _AdaptiveAvgPoolFn.apply(x, p)


def maxpool2x2_with_indices(x: torch.Tensor):
    return _MaxPool2x2IdxFn.apply(x)


def maxunpool2x2(x: torch.Tensor, idx: torch.Tensor, size=None) -> torch.Tensor:
    return _MaxUnpoolFn.apply(x, idx, size)


def taping(*tensors) -> bool:
    """True when a call with these inputs has to leave autograd nodes behind."""
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


class _FirstChannelsFn(torch.autograd.Function):
    """z[:, :c] of a channel-padded NHWC tensor as a VIEW; the gradient comes back as the padded buffer itself when the
    consumer (OhemCELoss) produced it in the same padded geometry — no zero-fill + copy of a full-resolution tensor."""

    @staticmethod
    def forward(ctx, z, c):
        ctx.shape = tuple(z.shape)
        return z[:, :c]

    @staticmethod
    def backward(ctx, g):
        n, cp, h, w = ctx.shape
        if (g.dim() == 4 and g.stride(1) == 1 and g.stride(3) == cp and g.stride(2) == w * cp and g.stride(0) == h * w * cp
                and g.storage_offset() == 0 and g.untyped_storage().nbytes() >= n * h * w * cp * g.element_size()):
            return torch.as_strided(g, (n, cp, h, w), (h * w * cp, 1, w * cp, cp)), None   # padding already zero
        full = torch.zeros((n, cp, h, w), dtype=g.dtype, device=g.device).contiguous(memory_format=torch.channels_last)
        full[:, :g.shape[1]] = g
        return full, None


class _PaddedBN:
    """BatchNorm2d stand-in with `pad` extra channels (weight 1, bias 0): what _BNActFn needs, running statistics copied back."""

    def __init__(self, bn, pad):
        self.bn, self.pad = bn, pad
        self.eps, self.momentum, self.track_running_stats = bn.eps, bn.momentum, bn.track_running_stats
        self.rtseg_sync = isinstance(bn, nn.SyncBatchNorm)
        dev = bn.weight.device
        self.running_mean = self.running_var = None
        if bn.running_mean is not None:
            self.running_mean = torch.cat([bn.running_mean.detach().float(), torch.zeros(pad, device=dev)])
            self.running_var = torch.cat([bn.running_var.detach().float(), torch.ones(pad, device=dev)])
        self.num_batches_tracked = bn.num_batches_tracked
        self.weight = torch.cat([bn.weight, torch.ones(pad, device=dev, dtype=bn.weight.dtype)])
        self.bias = torch.cat([bn.bias, torch.zeros(pad, device=dev, dtype=bn.bias.dtype)])

    def commit(self):
        if self.running_mean is not None:
            c = self.bn.running_mean.numel()
            with torch.no_grad():
                self.bn.running_mean.copy_(self.running_mean[:c])
                self.bn.running_var.copy_(self.running_var[:c])


def conv_bn_act_aligned(seq: nn.Module, x: torch.Tensor, pad_input: bool = False, keep_padding: bool = False) -> torch.Tensor:
    """conv_bn_act for an output channel count that is not a multiple of 16 (the 19-class logits layer) on bf16
    activations: the layer is run as if it had ceil16(cout) outputs — zero weights, BatchNorm weight 1 / bias 0 for the
    extra channels — so that forward, data gradient, weight gradient and the BatchNorm passes all take the aligned
    (tensor-core / vector) kernels; the result is the first cout channels of that buffer, as a view."""
    import torch.nn.functional as F
    conv, bn, act = seq[0], seq[1], seq[2].activation
    cout = conv.out_channels
    cp = (cout + 15) // 16 * 16
    transposed = bool(getattr(conv, "transposed", False))
    if cp == cout or x.dtype != torch.bfloat16 or conv.groups != 1 or bn.weight is None or bn.bias is None:
        return conv_bn_act(seq, x)
    pad = cp - cout
    cin_pad = 0
    if pad_input and not transposed and conv.in_channels % 16:
        # the image stem: 3 input channels become 16 (zero channels, zero weights) so that the forward and the weight
        # gradient run on the tensor cores instead of the 3-channel catch-all kernels
        cin_pad = (conv.in_channels + 15) // 16 * 16 - conv.in_channels
        xp = torch.zeros((x.shape[0], conv.in_channels + cin_pad, x.shape[2], x.shape[3]), dtype=x.dtype,
                         device=x.device).contiguous(memory_format=torch.channels_last)
        xp[:, :conv.in_channels] = x
        x = xp
    w_p = F.pad(conv.weight, (0, 0, 0, 0, 0, pad) if transposed else (0, 0, 0, 0, 0, cin_pad, 0, pad))
    b_p = F.pad(conv.bias, (0, pad)) if conv.bias is not None else None
    y = _ConvFn.apply(x, w_p, b_p, conv.stride[0], conv.dilation[0], tuple(conv.padding), transposed)
    pbn = _PaddedBN(bn, pad)
    z = _BNActFn.apply(y, None, pbn.weight, pbn.bias, _prelu_weight(act), pbn, act)
    pbn.commit()
    # keep_padding: the caller goes on in the padded geometry (e.g. a vectorised resize of the 32-channel logits buffer)
    # and takes the first cout channels itself with first_channels()
    return z if keep_padding else _FirstChannelsFn.apply(z, cout)


def first_channels(z: torch.Tensor, c: int) -> torch.Tensor:
    """z[:, :c] of a channel-padded NHWC tensor as a view whose gradient comes back in the padded geometry."""
    return z if z.shape[1] == c else _FirstChannelsFn.apply(z, c)
