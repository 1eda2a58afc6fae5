"""Drop-in replacements for the reference's building blocks (models/modules.py).

Same public names, constructor signatures, child layout and state-dict keys as
the reference classes, so that
  * `sys.modules['models.modules'] = rtseg_b200.modules` before `import models`
    makes all 36 unchanged model files of the reference build on these classes
    (SURVEY F5 — the reference module itself is never patched), and
  * README checkpoints, `deepcopy` (utils/model_ema.py:20), `state_dict()` zips
    (:30), `convert_sync_batchnorm` (utils/parallel.py:37) keep working.
What differs is `forward`: conv + eval-mode BatchNorm + activation run as ONE
sm_100a kernel launch through the C ABI instead of three ATen/cuDNN launches.

Inputs are NCHW-logical CUDA tensors (fp32 or bf16); outputs are NCHW-logical
tensors stored channels_last.  There is no CPU path and no cuDNN path behind
these modules: anything unsupported raises.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops

__all__ = ["conv3x3", "conv1x1", "channel_shuffle", "DSConvBNAct", "DWConvBNAct", "PWConvBNAct", "ConvBNAct",
           "DeConvBNAct", "Activation", "PyramidPoolingModule", "SegHead", "Conv2d"]


def _same_padding(kernel_size, dilation):
    """padding = (k-1)//2*dilation per dim — reference models/modules.py:49-52,76-79.

    Like the reference, anything that is neither int nor list/tuple is an error
    (there: UnboundLocalError; here a TypeError naming the cause).
    """
    if isinstance(kernel_size, (list, tuple)):
        return tuple((k - 1) // 2 * dilation for k in kernel_size[:2])
    if isinstance(kernel_size, int):
        return (kernel_size - 1) // 2 * dilation
    raise TypeError(f"kernel_size must be int, list or tuple, got {type(kernel_size).__name__}")


class _FoldCache:
    """Non-persistent cache of kernel-ready parameters, keyed on parameter versions.

    Never part of state_dict; rebuilt after load_state_dict / optimizer steps /
    .to() / .bfloat16() because those bump `_version` or change device/dtype.
    """

    def __init__(self):
        self.key = None
        self.value = None

    @staticmethod
    def _sig(tensors):
        return tuple((t.data_ptr(), t._version, t.device, t.dtype) if t is not None else None for t in tensors)

    def get(self, tensors, build):
        key = self._sig(tensors)
        if key != self.key:
            self.value = build()
            self.key = key
        return self.value


def _bn_tensors(bn):
    if bn is None:
        return []
    return [bn.weight, bn.bias, bn.running_mean, bn.running_var]


def _check_eval(mod: nn.Module, bn) -> None:
    if bn is not None and bn.training:
        raise NotImplementedError(
            f"{type(mod).__name__}: training-mode BatchNorm (batch statistics + backward) has no fused kernel for this "
            "call (depthwise blocks, slice / residual arguments); call .eval() — there is deliberately no cuDNN "
            "fallback on this path")


def _plain_call(res, res_mode, out, x_coff, y_coff, act2=0, r_coff=0) -> bool:
    return res is None and res_mode == ops.RES_NONE and out is None and x_coff == 0 and y_coff == 0 and act2 == 0 and r_coff == 0


class _FusedConvMixin:
    """forward() for Sequential(conv, bn, act)-shaped modules."""

    def _parts(self):
        return self[0], self[1], self[2].activation

    def _folded(self):
        conv, bn, act = self._parts()
        tensors = [conv.weight, conv.bias, *_bn_tensors(bn), getattr(act, "weight", None)]
        cache = self.__dict__.setdefault("_rtseg_cache", _FoldCache())
        return cache.get(tensors, lambda: ops.fold_conv(conv, bn, act, host=True))

    def forward(self, x, res=None, res_mode=ops.RES_NONE, out=None, x_coff=0, y_coff=0, act2=0, r_coff=0):
        if self[1].training and out is None and x_coff == 0 and y_coff == 0 and r_coff == 0 and act2 == 0 \
                and res_mode in (ops.RES_NONE, ops.RES_POST):
            from . import train   # batch statistics + autograd: Conv -> BN -> act as two tape nodes on rtseg kernels
            z = train.conv_bn_act(self, x)
            # a POST residual is the block's `x + conv(x)` (reference fastscnn.py:119, dabnet.py:95-97): a plain add node
            return z if res_mode == ops.RES_NONE else z + ops.to_nhwc(res)
        _check_eval(self, self[1])
        folded, post = self._folded()
        y = ops.conv(folded, ops.to_nhwc(x), res=res, res_mode=res_mode, out=out, x_coff=x_coff, y_coff=y_coff,
                     act2=act2, r_coff=r_coff)
        return post(y) if post is not None else y

    def __deepcopy__(self, memo):  # keep the cache out of EMA copies (utils/model_ema.py:20)
        import copy
        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k != "_rtseg_cache":
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new


class Conv2d(nn.Conv2d):
    """Bare convolution (what conv1x1 / conv3x3 return) routed to the same kernels."""

    def folded(self, bn=None, act=None):
        """(Folded, post module) of this convolution with an optional BatchNorm / activation behind it."""
        cache = self.__dict__.setdefault("_rtseg_cache", _FoldCache())
        tensors = [self.weight, self.bias, *_bn_tensors(bn), getattr(act, "weight", None)]
        return cache.get(tensors, lambda: ops.fold_conv(self, bn, act, host=True))

    def forward(self, x, res=None, res_mode=ops.RES_NONE, out=None, x_coff=0, y_coff=0, bn=None, act=None):
        cache = self.__dict__.setdefault("_rtseg_cache", _FoldCache())
        if self.training and (bn is None or bn.training) and out is None and x_coff == 0 and y_coff == 0 \
                and res_mode in (ops.RES_NONE, ops.RES_PRE) and torch.is_grad_enabled():
            from . import train
            y = train.conv(self, x)
            pre = res if res_mode == ops.RES_PRE else None     # act(bn(conv(x) + res)), reference fastscnn.py:90-92
            if bn is not None:
                return train.bn_act(bn, act, y, res=pre)
            if act is not None or pre is not None:
                return train.act(act, y, pre)                 # act(conv(x) + res), reference enet.py:165-181
            return y
        tensors = [self.weight, self.bias, *_bn_tensors(bn), getattr(act, "weight", None)]
        folded, post = cache.get(tensors, lambda: ops.fold_conv(self, bn, act, host=True))
        _check_eval(self, bn)
        y = ops.conv(folded, ops.to_nhwc(x), res=res, res_mode=res_mode, out=out, x_coff=x_coff, y_coff=y_coff)
        return post(y) if post is not None else y

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k != "_rtseg_cache":
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new


def conv3x3(in_channels, out_channels, stride=1, bias=False):
    """3x3 convolution, padding 1 (reference models/modules.py:7-9)."""
    return Conv2d(in_channels, out_channels, kernel_size=3, stride=stride, padding=1, bias=bias)


def conv1x1(in_channels, out_channels, stride=1, bias=False):
    """Pointwise convolution (reference models/modules.py:13-15)."""
    return Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, padding=0, bias=bias)


def channel_shuffle(x, groups=2):
    """ShuffleNet channel shuffle (reference models/modules.py:18-32): pure index permutation.

    With NHWC storage this is a view + one copy; the returned tensor is contiguous like the
    reference's `.contiguous()` result.
    """
    b, c, h, w = x.size()
    if c % groups:
        raise ValueError("channels must be divisible by groups")
    return x.view(b, groups, c // groups, h, w).transpose(1, 2).reshape(b, c, h, w)


class Activation(nn.Module):
    """Name -> nn activation hub; same 16 names as reference models/modules.py:114-123."""

    _HUB = {
        "relu": nn.ReLU, "relu6": nn.ReLU6, "leakyrelu": nn.LeakyReLU, "prelu": nn.PReLU, "celu": nn.CELU,
        "elu": nn.ELU, "hardswish": nn.Hardswish, "hardtanh": nn.Hardtanh, "gelu": nn.GELU, "glu": nn.GLU,
        "selu": nn.SELU, "silu": nn.SiLU, "sigmoid": nn.Sigmoid, "softmax": nn.Softmax, "tanh": nn.Tanh,
        "none": nn.Identity,
    }

    def __init__(self, act_type, **kwargs):
        super().__init__()
        name = act_type.lower()
        if name not in self._HUB:
            raise NotImplementedError(f"Unsupport activation type: {name}")
        self.activation = self._HUB[name](**kwargs)

    def forward(self, x):
        # Stand-alone use (e.g. lednet.py:86): a plain elementwise op on the caller's tensor.
        return self.activation(x)


class ConvBNAct(_FusedConvMixin, nn.Sequential):
    """Conv2d -> BatchNorm2d -> Activation as one kernel (reference models/modules.py:73-85)."""

    def __init__(self, in_channels, out_channels, kernel_size=3, stride=1, dilation=1, bias=False, act_type="relu",
                 **kwargs):
        padding = _same_padding(kernel_size, dilation)
        nn.Sequential.__init__(
            self,
            nn.Conv2d(in_channels, out_channels, kernel_size, stride, padding, dilation, bias=bias),
            nn.BatchNorm2d(out_channels),
            Activation(act_type, **kwargs),
        )


class DWConvBNAct(_FusedConvMixin, nn.Sequential):
    """Depthwise Conv2d(groups=in) -> BN -> act as one kernel (reference models/modules.py:46-59)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, dilation=1, act_type="relu", **kwargs):
        padding = _same_padding(kernel_size, dilation)
        nn.Sequential.__init__(
            self,
            nn.Conv2d(in_channels, out_channels, kernel_size, stride, padding, dilation=dilation, groups=in_channels,
                      bias=False),
            nn.BatchNorm2d(out_channels),
            Activation(act_type, **kwargs),
        )


class PWConvBNAct(_FusedConvMixin, nn.Sequential):
    """1x1 Conv2d(bias=True) -> BN -> act as one kernel (reference models/modules.py:63-69)."""

    def __init__(self, in_channels, out_channels, act_type="relu", bias=True, **kwargs):
        nn.Sequential.__init__(
            self,
            nn.Conv2d(in_channels, out_channels, 1, bias=bias),
            nn.BatchNorm2d(out_channels),
            Activation(act_type, **kwargs),
        )


class DSConvBNAct(nn.Sequential):
    """Depthwise-separable pair (reference models/modules.py:36-42)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, dilation=1, act_type="relu", **kwargs):
        super().__init__(
            DWConvBNAct(in_channels, in_channels, kernel_size, stride, dilation, act_type, **kwargs),
            PWConvBNAct(in_channels, out_channels, act_type, **kwargs),
        )


    def forward(self, x, out=None, y_coff=0):
        y = dw_then_pw(self[0], self[1]._folded() if not self[1][1].training else None, x, out=out, y_coff=y_coff)
        if y is not None:
            return y
        return self[1](self[0](x), out=out, y_coff=y_coff) if out is not None else self[1](self[0](x))


def dw_then_pw(dw: "DWConvBNAct", pw_folded, x, res=None, res_mode=ops.RES_NONE, out=None, y_coff=0, r_coff=0, act2=0):
    """DWConvBNAct -> (pointwise conv + BN + act [+ residual]) as ONE launch when the fused kernel applies (eval mode,
    bf16, 3x3 stride 1, channels a multiple of 64, 64 / 128 outputs); None otherwise — the caller then runs the two layers.
    `pw_folded` = (Folded, post module) of the pointwise layer, or None."""
    if pw_folded is None or dw[1].training or not isinstance(x, torch.Tensor) or not x.is_cuda or x.dtype != torch.bfloat16:
        return None
    if torch.is_grad_enabled() and x.requires_grad:
        return None
    fpw, post = pw_folded
    fdw, post_dw = dw._folded()
    if post is not None or post_dw is not None:
        return None
    x = ops.to_nhwc(x)
    if not ops.dwpw_supported(fdw, fpw, x, res_mode, res, out, 0, y_coff, r_coff, act2):
        return None
    return ops.dwpw(fdw, fpw, x, res=res, res_mode=res_mode, out=out, y_coff=y_coff, r_coff=r_coff, act2=act2)


class DeConvBNAct(nn.Module):
    """ConvTranspose2d -> BN -> act (reference models/modules.py:89-108); kernel for scale_factor 2."""

    def __init__(self, in_channels, out_channels, scale_factor=2, kernel_size=None, padding=None, act_type="relu",
                 **kwargs):
        super().__init__()
        if kernel_size is None:
            kernel_size = 2 * scale_factor - 1
        if padding is None:
            padding = (kernel_size - 1) // 2
        self.up_conv = nn.Sequential(
            nn.ConvTranspose2d(in_channels, out_channels, kernel_size=kernel_size, stride=scale_factor, padding=padding,
                               output_padding=scale_factor - 1),
            nn.BatchNorm2d(out_channels),
            Activation(act_type, **kwargs),
        )

    def _folded(self):
        conv, bn, act = self.up_conv[0], self.up_conv[1], self.up_conv[2].activation
        tensors = [conv.weight, conv.bias, *_bn_tensors(bn), getattr(act, "weight", None)]
        cache = self.__dict__.setdefault("_rtseg_cache", _FoldCache())
        return cache.get(tensors, lambda: ops.fold_conv(conv, bn, act, host=True))

    def forward(self, x, labels=None, want_logits=True):
        if self.up_conv[1].training and labels is None:
            from . import train
            return train.conv_bn_act_aligned(self.up_conv, x)   # == conv_bn_act unless cout needs channel padding
        _check_eval(self, self.up_conv[1])
        folded, post = self._folded()
        y = ops.conv(folded, ops.to_nhwc(x), labels=labels, want_logits=want_logits)
        return post(y) if (post is not None and y is not None) else y

    __deepcopy__ = Conv2d.__deepcopy__


class PyramidPoolingModule(nn.Module):
    """PSPNet pooling pyramid (reference models/modules.py:134-158).

    The reference runs 4x (adaptive pool, 1x1 conv, bilinear resize), a 5-way torch.cat and a
    PWConvBNAct.  Here each branch's resize writes straight into its channel slice of the
    concat buffer, so the cat never happens as a separate pass.
    """

    def __init__(self, in_channels, out_channels, act_type, pool_sizes=[1, 2, 4, 6], bias=False):
        super().__init__()
        assert len(pool_sizes) == 4, "Length of pool size should be 4.\n"
        hid_channels = int(in_channels // 4)
        for i, p in enumerate(pool_sizes, start=1):
            setattr(self, f"stage{i}", nn.Sequential(nn.AdaptiveAvgPool2d(p), conv1x1(in_channels, hid_channels)))
        self.conv = PWConvBNAct(2 * in_channels, out_channels, act_type=act_type, bias=bias)

    def concat_buffer(self, like: torch.Tensor, n: int, h: int, w: int) -> torch.Tensor:
        """The buffer the module's concat lives in: a producer may write the pass-through operand (channels [0, c)) straight
        into it and hand it to forward(cat=...) — the copy of x into the concat then never happens."""
        c = self.stage1[1].in_channels
        return ops.empty_nhwc(n, c + 4 * self.stage1[1].out_channels, h, w, like.dtype, like.device)

    def _stacked_branch_weights(self, convs):
        """fp32 [4, hid, c] (+ [4, hid] bias) of the four branch convolutions; rebuilt when a weight changes (_FoldCache)."""
        cache = self.__dict__.setdefault("_branch_cache", _FoldCache())

        def build():
            with torch.no_grad():
                wts = torch.stack([cv.weight.detach().float().reshape(cv.out_channels, cv.in_channels) for cv in convs]).contiguous()
                bias = None
                if any(cv.bias is not None for cv in convs):
                    bias = torch.stack([cv.bias.detach().float() if cv.bias is not None else
                                        torch.zeros(cv.out_channels, device=wts.device) for cv in convs]).contiguous()
            return wts, bias

        return cache.get([cv.weight for cv in convs] + [cv.bias for cv in convs], build)

    def forward(self, x, cat=None):
        """cat: a concat_buffer() whose first channels already hold x (then `x` is ignored)."""
        if cat is not None:
            c = self.stage1[1].in_channels
            n, _, h, w = cat.shape
            stages = [self.stage1, self.stage2, self.stage3, self.stage4]
            hid = stages[0][1].out_channels
            sizes = [int(st[0].output_size if isinstance(st[0].output_size, int) else st[0].output_size[0]) for st in stages]
            if sizes == [1, 2, 4, 6] and h % 4 == 0 and w % 4 == 0:
                pooled = ops.pyramid_avgpool(cat, 0, c)
                convs = [st[1] for st in stages]
                # the one-launch kernel keeps the 57 pooled pixels, the four weight matrices and its results in shared memory
                fits = 57 * hid * 4 + 57 * c * 2 + 4 * hid * (c + 8) * 2 <= 200 * 1024
                if (ops.USE_PPM_FUSION and cat.dtype == torch.bfloat16 and c % 8 == 0 and hid % 8 == 0 and fits and
                        all(type(cv) is Conv2d and cv.kernel_size == (1, 1) and cv.stride == (1, 1) and cv.groups == 1
                            and cv.out_channels == hid for cv in convs)):
                    wts, bias = self._stacked_branch_weights(convs)
                    ops.ppm_branches(pooled, wts, bias, cat, c)      # four conv1x1 + resize branches, one launch
                    return self.conv(cat)
            else:
                pooled = [ops.adaptive_avgpool(cat, p, 0, c) for p in sizes]
            for i, st in enumerate(stages):
                ops.upsample_bilinear(st[1](pooled[i]), (h, w), True, out=cat, y_coff=c + i * hid)
            return self.conv(cat)
        x = ops.to_nhwc(x)
        n, c, h, w = x.shape
        stages = [self.stage1, self.stage2, self.stage3, self.stage4]
        if self.conv[1].training or (torch.is_grad_enabled() and x.requires_grad):
            from . import train   # tape nodes with adjoint kernels; the concat is a copy node, as in the reference
            if not (self.conv[1].training and torch.is_grad_enabled()):
                raise NotImplementedError("PyramidPoolingModule: gradients need .train() (batch-statistics BatchNorm) — no "
                                          "silent partial autograd graph")
            feats = [x]
            for st in stages:
                p = st[0].output_size
                p = int(p if isinstance(p, int) else p[0])
                feats.append(train.upsample_bilinear(st[1](train.adaptive_avgpool(x, p)), (h, w), True))
            return self.conv(ops.to_nhwc(torch.cat(feats, dim=1)))
        hid = stages[0][1].out_channels
        cat = ops.empty_nhwc(n, c + 4 * hid, h, w, x.dtype, x.device)
        ops.copy_into_slice(x, cat)  # the pass-through operand of the concat
        return self.forward(None, cat=cat)


class SegHead(nn.Sequential):
    """3x3 ConvBNAct -> 1x1 classifier (reference models/modules.py:161-166)."""

    def __init__(self, in_channels, num_class, act_type, hid_channels=128):
        super().__init__(
            ConvBNAct(in_channels, hid_channels, 3, act_type=act_type),
            conv1x1(hid_channels, num_class),
        )


def interpolate_bilinear(x, size, align_corners=True):
    """F.interpolate(x, size, mode='bilinear') through the rtseg kernel (reference models/fastscnn.py:30)."""
    return ops.upsample_bilinear(x, tuple(size), align_corners)

