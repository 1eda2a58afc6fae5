"""The two-branch nets of BASELINE config 4 — BiSeNetv2 and DDRNet — on the fused blocks.

Same constructors, attribute names and state-dict keys as reference models/bisenetv2.py and models/ddrnet.py, so
a reference checkpoint loads with strict=True.  What is fused beyond the leaf modules:

* `GatherExpansionLayer` (bisenetv2.py:130-162): `act(x + left(x))` / `act(right(x) + left(x))` is the epilogue of
  the last pointwise convolution (POST residual + second activation) — no add, no activation kernel.
* `RB` / `RBB` (ddrnet.py:168-220): `act(conv2(.) + identity)` is the epilogue of the last convolution.
* `BilateralFusion` (ddrnet.py:223-243), `DAPPM` (:246-291), `BilateralGuidedAggregationLayer`
  (bisenetv2.py:172-203), `ContextEmbeddingBlock` (:153-170): every interpolate / sigmoid / multiply / add /
  activation chain between two convolutions is ONE `rtseg_combine_up` pass that resizes the low-resolution operand
  on the fly; DAPPM's five-way concat is five convolutions writing their own channel slice.
* `StemBlock` (bisenetv2.py:106-127): the concat is two slice writes.

Inference only (eval-mode BatchNorm folded into the convolutions); aux heads exist for state-dict parity and are
not evaluated, like the reference with is_training=False.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .modules import (Activation, ConvBNAct, DWConvBNAct, PWConvBNAct, SegHead, _bn_tensors, _check_eval, _FoldCache, conv1x1,
                      conv3x3)


def _param_dtype(m: nn.Module) -> torch.dtype:
    return next(m.parameters()).dtype


def _image_stem(block: ConvBNAct, x: torch.Tensor, act_dtype: Optional[torch.dtype]) -> torch.Tensor:
    """The first 3x3 stride-2 ConvBNAct on the NCHW (or uint8 HWC) image: the stem kernel when it applies."""
    x = ops.image_view(x)
    conv = block[0]
    dtype = act_dtype or _param_dtype(block)
    if conv.in_channels <= 4 and conv.out_channels <= 32 and tuple(conv.kernel_size) == (3, 3) and tuple(conv.stride) == (2, 2):
        _check_eval(block, block[1])
        folded, post = block._folded()
        if post is not None:
            raise NotImplementedError("non-pointwise activation in the stem")
        return ops.stem(folded, x, pool=False, out_dtype=dtype)
    if conv.in_channels <= 4 and conv.out_channels <= 64 and conv.out_channels % 16 == 0 \
            and tuple(conv.kernel_size) == (3, 3) and tuple(conv.stride) == (2, 2):
        # BiSeNetv2's 3 -> 64 detail stem (models/bisenetv2.py:47): two 32-output stem launches, each writing its
        # channel slice — the image (uint8 or fp32) is read by the stem kernel itself, never normalised into a copy
        import dataclasses
        _check_eval(block, block[1])
        folded, post = block._folded()
        if post is not None:
            raise NotImplementedError("non-pointwise activation in the stem")
        halves = block.__dict__.get("_rtseg_stem_halves")
        if halves is None or halves[0] is not folded:
            def cut(t, lo, hi):
                return None if t is None else t[lo:hi].contiguous()
            parts = []
            for lo in range(0, folded.cout, 32):
                hi = min(lo + 32, folded.cout)
                parts.append((lo, dataclasses.replace(folded, cout=hi - lo, w_f32=folded.w_f32[..., lo:hi].contiguous(), w_tc=None,
                                                      scale=cut(folded.scale, lo, hi), shift=cut(folded.shift, lo, hi),
                                                      slope=cut(folded.slope, lo, hi), w_tc_scaled=None)))
            halves = (folded, parts)
            block.__dict__["_rtseg_stem_halves"] = halves
        n, _, h, w = x.shape
        out = ops.empty_nhwc(n, folded.cout, (h - 1) // 2 + 1, (w - 1) // 2 + 1, dtype, x.device)
        for lo, part in halves[1]:
            ops.stem(part, x, pool=False, out_dtype=dtype, out=out, y_coff=lo)
        return out
    if x.dtype == torch.uint8:
        return block(ops.normalize_u8(x, dtype))
    return block(ops.to_nhwc(x, dtype))


def _act_code(act_module) -> int:
    code, _, _, prelu_w, post = ops.act_spec(act_module)
    if prelu_w is not None or post is not None:
        raise NotImplementedError("parametrised / non-pointwise block activation is not fused")
    return code


# ----------------------------------------------------------------------------- BiSeNetv2 ----
class DetailBranch(nn.Sequential):
    """reference models/bisenetv2.py:44-56"""

    act_dtype: Optional[torch.dtype] = None

    def __init__(self, in_channels, out_channels, act_type="relu"):
        super().__init__(
            ConvBNAct(in_channels, 64, 3, 2, act_type=act_type),
            ConvBNAct(64, 64, 3, 1, act_type=act_type),
            ConvBNAct(64, 64, 3, 2, act_type=act_type),
            ConvBNAct(64, 64, 3, 1, act_type=act_type),
            ConvBNAct(64, 128, 3, 1, act_type=act_type),
            ConvBNAct(128, 128, 3, 2, act_type=act_type),
            ConvBNAct(128, 128, 3, 1, act_type=act_type),
            ConvBNAct(128, out_channels, 3, 1, act_type=act_type),
        )

    def forward(self, x):
        x = _image_stem(self[0], x, self.act_dtype)
        for i in range(1, 8):
            x = self[i](x)
        return x


class StemBlock(nn.Module):
    """reference models/bisenetv2.py:106-127: cat[1x1 -> 3x3 s2, maxpool(3, 2, 1)] -> 3x3, the cat as slice writes."""

    act_dtype: Optional[torch.dtype] = None

    def __init__(self, in_channels, out_channels, act_type="relu"):
        super().__init__()
        self.conv_init = ConvBNAct(in_channels, out_channels, 3, 2, act_type=act_type)
        self.left_branch = nn.Sequential(ConvBNAct(out_channels, out_channels // 2, 1, act_type=act_type),
                                         ConvBNAct(out_channels // 2, out_channels, 3, 2, act_type=act_type))
        self.right_branch = nn.MaxPool2d(3, 2, 1)
        self.conv_last = ConvBNAct(out_channels * 2, out_channels, 3, 1, act_type=act_type)

    def forward(self, x):
        x = _image_stem(self.conv_init, x, self.act_dtype)
        n, c, h, w = x.shape
        cat = ops.empty_nhwc(n, 2 * c, (h - 1) // 2 + 1, (w - 1) // 2 + 1, x.dtype, x.device)
        self.left_branch[1](self.left_branch[0](x), out=cat, y_coff=0)
        ops.maxpool3x3s2(x, cat, y_coff=c)
        return self.conv_last(cat)


class GatherExpansionLayer(nn.Module):
    """reference models/bisenetv2.py:130-162.  The branch sum and the activation ride in the last pointwise
    convolution's epilogue; the x6 depthwise expansion is one channel-multiplier depthwise launch."""

    def __init__(self, in_channels, out_channels, stride, act_type="relu", expand_ratio=6):
        super().__init__()
        self.stride = stride
        hid = int(round(in_channels * expand_ratio))
        layers = [ConvBNAct(in_channels, in_channels, 3, act_type=act_type)]
        if stride == 2:
            layers.extend([DWConvBNAct(in_channels, hid, 3, 2, act_type="none"),
                           DWConvBNAct(hid, hid, 3, 1, act_type="none")])
            self.right_branch = nn.Sequential(DWConvBNAct(in_channels, in_channels, 3, 2, act_type="none"),
                                              PWConvBNAct(in_channels, out_channels, act_type="none"))
        else:
            layers.append(DWConvBNAct(in_channels, hid, 3, 1, act_type="none"))
        layers.append(PWConvBNAct(hid, out_channels, act_type="none"))
        self.left_branch = nn.Sequential(*layers)
        self.act = Activation(act_type)

    def forward(self, x):
        x = ops.to_nhwc(x)
        t = x
        for layer in list(self.left_branch)[:-1]:
            t = layer(t)
        other = self.right_branch(x) if self.stride == 2 else x
        return self.left_branch[-1](t, res=other, res_mode=ops.RES_POST, act2=_act_code(self.act.activation))


class ContextEmbeddingBlock(nn.Module):
    """reference models/bisenetv2.py:153-170: global pool -> BN -> 1x1 ConvBNAct, broadcast-added to x, then 3x3."""

    def __init__(self, in_channels, out_channels, act_type="relu"):
        super().__init__()
        self.pool = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.BatchNorm2d(in_channels))
        self.conv_mid = ConvBNAct(in_channels, in_channels, 1, act_type=act_type)
        self.conv_last = conv3x3(in_channels, out_channels)

    def forward(self, x):
        x = ops.to_nhwc(x)
        bn = self.pool[1]
        _check_eval(self, bn)
        cache = self.__dict__.setdefault("_rtseg_cache", _FoldCache())
        scale, shift = cache.get(_bn_tensors(bn), lambda: tuple(t.float().view(1, -1, 1, 1) for t in
                                                                ops.fold_affine(x.shape[1], None, bn)))
        pooled = ops.adaptive_avgpool(x, 1, out_dtype=torch.float32)          # [n, c, 1, 1] fp32
        pooled = torch.addcmul(shift, pooled, scale).to(x.dtype)                # eval BatchNorm on a 1x1 map: n * c values
        res = self.conv_mid(pooled.contiguous(memory_format=torch.channels_last))
        return self.conv_last(ops.combine_up(x, res, "add"))                     # res + x: a broadcast add


class SemanticBranch(nn.Sequential):
    """reference models/bisenetv2.py:59-103 (aux heads kept for the state dict, evaluated only on request)."""

    def __init__(self, in_channels, out_channels, num_class, act_type="relu", use_aux=False):
        super().__init__()
        self.use_aux = use_aux
        self.stage1to2 = StemBlock(in_channels, 16, act_type)
        self.stage3 = nn.Sequential(GatherExpansionLayer(16, 32, 2, act_type), GatherExpansionLayer(32, 32, 1, act_type))
        self.stage4 = nn.Sequential(GatherExpansionLayer(32, 64, 2, act_type), GatherExpansionLayer(64, 64, 1, act_type))
        self.stage5_1to4 = nn.Sequential(GatherExpansionLayer(64, 128, 2, act_type), GatherExpansionLayer(128, 128, 1, act_type),
                                         GatherExpansionLayer(128, 128, 1, act_type), GatherExpansionLayer(128, 128, 1, act_type))
        self.stage5_5 = ContextEmbeddingBlock(128, out_channels, act_type)
        if self.use_aux:
            self.seg_head2 = SegHead(16, num_class, act_type)
            self.seg_head3 = SegHead(32, num_class, act_type)
            self.seg_head4 = SegHead(64, num_class, act_type)
            self.seg_head5 = SegHead(128, num_class, act_type)

    def forward(self, x, with_aux: bool = False):
        x2 = self.stage1to2(x)
        x3 = self.stage3(x2)
        x4 = self.stage4(x3)
        x5 = self.stage5_1to4(x4)
        out = self.stage5_5(x5)
        if with_aux and self.use_aux:
            return out, self.seg_head2(x2), self.seg_head3(x3), self.seg_head4(x4), self.seg_head5(x5)
        return out


class BilateralGuidedAggregationLayer(nn.Module):
    """reference models/bisenetv2.py:172-203.  Three fused passes replace upsample, 2 x sigmoid, 2 x multiply,
    interpolate and add: x_high = d_high * sigmoid(up4(s_high)); x_low = d_low * sigmoid(s_low);
    res = x_high + up4(x_low)."""

    def __init__(self, in_channels, out_channels, act_type="relu"):
        super().__init__()
        self.detail_high = nn.Sequential(DWConvBNAct(in_channels, in_channels, 3, act_type=act_type),
                                         conv1x1(in_channels, in_channels))
        self.detail_low = nn.Sequential(DWConvBNAct(in_channels, in_channels, 3, 2, act_type=act_type),
                                        nn.AvgPool2d(3, 2, 1))
        self.semantic_high = nn.Sequential(ConvBNAct(in_channels, in_channels, 3, act_type=act_type),
                                           nn.Upsample(scale_factor=4, mode="bilinear", align_corners=True), nn.Sigmoid())
        self.semantic_low = nn.Sequential(DWConvBNAct(in_channels, in_channels, 3, act_type=act_type),
                                          conv1x1(in_channels, in_channels), nn.Sigmoid())
        self.conv_last = ConvBNAct(in_channels, out_channels, 3, act_type=act_type)

    def forward(self, x_d, x_s):
        x_d, x_s = ops.to_nhwc(x_d), ops.to_nhwc(x_s)
        if x_d.shape[2] != 4 * x_s.shape[2] or x_d.shape[3] != 4 * x_s.shape[3]:
            raise ValueError("the detail branch must be 4x the semantic branch (input sizes that are multiples of 32)")
        d_high = self.detail_high[1](self.detail_high[0](x_d))
        d_low = ops.avgpool3x3s2(self.detail_low[0](x_d))
        s_high = self.semantic_high[0](x_s)
        s_low = self.semantic_low[1](self.semantic_low[0](x_s))
        x_high = ops.combine_up(d_high, s_high, "mul", b_sigmoid=True)
        x_low = ops.combine_up(d_low, s_low, "mul", b_sigmoid=True)
        return self.conv_last(ops.combine_up(x_high, x_low, "add", out=x_high))


class BiSeNetv2(nn.Module):
    """BiSeNet V2 — reference models/bisenetv2.py:18-41."""

    HEAD_ALIGN_CORNERS = True

    def __init__(self, num_class=1, n_channel=3, act_type="relu", use_aux=True):
        super().__init__()
        self.use_aux = use_aux
        self.detail_branch = DetailBranch(n_channel, 128, act_type)
        self.semantic_branch = SemanticBranch(n_channel, 128, num_class, act_type, use_aux)
        self.bga_layer = BilateralGuidedAggregationLayer(128, 128, act_type)
        self.seg_head = SegHead(128, num_class, act_type)
        self.num_class = num_class

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        self.detail_branch.act_dtype = dtype
        self.semantic_branch.stage1to2.act_dtype = dtype
        return self

    def lowres_logits(self, x):
        """seg_head output at 1/8 resolution (before the bilinear x8)."""
        if self.training:
            raise NotImplementedError("BiSeNetv2: inference only on this path; call .eval()")
        x = ops.image_view(x)
        return self.seg_head(self.bga_layer(self.detail_branch(x), self.semantic_branch(x)))

    def forward(self, x, is_training=False):
        if is_training:
            raise NotImplementedError("BiSeNetv2: the training-mode aux outputs are not built on this path")
        x = ops.image_view(x)
        return ops.upsample_bilinear(self.lowres_logits(x), x.shape[2:], True)

    def predict(self, x, target: Optional[torch.Tensor] = None, confmat: Optional[torch.Tensor] = None,
                ignore_index: int = 255) -> torch.Tensor:
        x = ops.image_view(x)
        return ops.upsample_argmax(self.lowres_logits(x), x.shape[2:], True, target=target, confmat=confmat,
                                   ignore_index=ignore_index)


# -------------------------------------------------------------------------------- DDRNet ----
class RB(nn.Module):
    """Residual basic block — reference models/ddrnet.py:168-190; `out += identity; relu` in conv2's epilogue."""

    def __init__(self, in_channels, out_channels, stride=1, act_type="relu"):
        super().__init__()
        self.downsample = (stride > 1) or (in_channels != out_channels)
        self.conv1 = ConvBNAct(in_channels, out_channels, 3, stride, act_type=act_type)
        self.conv2 = ConvBNAct(out_channels, out_channels, 3, 1, act_type="none")
        if self.downsample:
            self.conv_down = ConvBNAct(in_channels, out_channels, 1, stride, act_type="none")
        self.act = nn.ReLU()

    def forward(self, x):
        x = ops.to_nhwc(x)
        identity = self.conv_down(x) if self.downsample else x
        return self.conv2(self.conv1(x), res=identity, res_mode=ops.RES_POST, act2=_act_code(self.act))


class RBB(nn.Module):
    """Residual bottleneck block — reference models/ddrnet.py:193-220."""

    def __init__(self, in_channels, out_channels, stride=1, act_type="relu"):
        super().__init__()
        self.downsample = (stride > 1) or (in_channels != out_channels)
        self.conv1 = ConvBNAct(in_channels, in_channels, 1, act_type=act_type)
        self.conv2 = ConvBNAct(in_channels, in_channels, 3, stride, act_type=act_type)
        self.conv3 = ConvBNAct(in_channels, out_channels, 1, act_type="none")
        if self.downsample:
            self.conv_down = ConvBNAct(in_channels, out_channels, 1, stride, act_type="none")
        self.act = Activation(act_type)

    def forward(self, x):
        x = ops.to_nhwc(x)
        identity = self.conv_down(x) if self.downsample else x
        return self.conv3(self.conv2(self.conv1(x)), res=identity, res_mode=ops.RES_POST, act2=_act_code(self.act.activation))


def build_blocks(block, in_channels, out_channels, stride, repeat_times, act_type):
    layers = [block(in_channels, out_channels, stride, act_type=act_type)]
    for _ in range(1, repeat_times):
        layers.append(block(out_channels, out_channels, 1, act_type=act_type))
    return nn.Sequential(*layers)


class BilateralFusion(nn.Module):
    """reference models/ddrnet.py:223-243.  x_low = act(x_low + conv_high(x_high)) is conv_high's epilogue;
    x_high = act(x_high + interpolate(conv_low(x_low))) is one combine pass."""

    def __init__(self, low_res_channels, high_res_channels, stride, act_type="relu"):
        super().__init__()
        self.conv_low = ConvBNAct(low_res_channels, high_res_channels, 1, act_type="none")
        self.conv_high = ConvBNAct(high_res_channels, low_res_channels, 3, stride, act_type="none")
        self.act = Activation(act_type)

    def forward(self, x_low, x_high):
        act = self.act.activation
        fuse_low = self.conv_low(x_low)
        new_low = self.conv_high(x_high, res=x_low, res_mode=ops.RES_POST, act2=_act_code(act))
        new_high = ops.combine_up(x_high, fuse_low, "add", act_module=act)
        return new_low, new_high


class DAPPM(nn.Module):
    """Deep aggregation pyramid pooling — reference models/ddrnet.py:246-291.  Each scale is pool -> 1x1 conv ->
    (resize + add, one pass) -> 3x3 conv that writes its slice of the concat buffer; `+ y0` is conv_last's epilogue."""

    def __init__(self, in_channels, out_channels, act_type="relu"):
        super().__init__()
        hid = int(in_channels // 4)
        self.conv0 = ConvBNAct(in_channels, out_channels, 1, act_type=act_type)
        self.conv1 = ConvBNAct(in_channels, hid, 1, act_type=act_type)
        self.pool2 = self._build_pool_layers(in_channels, hid, 5, 2)
        self.conv2 = ConvBNAct(hid, hid, 3, act_type=act_type)
        self.pool3 = self._build_pool_layers(in_channels, hid, 9, 4)
        self.conv3 = ConvBNAct(hid, hid, 3, act_type=act_type)
        self.pool4 = self._build_pool_layers(in_channels, hid, 17, 8)
        self.conv4 = ConvBNAct(hid, hid, 3, act_type=act_type)
        self.pool5 = self._build_pool_layers(in_channels, hid, -1, -1)
        self.conv5 = ConvBNAct(hid, hid, 3, act_type=act_type)
        self.conv_last = ConvBNAct(hid * 5, out_channels, 1, act_type=act_type)

    @staticmethod
    def _build_pool_layers(in_channels, out_channels, kernel_size, stride):
        layers = []
        if kernel_size == -1:
            layers.append(nn.AdaptiveAvgPool2d(1))
        else:
            layers.append(nn.AvgPool2d(kernel_size, stride, (kernel_size - 1) // 2))
        layers.append(conv1x1(in_channels, out_channels))
        return nn.Sequential(*layers)

    @staticmethod
    def _pooled(seq, x):
        pool = seq[0]
        if isinstance(pool, nn.AdaptiveAvgPool2d):
            p = ops.adaptive_avgpool(x, 1)
        else:
            k, s, pad = (v if isinstance(v, int) else v[0] for v in (pool.kernel_size, pool.stride, pool.padding))
            p = ops.avgpool2d(x, k, s, pad)
        return seq[1](p)

    def forward(self, x):
        x = ops.to_nhwc(x)
        n, _, h, w = x.shape
        hid = self.conv1[0].out_channels
        cat = ops.empty_nhwc(n, 5 * hid, h, w, x.dtype, x.device)
        y0 = self.conv0(x)
        self.conv1(x, out=cat, y_coff=0)
        for i, (pool, conv) in enumerate(((self.pool2, self.conv2), (self.pool3, self.conv3), (self.pool4, self.conv4),
                                          (self.pool5, self.conv5)), start=1):
            s = ops.combine_up(cat, self._pooled(pool, x), "add", a_coff=(i - 1) * hid, c=hid)   # y_{i-1} + resize(pool_i)
            conv(s, out=cat, y_coff=i * hid)
        return self.conv_last(cat, res=y0, res_mode=ops.RES_POST)


class Stage2(nn.Module):
    def __init__(self, init_channel, repeat_times, act_type="relu"):
        super().__init__()
        layers = [ConvBNAct(init_channel, init_channel, 3, 2, act_type=act_type)]
        for _ in range(repeat_times):
            layers.append(RB(init_channel, init_channel, 1, act_type))
        self.conv = nn.Sequential(*layers)

    def forward(self, x):
        return self.conv(x)


class Stage3(nn.Module):
    def __init__(self, init_channel, repeat_times, act_type="relu"):
        super().__init__()
        self.conv = build_blocks(RB, init_channel, init_channel * 2, 2, repeat_times, act_type)

    def forward(self, x):
        return self.conv(x)


class Stage4(nn.Module):
    """reference models/ddrnet.py:101-131"""

    def __init__(self, init_channel, repeat_times1, repeat_times2, act_type="relu"):
        super().__init__()
        cin, low, high = init_channel * 2, init_channel * 4, init_channel * 2
        if low < high:
            raise ValueError("Low resolution channel should be more than high resolution channel.\n")
        self.low_conv1 = build_blocks(RB, cin, low, 2, repeat_times1, act_type)
        self.high_conv1 = build_blocks(RB, cin, high, 1, repeat_times1, act_type)
        self.bilateral_fusion1 = BilateralFusion(low, high, 2)
        self.extra_conv = repeat_times2 > 0
        if self.extra_conv:
            self.low_conv2 = build_blocks(RB, low, low, 1, repeat_times2, act_type)
            self.high_conv2 = build_blocks(RB, high, high, 1, repeat_times2, act_type)
            self.bilateral_fusion2 = BilateralFusion(low, high, 2)

    def forward(self, x):
        x_low, x_high = self.bilateral_fusion1(self.low_conv1(x), self.high_conv1(x))
        if self.extra_conv:
            x_low, x_high = self.bilateral_fusion2(self.low_conv2(x_low), self.high_conv2(x_high))
        return x_low, x_high


class Stage5(nn.Module):
    """reference models/ddrnet.py:134-165"""

    def __init__(self, init_channel, repeat_times1, repeat_times2, act_type="relu"):
        super().__init__()
        low_in, high_in = init_channel * 4, init_channel * 2
        low1, high1, low2, high2 = init_channel * 8, init_channel * 2, init_channel * 16, init_channel * 4
        if low_in < high_in or low1 < high1 or low2 < high2:
            raise ValueError("Low resolution channel should be more than high resolution channel.\n")
        self.low_conv1 = build_blocks(RB, low_in, low1, 2, repeat_times1, act_type)
        self.high_conv1 = build_blocks(RB, high_in, high1, 1, repeat_times1, act_type)
        self.bilateral_fusion = BilateralFusion(low1, high1, 4)
        self.low_conv2 = build_blocks(RBB, low1, low2, 2, repeat_times2, act_type)
        self.high_conv2 = build_blocks(RBB, high1, high2, 1, repeat_times2, act_type)
        self.dappm = DAPPM(low2, high2)

    def forward(self, x_low, x_high):
        x_low, x_high = self.bilateral_fusion(self.low_conv1(x_low), self.high_conv1(x_high))
        x_low = self.dappm(self.low_conv2(x_low))
        x_high = self.high_conv2(x_high)
        return ops.combine_up(x_high, x_low, "add", out=x_high)     # high_conv2(x_high) + interpolate(x_low)


class DDRNet(nn.Module):
    """DDRNet — reference models/ddrnet.py:17-65 (the ONNX int8 argmax branch is `predict`)."""

    HEAD_ALIGN_CORNERS = True
    ARCH_HUB = {"DDRNet-23-slim": {"init_channel": 32, "repeat_times": [2, 2, 2, 0, 2, 1]},
                "DDRNet-23": {"init_channel": 64, "repeat_times": [2, 2, 2, 0, 2, 1]},
                "DDRNet-39": {"init_channel": 64, "repeat_times": [3, 4, 3, 3, 3, 1]}}
    act_dtype: Optional[torch.dtype] = None

    def __init__(self, num_class=1, n_channel=3, arch_type="DDRNet-23-slim", act_type="relu", use_aux=True):
        super().__init__()
        if arch_type not in self.ARCH_HUB:
            raise ValueError(f"Unsupport architecture type: {arch_type}.\n")
        init_channel = self.ARCH_HUB[arch_type]["init_channel"]
        rt = self.ARCH_HUB[arch_type]["repeat_times"]
        self.use_aux = use_aux
        self.conv1 = ConvBNAct(n_channel, init_channel, 3, 2, act_type=act_type)
        self.conv2 = Stage2(init_channel, rt[0], act_type)
        self.conv3 = Stage3(init_channel, rt[1], act_type)
        self.conv4 = Stage4(init_channel, rt[2], rt[3], act_type)
        self.conv5 = Stage5(init_channel, rt[4], rt[5], act_type)
        self.seg_head = SegHead(init_channel * 4, num_class, act_type)
        if self.use_aux:
            self.aux_head = SegHead(init_channel * 2, num_class, act_type)
        self.num_class = num_class

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        self.act_dtype = dtype
        return self

    def lowres_logits(self, x):
        if self.training:
            raise NotImplementedError("DDRNet: inference only on this path; call .eval()")
        x = _image_stem(self.conv1, x, self.act_dtype)
        x_low, x_high = self.conv4(self.conv3(self.conv2(x)))
        return self.seg_head(self.conv5(x_low, x_high))

    def forward(self, x, is_training=False):
        if is_training:
            raise NotImplementedError("DDRNet: the training-mode aux output is not built on this path")
        x = ops.image_view(x)
        return ops.upsample_bilinear(self.lowres_logits(x), x.shape[2:], True)

    def predict(self, x, target: Optional[torch.Tensor] = None, confmat: Optional[torch.Tensor] = None,
                ignore_index: int = 255) -> torch.Tensor:
        x = ops.image_view(x)
        return ops.upsample_argmax(self.lowres_logits(x), x.shape[2:], True, target=target, confmat=confmat,
                                   ignore_index=ignore_index)


__all__ = ["BiSeNetv2", "DDRNet", "GatherExpansionLayer", "BilateralGuidedAggregationLayer", "ContextEmbeddingBlock", "StemBlock",
           "DetailBranch", "SemanticBranch", "RB", "RBB", "BilateralFusion", "DAPPM"]
