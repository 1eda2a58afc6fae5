"""ERFNet and Fast-SCNN assembled from the fused blocks (configs 2, 3 and 5 of BASELINE.json).

The GPU box has no copy of the reference, so the two benchmark networks are
defined here.  Attribute names and child order reproduce the reference's
state-dict keys exactly (ERFNet: reference models/erfnet.py:15-46, 2,065,458
parameters; Fast-SCNN: models/fastscnn.py:16-121, 1,023,769 parameters — both
pinned by tests/golden/state_specs.json, generated from the reference), so a
reference checkpoint loads with `load_state_dict` and vice versa.

What differs from the reference is the execution plan of `forward`:
  * every conv+BN+act is one kernel; residual adds, the BN after an add, the
    InitialBlock max-pool + concat, PPM concat and FFM add are epilogue/slice
    fusions (no separate elementwise or cat passes);
  * activations stay NHWC bf16/fp32 between layers;
  * `predict(x)` ends in the fused argmax (ERFNet: inside the last transposed
    conv; Fast-SCNN: fused x8 bilinear upsample + argmax), so the
    19 x 1024 x 2048 logits are never written.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .modules import (_FoldCache, Activation, Conv2d, ConvBNAct, DeConvBNAct, DSConvBNAct, DWConvBNAct, PWConvBNAct,
                      PyramidPoolingModule, _bn_tensors, _check_eval, _FoldCache, conv1x1, dw_then_pw)


def _param_dtype(m: nn.Module) -> torch.dtype:
    return next(m.parameters()).dtype


def _pitch8(c: int) -> int:
    """Channel pitch that keeps every pixel of an NHWC bf16/fp32 buffer 16-byte aligned (TMA stores, vector I/O)."""
    return (c + 7) // 8 * 8


def _first_channels(t: torch.Tensor, c: int) -> torch.Tensor:
    """Channels [0, c) of a pitch-padded NHWC buffer as a dense NHWC tensor."""
    return t if t.shape[1] == c else t[:, :c].contiguous(memory_format=torch.channels_last)


class InitialBlock(nn.Module):
    """cat([ConvBNAct(in, out-in, 3, stride 2)(x), MaxPool2d(3,2,1)(x)]) — reference models/enet.py:38-48."""

    act_dtype: Optional[torch.dtype] = None  # storage type of the activations this stem emits (None: parameter dtype)

    def __init__(self, in_channels, out_channels, act_type, kernel_size=3, **kwargs):
        super().__init__()
        assert out_channels > in_channels, "out_channels should be larger than in_channels.\n"
        self.conv = ConvBNAct(in_channels, out_channels - in_channels, kernel_size, 2, act_type=act_type, **kwargs)
        self.pool = nn.MaxPool2d(3, 2, 1)

    def forward(self, x, out=None, y_coff=0):
        """`x` may be a wider NHWC buffer whose first in_channels channels are the input; `out` / `y_coff`
        place the result in a channel slice of a caller-owned buffer (DABNet's concatenations)."""
        x = ops.image_view(x)
        conv = self.conv[0]
        cin, cconv = conv.in_channels, conv.out_channels
        if self.conv[1].training and out is None and x.shape[1] == cin:
            from . import train   # training: conv branch and pool branch are tape nodes; the cat is a copy
            xn = ops.to_nhwc(x, self.act_dtype or (x.dtype if x.is_cuda and ops.is_nhwc(x) and x.dtype != torch.uint8
                                                    else _param_dtype(self.conv)))
            conv_out = train.conv_bn_act_aligned(self.conv, xn, pad_input=True) if cin <= 4 else train.conv_bn_act(self.conv, xn)
            return ops.to_nhwc(torch.cat([conv_out, train.maxpool3x3s2(xn)], dim=1))
        _check_eval(self, self.conv[1])
        folded, post = self.conv._folded()
        if post is not None:
            raise NotImplementedError("non-pointwise activation in InitialBlock")
        if cin <= 4 and cconv <= 32 and conv.kernel_size == (3, 3) and x.shape[1] == cin:
            return ops.stem(folded, x, pool=True, out_dtype=self.act_dtype or _param_dtype(self.conv), out=out,
                            y_coff=y_coff)
        x = ops.to_nhwc(x)
        n, _, h, w = x.shape
        if out is None:
            out = ops.empty_nhwc(n, cconv + cin, (h - 1) // 2 + 1, (w - 1) // 2 + 1, x.dtype, x.device)
        ops.conv(folded, x, out=out, y_coff=y_coff)
        ops.maxpool3x3s2(x, out, y_coff=y_coff + cconv, c=cin)
        return out


class NonBt1DBlock(nn.Module):
    """ERFNet non-bottleneck-1D block — reference models/erfnet.py:62-82.

    3x1 -> 1x3 -> 3x1(d) -> 1x3(d) -> (+x) -> BN -> act: four launches; the add,
    the trailing BatchNorm and the activation ride in the fourth conv's epilogue.
    The three inner activations are ReLU whatever `act_type` says (erfnet.py:66-68).
    """

    def __init__(self, channels, dilation=1, act_type="relu"):
        super().__init__()
        self.conv = nn.Sequential(
            ConvBNAct(channels, channels, (3, 1), inplace=True),
            ConvBNAct(channels, channels, (1, 3), inplace=True),
            ConvBNAct(channels, channels, (3, 1), dilation=dilation, inplace=True),
            Conv2d(channels, channels, (1, 3), dilation=dilation, padding=(0, dilation), bias=False),
        )
        self.bn_act = nn.Sequential(nn.BatchNorm2d(channels), Activation(act_type, inplace=True))

    def _pairs(self, x):
        """The block as two fused pair launches when the kernel covers it (bf16, undilated, 16 / 64 channels)."""
        if self.bn_act[0].training or any(self.conv[i][1].training for i in range(3)) or x.dtype != torch.bfloat16:
            return None
        fa, pa = self.conv[0]._folded()
        fb, pb = self.conv[1]._folded()
        fc, pc = self.conv[2]._folded()
        fd, pd = self.conv[3].folded(self.bn_act[0], self.bn_act[1].activation)
        if pa is not None or pb is not None or pc is not None or pd is not None:
            return None
        if not (ops.conv_pair_supported(fa, fb, x) and ops.conv_pair_supported(fc, fd, x, res=x)):
            return None
        return ops.conv_pair(fc, fd, ops.conv_pair(fa, fb, x), res=x)

    def forward(self, x):
        x = ops.to_nhwc(x)
        y = self._pairs(x)
        if y is not None:
            return y
        t = self.conv[0](x)
        t = self.conv[1](t)
        t = self.conv[2](t)
        if self.bn_act[0].training:
            from . import train   # conv -> (+x) -> BatchNorm(batch statistics) -> act, the add inside the BN node
            return train.bn_act(self.bn_act[0], self.bn_act[1].activation, train.conv(self.conv[3], t), res=x)
        return self.conv[3](t, res=x, res_mode=ops.RES_PRE, bn=self.bn_act[0], act=self.bn_act[1].activation)


def _blocks(block, channels, count, dilations=(), act_type="relu"):
    dilations = list(dilations) or [1] * count
    if len(dilations) != count:
        raise ValueError("Number of dilation should be equal to number of blocks")
    return nn.Sequential(*[block(channels, dilation=d, act_type=act_type) for d in dilations])


class ERFNet(nn.Module):
    """ERFNet — reference models/erfnet.py:15-46 (same layer names / state-dict keys)."""

    def __init__(self, num_class=1, n_channel=3, act_type="relu"):
        super().__init__()
        self.layer1 = InitialBlock(n_channel, 16, act_type=act_type)
        self.layer2 = InitialBlock(16, 64, act_type=act_type)
        self.layer3_7 = _blocks(NonBt1DBlock, 64, 5, act_type=act_type)
        self.layer8 = InitialBlock(64, 128, act_type=act_type)
        self.layer9_16 = _blocks(NonBt1DBlock, 128, 8, dilations=[2, 4, 8, 16, 2, 4, 8, 16], act_type=act_type)
        self.layer17 = DeConvBNAct(128, 64, act_type=act_type)
        self.layer18_19 = _blocks(NonBt1DBlock, 64, 2, act_type=act_type)
        self.layer20 = DeConvBNAct(64, 16, act_type=act_type)
        self.layer21_22 = _blocks(NonBt1DBlock, 16, 2, act_type=act_type)
        self.layer23 = DeConvBNAct(16, num_class, act_type=act_type)
        self.num_class = num_class

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        """bf16 activations with fp32 master parameters: the stem converts while it reads the fp32 image,
        BN is folded in fp32 and the weights are rounded to bf16 once (SURVEY §7 'hard parts')."""
        self.layer1.act_dtype = dtype
        return self

    def _trunk(self, x):
        x = self.layer1(ops.image_view(x))
        x = self.layer2(x)
        x = self.layer3_7(x)
        x = self.layer8(x)
        x = self.layer9_16(x)
        x = self.layer17(x)
        x = self.layer18_19(x)
        x = self.layer20(x)
        return self.layer21_22(x)

    def forward(self, x):
        """Logits [N, num_class, H, W] like the reference."""
        return self.layer23(self._trunk(x))

    def predict(self, x, labels: Optional[torch.Tensor] = None) -> torch.Tensor:
        """uint8 class map [N, H, W] == forward(x).max(dim=1)[1]; the logits are never materialised."""
        x = ops.image_view(x)
        n, _, h, w = x.shape
        if labels is None:
            labels = torch.empty((n, h, w), dtype=torch.uint8, device=x.device)
        self.layer23(self._trunk(x), labels=labels, want_logits=False)
        return labels


class InvertedResidual(nn.Module):
    """MobileNetV2 block of Fast-SCNN — reference models/fastscnn.py:105-121."""

    def __init__(self, in_channels, out_channels, stride, expand_ratio=6, act_type="relu"):
        super().__init__()
        hid = int(round(in_channels * expand_ratio))
        self.use_res_connect = stride == 1 and in_channels == out_channels
        self.conv = nn.Sequential(
            PWConvBNAct(in_channels, hid, act_type=act_type),
            DWConvBNAct(hid, hid, 3, stride, act_type=act_type),
            ConvBNAct(hid, out_channels, 1, act_type="none"),
        )

    def forward(self, x, out=None, y_coff=0):
        """out / y_coff (inference): write the result into a channel slice of a caller-owned buffer."""
        x = ops.to_nhwc(x)
        t = self.conv[0](x)
        kw = {} if out is None else {"out": out, "y_coff": y_coff}
        # depthwise 3x3 + linear projection (+ x) in one launch: the 6x-expanded depthwise output never reaches HBM
        proj = None if self.conv[2][1].training else self.conv[2]._folded()
        y = (dw_then_pw(self.conv[1], proj, t, res=x, res_mode=ops.RES_POST, **kw) if self.use_res_connect else
             dw_then_pw(self.conv[1], proj, t, **kw))
        if y is not None:
            return y
        t = self.conv[1](t)
        if self.use_res_connect:  # x + conv(x): the add rides in the projection's epilogue
            return self.conv[2](t, res=x, res_mode=ops.RES_POST, **kw)
        return self.conv[2](t, **kw)


class LearningToDownsample(nn.Sequential):
    """reference models/fastscnn.py:35-41"""

    act_dtype: Optional[torch.dtype] = None

    def __init__(self, in_channels, out_channels, hid_channels=[32, 48], act_type="relu"):
        super().__init__(
            ConvBNAct(in_channels, hid_channels[0], 3, 2, act_type=act_type),
            DSConvBNAct(hid_channels[0], hid_channels[1], 3, 2, act_type=act_type),
            DSConvBNAct(hid_channels[1], out_channels, 3, 2, act_type=act_type),
        )

    def forward(self, x):
        x = ops.image_view(x)
        stem = self[0]
        conv = stem[0]
        if stem[1].training:
            from . import train   # the stem as tape nodes (3 input channels padded to 16 for the tensor-core kernels)
            xn = ops.to_nhwc(x, self.act_dtype or _param_dtype(stem))
            x = train.conv_bn_act_aligned(stem, xn, pad_input=True) if conv.in_channels <= 4 else stem(xn)
        elif conv.in_channels <= 4 and conv.out_channels <= 32:
            folded, post = stem._folded()
            if post is not None:
                raise NotImplementedError("non-pointwise activation in the stem")
            x = ops.stem(folded, x, pool=False, out_dtype=self.act_dtype or _param_dtype(stem))
        else:
            x = stem(ops.to_nhwc(x, self.act_dtype or _param_dtype(stem)) if x.dtype == torch.uint8 else x)
        return self[2](self[1](x))


class GlobalFeatureExtractor(nn.Module):
    """reference models/fastscnn.py:44-70"""

    def __init__(self, in_channels, out_channels, act_type="relu"):
        super().__init__()
        feats = []
        for t, c, n, s in ([6, 64, 3, 2], [6, 96, 2, 2], [6, 128, 3, 1]):
            for i in range(n):
                feats.append(InvertedResidual(in_channels, c, s if i == 0 else 1, t, act_type=act_type))
                in_channels = c
        self.bottlenecks = nn.Sequential(*feats)
        self.ppm = PyramidPoolingModule(in_channels, out_channels, act_type=act_type, bias=True)

    def forward(self, x):
        if self.ppm.conv[1].training or (torch.is_grad_enabled() and isinstance(x, torch.Tensor) and x.requires_grad):
            return self.ppm(self.bottlenecks(x))
        # inference: the last bottleneck writes its output straight into the pyramid module's concat buffer
        blocks = list(self.bottlenecks)
        for blk in blocks[:-1]:
            x = blk(x)
        x = ops.to_nhwc(x)
        n, _, h, w = x.shape
        last = blocks[-1]
        stride = last.conv[1][0].stride[0]
        ho, wo = (h - 1) // stride + 1, (w - 1) // stride + 1
        cat = self.ppm.concat_buffer(x, n, ho, wo)
        last(x, out=cat, y_coff=0)
        return self.ppm(None, cat=cat)


class FeatureFusionModule(nn.Module):
    """reference models/fastscnn.py:73-93.  act(BN(conv_hi(hi) + conv_lo(DW(up(lo))))):
    the add, BN and activation are the epilogue of the low-resolution branch's 1x1 conv."""

    def __init__(self, higher_channels, lower_channels, out_channels, act_type="relu"):
        super().__init__()
        self.higher_res_conv = conv1x1(higher_channels, out_channels)
        self.lower_res_conv = nn.Sequential(
            DWConvBNAct(lower_channels, lower_channels, 3, 1, act_type=act_type),
            conv1x1(lower_channels, out_channels),
        )
        self.non_linear = nn.Sequential(nn.BatchNorm2d(out_channels), Activation(act_type))

    def forward(self, higher_res_feat, lower_res_feat):
        size = higher_res_feat.shape[2:]
        hi = self.higher_res_conv(higher_res_feat)
        if self.non_linear[0].training:
            from . import train
            lo = train.upsample_bilinear(lower_res_feat, size, True)
        else:
            lo = ops.upsample_bilinear(lower_res_feat, size, True)
        bn, act = self.non_linear[0], self.non_linear[1].activation
        if not bn.training:   # depthwise 3x3 -> 1x1 conv -> + hi -> BN -> act in one launch
            y = dw_then_pw(self.lower_res_conv[0], self.lower_res_conv[1].folded(bn, act), lo, res=hi, res_mode=ops.RES_PRE)
            if y is not None:
                return y
        lo = self.lower_res_conv[0](lo)
        return self.lower_res_conv[1](lo, res=hi, res_mode=ops.RES_PRE, bn=bn, act=act)


class Classifier(nn.Sequential):
    """reference models/fastscnn.py:96-102"""

    def __init__(self, in_channels, num_class, act_type="relu"):
        super().__init__(
            DSConvBNAct(in_channels, in_channels, 3, 1, act_type=act_type),
            DSConvBNAct(in_channels, in_channels, 3, 1, act_type=act_type),
            PWConvBNAct(in_channels, num_class, act_type=act_type),
        )


class FastSCNN(nn.Module):
    """Fast-SCNN — reference models/fastscnn.py:16-32 (same attribute names / state-dict keys)."""

    def __init__(self, num_class=1, n_channel=3, act_type="relu"):
        super().__init__()
        self.learning_to_downsample = LearningToDownsample(n_channel, 64, act_type=act_type)
        self.global_feature_extractor = GlobalFeatureExtractor(64, 128, act_type=act_type)
        self.feature_fusion = FeatureFusionModule(64, 128, 128, act_type=act_type)
        self.classifier = Classifier(128, num_class, act_type=act_type)
        self.num_class = num_class

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        self.learning_to_downsample.act_dtype = dtype
        return self

    def _features(self, x):
        hi = self.learning_to_downsample(ops.image_view(x))
        lo = self.global_feature_extractor(hi)
        return self.classifier[1](self.classifier[0](self.feature_fusion(hi, lo)))

    def lowres_logits(self, x):
        return self.classifier[2](self._features(x))

    def forward(self, x):
        """Full-resolution logits like the reference (bilinear x8, align_corners=True)."""
        x = ops.image_view(x)
        if self.classifier[2][1].training:
            from . import train
            # the 19-class layer runs with 32 outputs (zero weights for the pad), the x8 resize and its adjoint work on
            # whole 16-byte channel vectors of that buffer, and the loss reads the first 19 channels in place
            z = train.conv_bn_act_aligned(self.classifier[2], self._features(x), keep_padding=True)
            return train.first_channels(train.upsample_bilinear(z, x.shape[2:], True), self.num_class)
        return ops.upsample_bilinear(self.lowres_logits(x), x.shape[2:], True)

    def predict(self, x, target: Optional[torch.Tensor] = None, confmat: Optional[torch.Tensor] = None,
                ignore_index: int = 255) -> torch.Tensor:
        """uint8 class map via the fused upsample+argmax head (+ confusion-matrix update)."""
        x = ops.image_view(x)
        f = self._features(x)
        n, _, h, w = f.shape
        # the 19-class head writes into a 32-channel (16-byte aligned) buffer so that its output leaves the
        # tensor-core kernel through the TMA-store path; the fused head reads the first num_class channels
        cpad = (self.num_class + 15) // 16 * 16
        buf = ops.empty_nhwc(n, cpad, h, w, f.dtype, f.device)
        head = self.classifier[2]
        if f.dtype == torch.bfloat16 and not head[1].training and ops.USE_TENSOR_CORES and head[0].in_channels % 16 == 0:
            # the buffer is private: write all cpad channels (zero weights -> the pad channels hold act(0)), so that
            # every pixel leaves the kernel as one full 64-byte row instead of a 38-byte partial one
            ops.conv(self._padded_head(cpad), f, out=buf)
        else:
            head(f, out=buf)
        return ops.upsample_argmax(buf, x.shape[2:], True, target=target, confmat=confmat, ignore_index=ignore_index,
                                   num_classes=self.num_class)

    def _padded_head(self, cpad):
        """classifier[2] folded with its output channels padded to cpad (scale 1, shift 0, zero weights)."""
        import dataclasses
        head = self.classifier[2]
        folded, post = head._folded()
        if post is not None:
            raise NotImplementedError("non-pointwise activation in the classifier head")
        cached = self.__dict__.get("_rtseg_head_pad")
        if cached is None or cached[0] is not folded:
            def pad(t, fill):
                out = torch.full((cpad,), fill, dtype=torch.float32, device=folded.w_tc.device)
                if t is not None:
                    out[:folded.cout] = t
                return out
            fp = dataclasses.replace(folded, cout=cpad, scale=pad(folded.scale, 1.0), shift=pad(folded.shift, 0.0),
                                     slope=None if folded.slope is None else pad(folded.slope, 0.0), w_tc_scaled=None)
            cached = (folded, fp)
            self.__dict__["_rtseg_head_pad"] = cached
        return cached[1]


# ------------------------------------------------------------------------------ DABNet ----
class DABModule(nn.Module):
    """Depthwise asymmetric bottleneck — reference models/dabnet.py:71-98.

    3x3 (C -> C/2) -> [DW(3,1) -> DW(1,3)] || [DW(3,1,d) -> DW(1,3,d)] -> add -> 1x1 (C/2 -> C) -> += x:
    six launches; the branch sum rides in the second dilated depthwise conv's epilogue and the block
    residual in the 1x1 conv's epilogue (the reference needs 18 ATen kernels + 2 adds)."""

    def __init__(self, channels, dilation, act_type):
        super().__init__()
        assert channels % 2 == 0, "Input channel of DABModule should be multiple of 2.\n"
        hid = channels // 2
        self.init_conv = ConvBNAct(channels, hid, 3, act_type=act_type)
        self.left_branch = nn.Sequential(DWConvBNAct(hid, hid, (3, 1), act_type=act_type),
                                         DWConvBNAct(hid, hid, (1, 3), act_type=act_type))
        self.right_branch = nn.Sequential(DWConvBNAct(hid, hid, (3, 1), dilation=dilation, act_type=act_type),
                                          DWConvBNAct(hid, hid, (1, 3), dilation=dilation, act_type=act_type))
        self.last_conv = ConvBNAct(hid, channels, 1, act_type=act_type)

    def forward(self, x, x_coff=0, out=None, y_coff=0):
        """x_coff: the block's input is channels [x_coff, x_coff + C) of `x`; out / y_coff: write the result into
        a channel slice of a caller-owned buffer."""
        x = ops.to_nhwc(x)
        t = self.init_conv(x, x_coff=x_coff)
        left = self.left_branch[1](self.left_branch[0](t))
        s = self.right_branch[1](self.right_branch[0](t), res=left, res_mode=ops.RES_POST)   # x_left + x_right
        return self.last_conv(s, res=x, res_mode=ops.RES_POST, r_coff=x_coff, out=out, y_coff=y_coff)  # x += residual


def _dab_blocks(channels, count, dilation, act_type):
    return nn.Sequential(*[DABModule(channels, dilation, act_type=act_type) for _ in range(count)])


class DABNet(nn.Module):
    """DABNet — reference models/dabnet.py:16-61 (same layer names / state-dict keys).

    The three concatenations never run: each stage owns one NHWC buffer (channel pitch rounded up to 8 so that
    every slice is 16-byte aligned) and the producers — layer3 / layer4 / layer8, the last DABModule of a stage,
    the average-pool pyramid — write their channel slices of it directly.  The 35-, 131- and 259-channel
    convolutions that read those buffers run on the tensor cores (K zero-filled by the TMA unit)."""

    def __init__(self, num_class=1, n_channel=3, act_type="prelu"):
        super().__init__()
        self.layer1 = ConvBNAct(n_channel, 32, 3, 2, act_type=act_type)
        self.layer2 = ConvBNAct(32, 32, 3, 1, act_type=act_type)
        self.layer3 = ConvBNAct(32, 32, 3, 1, act_type=act_type)
        self.layer4 = InitialBlock(32 + n_channel, 64, act_type=act_type)
        self.layer5_7 = _dab_blocks(64, 3, 2, act_type)
        self.layer8 = ConvBNAct(64 * 2 + n_channel, 128, 3, 2, act_type=act_type)
        self.layer9_10 = _dab_blocks(128, 2, 4, act_type)
        self.layer11_12 = _dab_blocks(128, 2, 8, act_type)
        self.layer13_14 = _dab_blocks(128, 2, 16, act_type)
        self.layer15 = conv1x1(128 * 2 + n_channel, num_class)
        self.num_class = num_class
        self.act_dtype: Optional[torch.dtype] = None

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        self.act_dtype = dtype
        return self

    @staticmethod
    def _stage(blocks, buf, x_coff):
        """Runs a chain of DABModules whose input is buf[:, x_coff : x_coff + C]; the last one writes buf[:, :C]."""
        y, coff = buf, x_coff
        for i, m in enumerate(blocks):
            last = i == len(blocks) - 1
            y = m(y, x_coff=coff, out=buf if last else None, y_coff=0)
            coff = 0
        return buf

    def _layer8(self, buf2, buf3, c3, cin_img):
        """layer8 = ConvBNAct(2*64 + 3 -> 128, 3, stride 2).  On the bf16 path K is split across two launches: the
        n_channel-wide tail (x_d4) runs first as a bare 3x3 stride-2 convolution whose sums enter the 128-channel
        main convolution as a PRE residual (acc + res -> BN -> act).  The main launch then has K = 128 = 2 x 64:
        full 128-byte TMA rows and 18 pipeline stages per tile instead of 81 32-byte-row stages for K = 131."""
        conv, bn, act = self.layer8._parts()
        cmain = conv.in_channels - cin_img
        if buf2.dtype != torch.bfloat16 or cmain % 64 or not ops.USE_TENSOR_CORES:
            return self.layer8(buf2, out=buf3, y_coff=c3)
        _check_eval(self, bn)
        tensors = [conv.weight, conv.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var, getattr(act, "weight", None)]
        cache = self.__dict__.setdefault("_rtseg_cache8", _FoldCache())

        def build():
            main, post = ops.fold_conv(conv, bn, act, in_slice=(0, cmain), host=True)
            side, _ = ops.fold_conv(conv, None, None, in_slice=(cmain, conv.in_channels), tc_small_cin=True, host=True)
            side.shift = None  # the bias (if any) is applied once, by the main launch
            return main, side, post
        main, side, post = cache.get(tensors, build)
        if post is not None:
            return self.layer8(buf2, out=buf3, y_coff=c3)
        tail = ops.conv(side, buf2, x_coff=cmain)
        return ops.conv(main, buf2, res=tail, res_mode=ops.RES_PRE, out=buf3, y_coff=c3)

    def __deepcopy__(self, memo):  # keep the fold cache out of EMA copies (utils/model_ema.py:20)
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k != "_rtseg_cache8":
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new

    def _lowres_padded(self, x):
        x = ops.image_view(x)
        dt = self.act_dtype or _param_dtype(self.layer1)
        n, cin, h, w = x.shape
        half = lambda v: (v - 1) // 2 + 1
        pitch = _pitch8
        h2, w2, h4, w4, h8, w8 = half(h), half(w), half(half(h)), half(half(w)), half(half(half(h))), half(half(half(w)))
        c1, c2, c3 = self.layer3[0].out_channels, self.layer4.conv[0].out_channels + self.layer4.conv[0].in_channels, \
            self.layer8[0].out_channels
        buf1 = ops.empty_nhwc(n, pitch(c1 + cin), h2, w2, dt, x.device)          # [layer3 | x_d2]
        buf2 = ops.empty_nhwc(n, pitch(2 * c2 + cin), h4, w4, dt, x.device)      # [layer5_7 | layer4 | x_d4]
        buf3 = ops.empty_nhwc(n, pitch(2 * c3 + cin), h8, w8, dt, x.device)      # [layer13_14 | layer8 | x_d8]
        ops.avgpool3x3s2(x, out=buf1, y_coff=c1)
        ops.avgpool3x3s2(buf1, out=buf2, y_coff=2 * c2, x_coff=c1, c=cin)
        ops.avgpool3x3s2(buf2, out=buf3, y_coff=2 * c3, x_coff=2 * c2, c=cin)
        stem, conv = self.layer1, self.layer1[0]
        _check_eval(self, stem[1])
        folded, post = stem._folded()
        if post is None and conv.in_channels <= 4:
            y = ops.stem(folded, x, pool=False, out_dtype=dt)
        else:
            y = stem(ops.to_nhwc(x, dt))
        self.layer3(self.layer2(y), out=buf1, y_coff=0)
        self.layer4(buf1, out=buf2, y_coff=c2)
        self._stage(self.layer5_7, buf2, c2)
        self._layer8(buf2, buf3, c3, cin)
        self._stage([*self.layer9_10, *self.layer11_12, *self.layer13_14], buf3, c3)
        logits = ops.empty_nhwc(n, _pitch8(self.num_class), h8, w8, dt, x.device)   # pitch 24 for 19 classes
        return self.layer15(buf3, out=logits)

    def lowres_logits(self, x):
        return _first_channels(self._lowres_padded(x), self.num_class)

    def _train_forward(self, x):
        """The reference's own graph (models/dabnet.py:36-61) as tape nodes: the three concatenations are torch copy
        nodes here, exactly as in the reference; the average-pool pyramid of the image needs no gradient."""
        from . import train
        dt = self.act_dtype or _param_dtype(self.layer1)
        cat = lambda ts: ops.to_nhwc(torch.cat(ts, dim=1))
        with torch.no_grad():
            x_d2 = ops.avgpool3x3s2(x, out_dtype=dt)
            x_d4 = ops.avgpool3x3s2(x_d2)
            x_d8 = ops.avgpool3x3s2(x_d4)
        stem = self.layer1
        xn = ops.to_nhwc(x, dt)
        y = train.conv_bn_act_aligned(stem, xn, pad_input=True) if stem[0].in_channels <= 4 else stem(xn)
        y = self.layer3(self.layer2(y))
        b2 = self.layer4(cat([y, x_d2]))
        y = self.layer5_7(b2)
        b3 = self.layer8(cat([y, b2, x_d4]))
        y = self.layer13_14(self.layer11_12(self.layer9_10(b3)))
        y = self.layer15(cat([y, b3, x_d8]))
        return train.upsample_bilinear(y, x.shape[2:], True)

    def forward(self, x):
        x = ops.image_view(x)
        if self.layer1[1].training:
            return self._train_forward(x)
        return ops.upsample_bilinear(self._lowres_padded(x), x.shape[2:], True, c=self.num_class)

    def predict(self, x, target: Optional[torch.Tensor] = None, confmat: Optional[torch.Tensor] = None,
                ignore_index: int = 255) -> torch.Tensor:
        x = ops.image_view(x)
        return ops.upsample_argmax(self._lowres_padded(x), x.shape[2:], True, target=target, confmat=confmat,
                                   ignore_index=ignore_index, num_classes=self.num_class)


# ------------------------------------------------------------------------------ LEDNet ----
class SSnbtUnit(nn.Module):
    """Split-shuffle non-bottleneck unit — reference models/lednet.py:52-88.

    The two halves are channel SLICES of the input (no clone), each branch's last conv writes its
    half of the output with `+ x` and the activation in its epilogue (no cat, no separate add);
    the channel shuffle is one permutation kernel."""

    def __init__(self, channels, dilation, act_type):
        super().__init__()
        assert channels % 2 == 0, "Input channel should be multiple of 2.\n"
        sc = channels // 2
        self.split_channels = sc
        self.left_branch = nn.Sequential(
            Conv2d(sc, sc, (3, 1), padding=(1, 0)), Activation(act_type),
            ConvBNAct(sc, sc, (1, 3), act_type=act_type),
            Conv2d(sc, sc, (3, 1), padding=(dilation, 0), dilation=dilation), Activation(act_type),
            ConvBNAct(sc, sc, (1, 3), dilation=dilation, act_type=act_type),
        )
        self.right_branch = nn.Sequential(
            Conv2d(sc, sc, (1, 3), padding=(0, 1)), Activation(act_type),
            ConvBNAct(sc, sc, (3, 1), act_type=act_type),
            Conv2d(sc, sc, (1, 3), padding=(0, dilation), dilation=dilation), Activation(act_type),
            ConvBNAct(sc, sc, (3, 1), dilation=dilation, act_type=act_type),
        )
        self.act = Activation(act_type)

    def _branch(self, br, x, coff, out):
        t = br[0](x, x_coff=coff, act=br[1].activation)
        t = br[2](t)
        t = br[3](t, act=br[4].activation)
        act2 = ops.act_spec(self.act.activation)
        if act2[3] is not None or act2[4] is not None:
            raise NotImplementedError("SSnbtUnit: parametrised final activation is not fused")
        br[5](t, res=x, res_mode=ops.RES_POST, out=out, y_coff=coff, act2=act2[0], r_coff=coff)

    def forward(self, x):
        x = ops.to_nhwc(x)
        out = torch.empty_like(x, memory_format=torch.channels_last)
        self._branch(self.left_branch, x, 0, out)
        self._branch(self.right_branch, x, self.split_channels, out)
        return ops.channel_shuffle(out, 2)


class AttentionPyramidNetwork(nn.Module):
    """LEDNet decoder — reference models/lednet.py:91-136.  Convolutions, pooling and resizes are rtseg
    kernels; the two pyramid additions are conv epilogues; the attention product and the final sum are
    element-wise torch ops on 1/8-resolution 19-channel maps (0.1 % of the net's bytes)."""

    def __init__(self, in_channels, out_channels, act_type):
        super().__init__()
        self.left_conv1_1 = ConvBNAct(in_channels, in_channels, 3, 2, act_type=act_type)
        self.left_conv1_2 = ConvBNAct(in_channels, out_channels, 3, act_type=act_type)
        self.left_conv2_1 = ConvBNAct(in_channels, in_channels, 3, 2, act_type=act_type)
        self.left_conv2_2 = ConvBNAct(in_channels, out_channels, 3, act_type=act_type)
        self.left_conv3 = nn.Sequential(ConvBNAct(in_channels, in_channels, 3, 2, act_type=act_type),
                                        ConvBNAct(in_channels, out_channels, 3, act_type=act_type))
        self.mid_branch = ConvBNAct(in_channels, out_channels, act_type=act_type)
        self.right_branch = nn.Sequential(nn.AdaptiveAvgPool2d(1), ConvBNAct(in_channels, out_channels, act_type=act_type))

    def forward(self, x, padded=False):
        """padded=True returns the [n, pitch8(out_channels), h, w] buffer the kernels write (channels beyond
        out_channels are undefined); the default returns the dense tensor the reference returns."""
        x = ops.to_nhwc(x)
        n, cout = x.shape[0], self.mid_branch[0].out_channels
        # every out_channels-wide map lives in a pitch-8 buffer: TMA stores / ring residuals need 16-byte pixels
        buf = lambda size: ops.empty_nhwc(n, _pitch8(cout), int(size[0]), int(size[1]), x.dtype, x.device)
        up = lambda t, size: ops.upsample_bilinear(t, size, True, out=buf(size), c=cout)
        size0 = x.shape[2:]
        l1 = self.left_conv1_1(x)
        size1 = l1.shape[2:]
        l2 = self.left_conv2_1(l1)
        size2 = l2.shape[2:]
        l3 = self.left_conv3[0](l2)
        l3 = up(self.left_conv3[1](l3, out=buf(l3.shape[2:])), size2)
        l2 = up(self.left_conv2_2(l2, res=l3, res_mode=ops.RES_POST, out=buf(size2)), size1)    # x_left2 += x_left3
        l1 = up(self.left_conv1_2(l1, res=l2, res_mode=ops.RES_POST, out=buf(size1)), size0)    # x_left += x_left2
        mid = self.mid_branch(x, out=buf(size0))
        right = self.right_branch[1](ops.adaptive_avgpool(x, 1), out=buf((1, 1)))
        out = torch.addcmul(up(right, size0), mid, l1)                                           # mid * left + right
        return out if padded else _first_channels(out, cout)


class LEDNet(nn.Module):
    """LEDNet — reference models/lednet.py:16-50 (same attribute names / state-dict keys)."""

    def __init__(self, num_class=1, n_channel=3, act_type="relu"):
        super().__init__()
        dils = [1, 2, 5, 9, 2, 5, 9, 17]
        self.encoder = nn.Sequential(
            InitialBlock(n_channel, 32, act_type),
            *[SSnbtUnit(32, 1, act_type=act_type) for _ in range(3)],
            InitialBlock(32, 64, act_type),
            *[SSnbtUnit(64, 1, act_type=act_type) for _ in range(2)],
            InitialBlock(64, 128, act_type),
            *[SSnbtUnit(128, d, act_type=act_type) for d in dils],
        )
        self.apn = AttentionPyramidNetwork(128, num_class, act_type)
        self.num_class = num_class

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        self.encoder[0].act_dtype = dtype
        return self

    def _lowres_padded(self, x):
        return self.apn(self.encoder(ops.image_view(x)), padded=True)

    def lowres_logits(self, x):
        return _first_channels(self._lowres_padded(x), self.num_class)

    def forward(self, x):
        x = ops.image_view(x)
        return ops.upsample_bilinear(self._lowres_padded(x), x.shape[2:], True, c=self.num_class)

    def predict(self, x, target: Optional[torch.Tensor] = None, confmat: Optional[torch.Tensor] = None,
                ignore_index: int = 255) -> torch.Tensor:
        x = ops.image_view(x)
        return ops.upsample_argmax(self._lowres_padded(x), x.shape[2:], True, target=target, confmat=confmat,
                                   ignore_index=ignore_index, num_classes=self.num_class)


# -------------------------------------------------------------------------------- ENet ----
class Upsample(nn.Module):
    """reference models/enet.py:184-204: a bare ConvTranspose2d(k, s = 2, bias=False) ('deconvolution') or
    ConvBNAct 1x1 + bilinear nn.Upsample (any other type)."""

    def __init__(self, in_channels, out_channels, scale_factor=2, kernel_size=None, upsample_type=None, act_type="relu"):
        super().__init__()
        if upsample_type == "deconvolution":
            if kernel_size is None:
                kernel_size = 2 * scale_factor - 1
            self.up_conv = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=kernel_size, stride=scale_factor,
                                              padding=(kernel_size - 1) // 2, output_padding=1, bias=False)
        else:
            self.up_conv = nn.Sequential(ConvBNAct(in_channels, out_channels, 1, act_type=act_type),
                                         nn.Upsample(scale_factor=scale_factor, mode="bilinear"))

    def _folded(self):
        cache = self.__dict__.setdefault("_rtseg_cache", _FoldCache())
        return cache.get([self.up_conv.weight], lambda: ops.fold_conv(self.up_conv, None, None, host=True))

    def forward(self, x, labels=None, want_logits=True):
        x = ops.to_nhwc(x)
        deconv = isinstance(self.up_conv, nn.ConvTranspose2d)
        if self.training and labels is None:
            from . import train   # tape nodes: the transposed conv's gradients are its adjoint stride-2 convolution
            if deconv:
                return train.conv(self.up_conv, x)
            y = self.up_conv[0](x)
            sf = self.up_conv[1].scale_factor
            return train.upsample_bilinear(y, (int(y.shape[2] * sf), int(y.shape[3] * sf)), False)
        if deconv:
            return ops.conv(self._folded()[0], x, labels=labels, want_logits=want_logits)
        y = self.up_conv[0](x)
        sf = self.up_conv[1].scale_factor
        return ops.upsample_bilinear(y, (int(y.shape[2] * sf), int(y.shape[3] * sf)), False)

    __deepcopy__ = Conv2d.__deepcopy__


class Bottleneck(nn.Module):
    """ENet bottleneck — reference models/enet.py:119-181.  The shortcut add (or the pooled / unpooled left branch)
    and the PReLU are the epilogue of the last 1x1 convolution; MaxPool2d(2, 2, return_indices) / MaxUnpool2d are the
    rtseg pooling pair (one byte per index)."""

    def __init__(self, in_channels, out_channels, conv_type, act_type="prelu", upsample_type="regular", dilation=1,
                 drop_p=0.1, shrink_ratio=0.25):
        super().__init__()
        self.conv_type = conv_type
        hid = int(in_channels * shrink_ratio)
        if conv_type == "regular":
            self.right_init_conv = nn.Sequential(ConvBNAct(in_channels, hid, 1), ConvBNAct(hid, hid))
        elif conv_type == "downsampling":
            self.left_pool = nn.MaxPool2d(2, 2, return_indices=True)
            self.left_conv = ConvBNAct(in_channels, out_channels, 1)
            self.right_init_conv = nn.Sequential(ConvBNAct(in_channels, hid, 3, 2), ConvBNAct(hid, hid))
        elif conv_type == "upsampling":
            self.left_conv = ConvBNAct(in_channels, out_channels, 1)
            self.left_pool = nn.MaxUnpool2d(2, 2)
            self.right_init_conv = nn.Sequential(ConvBNAct(in_channels, hid, 1),
                                                 Upsample(hid, hid, scale_factor=2, kernel_size=3, upsample_type=upsample_type))
        elif conv_type == "dilate":
            self.right_init_conv = nn.Sequential(ConvBNAct(in_channels, hid, 1), ConvBNAct(hid, hid, dilation=dilation))
        elif conv_type == "asymmetric":
            self.right_init_conv = nn.Sequential(ConvBNAct(in_channels, hid, 1), ConvBNAct(hid, hid, (5, 1)),
                                                 ConvBNAct(hid, hid, (1, 5)))
        else:
            raise ValueError(f"[!] Unsupport convolution type: {conv_type}")
        self.right_last_conv = nn.Sequential(conv1x1(hid, out_channels), nn.Dropout(drop_p))
        self.act = Activation(act_type)

    def forward(self, x, indices=None):
        x = ops.to_nhwc(x)
        last, act = self.right_last_conv[0], self.act.activation
        if self.training:
            from . import train   # Dropout is torch's own elementwise mask node (reference enet.py:160)
            right = self.right_last_conv[1](last(self.right_init_conv(x)))
            if self.conv_type == "downsampling":
                pooled, idx = train.maxpool2x2_with_indices(x)       # adjoint = the unpooling kernel
                return train.act(act, self.left_conv(pooled), right), idx
            if self.conv_type == "upsampling":
                if indices is None:
                    raise ValueError("Upsampling-type conv needs pooling indices.")
                return train.act(act, train.maxunpool2x2(self.left_conv(x), indices), right)
            return train.act(act, right, x)
        right = self.right_init_conv(x)
        if self.conv_type == "downsampling":
            pooled, idx = ops.maxpool2x2_with_indices(x)
            left = self.left_conv(pooled)
            return last(right, res=left, res_mode=ops.RES_PRE, act=act), idx      # act(x_left + x_right)
        if self.conv_type == "upsampling":
            if indices is None:
                raise ValueError("Upsampling-type conv needs pooling indices.")
            left = ops.maxunpool2x2(self.left_conv(x), indices)
            return last(right, res=left, res_mode=ops.RES_PRE, act=act)
        return last(right, res=x, res_mode=ops.RES_PRE, act=act)                   # act(x + x_right)


class BottleNeck1(nn.Module):
    """reference models/enet.py:51-66"""

    def __init__(self, in_channels, out_channels, act_type="prelu", drop_p=0.01):
        super().__init__()
        self.conv_pool = Bottleneck(in_channels, out_channels, "downsampling", act_type, drop_p=drop_p)
        self.conv_regular = nn.Sequential(*[Bottleneck(out_channels, out_channels, "regular", act_type, drop_p=drop_p)
                                            for _ in range(4)])

    def forward(self, x):
        x, indices = self.conv_pool(x)
        return self.conv_regular(x), indices


class BottleNeck23(nn.Module):
    """reference models/enet.py:69-95"""

    def __init__(self, in_channels, out_channels, act_type="prelu", downsample=True):
        super().__init__()
        self.downsample = downsample
        if downsample:
            self.conv_pool = Bottleneck(in_channels, out_channels, "downsampling", act_type=act_type)
        c = out_channels
        self.conv_regular = nn.Sequential(
            Bottleneck(c, c, "regular", act_type), Bottleneck(c, c, "dilate", act_type, dilation=2),
            Bottleneck(c, c, "asymmetric", act_type), Bottleneck(c, c, "dilate", act_type, dilation=4),
            Bottleneck(c, c, "regular", act_type), Bottleneck(c, c, "dilate", act_type, dilation=8),
            Bottleneck(c, c, "asymmetric", act_type), Bottleneck(c, c, "dilate", act_type, dilation=16))

    def forward(self, x):
        indices = None
        if self.downsample:
            x, indices = self.conv_pool(x)
        x = self.conv_regular(x)
        return (x, indices) if self.downsample else x


class BottleNeck45(nn.Module):
    """reference models/enet.py:98-116"""

    def __init__(self, in_channels, out_channels, act_type="prelu", upsample_type=None, extra_conv=False):
        super().__init__()
        self.extra_conv = extra_conv
        self.conv_unpool = Bottleneck(in_channels, out_channels, "upsampling", act_type, upsample_type)
        self.conv_regular = Bottleneck(out_channels, out_channels, "regular", act_type)
        if extra_conv:
            self.conv_extra = Bottleneck(out_channels, out_channels, "regular", act_type)

    def forward(self, x, indices):
        x = self.conv_regular(self.conv_unpool(x, indices))
        return self.conv_extra(x) if self.extra_conv else x


class ENet(nn.Module):
    """ENet — reference models/enet.py:14-35 (same attribute names / state-dict keys; BASELINE configs[0])."""

    def __init__(self, num_class=1, n_channel=3, act_type="prelu", upsample_type="deconvolution"):
        super().__init__()
        self.initial = InitialBlock(n_channel, 16, act_type)
        self.bottleneck1 = BottleNeck1(16, 64, act_type)
        self.bottleneck2 = BottleNeck23(64, 128, act_type, True)
        self.bottleneck3 = BottleNeck23(128, 128, act_type, False)
        self.bottleneck4 = BottleNeck45(128, 64, act_type, upsample_type, True)
        self.bottleneck5 = BottleNeck45(64, 16, act_type, upsample_type, False)
        # (the reference does not hand upsample_type to this one, models/enet.py:24: the head is always
        # ConvBNAct 1x1 + bilinear x2, align_corners=False)
        self.fullconv = Upsample(16, num_class, scale_factor=2, act_type=act_type)
        self.num_class = num_class

    def set_act_dtype(self, dtype: Optional[torch.dtype]):
        self.initial.act_dtype = dtype
        return self

    def _trunk(self, x):
        x = self.initial(ops.image_view(x))
        x, indices1 = self.bottleneck1(x)
        x, indices2 = self.bottleneck2(x)
        x = self.bottleneck3(x)
        x = self.bottleneck4(x, indices2)
        return self.bottleneck5(x, indices1)

    def forward(self, x):
        return self.fullconv(self._trunk(x))

    HEAD_ALIGN_CORNERS = False   # nn.Upsample(scale_factor=2, mode='bilinear') of the head (the other nets: True)

    def lowres_logits(self, x):
        """The head's 1x1 ConvBNAct output (half resolution), before its bilinear x2."""
        if isinstance(self.fullconv.up_conv, nn.ConvTranspose2d):
            raise NotImplementedError("a transposed-conv head has no low-resolution logits")
        return self.fullconv.up_conv[0](self._trunk(x))

    def predict(self, x, target: Optional[torch.Tensor] = None, confmat: Optional[torch.Tensor] = None,
                ignore_index: int = 255) -> torch.Tensor:
        """uint8 class map: the head's bilinear x2 and the argmax run as the fused upsample+argmax kernel."""
        x = ops.image_view(x)
        f = self._trunk(x)
        if isinstance(self.fullconv.up_conv, nn.ConvTranspose2d):
            n, _, h, w = x.shape
            labels = torch.empty((n, h, w), dtype=torch.uint8, device=x.device)
            self.fullconv(f, labels=labels, want_logits=False)
            return labels
        low = self.fullconv.up_conv[0](f)
        return ops.upsample_argmax(low, (2 * low.shape[2], 2 * low.shape[3]), False, target=target, confmat=confmat,
                                   ignore_index=ignore_index)


from .nets_dual import BiSeNetv2, DDRNet  # noqa: E402  (config 4; they use this module's leaf blocks only via modules.py)

MODEL_HUB = {"erfnet": ERFNet, "fastscnn": FastSCNN, "dabnet": DABNet, "lednet": LEDNet, "enet": ENet,
             "bisenetv2": BiSeNetv2, "ddrnet": DDRNet}
AUX_MODELS = ("bisenetv2", "ddrnet")   # reference models/__init__.py:13-14


def get_model(config) -> nn.Module:
    """Subset of reference models/__init__.py:47-99 for the nets defined in this package."""
    name = config.model
    if name not in MODEL_HUB:
        raise NotImplementedError(f"Unsupport model type: {name}")
    if getattr(config, "use_detail_head", False):
        raise ValueError(f"Model {name} does not support detail heads.\n")
    if name in AUX_MODELS:
        return MODEL_HUB[name](num_class=config.num_class, use_aux=bool(getattr(config, "use_aux", False)))
    if getattr(config, "use_aux", False):
        raise ValueError(f"Model {name} does not support auxiliary heads.\n")
    return MODEL_HUB[name](num_class=config.num_class)


__all__ = ["ERFNet", "FastSCNN", "DABNet", "LEDNet", "ENet", "BiSeNetv2", "DDRNet", "Bottleneck", "InitialBlock", "NonBt1DBlock", "InvertedResidual", "DABModule",
           "SSnbtUnit", "get_model", "MODEL_HUB"]
_ = (_bn_tensors, _FoldCache)
