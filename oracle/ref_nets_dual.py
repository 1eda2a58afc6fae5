"""TEST INFRASTRUCTURE (oracle): functional fp32 restatement of the two-branch nets of BASELINE config 4.

BiSeNetv2 (reference models/bisenetv2.py) and DDRNet (reference models/ddrnet.py) in eval mode, written with
torch.nn.functional on a state dict, op for op as the reference modules run them.  Pinned to the REAL reference by
tests/golden/cfg4_outputs.npz (tests/golden/make_golden_cfg4.py).  `B.q()` marks the tensors the CUDA path
materialises (storage emulation for the bf16 tests); it is the identity in fp32.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import ref_blocks as B


def _up(x, size):
    return F.interpolate(x, size, mode="bilinear", align_corners=True)


def seg_head(x, sd, prefix, act_type="relu"):
    """SegHead (reference models/modules.py:161-166): 3x3 ConvBNAct -> 1x1 conv."""
    y = B.conv_bn_act(x, sd, prefix + "0.", 3, 1, 1, act_type)
    return B.q(F.conv2d(y, B.qw(sd[prefix + "1.weight"]), sd.get(prefix + "1.bias")))


# ----------------------------------------------------------------------------- BiSeNetv2 ----
def gather_expansion(x, sd, prefix, stride, act_type="relu"):
    """GatherExpansionLayer (reference models/bisenetv2.py:130-162)."""
    p = prefix + "left_branch."
    y = B.conv_bn_act(x, sd, p + "0.", 3, 1, 1, act_type)
    if stride == 2:
        y = B.dw_conv_bn_act(y, sd, p + "1.", 3, 2, 1, "none")
        y = B.dw_conv_bn_act(y, sd, p + "2.", 3, 1, 1, "none")
        other = B.dw_conv_bn_act(x, sd, prefix + "right_branch.0.", 3, 2, 1, "none")
        other = B.pw_conv_bn_act(other, sd, prefix + "right_branch.1.", "none")
        return B.pw_conv_bn_act(y, sd, p + "3.", "none", post_add=other, post_act=act_type)
    y = B.dw_conv_bn_act(y, sd, p + "1.", 3, 1, 1, "none")
    return B.pw_conv_bn_act(y, sd, p + "2.", "none", post_add=x, post_act=act_type)


def stem_block(x, sd, prefix, act_type="relu"):
    """StemBlock (reference models/bisenetv2.py:106-127)."""
    x = B.conv_bn_act(x, sd, prefix + "conv_init.", 3, 2, 1, act_type)
    left = B.conv_bn_act(x, sd, prefix + "left_branch.0.", 1, 1, 1, act_type)
    left = B.conv_bn_act(left, sd, prefix + "left_branch.1.", 3, 2, 1, act_type)
    right = F.max_pool2d(x, 3, 2, 1)
    return B.conv_bn_act(torch.cat([left, right], dim=1), sd, prefix + "conv_last.", 3, 1, 1, act_type)


def context_embedding(x, sd, prefix, act_type="relu"):
    """ContextEmbeddingBlock (reference models/bisenetv2.py:153-170)."""
    res = B.q(B.batch_norm(F.adaptive_avg_pool2d(x, 1), sd, prefix + "pool.1."))
    res = B.conv_bn_act(res, sd, prefix + "conv_mid.", 1, 1, 1, act_type)
    return B.q(F.conv2d(B.q(res + x), B.qw(sd[prefix + "conv_last.weight"]), sd.get(prefix + "conv_last.bias"), 1, 1))


def bilateral_guided_aggregation(x_d, x_s, sd, prefix, act_type="relu"):
    """BilateralGuidedAggregationLayer (reference models/bisenetv2.py:172-203)."""
    d_high = B.dw_conv_bn_act(x_d, sd, prefix + "detail_high.0.", 3, 1, 1, act_type)
    d_high = B.q(F.conv2d(d_high, B.qw(sd[prefix + "detail_high.1.weight"])))
    d_low = B.dw_conv_bn_act(x_d, sd, prefix + "detail_low.0.", 3, 2, 1, act_type)
    d_low = B.q(F.avg_pool2d(d_low, 3, 2, 1))
    s_high = B.conv_bn_act(x_s, sd, prefix + "semantic_high.0.", 3, 1, 1, act_type)
    s_high = torch.sigmoid(F.interpolate(s_high, scale_factor=4, mode="bilinear", align_corners=True))
    s_low = B.dw_conv_bn_act(x_s, sd, prefix + "semantic_low.0.", 3, 1, 1, act_type)
    s_low = torch.sigmoid(B.q(F.conv2d(s_low, B.qw(sd[prefix + "semantic_low.1.weight"]))))
    x_high = B.q(d_high * s_high)
    x_low = B.q(d_low * s_low)
    res = B.q(x_high + _up(x_low, x_high.shape[2:]))
    return B.conv_bn_act(res, sd, prefix + "conv_last.", 3, 1, 1, act_type)


def bisenetv2_lowres_logits(sd, x, act_type="relu"):
    d = x
    for i, stride in enumerate((2, 1, 2, 1, 1, 2, 1, 1)):          # DetailBranch (:44-56)
        d = B.conv_bn_act(d, sd, f"detail_branch.{i}.", 3, stride, 1, act_type)
    p = "semantic_branch."
    s = stem_block(x, sd, p + "stage1to2.", act_type)
    for stage, strides in (("stage3", (2, 1)), ("stage4", (2, 1)), ("stage5_1to4", (2, 1, 1, 1))):
        for i, stride in enumerate(strides):
            s = gather_expansion(s, sd, f"{p}{stage}.{i}.", stride, act_type)
    s = context_embedding(s, sd, p + "stage5_5.", act_type)
    y = bilateral_guided_aggregation(d, s, sd, "bga_layer.", act_type)
    return seg_head(y, sd, "seg_head.", act_type)


def bisenetv2_forward(sd, x, act_type="relu"):
    """BiSeNetv2.forward with is_training=False (reference models/bisenetv2.py:27-41)."""
    return _up(bisenetv2_lowres_logits(sd, x, act_type), x.shape[2:])


# -------------------------------------------------------------------------------- DDRNet ----
def rb(x, sd, prefix, cin, cout, stride, act_type="relu"):
    """RB (reference models/ddrnet.py:168-190); its own activation is nn.ReLU whatever act_type is."""
    identity = x
    if stride > 1 or cin != cout:
        identity = B.conv_bn_act(x, sd, prefix + "conv_down.", 1, stride, 1, "none")
    y = B.conv_bn_act(x, sd, prefix + "conv1.", 3, stride, 1, act_type)
    return B.conv_bn_act(y, sd, prefix + "conv2.", 3, 1, 1, "none", post_add=identity, post_act="relu")


def rbb(x, sd, prefix, cin, cout, stride, act_type="relu"):
    """RBB (reference models/ddrnet.py:193-220)."""
    identity = x
    if stride > 1 or cin != cout:
        identity = B.conv_bn_act(x, sd, prefix + "conv_down.", 1, stride, 1, "none")
    y = B.conv_bn_act(x, sd, prefix + "conv1.", 1, 1, 1, act_type)
    y = B.conv_bn_act(y, sd, prefix + "conv2.", 3, stride, 1, act_type)
    return B.conv_bn_act(y, sd, prefix + "conv3.", 1, 1, 1, "none", post_add=identity, post_act=act_type)


def _blocks(block, x, sd, prefix, cin, cout, stride, repeat, act_type):
    for i in range(repeat):
        x = block(x, sd, f"{prefix}{i}.", cin if i == 0 else cout, cout, stride if i == 0 else 1, act_type)
    return x


def bilateral_fusion(x_low, x_high, sd, prefix, stride, act_type="relu"):
    """BilateralFusion (reference models/ddrnet.py:223-243)."""
    fuse_low = B.conv_bn_act(x_low, sd, prefix + "conv_low.", 1, 1, 1, "none")
    new_low = B.conv_bn_act(x_high, sd, prefix + "conv_high.", 3, stride, 1, "none", post_add=x_low, post_act=act_type)
    new_high = B.q(B.activation(x_high + _up(fuse_low, x_high.shape[2:]), act_type))
    return new_low, new_high


def dappm(x, sd, prefix, act_type="relu"):
    """DAPPM (reference models/ddrnet.py:246-291)."""
    size = x.shape[2:]
    y0 = B.conv_bn_act(x, sd, prefix + "conv0.", 1, 1, 1, act_type)
    ys = [B.conv_bn_act(x, sd, prefix + "conv1.", 1, 1, 1, act_type)]
    for i, (k, s) in enumerate(((5, 2), (9, 4), (17, 8), (-1, -1)), start=2):
        pooled = F.adaptive_avg_pool2d(x, 1) if k == -1 else F.avg_pool2d(x, k, s, (k - 1) // 2)
        pooled = B.q(F.conv2d(B.q(pooled), B.qw(sd[f"{prefix}pool{i}.1.weight"])))
        ys.append(B.conv_bn_act(B.q(ys[-1] + _up(pooled, size)), sd, f"{prefix}conv{i}.", 3, 1, 1, act_type))
    return B.conv_bn_act(torch.cat(ys, dim=1), sd, prefix + "conv_last.", 1, 1, 1, act_type, post_add=y0)


ARCH_HUB = {"DDRNet-23-slim": (32, [2, 2, 2, 0, 2, 1]), "DDRNet-23": (64, [2, 2, 2, 0, 2, 1]),
            "DDRNet-39": (64, [3, 4, 3, 3, 3, 1])}


def ddrnet_lowres_logits(sd, x, act_type="relu", arch_type="DDRNet-23-slim"):
    c, rt = ARCH_HUB[arch_type]
    x = B.conv_bn_act(x, sd, "conv1.", 3, 2, 1, act_type)
    x = B.conv_bn_act(x, sd, "conv2.conv.0.", 3, 2, 1, act_type)                       # Stage2 (:68-82)
    for i in range(rt[0]):
        x = rb(x, sd, f"conv2.conv.{i + 1}.", c, c, 1, act_type)
    x = _blocks(rb, x, sd, "conv3.conv.", c, 2 * c, 2, rt[1], act_type)                # Stage3 (:92-98)
    p = "conv4."                                                                       # Stage4 (:101-131)
    low = _blocks(rb, x, sd, p + "low_conv1.", 2 * c, 4 * c, 2, rt[2], act_type)
    high = _blocks(rb, x, sd, p + "high_conv1.", 2 * c, 2 * c, 1, rt[2], act_type)
    low, high = bilateral_fusion(low, high, sd, p + "bilateral_fusion1.", 2)
    if rt[3] > 0:
        low = _blocks(rb, low, sd, p + "low_conv2.", 4 * c, 4 * c, 1, rt[3], act_type)
        high = _blocks(rb, high, sd, p + "high_conv2.", 2 * c, 2 * c, 1, rt[3], act_type)
        low, high = bilateral_fusion(low, high, sd, p + "bilateral_fusion2.", 2)
    p = "conv5."                                                                       # Stage5 (:134-165)
    low = _blocks(rb, low, sd, p + "low_conv1.", 4 * c, 8 * c, 2, rt[4], act_type)
    high = _blocks(rb, high, sd, p + "high_conv1.", 2 * c, 2 * c, 1, rt[4], act_type)
    low, high = bilateral_fusion(low, high, sd, p + "bilateral_fusion.", 4)
    low = _blocks(rbb, low, sd, p + "low_conv2.", 8 * c, 16 * c, 2, rt[5], act_type)
    low = dappm(low, sd, p + "dappm.")
    high = _blocks(rbb, high, sd, p + "high_conv2.", 2 * c, 4 * c, 1, rt[5], act_type)
    return seg_head(B.q(high + _up(low, high.shape[2:])), sd, "seg_head.", act_type)


def ddrnet_forward(sd, x, act_type="relu"):
    """DDRNet.forward with is_training=False (reference models/ddrnet.py:40-65)."""
    return _up(ddrnet_lowres_logits(sd, x, act_type), x.shape[2:])
