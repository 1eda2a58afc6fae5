"""Functional fp32 restatement of the reference's building blocks (reference models/modules.py).

Every function takes the network's state dict `sd` and the key prefix of the module it restates,
and evaluates exactly the op sequence the reference module runs in eval mode:
conv (F.conv2d) -> batch_norm with running statistics -> activation.  Unfused on purpose.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm2d default, never overridden in the reference

# Optional emulation of reduced-precision STORAGE (not arithmetic): when set to torch.bfloat16 every
# fused layer's output, the weights of every dense conv (groups == 1: all of them run on the tensor
# cores) and the image as the stem reads it are rounded to bf16 exactly where the CUDA path rounds them; all arithmetic stays
# fp32.  Used by the bf16 parity tests to separate "the kernel is wrong" from "bf16 has 8 bits".
STORAGE = None


def q(t):
    """Round a layer output to the emulated storage type."""
    return t if STORAGE is None else t.to(STORAGE).float()


def qw(w, groups=1, transposed=False):
    cin = w.shape[0] if transposed else w.shape[1] * groups
    if STORAGE is None or groups != 1:   # depthwise weights stay fp32 in the kernels
        return w
    return w.to(STORAGE).float()


def activation(x, act_type: str, sd=None, prefix: str = "", **kw):
    """Activation hub of reference models/modules.py:111-131 (the pointwise members)."""
    a = act_type.lower()
    if a == "none":
        return x
    if a == "relu":
        return F.relu(x)
    if a == "relu6":
        return F.relu6(x)
    if a == "leakyrelu":
        return F.leaky_relu(x, kw.get("negative_slope", 0.01))
    if a == "prelu":
        return F.prelu(x, sd[prefix + "activation.weight"])
    if a == "elu":
        return F.elu(x, kw.get("alpha", 1.0))
    if a == "celu":
        return F.celu(x, kw.get("alpha", 1.0))
    if a == "selu":
        return F.selu(x)
    if a == "hardswish":
        return F.hardswish(x)
    if a == "hardtanh":
        return F.hardtanh(x, kw.get("min_val", -1.0), kw.get("max_val", 1.0))
    if a == "gelu":
        return F.gelu(x)
    if a == "silu":
        return F.silu(x)
    if a == "sigmoid":
        return torch.sigmoid(x)
    if a == "tanh":
        return torch.tanh(x)
    raise NotImplementedError(f"Unsupport activation type: {a}")


# Training restatement: with TRAINING = True every BatchNorm uses batch statistics (and updates the running
# statistics in `sd` in place, momentum 0.1) like the reference's modules in .train(); torch autograd on the
# tensors of `sd` then yields the reference gradients.
TRAINING = False


def batch_norm(x, sd, prefix: str):
    """nn.BatchNorm2d: eval mode (running statistics) or, with TRAINING, batch statistics."""
    return F.batch_norm(x, sd[prefix + "running_mean"], sd[prefix + "running_var"], sd[prefix + "weight"],
                        sd[prefix + "bias"], TRAINING, 0.1 if TRAINING else 0.0, BN_EPS)


def _pad(kernel_size, dilation):
    # reference models/modules.py:49-52 / :76-79
    if isinstance(kernel_size, (tuple, list)):
        return ((kernel_size[0] - 1) // 2 * dilation, (kernel_size[1] - 1) // 2 * dilation)
    return (kernel_size - 1) // 2 * dilation


def conv_bn_act(x, sd, prefix, kernel_size=3, stride=1, dilation=1, act_type="relu", groups=1, post_add=None,
                post_act="none", **kw):
    """ConvBNAct (reference models/modules.py:73-85); children `0` conv, `1` bn, `2` activation.

    `post_add` / `post_act` restate a caller's `y += t; y = act(y)` that follows the block (DABModule,
    SSnbtUnit, AttentionPyramidNetwork) so that storage emulation rounds once, where the fused kernel does."""
    if groups == 1 and x.shape[1] <= 4:
        x = q(x)   # the tensor-core stem converts the fp32 image to bf16 while it builds its im2col rows
    y = F.conv2d(x, qw(sd[prefix + "0.weight"], groups), sd.get(prefix + "0.bias"), stride, _pad(kernel_size, dilation),
                 dilation, groups)
    y = batch_norm(y, sd, prefix + "1.")
    y = activation(y, act_type, sd, prefix + "2.", **kw)
    if post_add is not None:
        y = activation(y + post_add, post_act)
    return q(y)


def dw_conv_bn_act(x, sd, prefix, kernel_size, stride=1, dilation=1, act_type="relu", **kw):
    """DWConvBNAct (reference models/modules.py:46-59): groups = in_channels."""
    return conv_bn_act(x, sd, prefix, kernel_size, stride, dilation, act_type, groups=x.shape[1], **kw)


def pw_conv_bn_act(x, sd, prefix, act_type="relu", **kw):
    """PWConvBNAct (reference models/modules.py:63-69)."""
    return conv_bn_act(x, sd, prefix, 1, 1, 1, act_type, **kw)


def ds_conv_bn_act(x, sd, prefix, kernel_size, stride=1, dilation=1, act_type="relu", **kw):
    """DSConvBNAct (reference models/modules.py:36-42): children `0` depthwise, `1` pointwise."""
    y = dw_conv_bn_act(x, sd, prefix + "0.", kernel_size, stride, dilation, act_type, **kw)
    return pw_conv_bn_act(y, sd, prefix + "1.", act_type, **kw)


def deconv_bn_act(x, sd, prefix, scale_factor=2, act_type="relu", **kw):
    """DeConvBNAct (reference models/modules.py:89-108): k = 2s-1, p = (k-1)//2, output_padding = s-1."""
    k = 2 * scale_factor - 1
    y = F.conv_transpose2d(x, qw(sd[prefix + "up_conv.0.weight"], 1, True), sd.get(prefix + "up_conv.0.bias"),
                           scale_factor, (k - 1) // 2, scale_factor - 1)
    y = batch_norm(y, sd, prefix + "up_conv.1.")
    return q(activation(y, act_type, sd, prefix + "up_conv.2.", **kw))


def initial_block(x, sd, prefix, act_type="relu"):
    """InitialBlock (reference models/enet.py:38-48): cat[conv3x3 s2 branch, maxpool(3,2,1) of the input]."""
    a = conv_bn_act(x, sd, prefix + "conv.", 3, 2, 1, act_type)
    b = q(F.max_pool2d(x, 3, 2, 1))
    return torch.cat([a, b], dim=1)


def pyramid_pooling(x, sd, prefix, act_type="relu", pool_sizes=(1, 2, 4, 6)):
    """PyramidPoolingModule (reference models/modules.py:134-158)."""
    size = x.shape[2:]
    feats = [x]
    for i, p in enumerate(pool_sizes, start=1):
        z = q(F.adaptive_avg_pool2d(x, p))
        z = q(F.conv2d(z, qw(sd[f"{prefix}stage{i}.1.weight"])))
        feats.append(q(F.interpolate(z, size, mode="bilinear", align_corners=True)))
    return pw_conv_bn_act(torch.cat(feats, dim=1), sd, prefix + "conv.", act_type)
