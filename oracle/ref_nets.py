"""Functional fp32 restatement of the benchmark networks and their seeded weight recipe."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import ref_blocks as B


# ----------------------------------------------------------------------------- ERFNet ----
def non_bt_1d(x, sd, prefix, dilation=1, act_type="relu"):
    """NonBt1DBlock (reference models/erfnet.py:62-82); inner activations are always ReLU (:66-68)."""
    y = B.conv_bn_act(x, sd, prefix + "conv.0.", (3, 1), 1, 1, "relu")
    y = B.conv_bn_act(y, sd, prefix + "conv.1.", (1, 3), 1, 1, "relu")
    y = B.conv_bn_act(y, sd, prefix + "conv.2.", (3, 1), 1, dilation, "relu")
    y = F.conv2d(y, B.qw(sd[prefix + "conv.3.weight"]), None, 1, (0, dilation), dilation)
    y = y + x
    y = B.batch_norm(y, sd, prefix + "bn_act.0.")
    return B.q(B.activation(y, act_type, sd, prefix + "bn_act.1."))


def erfnet_forward(sd, x, act_type="relu"):
    """ERFNet.forward (reference models/erfnet.py:35-46)."""
    x = B.initial_block(x, sd, "layer1.", act_type)
    x = B.initial_block(x, sd, "layer2.", act_type)
    for i in range(5):
        x = non_bt_1d(x, sd, f"layer3_7.{i}.", 1, act_type)
    x = B.initial_block(x, sd, "layer8.", act_type)
    for i, d in enumerate([2, 4, 8, 16, 2, 4, 8, 16]):
        x = non_bt_1d(x, sd, f"layer9_16.{i}.", d, act_type)
    x = B.deconv_bn_act(x, sd, "layer17.", 2, act_type)
    for i in range(2):
        x = non_bt_1d(x, sd, f"layer18_19.{i}.", 1, act_type)
    x = B.deconv_bn_act(x, sd, "layer20.", 2, act_type)
    for i in range(2):
        x = non_bt_1d(x, sd, f"layer21_22.{i}.", 1, act_type)
    return B.deconv_bn_act(x, sd, "layer23.", 2, act_type)


# -------------------------------------------------------------------------- Fast-SCNN ----
_IR_SETTING = [(6, 64, 3, 2), (6, 96, 2, 2), (6, 128, 3, 1)]  # t, c, n, s — reference models/fastscnn.py:47-52


def inverted_residual(x, sd, prefix, cin, cout, stride, act_type="relu"):
    """InvertedResidual (reference models/fastscnn.py:105-121)."""
    y = B.pw_conv_bn_act(x, sd, prefix + "conv.0.", act_type)
    y = B.dw_conv_bn_act(y, sd, prefix + "conv.1.", 3, stride, 1, act_type)
    y = B.conv_bn_act(y, sd, prefix + "conv.2.", 1, 1, 1, "none")
    return B.q(x + y) if (stride == 1 and cin == cout) else y


def fastscnn_lowres_logits(sd, x, act_type="relu"):
    # LearningToDownsample (reference models/fastscnn.py:35-41)
    p = "learning_to_downsample."
    hi = B.conv_bn_act(x, sd, p + "0.", 3, 2, 1, act_type)
    hi = B.ds_conv_bn_act(hi, sd, p + "1.", 3, 2, 1, act_type)
    hi = B.ds_conv_bn_act(hi, sd, p + "2.", 3, 2, 1, act_type)
    # GlobalFeatureExtractor (:44-70)
    lo, cin, idx = hi, 64, 0
    for _t, c, n, s in _IR_SETTING:
        for i in range(n):
            lo = inverted_residual(lo, sd, f"global_feature_extractor.bottlenecks.{idx}.", cin, c, s if i == 0 else 1,
                                   act_type)
            cin, idx = c, idx + 1
    lo = B.pyramid_pooling(lo, sd, "global_feature_extractor.ppm.", act_type)
    # FeatureFusionModule (:73-93)
    p = "feature_fusion."
    h = B.q(F.conv2d(hi, B.qw(sd[p + "higher_res_conv.weight"])))
    l = B.q(F.interpolate(lo, hi.shape[2:], mode="bilinear", align_corners=True))
    l = B.dw_conv_bn_act(l, sd, p + "lower_res_conv.0.", 3, 1, 1, act_type)
    l = F.conv2d(l, B.qw(sd[p + "lower_res_conv.1.weight"]))
    y = B.q(B.activation(B.batch_norm(h + l, sd, p + "non_linear.0."), act_type, sd, p + "non_linear.1."))
    # Classifier (:96-102)
    p = "classifier."
    y = B.ds_conv_bn_act(y, sd, p + "0.", 3, 1, 1, act_type)
    y = B.ds_conv_bn_act(y, sd, p + "1.", 3, 1, 1, act_type)
    return B.pw_conv_bn_act(y, sd, p + "2.", act_type)


def fastscnn_forward(sd, x, act_type="relu"):
    """FastSCNN.forward (reference models/fastscnn.py:24-32): x8 bilinear, align_corners=True."""
    return F.interpolate(fastscnn_lowres_logits(sd, x, act_type), x.shape[2:], mode="bilinear", align_corners=True)


# ----------------------------------------------------------------------------- DABNet ----
def dab_module(x, sd, prefix, dilation, act_type="prelu"):
    """DABModule (reference models/dabnet.py:71-98)."""
    t = B.conv_bn_act(x, sd, prefix + "init_conv.", 3, 1, 1, act_type)
    left = B.dw_conv_bn_act(t, sd, prefix + "left_branch.0.", (3, 1), 1, 1, act_type)
    left = B.dw_conv_bn_act(left, sd, prefix + "left_branch.1.", (1, 3), 1, 1, act_type)
    right = B.dw_conv_bn_act(t, sd, prefix + "right_branch.0.", (3, 1), 1, dilation, act_type)
    both = B.dw_conv_bn_act(right, sd, prefix + "right_branch.1.", (1, 3), 1, dilation, act_type, post_add=left)
    return B.conv_bn_act(both, sd, prefix + "last_conv.", 1, 1, 1, act_type, post_add=x)


def dabnet_lowres_logits(sd, x, act_type="prelu"):
    """DABNet.forward up to layer15 (reference models/dabnet.py:30-58)."""
    x_d2 = B.q(F.avg_pool2d(x, 3, 2, 1))
    x_d4 = B.q(F.avg_pool2d(x_d2, 3, 2, 1))
    x_d8 = B.q(F.avg_pool2d(x_d4, 3, 2, 1))
    y = B.conv_bn_act(x, sd, "layer1.", 3, 2, 1, act_type)
    y = B.conv_bn_act(y, sd, "layer2.", 3, 1, 1, act_type)
    y = B.conv_bn_act(y, sd, "layer3.", 3, 1, 1, act_type)
    y = B.initial_block(torch.cat([y, x_d2], dim=1), sd, "layer4.", act_type)
    block1 = y
    for i in range(3):
        y = dab_module(y, sd, f"layer5_7.{i}.", 2, act_type)
    y = B.conv_bn_act(torch.cat([y, block1, x_d4], dim=1), sd, "layer8.", 3, 2, 1, act_type)
    block2 = y
    for name, d in (("layer9_10", 4), ("layer11_12", 8), ("layer13_14", 16)):
        for i in range(2):
            y = dab_module(y, sd, f"{name}.{i}.", d, act_type)
    y = torch.cat([y, block2, x_d8], dim=1)
    return B.q(F.conv2d(y, B.qw(sd["layer15.weight"])))


def dabnet_forward(sd, x, act_type="prelu"):
    return F.interpolate(dabnet_lowres_logits(sd, x, act_type), x.shape[2:], mode="bilinear", align_corners=True)


# ----------------------------------------------------------------------------- LEDNet ----
def channel_shuffle(x, groups=2):
    """reference models/modules.py:18-32: out[:, i*groups + g] = x[:, g*(C/groups) + i]."""
    n, c, h, w = x.shape
    return x.reshape(n, groups, c // groups, h, w).transpose(1, 2).reshape(n, c, h, w)


def ss_nbt_unit(x, sd, prefix, dilation, act_type="relu"):
    """SSnbtUnit (reference models/lednet.py:52-88)."""
    sc = x.shape[1] // 2
    halves = []
    for name, first, second, xs in (("left_branch.", (3, 1), (1, 3), x[:, :sc]), ("right_branch.", (1, 3), (3, 1), x[:, sc:])):
        p = prefix + name
        t = F.conv2d(xs, B.qw(sd[p + "0.weight"]), sd[p + "0.bias"], 1, B._pad(first, 1), 1)
        t = B.q(B.activation(t, act_type, sd, p + "1."))
        t = B.conv_bn_act(t, sd, p + "2.", second, 1, 1, act_type)
        t = F.conv2d(t, B.qw(sd[p + "3.weight"]), sd[p + "3.bias"], 1, B._pad(first, dilation), dilation)
        t = B.q(B.activation(t, act_type, sd, p + "4."))
        halves.append(B.conv_bn_act(t, sd, p + "5.", second, 1, dilation, act_type, post_add=xs, post_act=act_type))
    return channel_shuffle(torch.cat(halves, dim=1))


def attention_pyramid(x, sd, prefix, act_type="relu"):
    """AttentionPyramidNetwork (reference models/lednet.py:91-136)."""
    up = lambda t, size: B.q(F.interpolate(t, size, mode="bilinear", align_corners=True))
    size0 = x.shape[2:]
    l1 = B.conv_bn_act(x, sd, prefix + "left_conv1_1.", 3, 2, 1, act_type)
    l2 = B.conv_bn_act(l1, sd, prefix + "left_conv2_1.", 3, 2, 1, act_type)
    l3 = B.conv_bn_act(l2, sd, prefix + "left_conv3.0.", 3, 2, 1, act_type)
    l3 = up(B.conv_bn_act(l3, sd, prefix + "left_conv3.1.", 3, 1, 1, act_type), l2.shape[2:])
    l2 = up(B.conv_bn_act(l2, sd, prefix + "left_conv2_2.", 3, 1, 1, act_type, post_add=l3), l1.shape[2:])
    l1 = up(B.conv_bn_act(l1, sd, prefix + "left_conv1_2.", 3, 1, 1, act_type, post_add=l2), size0)
    mid = B.conv_bn_act(x, sd, prefix + "mid_branch.", 3, 1, 1, act_type)
    right = B.conv_bn_act(B.q(F.adaptive_avg_pool2d(x, 1)), sd, prefix + "right_branch.1.", 3, 1, 1, act_type)
    return B.q(mid * l1 + up(right, size0))   # one rounding: the CUDA path evaluates mul + add in a single pass


def lednet_lowres_logits(sd, x, act_type="relu"):
    """LEDNet encoder + APN (reference models/lednet.py:16-50)."""
    idx = 0
    for cout, dils in ((32, [1, 1, 1]), (64, [1, 1]), (128, [1, 2, 5, 9, 2, 5, 9, 17])):
        x = B.initial_block(x, sd, f"encoder.{idx}.", act_type)
        idx += 1
        for d in dils:
            x = ss_nbt_unit(x, sd, f"encoder.{idx}.", d, act_type)
            idx += 1
    return attention_pyramid(x, sd, "apn.", act_type)


def lednet_forward(sd, x, act_type="relu"):
    return F.interpolate(lednet_lowres_logits(sd, x, act_type), x.shape[2:], mode="bilinear", align_corners=True)


# ------------------------------------------------------------------------------- ENet ----
def enet_bottleneck(x, sd, prefix, conv_type, dilation=1, indices=None, act_type="prelu"):
    """Bottleneck (reference models/enet.py:119-181), eval mode (Dropout = identity).  Inner ConvBNActs are ReLU
    (their default act_type); the block's own activation is `act_type` applied after the add."""
    p = prefix + "right_init_conv."
    if conv_type == "downsampling":
        r = B.conv_bn_act(x, sd, p + "0.", 3, 2, 1, "relu")
        r = B.conv_bn_act(r, sd, p + "1.", 3, 1, 1, "relu")
    elif conv_type == "upsampling":
        r = B.conv_bn_act(x, sd, p + "0.", 1, 1, 1, "relu")
        r = B.q(F.conv_transpose2d(r, B.qw(sd[p + "1.up_conv.weight"], 1, True), None, 2, 1, 1))
    elif conv_type == "asymmetric":
        r = B.conv_bn_act(x, sd, p + "0.", 1, 1, 1, "relu")
        r = B.conv_bn_act(r, sd, p + "1.", (5, 1), 1, 1, "relu")
        r = B.conv_bn_act(r, sd, p + "2.", (1, 5), 1, 1, "relu")
    else:  # regular / dilate
        r = B.conv_bn_act(x, sd, p + "0.", 1, 1, 1, "relu")
        r = B.conv_bn_act(r, sd, p + "1.", 3, 1, dilation, "relu")
    r = F.conv2d(r, B.qw(sd[prefix + "right_last_conv.0.weight"]))
    if conv_type == "downsampling":
        left, idx = F.max_pool2d(x, 2, 2, return_indices=True)
        left = B.conv_bn_act(B.q(left), sd, prefix + "left_conv.", 1, 1, 1, "relu")
        return B.q(B.activation(left + r, act_type, sd, prefix + "act.")), idx
    if conv_type == "upsampling":
        left = B.conv_bn_act(x, sd, prefix + "left_conv.", 1, 1, 1, "relu")
        left = F.max_unpool2d(left, indices, 2, 2)
        return B.q(B.activation(left + r, act_type, sd, prefix + "act."))
    return B.q(B.activation(x + r, act_type, sd, prefix + "act."))


_ENET_STAGE23 = [("regular", 1), ("dilate", 2), ("asymmetric", 1), ("dilate", 4), ("regular", 1), ("dilate", 8),
                 ("asymmetric", 1), ("dilate", 16)]   # reference models/enet.py:76-85


def enet_forward(sd, x, act_type="prelu"):
    """ENet.forward (reference models/enet.py:25-35), upsample_type='deconvolution' (the default)."""
    x = B.initial_block(x, sd, "initial.", act_type)
    x, idx1 = enet_bottleneck(x, sd, "bottleneck1.conv_pool.", "downsampling", act_type=act_type)
    for i in range(4):
        x = enet_bottleneck(x, sd, f"bottleneck1.conv_regular.{i}.", "regular", act_type=act_type)
    x, idx2 = enet_bottleneck(x, sd, "bottleneck2.conv_pool.", "downsampling", act_type=act_type)
    for stage in ("bottleneck2", "bottleneck3"):
        for i, (kind, d) in enumerate(_ENET_STAGE23):
            x = enet_bottleneck(x, sd, f"{stage}.conv_regular.{i}.", kind, d, act_type=act_type)
    x = enet_bottleneck(x, sd, "bottleneck4.conv_unpool.", "upsampling", indices=idx2, act_type=act_type)
    x = enet_bottleneck(x, sd, "bottleneck4.conv_regular.", "regular", act_type=act_type)
    x = enet_bottleneck(x, sd, "bottleneck4.conv_extra.", "regular", act_type=act_type)
    x = enet_bottleneck(x, sd, "bottleneck5.conv_unpool.", "upsampling", indices=idx1, act_type=act_type)
    x = enet_bottleneck(x, sd, "bottleneck5.conv_regular.", "regular", act_type=act_type)
    # fullconv = Upsample(16, num_class, scale_factor=2, act_type) WITHOUT upsample_type (models/enet.py:24), i.e. the
    # ConvBNAct 1x1 + nn.Upsample(scale_factor=2, mode='bilinear') branch (:197-201), align_corners=False
    x = B.conv_bn_act(x, sd, "fullconv.up_conv.0.", 1, 1, 1, act_type)
    return B.q(F.interpolate(x, scale_factor=2, mode="bilinear"))


from .ref_nets_dual import bisenetv2_forward, ddrnet_forward  # noqa: E402  (the two-branch nets of config 4)

FORWARD = {"erfnet": erfnet_forward, "fastscnn": fastscnn_forward, "dabnet": dabnet_forward, "lednet": lednet_forward,
           "enet": enet_forward, "bisenetv2": bisenetv2_forward, "ddrnet": ddrnet_forward}


# --------------------------------------------------------------------- seeded weights ----
def make_state(spec, seed: int, perturb_bn: bool = True):
    """Deterministic state dict for `spec` = [(key, shape, dtype_str), ...] (reference key order).

    perturb_bn=False reproduces the information content of a default init (BatchNorm = identity,
    SURVEY F6); perturb_bn=True draws running statistics / affine terms so that a wrong BN fold,
    a dropped bias or a wrong PReLU slope cannot hide.  CPU generator => identical on every host.
    """
    keys = {k for k, _, _ in spec}
    sd = {}
    for idx, (key, shape, dtype) in enumerate(spec):
        g = torch.Generator().manual_seed(seed * 1000003 + idx)
        leaf = key.rsplit(".", 1)[-1]
        stem = key[: -len(leaf)]
        is_bn = (stem + "running_mean") in keys
        if leaf == "num_batches_tracked":
            t = torch.zeros(shape, dtype=torch.int64)
        elif len(shape) == 4:  # conv / deconv weight, He-style so activations stay O(1)
            fan_in = shape[1] * shape[2] * shape[3]
            t = torch.randn(shape, generator=g) * (2.0 / max(fan_in, 1)) ** 0.5
        elif is_bn:
            if not perturb_bn:
                t = {"weight": torch.ones, "bias": torch.zeros, "running_mean": torch.zeros,
                     "running_var": torch.ones}[leaf](shape)
            elif leaf == "weight":
                t = 1.0 + 0.2 * torch.randn(shape, generator=g)
            elif leaf == "running_var":
                t = 0.5 + torch.rand(shape, generator=g)
            else:  # bias, running_mean
                t = 0.1 * torch.randn(shape, generator=g)
        elif "activation" in key:  # PReLU slope
            t = 0.25 + (0.05 * torch.randn(shape, generator=g) if perturb_bn else torch.zeros(shape))
        else:  # conv bias
            t = 0.1 * torch.randn(shape, generator=g)
        sd[key] = t
    return sd


def make_input(seed: int, n: int, h: int, w: int, c: int = 3):
    g = torch.Generator().manual_seed(seed)
    return torch.randn((n, c, h, w), generator=g)
