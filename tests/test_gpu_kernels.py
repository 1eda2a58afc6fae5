"""GPU parity tests, kernel by kernel: librtseg (through the C ABI) against the CPU oracle.

Tolerances are BASELINE.json's: fp32 <= 1e-4, bf16 <= 2e-2 (max-abs error / max-abs reference);
integer outputs (labels away from ties, confusion matrices) are bit-exact.
"""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import ref_blocks as B
from oracle import ref_head, ref_nets, ref_ohem
from tests.helpers import GOLDEN, TOL_BF16, TOL_FP32, rel_err, seeded

pytestmark = pytest.mark.gpu

DTYPES = [torch.float32, torch.bfloat16]


def _tol(dtype):
    return TOL_FP32 if dtype == torch.float32 else TOL_BF16


def _randomise(module, seed=0):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, v in module.state_dict().items():
            if k.endswith("running_var"):
                v.copy_(0.5 + torch.rand(v.shape, generator=g))
            elif k.endswith("num_batches_tracked"):
                continue
            elif v.dim() == 4:
                fan = v.shape[1] * v.shape[2] * v.shape[3]
                v.copy_(torch.randn(v.shape, generator=g) * (1.0 / max(fan, 1)) ** 0.5)
            else:
                v.copy_(torch.randn(v.shape, generator=g) * 0.3 + (1.0 if k.endswith("1.weight") else 0.0))
    return module.eval()


def _sd(module, prefix="m."):
    return {prefix + k: v.detach().clone().float().cpu() for k, v in module.state_dict().items()}


def _run(module, x_cpu, dtype, **kw):
    import rtseg_b200.ops as ops
    m = module.to("cuda")  # fp32 master parameters; the activation dtype follows the input
    with torch.no_grad():
        y = m(x_cpu.to("cuda", dtype), **kw)
    assert ops.tc_timeout_flag() == 0, "a tensor-core kernel hit its barrier watchdog"
    return y.float().cpu()


# ------------------------------------------------------------------------------ layout ----
@pytest.mark.parametrize("shape", [(2, 3, 17, 33), (1, 19, 8, 40), (2, 64, 5, 7), (1, 1, 4, 4)])
def test_layout_roundtrip(shape):
    from rtseg_b200 import ops
    x = seeded(shape, 0).cuda()
    y = ops.to_nhwc(x)
    assert y.is_contiguous(memory_format=torch.channels_last) and torch.equal(y, x)
    assert torch.equal(ops.to_nchw(y), x) and ops.to_nchw(y).is_contiguous()
    yb = ops.to_nhwc(x, torch.bfloat16)
    assert torch.equal(yb, x.bfloat16())
    assert torch.equal(ops.to_nchw(yb, torch.float32), x.bfloat16().float())


# --------------------------------------------------------------------------- depthwise ----
DW_CASES = [
    # cin, cout, kernel, stride, dil, h, w, act
    (32, 32, 3, 2, 1, 36, 70, "relu"),        # Fast-SCNN DS1.dw
    (48, 48, 3, 2, 1, 17, 33, "relu"),
    (128, 128, 3, 1, 1, 20, 96, "relu"),      # PX=8 window kernel
    (384, 384, 3, 1, 1, 9, 13, "relu6"),
    (64, 64, (3, 1), 1, 4, 24, 40, "prelu"),  # DABNet branches
    (64, 64, (1, 3), 1, 16, 24, 40, "prelu"),
    (32, 32, (1, 3), 1, 1, 11, 75, "relu"),
    (32, 32, (3, 1), 2, 1, 20, 30, "relu6"),  # ADSCNet: (3,1) with stride 2 halves both dims
    (19, 19, 3, 1, 1, 10, 12, "none"),        # odd channels -> scalar kernel
    (16, 96, 3, 2, 1, 12, 20, "relu"),        # channel multiplier 6 (BiSeNetv2): vector kernel, two input channels per vector
    (32, 192, 3, 1, 1, 13, 21, "none"),
    (128, 768, 3, 1, 1, 8, 9, "prelu"),
    (64, 256, 3, 1, 1, 10, 12, "relu"),       # multiplier 4
    (12, 24, 3, 2, 1, 10, 12, "relu"),        # multiplier 2: a vector would span four input channels -> scalar kernel
    (8, 40, (1, 3), 1, 2, 9, 17, "relu6"),    # multiplier 5, dilated 1x3
    (24, 24, 5, 1, 1, 14, 18, "hardswish"),
]


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("case", DW_CASES)
def test_depthwise_conv(case, dtype):
    from rtseg_b200 import modules
    cin, cout, k, s, d, h, w, act = case
    m = _randomise(modules.DWConvBNAct(cin, cout, k, s, d, act_type=act), 1)
    x = seeded((2, cin, h, w), 2)
    with torch.no_grad():
        ref = B.conv_bn_act(x, _sd(m), "m.", k, s, d, act, groups=cin)
    out = _run(m, x, dtype)
    assert out.shape == ref.shape
    assert rel_err(out, ref) < _tol(dtype)


# ------------------------------------------------------------------ dense, CUDA cores ----
DENSE_CASES = [
    # cin, cout, kernel, stride, dil, bias, act, h, w
    (16, 48, 3, 2, 1, False, "relu", 20, 36),
    (64, 64, (3, 1), 1, 2, False, "relu", 16, 24),
    (64, 64, (1, 3), 1, 3, False, "prelu", 16, 24),
    (35, 29, 3, 2, 1, False, "prelu", 15, 21),     # DABNet layer4 (odd channels)
    (131, 128, 1, 1, 1, True, "relu", 9, 11),
    (259, 19, 1, 1, 1, False, "none", 8, 8),       # DABNet classifier
    (16, 16, (5, 1), 1, 1, False, "relu", 12, 10),  # ENet asymmetric
    (3, 32, 3, 1, 1, False, "leakyrelu", 10, 14),
    (128, 128, 3, 1, 1, False, "elu", 6, 10),
]


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("case", DENSE_CASES)
def test_dense_conv_cuda_cores(case, dtype, monkeypatch):
    from rtseg_b200 import modules, ops
    monkeypatch.setattr(ops, "USE_TENSOR_CORES", False)
    cin, cout, k, s, d, bias, act, h, w = case
    m = _randomise(modules.ConvBNAct(cin, cout, k, s, d, bias=bias, act_type=act), 3)
    x = seeded((2, cin, h, w), 4)
    with torch.no_grad():
        ref = B.conv_bn_act(x, _sd(m), "m.", k, s, d, act)
    out = _run(m, x, dtype)
    assert out.shape == ref.shape and rel_err(out, ref) < _tol(dtype)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("mode", ["pre", "post"])
def test_dense_conv_residual_modes(mode, dtype, monkeypatch):
    from rtseg_b200 import modules, ops
    monkeypatch.setattr(ops, "USE_TENSOR_CORES", False)
    m = _randomise(modules.ConvBNAct(24, 24, 1, act_type="relu"), 5)
    x, r = seeded((1, 24, 9, 13), 6), seeded((1, 24, 9, 13), 7)
    sd = _sd(m)
    with torch.no_grad():
        conv = F.conv2d(x, sd["m.0.weight"])
        if mode == "pre":
            ref = F.relu(B.batch_norm(conv + r, sd, "m.1."))
        else:
            ref = F.relu(B.batch_norm(conv, sd, "m.1.")) + r
    res = ops.to_nhwc(r.to("cuda", dtype))
    out = _run(m, x, dtype, res=res, res_mode=ops.RES_PRE if mode == "pre" else ops.RES_POST)
    assert rel_err(out, ref) < _tol(dtype)


# ------------------------------------------------------------------ dense, tensor cores ----
TC_CASES = [
    # cin, cout, kernel, stride, dil, bias, act, n, h, w
    (64, 64, (3, 1), 1, 1, False, "relu", 2, 16, 128),     # ERFNet layer3_7
    (64, 64, (1, 3), 1, 1, False, "relu", 2, 16, 128),
    (128, 128, (3, 1), 1, 16, False, "relu", 1, 40, 128),  # ERFNet layer9_16, dilation 16
    (128, 128, (1, 3), 1, 8, False, "relu", 1, 8, 256),
    (16, 16, (3, 1), 1, 1, False, "relu", 1, 12, 256),     # ERFNet layer21_22 (32-byte swizzle)
    (16, 16, (1, 3), 1, 1, False, "relu", 1, 12, 256),
    (32, 48, 1, 1, 1, True, "relu", 1, 16, 64),            # Fast-SCNN pointwise (64-byte swizzle)
    (48, 64, 1, 1, 1, True, "relu", 1, 16, 64),
    (64, 384, 1, 1, 1, True, "relu", 1, 16, 32),           # two N tiles of 192
    (384, 64, 1, 1, 1, False, "none", 1, 16, 32),
    (768, 128, 1, 1, 1, False, "none", 1, 8, 16),          # K = 768, N split for shared memory
    (128, 19, 1, 1, 1, True, "relu", 1, 16, 32),           # classifier head, cout padded to 32
    (64, 64, 3, 1, 1, False, "prelu", 1, 10, 24),          # partial tiles in w and h
    (16, 48, 3, 2, 1, False, "relu", 2, 16, 64),           # stride-2 downsampler via parity views
    (64, 64, 3, 2, 1, False, "relu", 1, 12, 40),
    (128, 128, (3, 1), 1, 2, False, "relu", 3, 9, 37),     # ragged everything
]


@pytest.mark.parametrize("case", TC_CASES)
def test_dense_conv_tensor_cores(case):
    from rtseg_b200 import _lib, modules, ops
    cin, cout, k, s, d, bias, act, n, h, w = case
    m = _randomise(modules.ConvBNAct(cin, cout, k, s, d, bias=bias, act_type=act), 8)
    x = seeded((n, cin, h, w), 9)
    with torch.no_grad():
        ref = B.conv_bn_act(x, _sd(m), "m.", k, s, d, act)
    before = _lib.launch_count()
    out = _run(m, x, torch.bfloat16)
    assert _lib.launch_count() - before >= 1
    assert out.shape == ref.shape
    err = rel_err(out, ref)
    assert err < TOL_BF16, f"tensor-core conv off by {err:.3e}"
    # and it must agree with the CUDA-core kernel on the same bf16 inputs much more tightly
    ops.USE_TENSOR_CORES = False
    try:
        simt = _run(m, x, torch.bfloat16)
    finally:
        ops.USE_TENSOR_CORES = True
    assert rel_err(out, simt) < 1e-2


@pytest.mark.parametrize("mode", ["pre", "post"])
def test_tensor_core_residual_and_slices(mode):
    """ERFNet's fused tail (add -> BN -> act) and writing into a channel slice of a wider buffer."""
    from rtseg_b200 import modules, ops
    m = _randomise(modules.ConvBNAct(64, 64, (1, 3), dilation=2, act_type="relu"), 10).to("cuda", torch.bfloat16)
    x, r = seeded((2, 64, 8, 128), 11), seeded((2, 64, 8, 128), 12)
    sd = _sd(m)
    with torch.no_grad():
        conv = F.conv2d(x, sd["m.0.weight"], None, 1, (0, 2), 2)
        ref = F.relu(B.batch_norm(conv + r, sd, "m.1.")) if mode == "pre" else F.relu(B.batch_norm(conv, sd, "m.1.")) + r
    xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
    rb = ops.to_nhwc(r.to("cuda", torch.bfloat16))
    wide = torch.zeros((2, 128, 8, 128), dtype=torch.bfloat16, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        m(xb, res=rb, res_mode=ops.RES_PRE if mode == "pre" else ops.RES_POST, out=wide, y_coff=64)
    assert ops.tc_timeout_flag() == 0
    assert rel_err(wide[:, 64:], ref) < TOL_BF16
    assert (wide[:, :64] == 0).all(), "the kernel wrote outside its channel slice"


# ---------------------------------------------------------------------------------- stem ----
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("nhwc_input", [False, True])
def test_initial_block_stem(dtype, nhwc_input):
    from rtseg_b200 import nets
    m = _randomise(nets.InitialBlock(3, 16, "relu"), 13)
    x = seeded((2, 3, 37, 50), 14)
    with torch.no_grad():
        ref = B.initial_block(x, _sd(m), "m.", "relu")
    xin = x.cuda()  # the caller's fp32 NCHW tensor, read directly by the kernel
    if nhwc_input:
        xin = xin.contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        out = m.to("cuda", dtype)(xin)
    assert out.dtype == dtype and out.shape == ref.shape
    assert rel_err(out, ref) < _tol(dtype)
    # pool branch is exact in fp32
    if dtype == torch.float32:
        assert torch.equal(out[:, 13:].cpu(), ref[:, 13:])


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("cin,cout", [(16, 64), (64, 128)])
def test_initial_block_deep(cin, cout, dtype):
    from rtseg_b200 import nets
    m = _randomise(nets.InitialBlock(cin, cout, "relu"), 15)
    x = seeded((1, cin, 16, 64), 16)
    with torch.no_grad():
        ref = B.initial_block(x, _sd(m), "m.", "relu")
    out = _run(m, x, dtype)
    assert rel_err(out, ref) < _tol(dtype)


def test_fastscnn_stem_without_pool():
    from rtseg_b200 import nets
    m = _randomise(nets.LearningToDownsample(3, 64), 17)
    x = seeded((1, 3, 64, 96), 18)
    sd = _sd(m)
    with torch.no_grad():
        ref = B.conv_bn_act(x, sd, "m.0.", 3, 2, 1, "relu")
        ref = B.ds_conv_bn_act(ref, sd, "m.1.", 3, 2, 1, "relu")
        ref = B.ds_conv_bn_act(ref, sd, "m.2.", 3, 2, 1, "relu")
    for dtype in DTYPES:
        with torch.no_grad():
            out = m.to("cuda", dtype)(x.cuda())
        assert rel_err(out, ref) < _tol(dtype)


# -------------------------------------------------------------------------------- deconv ----
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("cin,cout,h,w", [(128, 64, 8, 16), (64, 16, 6, 128), (16, 19, 9, 130)])
def test_deconv(cin, cout, h, w, dtype):
    from rtseg_b200 import modules, ops
    m = _randomise(modules.DeConvBNAct(cin, cout, act_type="relu"), 19)
    x = seeded((2, cin, h, w), 20)
    with torch.no_grad():
        ref = B.deconv_bn_act(x, _sd(m), "m.", 2, "relu")
    out = _run(m, x, dtype)
    assert out.shape == ref.shape == (2, cout, 2 * h, 2 * w)
    assert rel_err(out, ref) < _tol(dtype)
    # fused argmax: identical to the argmax of the kernel's own logits
    labels = torch.empty((2, 2 * h, 2 * w), dtype=torch.uint8, device="cuda")
    with torch.no_grad():
        m(x.to("cuda", dtype), labels=labels, want_logits=False)
    assert ops.tc_timeout_flag() == 0
    own = out.to(dtype).float().max(dim=1)[1]
    assert (labels.cpu().long() == own).float().mean().item() == 1.0


@pytest.mark.parametrize("act", ["relu", "relu6", "none", "hardtanh"])
def test_deconv_labels_head_matches_torch_max_on_ties_nan_and_inf(act):
    """The fused head compares bf16-rounded logits; ties (ReLU zeros, ReLU6 saturation) take the first class and a
    NaN logit wins like torch.max — checked against the argmax of the kernel's own materialised logits."""
    from rtseg_b200 import modules, ops
    m = _randomise(modules.DeConvBNAct(16, 19, act_type=act), 61)
    with torch.no_grad():
        m.up_conv[0].weight.mul_(6.0)             # push many logits past 6 so that relu6 / hardtanh saturate
    x = seeded((1, 16, 12, 140), 62)
    x[0, :, 3, 5] = float("nan")
    x[0, 2, 7, 9] = float("inf")
    xb = x.to("cuda", torch.bfloat16)
    labels = torch.empty((1, 24, 280), dtype=torch.uint8, device="cuda")
    with torch.no_grad():
        logits = m.cuda()(xb)
        m(xb, labels=labels, want_logits=False)
    assert ops.tc_timeout_flag() == 0
    own = logits.float().max(dim=1)[1]
    assert torch.isnan(logits).any()
    assert torch.equal(labels.long(), own)


# -------------------------------------------------------------------------------- resize ----
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("align", [True, False])
@pytest.mark.parametrize("c,hi,wi,ho,wo", [(128, 8, 16, 32, 64), (19, 16, 32, 128, 256), (32, 1, 1, 8, 16),
                                           (24, 7, 9, 14, 18)])
def test_bilinear_resize(c, hi, wi, ho, wo, align, dtype):
    from rtseg_b200 import ops
    x = seeded((2, c, hi, wi), 21)
    ref = F.interpolate(x, (ho, wo), mode="bilinear", align_corners=align)
    out = ops.upsample_bilinear(x.to("cuda", dtype), (ho, wo), align)
    assert rel_err(out, ref) < (1e-5 if dtype == torch.float32 else TOL_BF16)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("align,c,hi,wi,ho,wo", [(True, 19, 16, 32, 128, 256), (False, 19, 24, 40, 48, 80),
                                                 (True, 7, 5, 9, 37, 61), (True, 32, 8, 8, 64, 64)])
def test_fused_upsample_argmax_and_confmat(align, c, hi, wi, ho, wo, dtype):
    from rtseg_b200 import ops
    x = seeded((2, c, hi, wi), 22).to(dtype)
    xf = x.float()
    up = F.interpolate(xf, (ho, wo), mode="bilinear", align_corners=align)
    ref = up.max(dim=1)[1]
    rng = np.random.default_rng(5)
    target = torch.from_numpy(rng.integers(0, c, size=(2, ho, wo))).long()
    target[torch.from_numpy(rng.random((2, ho, wo)) < 0.1)] = 255
    cm = torch.zeros((c, c), dtype=torch.int64, device="cuda")
    labels = ops.upsample_argmax(x.cuda(), (ho, wo), align, target=target.cuda(), confmat=cm, ignore_index=255)
    assert labels.dtype == torch.uint8 and labels.shape == (2, ho, wo)
    got = labels.cpu().long()
    # pixels whose top-2 margin is above fp32 interpolation noise must agree exactly
    top2 = up.topk(2, dim=1)[0]
    clear = (top2[:, 0] - top2[:, 1]) > 1e-4 * up.abs().max()
    assert (got[clear] == ref[clear]).all()
    assert (got == ref).float().mean().item() >= 0.999
    # confusion matrix: bit-exact for the labels the kernel itself produced
    want = ref_head.confusion_matrix(got.numpy(), target.numpy(), c, 255)
    assert np.array_equal(cm.cpu().numpy(), want)
    # int64 label variant, no confusion matrix
    l64 = ops.upsample_argmax(x.cuda(), (ho, wo), align, out_dtype=torch.int64)
    assert torch.equal(l64.cpu(), got)


def test_argmax_semantics_first_max_and_nan():
    from rtseg_b200 import ops
    x = torch.zeros(1, 5, 2, 3)
    x[0, 3, 0, 0] = 1.0
    x[0, 2, 0, 1] = 1.0
    x[0, 4, 0, 1] = 1.0            # tie -> first index (2)
    x[0, 1, 1, 1] = float("nan")   # NaN wins (torch.max semantics)
    ref = x.max(dim=1)[1]
    for t in (x.cuda(), x.cuda().contiguous(memory_format=torch.channels_last)):
        assert torch.equal(ops.argmax_channels(t).cpu(), ref)
        assert torch.equal(ops.argmax_channels(t, torch.uint8).cpu().long(), ref)
    assert ref[0, 0, 2] == 0 and ref[0, 0, 1] == 2 and ref[0, 1, 1] == 1


# ---------------------------------------------------------------------- confusion matrix ----
@pytest.mark.parametrize("pred_dtype", [torch.uint8, torch.int64])
@pytest.mark.parametrize("npix", [0, 1, 1000, 300001])
def test_confusion_matrix_bit_exact(npix, pred_dtype):
    from rtseg_b200 import ops
    rng = np.random.default_rng(npix)
    pred = rng.integers(0, 19, size=npix)
    target = rng.integers(0, 19, size=npix)
    target[rng.random(npix) < 0.1] = 255
    cm = torch.zeros((19, 19), dtype=torch.int64, device="cuda")
    for _ in range(2):  # accumulates across updates like the metric object
        ops.confmat_update(cm, torch.from_numpy(pred).to(pred_dtype).cuda(), torch.from_numpy(target).cuda(), 255)
    want = 2 * ref_head.confusion_matrix(pred, target, 19, 255)
    assert cm.dtype == torch.int64 and np.array_equal(cm.cpu().numpy(), want)


def test_confusion_matrix_edge_cases():
    from rtseg_b200 import metrics, ops
    cm = torch.zeros((19, 19), dtype=torch.int64, device="cuda")
    pred = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    ops.confmat_update(cm, pred, torch.full((4096,), 255, device="cuda"), 255)   # everything ignored
    assert cm.sum().item() == 0
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    ops.confmat_update(cm, pred, torch.full((4096,), 77, device="cuda"), 255, bad_count=bad)  # out-of-range targets
    assert cm.sum().item() == 0 and bad.item() == 4096
    ops.confmat_update(cm, pred, torch.full((4096,), 3, device="cuda"), 255)     # one hot bin, 4096 collisions
    assert cm[3, 0].item() == 4096 and cm.sum().item() == 4096
    m = metrics.SegMetrics(19, 255).to("cuda")
    logits = seeded((2, 19, 8, 8), 30).cuda()
    target = torch.randint(0, 19, (2, 8, 8), generator=torch.Generator().manual_seed(1)).cuda()
    m.update(logits, target)
    want = ref_head.confusion_matrix(logits.cpu().max(1)[1].numpy(), target.cpu().numpy(), 19)
    assert np.array_equal(m.confmat.cpu().numpy(), want)
    assert np.array_equal(m.compute().cpu().numpy(), ref_head.iou_from_confmat(want))
    m.reset()
    assert m.confmat.sum().item() == 0


# ------------------------------------------------------------------------------------ OHEM ----
def _ohem_case(case, margin, ignore_frac, shape=(2, 19, 24, 40)):
    g = torch.Generator().manual_seed(100 + case)
    n, c, h, w = shape
    logits = torch.randn(shape, generator=g)
    labels = torch.randint(0, c, (n, h, w), generator=g)
    if margin:
        logits.scatter_add_(1, labels.unsqueeze(1), torch.full((n, 1, h, w), margin))
    labels[torch.rand((n, h, w), generator=g) < ignore_frac] = 255
    return logits, labels


def test_ohem_matches_reference_golden_cases():
    import rtseg_b200
    with open(os.path.join(GOLDEN, "ohem_cases.json")) as f:
        cases = json.load(f)
    crit = rtseg_b200.OhemCELoss(0.7, 255)
    for c in cases:
        logits, labels = _ohem_case(c["case"], c["margin"], c["ignore_frac"])
        lg = logits.cuda().requires_grad_(True)
        loss = crit(lg, labels.cuda())
        loss.backward()
        assert abs(loss.item() - c["loss"]) <= 2e-6 * max(1.0, abs(c["loss"])), c
        assert abs(lg.grad.abs().sum().item() - c["grad_abs_sum"]) <= 1e-4 * c["grad_abs_sum"], c


@pytest.mark.parametrize("layout", ["nchw", "nhwc"])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("margin,ignore_frac", [(0.0, 0.1), (12.0, 0.1), (12.0, 0.0)])
def test_ohem_forward_backward_vs_oracle(margin, ignore_frac, dtype, layout):
    import rtseg_b200
    logits, labels = _ohem_case(7, margin, ignore_frac, (2, 19, 33, 47))
    logits = logits.to(dtype)
    ref_in = logits.float().clone().requires_grad_(True)
    ref_loss, branch = ref_ohem.ohem_ce_loss(ref_in, labels, 0.7, 255)
    ref_loss.backward()
    lg = logits.detach().cuda()
    if layout == "nhwc":
        lg = lg.contiguous(memory_format=torch.channels_last)
    lg = lg.detach().requires_grad_(True)
    loss = rtseg_b200.OhemCELoss(0.7, 255)(lg, labels.cuda())
    (2.0 * loss).backward()  # upstream gradient != 1
    assert abs(loss.item() - ref_loss.item()) <= 1e-5 * max(1.0, abs(ref_loss.item()))
    assert rel_err(lg.grad, 2.0 * ref_in.grad) < (1e-4 if dtype == torch.float32 else 1e-2)
    assert (branch == 1) == (margin > 0)


def test_ohem_all_ignored_is_nan_and_workspace_is_reset():
    import rtseg_b200
    crit = rtseg_b200.OhemCELoss(0.7, 255)
    logits = seeded((1, 19, 8, 8), 40).cuda()
    assert torch.isnan(crit(logits, torch.full((1, 8, 8), 255, device="cuda")))
    a = crit(logits, torch.zeros((1, 8, 8), dtype=torch.long, device="cuda")).item()
    b = crit(logits, torch.zeros((1, 8, 8), dtype=torch.long, device="cuda")).item()
    assert a == b and np.isfinite(a)


# ------------------------------------------------------------------------------ pooling ----
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("p", [1, 2, 4, 6])
def test_adaptive_avgpool(p, dtype):
    from rtseg_b200 import ops
    x = seeded((2, 40, 13, 21), 41)
    ref = F.adaptive_avg_pool2d(x, p)
    out = ops.adaptive_avgpool(ops.to_nhwc(x.to("cuda", dtype)), p)
    assert out.shape == ref.shape and rel_err(out, ref) < (1e-5 if dtype == torch.float32 else TOL_BF16)


@pytest.mark.parametrize("dtype", DTYPES)
def test_pyramid_pooling_module(dtype):
    from rtseg_b200 import modules
    m = _randomise(modules.PyramidPoolingModule(128, 128, "relu", bias=True), 42)
    x = seeded((1, 128, 8, 16), 43)
    with torch.no_grad():
        ref = B.pyramid_pooling(x, _sd(m), "m.", "relu")
    assert rel_err(_run(m, x, dtype), ref) < _tol(dtype)


def test_c_abi_host_buffer_entry_point():
    """rtseg_upsample_argmax_host: HOST pointers in and out, copies inside the call."""
    import ctypes
    from rtseg_b200 import _lib
    lib = _lib.load()
    _lib.sync_device(0)
    x = seeded((1, 16, 32, 19), 44).contiguous()  # NHWC fp32 on the host
    labels = torch.empty((1, 128, 256), dtype=torch.uint8)
    scratch = torch.empty(x.numel() * 4 + 256 + labels.numel(), dtype=torch.uint8, device="cuda")
    rc = lib.rtseg_upsample_argmax_host(1, 16, 32, 19, 128, 256, 1, 0, x.data_ptr(), labels.data_ptr(),
                                        scratch.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _lib.last_error()
    ref = ref_head.upsample_argmax(x.permute(0, 3, 1, 2), (128, 256), True)
    assert (labels.long() == ref).float().mean().item() >= 0.999
    assert ctypes.sizeof(ctypes.c_void_p) == 8


def test_c_abi_host_buffer_scratch_is_exactly_the_documented_size():
    """A caller that sizes the scratch as include/rtseg.h says (logits bytes + n*ho*wo) must not be written past:
    logits size not a multiple of 256 bytes, guard bytes behind the scratch stay untouched."""
    from rtseg_b200 import _lib
    lib = _lib.load()
    _lib.sync_device(0)
    n, hi, wi, c, ho, wo = 1, 7, 9, 19, 28, 36            # 7*9*19*4 = 4788 bytes, 4788 % 256 = 180
    need = lib.rtseg_upsample_argmax_host_scratch_bytes(n, hi, wi, c, ho, wo, 0)
    assert need == n * hi * wi * c * 4 + n * ho * wo and (n * hi * wi * c * 4) % 256 != 0
    x = seeded((n, hi, wi, c), 46).contiguous()
    labels = torch.empty((n, ho, wo), dtype=torch.uint8)
    buf = torch.full((need + 512,), 0xA5, dtype=torch.uint8, device="cuda")
    rc = lib.rtseg_upsample_argmax_host(n, hi, wi, c, ho, wo, 1, 0, x.data_ptr(), labels.data_ptr(), buf.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _lib.last_error()
    assert (buf[need:] == 0xA5).all(), "the call wrote behind the documented scratch size"
    ref = ref_head.upsample_argmax(x.permute(0, 3, 1, 2), (ho, wo), True)
    assert (labels.long() == ref).float().mean().item() >= 0.999


@pytest.mark.parametrize("align", [True, False])
def test_fused_head_with_infinite_logits_matches_torch(align):
    """+inf / -inf logits: torch interpolates them to +-inf (or NaN where inf meets -inf or a zero weight) and max picks
    accordingly; a pixel whose logits are all -inf keeps class 0 and never lands outside [0, c) in the histogram."""
    from rtseg_b200 import ops
    c, hi, wi, ho, wo = 19, 12, 16, 96, 128
    x = seeded((2, c, hi, wi), 47)
    x[0, 3, 2, 3] = float("inf")
    x[0, 5, 2, 4] = float("-inf")
    x[0, 7, 6:8, 6:8] = float("inf")
    x[1, :, 4, 5] = float("-inf")           # every class -inf at one source pixel
    x[1, :, 4, 6] = float("-inf")
    x[1, :, 5, 5:7] = float("-inf")
    x[1, 2, 9, 9] = float("nan")
    up = F.interpolate(x, (ho, wo), mode="bilinear", align_corners=align)
    ref = up.max(dim=1)[1]
    target = torch.randint(0, c, (2, ho, wo), generator=torch.Generator().manual_seed(3))
    cm = torch.zeros((c, c), dtype=torch.int64, device="cuda")
    got = ops.upsample_argmax(x.cuda(), (ho, wo), align, target=target.cuda(), confmat=cm).cpu().long()
    assert int(got.max()) < c
    finite = torch.isfinite(up).all(dim=1)
    top2 = up.topk(2, dim=1)[0]
    clear = ~finite | ((top2[:, 0] - top2[:, 1]) > 1e-4)
    assert (got[clear] == ref[clear]).all(), "non-finite pixels must follow torch.max exactly"
    assert cm.sum().item() == target.numel()
    assert np.array_equal(cm.cpu().numpy(), ref_head.confusion_matrix(got.numpy(), target.numpy(), c, 255))


def test_labels_head_accepts_an_odd_label_address():
    """The 16-bit paired label store needs a 2-byte aligned class map; an odd address takes the general path."""
    from rtseg_b200 import modules, ops
    m = _randomise(modules.DeConvBNAct(16, 19, act_type="relu"), 63).cuda()
    x = seeded((1, 16, 20, 72), 64).to("cuda", torch.bfloat16)
    raw = torch.zeros(40 * 144 + 2, dtype=torch.uint8, device="cuda")
    even, odd = raw[:40 * 144].view(1, 40, 144), raw[1:40 * 144 + 1].view(1, 40, 144)
    with torch.no_grad():
        m(x, labels=even, want_logits=False)
        want = even.clone()
        raw.zero_()
        m(x, labels=odd, want_logits=False)
    assert ops.tc_timeout_flag() == 0
    assert odd.data_ptr() % 2 == 1 and torch.equal(odd, want)


def test_raw_ops_refuse_tensors_on_the_autograd_tape():
    from rtseg_b200 import modules, ops
    x = torch.randn(1, 8, 8, 8, device="cuda", requires_grad=True)
    for fn in (lambda: ops.upsample_bilinear(x, (16, 16), True), lambda: ops.adaptive_avgpool(x, 2),
               lambda: ops.channel_shuffle(x, 2), lambda: ops.avgpool3x3s2(x)):
        with pytest.raises(NotImplementedError, match="no backward"):
            fn()
    ppm = modules.PyramidPoolingModule(8, 8, "relu").cuda()       # training mode: tape nodes with adjoint kernels (row T)
    out = ppm(x)
    out.sum().backward()
    assert x.grad is not None and all(p.grad is not None for p in ppm.parameters())
    with pytest.raises(NotImplementedError, match="PyramidPoolingModule"):
        ppm.eval()(x)                                              # eval-mode BatchNorm with a taped input: refused
    with torch.no_grad():
        assert ppm.eval()(x).shape == (1, 8, 8, 8)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("c,h,w,ct,coff", [(128, 32, 64, 128, 0), (128, 32, 64, 256, 0), (48, 8, 12, 64, 16), (19, 4, 4, 19, 0), (64, 16, 20, 64, 0)])
def test_pyramid_avgpool_matches_four_adaptive_pools(c, h, w, ct, coff, dtype):
    """AdaptiveAvgPool2d(1 / 2 / 4 / 6) in one pass (PyramidPoolingModule, reference models/modules.py:145-149); the
    result is the same from launch to launch (no atomics)."""
    from rtseg_b200 import ops
    x = seeded((2, ct, h, w), 60).to(dtype)
    xg = x.cuda().contiguous(memory_format=torch.channels_last)
    outs = ops.pyramid_avgpool(xg, coff, c)
    again = ops.pyramid_avgpool(xg, coff, c)
    tol = 2e-6 if dtype == torch.float32 else TOL_BF16 / 4
    for p, o, o2 in zip((1, 2, 4, 6), outs, again):
        ref = F.adaptive_avg_pool2d(x[:, coff:coff + c].float(), p)
        assert tuple(o.shape) == (2, c, p, p) and rel_err(o, ref) < tol and torch.equal(o, o2)


@pytest.mark.parametrize("n,c,hid,h,w,ct", [(2, 128, 32, 32, 64, 256), (1, 64, 16, 8, 12, 128), (3, 32, 8, 4, 4, 64)])
def test_ppm_branches_match_conv_and_resize(n, c, hid, h, w, ct):
    """conv1x1 + F.interpolate(bilinear, align_corners=True) of the four pooled maps into the concat buffer in ONE launch
    (PyramidPoolingModule, reference models/modules.py:145-156): against torch on the bf16-rounded operands, and against
    the unfused rtseg path (tensor-core conv1x1 + resize kernel per branch) through the module."""
    from rtseg_b200 import modules, ops
    x = seeded((n, c, h, w), 71).to(torch.bfloat16).cuda().contiguous(memory_format=torch.channels_last)
    pooled = ops.pyramid_avgpool(x, 0, c)
    wts = (seeded((4, hid, c), 72) * 0.2).cuda().contiguous()
    bias = seeded((4, hid), 73).cuda().contiguous()
    cat = torch.full((n, ct, h, w), 7.0, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ops.ppm_branches(pooled, wts, bias, cat, ct - 4 * hid)
    assert torch.all(cat[:, :ct - 4 * hid] == 7.0)                       # channels outside the slice untouched
    for b, p in enumerate((1, 2, 4, 6)):
        z = F.conv2d(pooled[b].float(), wts[b].to(torch.bfloat16).float()[:, :, None, None], bias[b])
        ref = F.interpolate(z.to(torch.bfloat16).float(), (h, w), mode="bilinear", align_corners=True)
        got = cat[:, ct - 4 * hid + b * hid: ct - 4 * hid + (b + 1) * hid].float()
        assert rel_err(got, ref) < TOL_BF16, (b, rel_err(got, ref))
    if c == 128:   # the module, fused against unfused
        m = _randomise(modules.PyramidPoolingModule(c, c, "relu"), 74).eval().cuda()
        with torch.no_grad():
            fused = m(x)
            ops.USE_PPM_FUSION = False
            try:
                plain = m(x)
            finally:
                ops.USE_PPM_FUSION = True
        assert rel_err(fused, plain) < TOL_BF16


# --------------------------------------------------- DABNet / LEDNet building blocks ----
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("shape", [(2, 3, 32, 64), (1, 3, 17, 33), (1, 8, 9, 7)])
def test_avgpool3x3s2_matches_avg_pool2d(shape, dtype):
    """F.avg_pool2d(x, 3, 2, 1): zero padding counted in the divisor (reference models/dabnet.py:32-34)."""
    from rtseg_b200 import ops
    x = seeded(shape, 50)
    ref = F.avg_pool2d(x, 3, 2, 1)
    tol = 1e-6 if dtype == torch.float32 else TOL_BF16
    out = ops.avgpool3x3s2(x.cuda(), out_dtype=dtype)                       # NCHW fp32 image in, NHWC out
    assert out.dtype == dtype and out.shape == ref.shape and ops.is_nhwc(out)
    assert rel_err(out, ref) < tol
    out2 = ops.avgpool3x3s2(ops.to_nhwc(x.to("cuda", dtype)))               # NHWC in, chained like x_d4 / x_d8
    assert rel_err(out2, ref) < tol
    # into a channel slice of a wider buffer
    wide = torch.zeros((shape[0], shape[1] + 5, ref.shape[2], ref.shape[3]), device="cuda", dtype=dtype).contiguous(
        memory_format=torch.channels_last)
    ops.avgpool3x3s2(x.cuda(), out=wide, y_coff=4)
    assert rel_err(wide[:, 4:4 + shape[1]], ref) < tol
    assert wide[:, :4].abs().sum().item() == 0 and wide[:, 4 + shape[1]:].abs().sum().item() == 0


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("c,groups", [(32, 2), (128, 2), (24, 3), (6, 2)])
def test_channel_shuffle_is_the_reference_permutation(c, groups, dtype):
    from rtseg_b200 import ops
    x = seeded((2, c, 5, 9), 51).to(dtype)
    ref = ref_nets.channel_shuffle(x.float(), groups)
    out = ops.channel_shuffle(x.cuda(), groups)
    assert torch.equal(out.float().cpu(), ref)                              # a permutation: bit-exact


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("channels,dilation", [(64, 2), (128, 16)])
def test_dab_module(channels, dilation, dtype):
    from rtseg_b200 import nets
    m = _randomise(nets.DABModule(channels, dilation, "prelu"), 52)
    x = seeded((1, channels, 24, 40), 53)
    with torch.no_grad():
        ref = ref_nets.dab_module(x, _sd(m), "m.", dilation, "prelu")
    assert rel_err(_run(m, x, dtype), ref) < _tol(dtype)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("channels,dilation", [(32, 1), (64, 1), (128, 5), (128, 17)])
def test_ss_nbt_unit(channels, dilation, dtype):
    from rtseg_b200 import nets
    m = _randomise(nets.SSnbtUnit(channels, dilation, "relu"), 54)
    x = seeded((1, channels, 24, 40), 55)
    with torch.no_grad():
        ref = ref_nets.ss_nbt_unit(x, _sd(m), "m.", dilation, "relu")
    assert rel_err(_run(m, x, dtype), ref) < _tol(dtype)


@pytest.mark.parametrize("dtype", DTYPES)
def test_attention_pyramid_network(dtype):
    from rtseg_b200 import nets
    m = _randomise(nets.AttentionPyramidNetwork(128, 19, "relu"), 56)
    x = seeded((1, 128, 16, 32), 57)
    with torch.no_grad():
        ref = ref_nets.attention_pyramid(x, _sd(m), "m.", "relu")
    assert rel_err(_run(m, x, dtype), ref) < _tol(dtype)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("p", [1, 2])
def test_adaptive_avgpool_large_map_two_stage(p, dtype):
    """LEDNet's APN pools a 1/8-resolution map to 1x1 (reference models/lednet.py:104): two exact stages."""
    from rtseg_b200 import ops
    x = seeded((2, 24, 128, 256), 58).to(dtype).float()
    ref = F.adaptive_avg_pool2d(x, p)
    out = ops.adaptive_avgpool(ops.to_nhwc(x.to("cuda", dtype)), p)
    assert out.shape == ref.shape and out.dtype == dtype
    assert (out.float().cpu() - ref).abs().max().item() < (1e-5 if dtype == torch.float32 else 2e-3)


@pytest.mark.parametrize("cin,cout,k,stride", [(35, 29, 3, 2), (131, 128, 3, 2), (259, 19, 1, 1), (40, 24, 3, 1)])
def test_tensor_core_conv_with_ragged_k(cin, cout, k, stride):
    """cin not a multiple of 16 (DABNet's concatenations): K is zero-filled by the TMA unit; channels of the
    buffer beyond cin are poisoned with NaN to prove they are never read."""
    from rtseg_b200 import modules
    m = _randomise(modules.ConvBNAct(cin, cout, k, stride, act_type="prelu"), 59)
    x = seeded((1, cin, 32, 48), 60)
    with torch.no_grad():
        ref = B.conv_bn_act(x, _sd(m), "m.", k, stride, 1, "prelu")
    pitch = (cin + 7) // 8 * 8 + 8
    buf = torch.full((1, pitch, 32, 48), float("nan"), device="cuda", dtype=torch.bfloat16).contiguous(
        memory_format=torch.channels_last)
    buf[:, :cin] = x.cuda().to(torch.bfloat16)
    with torch.no_grad():
        out = m.cuda()(buf)
    assert out.shape == ref.shape and torch.isfinite(out).all()
    assert rel_err(out, ref) < TOL_BF16


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("act", ["relu", "relu6", "prelu", "hardtanh"])
def test_activations_propagate_nan_like_torch(act, dtype):
    """torch's relu / relu6 / hardtanh / prelu / max_pool2d hand a NaN through; fmaxf-style clamps would hide it."""
    from rtseg_b200 import modules, nets
    x = seeded((1, 32, 10, 24), 63)
    x[0, 5, 4, 7] = float("nan")
    for m, fn in ((modules.ConvBNAct(32, 32, 3, act_type=act), lambda sd, t: B.conv_bn_act(t, sd, "m.", 3, 1, 1, act)),
                  (modules.DWConvBNAct(32, 32, 3, act_type=act), lambda sd, t: B.dw_conv_bn_act(t, sd, "m.", 3, 1, 1, act))):
        m = _randomise(m, 64)
        with torch.no_grad():
            ref = fn(_sd(m), x)
        out = _run(m, x, dtype)
        assert torch.equal(torch.isnan(out), torch.isnan(ref)) and torch.isnan(ref).any()
    img = seeded((1, 3, 12, 20), 65)
    img[0, 1, 6, 9] = float("nan")
    blk = _randomise(nets.InitialBlock(3, 16, act), 66)
    with torch.no_grad():
        ref = B.initial_block(img, _sd(blk), "m.", act)
    blk.act_dtype = dtype
    with torch.no_grad():
        out = blk.cuda()(img.cuda()).float().cpu()
    assert torch.equal(torch.isnan(out), torch.isnan(ref)) and torch.isnan(ref[:, 13:]).any()


@pytest.mark.parametrize("shape", [(2, 64, 128), (1, 37, 53), (1, 3, 1)])
def test_colorize_blend_matches_pillow_restatement(shape):
    """Predict epilogue (reference core/seg_trainer.py:172-191): LUT gather + Pillow blend, bit-exact (uint8)."""
    from rtseg_b200 import ops
    rng = np.random.default_rng(12)
    labels = rng.integers(0, 19, size=shape).astype(np.uint8)
    image = rng.integers(0, 256, size=(*shape, 3)).astype(np.uint8)
    lab = torch.from_numpy(labels).cuda()
    out = ops.colorize(lab)
    assert out.dtype == torch.uint8 and tuple(out.shape) == (*shape, 3)
    assert np.array_equal(out.cpu().numpy(), ref_head.colorize_blend(labels, ops.CITYSCAPES_COLORMAP))
    for alpha in (0.0, 0.3, 0.5, 0.77, 1.0):
        got = ops.colorize(lab, image=torch.from_numpy(image).cuda(), alpha=alpha).cpu().numpy()
        assert np.array_equal(got, ref_head.colorize_blend(labels, ops.CITYSCAPES_COLORMAP, image, alpha)), alpha
    with pytest.raises(RuntimeError):
        ops.colorize(lab, image=torch.from_numpy(image).cuda(), alpha=1.5)


# ------------------------------------------------------------------ auxiliary-head path ----
@pytest.mark.parametrize("hi,wi,ho,wo", [(64, 128, 8, 16), (37, 53, 19, 27), (16, 16, 64, 48), (1024, 1024, 128, 128), (9, 7, 9, 7)])
def test_nearest_resize_of_label_maps_matches_interpolate(hi, wi, ho, wo):
    """reference core/seg_trainer.py:58-62: masks.unsqueeze(1).float() -> F.interpolate(nearest) -> .long()"""
    from rtseg_b200 import ops
    g = torch.Generator().manual_seed(80)
    masks = torch.randint(0, 19, (2, hi, wi), generator=g)
    masks[torch.rand((2, hi, wi), generator=g) < 0.1] = 255
    want = F.interpolate(masks.unsqueeze(1).float(), (ho, wo), mode="nearest").squeeze(1).long()
    got = ops.resize_labels_nearest(masks.cuda(), (ho, wo))
    assert got.dtype == torch.int64 and torch.equal(got.cpu(), want)


def test_loss_with_aux_heads_matches_reference_composition():
    """loss + sum_i coef_i * loss(aux_i, nearest(masks -> aux_i size)) with the OHEM loss (both restated by the oracle)."""
    import rtseg_b200
    g = torch.Generator().manual_seed(81)
    masks = torch.randint(0, 19, (2, 64, 96), generator=g)
    masks[torch.rand((2, 64, 96), generator=g) < 0.1] = 255
    preds = torch.randn((2, 19, 64, 96), generator=g)
    aux = [torch.randn((2, 19, 8, 12), generator=g), torch.randn((2, 19, 16, 24), generator=g)]
    coef = [0.4, 0.7]
    want = ref_ohem.ohem_ce_loss(preds, masks, 0.7)[0]
    for c, a in zip(coef, aux):
        m = F.interpolate(masks.unsqueeze(1).float(), a.shape[2:], mode="nearest").squeeze(1).long()
        want = want + c * ref_ohem.ohem_ce_loss(a, m, 0.7)[0]
    fn = rtseg_b200.OhemCELoss(0.7)
    got = rtseg_b200.loss_with_aux_heads(fn, preds.cuda().requires_grad_(True), [a.cuda() for a in aux], masks.cuda(), coef)
    assert abs(got.item() - want.item()) < 1e-5 * max(1.0, abs(want.item()))
    got.backward()
