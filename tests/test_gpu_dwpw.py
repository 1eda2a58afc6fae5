"""GPU parity of the fused depthwise 3x3 -> pointwise 1x1 kernel (rtseg_dwpw_tc): DSConvBNAct (reference
models/modules.py:36-42), the tail of InvertedResidual (models/fastscnn.py:113-121) and FeatureFusionModule's
low-resolution branch (:80-92), against the oracle and against the two-launch path it replaces."""
import pytest
import torch

from oracle import ref_blocks as B
from oracle import ref_nets
from tests.helpers import TOL_BF16, rel_err, seeded
from tests.test_gpu_train import _randomise

pytestmark = pytest.mark.gpu


def _nhwc16(t):
    return t.to("cuda", torch.bfloat16).contiguous(memory_format=torch.channels_last)


def _sd(m):
    return {"m." + k: v.detach().float().cpu() for k, v in m.state_dict().items()}


def _both(m, *inputs):
    """(fused output, two-launch output, librtseg launches of the fused call)"""
    from rtseg_b200 import _lib, ops
    m = m.eval().cuda()
    with torch.no_grad():
        ops.USE_DWPW_FUSION = False
        try:
            two = m(*inputs)
        finally:
            ops.USE_DWPW_FUSION = True
        m(*inputs)
        before = _lib.launch_count()
        one = m(*inputs)
        launches = _lib.launch_count() - before
    assert ops.tc_timeout_flag() == 0
    return one, two, launches


def _emulated(fn):
    B.STORAGE = torch.bfloat16
    try:
        with torch.no_grad():
            return fn()
    finally:
        B.STORAGE = None


@pytest.mark.parametrize("shape", [(2, 16, 32), (2, 13, 27), (1, 4, 8), (3, 40, 72), (1, 9, 130)])
@pytest.mark.parametrize("cin,cout,act", [(128, 128, "relu"), (64, 64, "relu6"), (128, 64, "none"), (256, 128, "relu")])
def test_ds_conv_fused_matches_oracle_and_two_launch_path(cin, cout, act, shape):
    from rtseg_b200 import modules
    n, h, w = shape
    m = _randomise(modules.DSConvBNAct(cin, cout, 3, 1, act_type=act), 300)
    x = seeded((n, cin, h, w), 301)
    xq = x.to(torch.bfloat16).float()
    ref = _emulated(lambda: B.ds_conv_bn_act(xq, _sd(m), "m.", 3, 1, 1, act))
    one, two, launches = _both(m, _nhwc16(x))
    assert launches == 1
    assert one.shape == ref.shape and rel_err(one, ref) < TOL_BF16 and rel_err(one, two) < TOL_BF16 / 2


def test_ds_conv_fused_writes_a_channel_slice_of_a_wider_buffer():
    from rtseg_b200 import modules, ops
    m = _randomise(modules.DSConvBNAct(128, 64, 3, 1), 302).eval().cuda()
    x = _nhwc16(seeded((2, 128, 12, 20), 303))
    wide = torch.full((2, 192, 12, 20), 7.0, dtype=torch.bfloat16, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        alone = m(x)
        m(x, out=wide, y_coff=64)
    assert torch.equal(wide[:, 64:128], alone) and (wide[:, :64] == 7).all() and (wide[:, 128:] == 7).all()
    assert ops.tc_timeout_flag() == 0


@pytest.mark.parametrize("cin,cout,stride,shape", [(64, 64, 1, (2, 16, 32)), (64, 64, 1, (1, 11, 19)), (64, 96, 2, (2, 16, 32)),
                                                   (96, 128, 1, (2, 8, 16))])
def test_inverted_residual_tail_fused(cin, cout, stride, shape):
    """384 -> 64 (+ x) runs fused; 576 / 768 hidden channels and stride 2 keep the two-launch path (same results)."""
    from rtseg_b200 import nets
    n, h, w = shape
    m = _randomise(nets.InvertedResidual(cin, cout, stride), 304)
    x = seeded((n, cin, h, w), 305)
    xq = x.to(torch.bfloat16).float()
    ref = _emulated(lambda: ref_nets.inverted_residual(xq, _sd(m), "m.", cin, cout, stride))
    one, two, launches = _both(m, _nhwc16(x))
    assert launches == (2 if (cin * 6 == 384 and stride == 1) else 3)
    assert rel_err(one, ref) < TOL_BF16 and rel_err(one, two) < TOL_BF16 / 2


@pytest.mark.parametrize("shape", [(2, 16, 32), (1, 12, 20)])
def test_feature_fusion_lower_branch_fused(shape):
    """upsample -> [DW 3x3 -> 1x1 conv -> + conv_hi(hi) -> BN -> relu] (reference models/fastscnn.py:73-93)."""
    from rtseg_b200 import nets
    n, h, w = shape
    m = _randomise(nets.FeatureFusionModule(64, 128, 128), 306)
    hi, lo = seeded((n, 64, h, w), 307), seeded((n, 128, h // 4, w // 4), 308)
    one, two, launches = _both(m, _nhwc16(hi), _nhwc16(lo))
    assert launches == 3      # conv_hi, bilinear resize, fused lower branch
    assert rel_err(one, two) < TOL_BF16 / 2
    # oracle: the same ops unfused, bf16 storage
    import torch.nn.functional as F
    sd = _sd(m)

    def ref_fn():
        hq, lq = hi.to(torch.bfloat16).float(), lo.to(torch.bfloat16).float()
        hh = B.q(F.conv2d(hq, B.qw(sd["m.higher_res_conv.weight"])))
        ll = B.q(F.interpolate(lq, hq.shape[2:], mode="bilinear", align_corners=True))
        ll = B.dw_conv_bn_act(ll, sd, "m.lower_res_conv.0.", 3, 1, 1, "relu")
        ll = F.conv2d(ll, B.qw(sd["m.lower_res_conv.1.weight"]))
        return B.q(torch.relu(B.batch_norm(hh + ll, sd, "m.non_linear.0.")))
    assert rel_err(one, _emulated(ref_fn)) < TOL_BF16


@pytest.mark.parametrize("shape", [(2, 16, 32), (2, 34, 50), (1, 8, 16), (3, 64, 96)])
@pytest.mark.parametrize("cin,cout,act", [(32, 48, "relu"), (48, 64, "relu"), (16, 16, "relu6"), (64, 32, "none")])
def test_ds_conv_stride2_fused_matches_oracle_and_two_launch_path(cin, cout, act, shape):
    """The stride-2 narrow-layer variant: Fast-SCNN's learning-to-downsample pairs (reference models/fastscnn.py:38-40)."""
    from rtseg_b200 import modules
    n, h, w = shape
    m = _randomise(modules.DSConvBNAct(cin, cout, 3, 2, act_type=act), 320)
    x = seeded((n, cin, h, w), 321)
    xq = x.to(torch.bfloat16).float()
    ref = _emulated(lambda: B.ds_conv_bn_act(xq, _sd(m), "m.", 3, 2, 1, act))
    one, two, launches = _both(m, _nhwc16(x))
    assert launches == 1
    assert one.shape == ref.shape and rel_err(one, ref) < TOL_BF16 and rel_err(one, two) < TOL_BF16 / 2


def test_fused_path_declines_what_it_does_not_cover():
    from rtseg_b200 import _lib, modules
    for args in ((32, 48, 3, 2, 2), (384, 64, 3, 2), (128, 128, 3, 1, 2), (128, 19, 3, 1)):   # dilation 2, wide stride 2, 19 outputs
        m = _randomise(modules.DSConvBNAct(*args), 309).eval().cuda()
        x = _nhwc16(seeded((1, args[0], 16, 24), 310))
        with torch.no_grad():
            m(x)
            before = _lib.launch_count()
            m(x)
        assert _lib.launch_count() - before == 2


def test_fused_kernel_in_the_timed_regime_many_tiles_per_cta_two_streams():
    """The shape bench.py times (Fast-SCNN's classifier at batch 32: 8,192 tiles, 55 per CTA, every ring slot and both
    accumulators reused many times): equal to the two-launch path, bit-identical from launch to launch, and unchanged
    when two streams run it concurrently (the engine's two slots do)."""
    from rtseg_b200 import modules, ops
    m = _randomise(modules.DSConvBNAct(128, 128, 3, 1), 311).eval().cuda()
    xs = [_nhwc16(seeded((32, 128, 128, 256), 312 + i)) for i in range(2)]
    with torch.no_grad():
        ops.USE_DWPW_FUSION = False
        try:
            two = [m(x) for x in xs]
        finally:
            ops.USE_DWPW_FUSION = True
        one = [m(x) for x in xs]
        for a, b in zip(one, two):
            assert rel_err(a, b) < TOL_BF16 / 2
        for _ in range(10):
            assert torch.equal(m(xs[0]), one[0])
        streams = [torch.cuda.Stream() for _ in range(2)]
        torch.cuda.synchronize()
        outs = [[], []]
        for _ in range(8):
            for i, st in enumerate(streams):
                with torch.cuda.stream(st):
                    outs[i].append(m(xs[i]))
        torch.cuda.synchronize()
        for i in range(2):
            for o in outs[i]:
                assert torch.equal(o, one[i])
    assert ops.tc_timeout_flag() == 0
