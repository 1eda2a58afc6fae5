"""GPU parity tests of the fused 3x1 -> 1x3 pair kernel (rtseg_conv_pair_tc, csrc/pair_tc.cu) against the CPU oracle.

The pair replaces two ConvBNAct launches of a NonBt1D block (reference models/erfnet.py:65-71) or its tail
(conv -> `+= x` -> BatchNorm -> act, :77-81); the intermediate tensor lives only in shared memory, rounded to bf16
exactly where the two-launch path rounds it.  Tolerance: bf16 <= 2e-2 of the largest reference magnitude; against
the two-launch rtseg path (same rounding points, different tiling) the results must agree to a few bf16 ulps.
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import ref_blocks as B
from oracle import ref_nets
from tests.helpers import TOL_BF16, rel_err, seeded
from tests.test_gpu_kernels import _randomise, _sd

pytestmark = pytest.mark.gpu

CASES = [
    # c, n, h, w
    (64, 1, 8, 30),          # exactly two tiles high, one wide
    (64, 2, 19, 45),         # ragged: partial tiles right and bottom
    (64, 1, 5, 8),           # narrower than a tile
    (64, 3, 64, 128),
    (16, 1, 16, 62),         # one 16-channel tile
    (16, 2, 37, 150),
    (16, 1, 3, 9),
    (64, 4, 256, 512),       # ERFNet layer3_7 at its own size: ~125 tiles per CTA
    (16, 4, 512, 1024),      # ERFNet layer21_22 at its own size
]


def _oracle_pair(x, blk_sd, res):
    """conv[0] -> conv[1] (res=False) or conv[2] -> conv[3] -> += x -> bn_act (res=True) of a NonBt1D block, fp32,
    on the bf16-rounded input; the intermediate is rounded to bf16 like the kernel's shared-memory tile."""
    xq = x.bfloat16().float()
    if not res:
        t = B.conv_bn_act(xq, blk_sd, "m.conv.0.", (3, 1), 1, 1, "relu").bfloat16().float()
        return B.conv_bn_act(t, blk_sd, "m.conv.1.", (1, 3), 1, 1, "relu")
    t = B.conv_bn_act(xq, blk_sd, "m.conv.2.", (3, 1), 1, 1, "relu").bfloat16().float()
    y = F.conv2d(t, blk_sd["m.conv.3.weight"], None, 1, (0, 1), 1)
    return F.relu(B.batch_norm(y + xq, blk_sd, "m.bn_act.0."))


@pytest.mark.parametrize("fold", [True, False], ids=["bn-in-mma", "bn-in-epilogue"])
@pytest.mark.parametrize("res", [False, True])
@pytest.mark.parametrize("c,n,h,w", CASES)
def test_pair_kernel_vs_oracle(c, n, h, w, res, fold, monkeypatch):
    """fold=True: weights carry the BatchNorm scale and the shift is added by the bias MMA (the default);
    fold=False: scale and shift applied by the epilogue like conv_tc.cu does."""
    from rtseg_b200 import _lib, nets, ops
    monkeypatch.setattr(ops, "PAIR_FOLD_BN", fold)
    blk = _randomise(nets.NonBt1DBlock(c, 1), 70 + c).cuda()
    sd = _sd(blk)
    x = seeded((n, c, h, w), 71)
    xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
    i0 = 2 if res else 0
    fa, _ = blk.conv[i0]._folded()
    fb, _ = blk.conv[3].folded(blk.bn_act[0], blk.bn_act[1].activation) if res else blk.conv[1]._folded()
    assert ops.conv_pair_supported(fa, fb, xb, res=xb if res else None)
    with torch.no_grad():
        ref = _oracle_pair(x, sd, res)
        before = _lib.launch_count()
        first = ops.conv_pair(fa, fb, xb, res=xb if res else None).clone()
        for _ in range(3):     # bit-identical run to run
            assert torch.equal(first, ops.conv_pair(fa, fb, xb, res=xb if res else None))
        assert _lib.launch_count() - before == 4      # one launch per pair
        # the two-launch path: same operands, same rounding points
        t = blk.conv[i0](xb)
        two = blk.conv[3](t, res=xb, res_mode=ops.RES_PRE, bn=blk.bn_act[0], act=blk.bn_act[1].activation) if res else blk.conv[1](t)
    assert ops.tc_timeout_flag() == 0
    assert first.shape == ref.shape and first.dtype == torch.bfloat16
    err = rel_err(first, ref)
    assert err < TOL_BF16, f"pair kernel off by {err:.3e}"
    # against the two-launch kernels only accumulation order differs: a few bf16 ulps of the largest value
    assert rel_err(first, two) < 1.2e-2
    # the image border columns are where a missing zero-padding of the intermediate would show
    for col in (0, w - 1):
        assert rel_err(first[..., col], ref[..., col]) < TOL_BF16


def test_pair_output_into_a_channel_slice_and_residual_from_a_slice():
    from rtseg_b200 import nets, ops
    blk = _randomise(nets.NonBt1DBlock(64, 1), 72).cuda()
    x, r = seeded((2, 64, 12, 40), 73), seeded((2, 64, 12, 40), 74)
    wide_x = torch.full((2, 96, 12, 40), float("nan"), dtype=torch.bfloat16, device="cuda").contiguous(memory_format=torch.channels_last)
    wide_x[:, :64] = x.to("cuda", torch.bfloat16)       # the input occupies channels [0, 64) of a wider buffer
    fa, _ = blk.conv[2]._folded()
    fb, _ = blk.conv[3].folded(blk.bn_act[0], blk.bn_act[1].activation)
    rb = ops.to_nhwc(r.to("cuda", torch.bfloat16))
    sd = _sd(blk)
    with torch.no_grad():
        t = B.conv_bn_act(x.bfloat16().float(), sd, "m.conv.2.", (3, 1), 1, 1, "relu").bfloat16().float()
        ref = F.relu(B.batch_norm(F.conv2d(t, sd["m.conv.3.weight"], None, 1, (0, 1)) + r.bfloat16().float(), sd, "m.bn_act.0."))
        out = ops.conv_pair(fa, fb, wide_x, res=rb)
    assert rel_err(out, ref) < TOL_BF16


@pytest.mark.parametrize("c,h,w", [(64, 64, 96), (16, 40, 200)])
def test_nonbt1d_block_takes_the_pair_path_and_matches_the_oracle(c, h, w, monkeypatch):
    from rtseg_b200 import _lib, nets, ops
    blk = _randomise(nets.NonBt1DBlock(c, 1), 75).cuda()
    x = seeded((2, c, h, w), 76)
    xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
    with torch.no_grad():
        ref = ref_nets.non_bt_1d(x.bfloat16().float(), _sd(blk), "m.", 1, "relu")
        before = _lib.launch_count()
        y = blk(xb)
        assert _lib.launch_count() - before == 2, "an undilated bf16 block is two pair launches"
        monkeypatch.setattr(ops, "USE_PAIR_FUSION", False)
        before = _lib.launch_count()
        y4 = blk(xb)
        assert _lib.launch_count() - before == 4
    assert ops.tc_timeout_flag() == 0
    assert rel_err(y, ref) < TOL_BF16 and rel_err(y4, ref) < TOL_BF16
    dil = _randomise(nets.NonBt1DBlock(c, 2), 77).cuda()          # dilated blocks keep the four-launch path
    before = _lib.launch_count()
    with torch.no_grad():
        dil(xb)
    assert _lib.launch_count() - before == 4
