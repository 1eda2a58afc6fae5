"""GPU parity tests for ENet (BASELINE configs[0]): the pooling pair with indices and the network.

Oracle: torch's own max_pool2d(return_indices) / max_unpool2d for the kernels (the reference calls exactly these through
nn.MaxPool2d / nn.MaxUnpool2d, models/enet.py:130-139); oracle/ref_nets.enet_forward — pinned to the real reference's
outputs in tests/golden/enet_outputs.npz — for the network.
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import ref_nets
from tests.helpers import ARGMAX_AGREE_FP32, TOL_BF16, TOL_FP32, load_net_outputs, load_specs, rel_err, seeded

pytestmark = pytest.mark.gpu


def _torch_indices(idx_u8, w):
    """uint8 window positions [N, hp, wp, C] -> torch's flat int64 indices [N, C, hp, wp] into the H x W plane."""
    n, hp, wp, c = idx_u8.shape
    pos = idx_u8.permute(0, 3, 1, 2).long()
    rows = torch.arange(hp, device=pos.device).view(1, 1, hp, 1) * 2 + pos // 2
    cols = torch.arange(wp, device=pos.device).view(1, 1, 1, wp) * 2 + pos % 2
    return rows * w + cols


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("c,h,w", [(16, 32, 64), (64, 17, 23), (128, 8, 8), (3, 10, 14), (20, 6, 9)])
def test_maxpool2x2_with_indices_and_unpool_match_torch(c, h, w, dtype):
    from rtseg_b200 import ops
    x = seeded((2, c, h, w), 90).to(dtype)
    x[0, 0, 0, 0] = x[0, 0, 0, 1]            # a tie: the first maximum in row-major order wins
    x[1, c - 1, 2, 3] = float("nan")         # a NaN is always taken
    ref, ref_idx = F.max_pool2d(x.float(), 2, 2, return_indices=True)
    pooled, idx = ops.maxpool2x2_with_indices(x.cuda())
    assert pooled.dtype == dtype and idx.dtype == torch.uint8 and tuple(idx.shape) == (2, h // 2, w // 2, c)
    assert torch.equal(torch.nan_to_num(pooled.float().cpu(), nan=-7.0), torch.nan_to_num(ref, nan=-7.0))
    assert torch.equal(_torch_indices(idx, w).cpu(), ref_idx)
    # unpool: values back at their positions, zeros elsewhere (odd sizes: the last row / column stays zero)
    v = seeded(tuple(ref.shape), 91).to(dtype)
    want = F.max_unpool2d(v.float(), ref_idx, 2, 2, output_size=(h, w))
    got = ops.maxunpool2x2(v.cuda(), idx, (h, w))
    assert got.dtype == dtype and torch.equal(got.float().cpu(), want)
    if h % 2 == 0 and w % 2 == 0:                 # default size 2h' x 2w' (torch's flat indices need the true width)
        got2 = ops.maxunpool2x2(v.cuda(), idx)
        assert torch.equal(got2.float().cpu(), F.max_unpool2d(v.float(), ref_idx, 2, 2))


def _model(seed=1, perturb=True):
    from rtseg_b200 import nets
    sd = ref_nets.make_state(load_specs()["enet"]["spec"], seed, perturb)
    m = nets.ENet(num_class=19)
    m.load_state_dict(sd, strict=True)
    return m.eval().cuda(), sd


@pytest.mark.parametrize("tag,perturb", [("perturbed", True), ("identity_bn", False)])
def test_enet_matches_reference_golden_vectors_at_a_second_size(tag, perturb):
    """64 x 96 input (three pooling levels with non-power-of-two width): the real reference's logits, fp32 <= 1e-4."""
    m, _ = _model(1, perturb)
    x = ref_nets.make_input(7, 1, 64, 96)
    with torch.no_grad():
        y = m(x.cuda())
    ref = torch.from_numpy(load_net_outputs()[f"enet.{tag}.64x96"])
    assert y.shape == ref.shape and rel_err(y, ref) < TOL_FP32
    assert (y.float().cpu().max(1)[1] == ref.max(1)[1]).float().mean().item() >= ARGMAX_AGREE_FP32


def test_enet_config1_shape_fp32_and_bf16():
    """BASELINE configs[0]: ENet, 1 x 3 x 512 x 1024, 19 classes (reference tools/test_speed.py: ratio 0.5 of 1024 x 2048).
    fp32 <= 1e-4 and argmax >= 99.9 % against the oracle; bf16 within the format's own error; fused head = oracle head."""
    from oracle import ref_blocks, ref_head
    from rtseg_b200 import ops
    m, sd = _model(2)
    x = ref_nets.make_input(8, 1, 512, 1024)
    with torch.no_grad():
        ref = ref_nets.enet_forward(sd, x)
        y32 = m(x.cuda())
        labels32 = m.predict(x.cuda())
        ref_blocks.STORAGE = torch.bfloat16
        try:
            emu = ref_nets.enet_forward(sd, x)
        finally:
            ref_blocks.STORAGE = None
        m.set_act_dtype(torch.bfloat16)
        y16 = m(x.cuda())
        low = m.lowres_logits(x.cuda())
        labels16 = m.predict(x.cuda())
    assert ops.tc_timeout_flag() == 0
    from tests.helpers import rel_err_robust
    assert tuple(y32.shape) == (1, 19, 512, 1024) and rel_err_robust(y32, ref, 1e-5)[0] < TOL_FP32
    want = ref.max(1)[1]
    assert (y32.float().cpu().max(1)[1] == want).float().mean().item() >= ARGMAX_AGREE_FP32
    assert (labels32.cpu().long() == want).float().mean().item() >= ARGMAX_AGREE_FP32
    err_fmt = rel_err(emu, ref)
    assert rel_err(y16, emu) < 2.0 * err_fmt and rel_err(y16, ref) < max(TOL_BF16, 1.5 * err_fmt)
    head = ref_head.upsample_argmax(low.float().cpu(), (512, 1024), False)
    assert (labels16.cpu().long() == head).float().mean().item() >= ARGMAX_AGREE_FP32
    print(f"enet 1x3x512x1024: fp32 {rel_err(y32, ref):.2e}; bf16 vs emulation {rel_err(y16, emu):.2e} (format {err_fmt:.2e}); "
          f"bf16 argmax agreement with fp32 oracle {(labels16.cpu().long() == want).float().mean().item():.5f}")
