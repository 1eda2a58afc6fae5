"""GPU parity tests for BASELINE config 4: the fused glue ops, the blocks of BiSeNetv2 / DDRNet and
the two networks.

Oracle: torch's own interpolate / avg_pool2d for the two glue kernels; oracle/ref_nets_dual.py — pinned to the real
reference's outputs in tests/golden/cfg4_outputs.npz — for blocks and networks.
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import ref_blocks, ref_head, ref_nets, ref_nets_dual as D
from tests.helpers import ARGMAX_AGREE_FP32, TOL_BF16, TOL_FP32, load_net_outputs, load_specs, rel_err, seeded
from tests.test_gpu_train import _randomise

pytestmark = pytest.mark.gpu
DTYPES = [torch.float32, torch.bfloat16]


def _nhwc(t, dtype):
    return t.to("cuda", dtype).contiguous(memory_format=torch.channels_last)


# ------------------------------------------------------------------------------ glue kernels ----
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("c,h,w,hb,wb,op,sig,act", [
    (128, 16, 32, 4, 8, "mul", True, None), (128, 8, 16, 8, 16, "mul", True, None), (128, 16, 32, 4, 8, "add", False, None),
    (128, 8, 12, 1, 1, "add", False, None), (64, 16, 24, 4, 6, "add", False, "relu"), (19, 9, 13, 3, 5, "add", False, "relu"),
    (32, 6, 10, 1, 2, "add", False, None)])
def test_combine_up_matches_torch(c, h, w, hb, wb, op, sig, act, dtype):
    """act(a (+|*) g(interpolate(b, align_corners=True))) — reference models/bisenetv2.py:190-201, ddrnet.py:238-243."""
    from rtseg_b200 import ops
    a, b = seeded((2, c, h, w), 200).to(dtype), seeded((2, c, hb, wb), 201).to(dtype)
    up = F.interpolate(b.float(), (h, w), mode="bilinear", align_corners=True) if (hb, wb) != (h, w) else b.float()
    g = torch.sigmoid(up) if sig else up
    ref = a.float() * g if op == "mul" else a.float() + g
    ref = F.relu(ref) if act else ref
    out = ops.combine_up(_nhwc(a, dtype), _nhwc(b, dtype), op, b_sigmoid=sig, act_module=torch.nn.ReLU() if act else None)
    assert out.dtype == dtype and rel_err(out, ref) < (1e-6 if dtype == torch.float32 else TOL_BF16 / 4)


@pytest.mark.parametrize("dtype", DTYPES)
def test_combine_up_reads_and_writes_channel_slices(dtype):
    from rtseg_b200 import ops
    a, b = seeded((1, 96, 8, 12), 202).to(dtype), seeded((1, 32, 2, 3), 203).to(dtype)
    buf = _nhwc(a, dtype)
    keep = buf.clone()
    out = ops.combine_up(buf, _nhwc(b, dtype), "add", a_coff=32, c=32, out=buf, y_coff=64)      # buf[64:96] = buf[32:64] + up(b)
    ref = a[:, 32:64].float() + F.interpolate(b.float(), (8, 12), mode="bilinear", align_corners=True)
    assert out is buf and rel_err(buf[:, 64:], ref) < (1e-6 if dtype == torch.float32 else TOL_BF16 / 4)
    assert torch.equal(buf[:, :64], keep[:, :64])


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("c,h,w,k,s,p", [(512, 16, 32, 5, 2, 2), (512, 16, 32, 9, 4, 4), (512, 16, 32, 17, 8, 8), (128, 9, 14, 3, 2, 1),
                                         (64, 1, 2, 17, 8, 8), (19, 4, 6, 5, 2, 2)])
def test_avgpool2d_matches_torch(c, h, w, k, s, p, dtype):
    """nn.AvgPool2d(k, s, p), count_include_pad=True — DAPPM's pools, reference models/ddrnet.py:266-275."""
    from rtseg_b200 import ops
    x = seeded((2, c, h, w), 204).to(dtype)
    ref = F.avg_pool2d(x.float(), k, s, p)
    out = ops.avgpool2d(_nhwc(x, dtype), k, s, p)
    assert tuple(out.shape) == tuple(ref.shape) and rel_err(out, ref) < (1e-6 if dtype == torch.float32 else TOL_BF16 / 4)


# ------------------------------------------------------------------------------------ blocks ----
def _block_check(m, fn, x, dtype, extra=()):
    m = m.eval().cuda()
    sd = {"m." + k: v.detach().float().cpu() for k, v in m.state_dict().items()}
    with torch.no_grad():
        ref = fn(sd, x, *extra)
        out = m(_nhwc(x, dtype), *[_nhwc(e, dtype) for e in extra])
    tol = TOL_FP32 if dtype == torch.float32 else TOL_BF16
    if isinstance(ref, tuple):
        for o, r in zip(out, ref):
            assert rel_err(o, r) < tol
    else:
        assert tuple(out.shape) == tuple(ref.shape) and rel_err(out, ref) < tol


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("cin,cout,stride", [(16, 32, 2), (32, 32, 1), (64, 128, 2), (128, 128, 1)])
def test_gather_expansion_layer(cin, cout, stride, dtype):
    from rtseg_b200 import nets_dual
    m = _randomise(nets_dual.GatherExpansionLayer(cin, cout, stride), 210)
    _block_check(m, lambda sd, t: D.gather_expansion(t, sd, "m.", stride), seeded((2, cin, 16, 24), 211), dtype)


@pytest.mark.parametrize("dtype", DTYPES)
def test_stem_block_and_context_embedding(dtype):
    from rtseg_b200 import nets_dual
    m = _randomise(nets_dual.StemBlock(3, 16), 212)
    m.act_dtype = dtype
    m = m.eval().cuda()
    sd = {"m." + k: v.detach().float().cpu() for k, v in m.state_dict().items()}
    x = seeded((2, 3, 32, 48), 213)
    with torch.no_grad():
        assert rel_err(m(x.cuda()), D.stem_block(x, sd, "m.")) < (TOL_FP32 if dtype == torch.float32 else TOL_BF16)
    m = _randomise(nets_dual.ContextEmbeddingBlock(128, 128), 214)
    _block_check(m, lambda sd, t: D.context_embedding(t, sd, "m."), seeded((2, 128, 4, 8), 215), dtype)


@pytest.mark.parametrize("dtype", DTYPES)
def test_bilateral_guided_aggregation_layer(dtype):
    from rtseg_b200 import nets_dual
    m = _randomise(nets_dual.BilateralGuidedAggregationLayer(128, 128), 216)
    x_d, x_s = seeded((2, 128, 16, 32), 217), seeded((2, 128, 4, 8), 218)
    _block_check(m, lambda sd, d, s: D.bilateral_guided_aggregation(d, s, sd, "m."), x_d, dtype, extra=(x_s,))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("block,cin,cout,stride", [("RB", 32, 32, 1), ("RB", 32, 64, 2), ("RB", 64, 128, 1), ("RBB", 64, 128, 1),
                                                   ("RBB", 256, 512, 2)])
def test_residual_blocks(block, cin, cout, stride, dtype):
    from rtseg_b200 import nets_dual
    m = _randomise(getattr(nets_dual, block)(cin, cout, stride), 219)
    fn = D.rb if block == "RB" else D.rbb
    _block_check(m, lambda sd, t: fn(t, sd, "m.", cin, cout, stride), seeded((2, cin, 12, 20), 220), dtype)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("low,high,stride", [(128, 64, 2), (256, 64, 4)])
def test_bilateral_fusion(low, high, stride, dtype):
    from rtseg_b200 import nets_dual
    m = _randomise(nets_dual.BilateralFusion(low, high, stride), 221)
    x_low, x_high = seeded((2, low, 4, 8), 222), seeded((2, high, 4 * stride, 8 * stride), 223)
    _block_check(m, lambda sd, lo, hi: D.bilateral_fusion(lo, hi, sd, "m.", stride), x_low, dtype, extra=(x_high,))


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("h,w", [(16, 32), (1, 2), (5, 7)])
def test_dappm(h, w, dtype):
    from rtseg_b200 import nets_dual
    m = _randomise(nets_dual.DAPPM(512, 128), 224)
    _block_check(m, lambda sd, t: D.dappm(t, sd, "m."), seeded((2, 512, h, w), 225), dtype)


# ---------------------------------------------------------------------------------- networks ----
def _model(name, seed=1, perturb=True):
    from rtseg_b200 import nets
    sd = ref_nets.make_state(load_specs()[name]["spec"], seed, perturb)
    m = nets.MODEL_HUB[name](num_class=19, use_aux=False)       # get_model builds them like this (use_aux False)
    m.load_state_dict(sd, strict=True)
    return m.eval().cuda(), sd


@pytest.mark.parametrize("name", ["bisenetv2", "ddrnet"])
@pytest.mark.parametrize("tag,perturb", [("perturbed", True), ("identity_bn", False)])
def test_config4_nets_match_reference_golden_vectors(name, tag, perturb):
    """The REAL reference's logits on a 1 x 3 x 64 x 128 input: fp32 <= 1e-4, class maps >= 99.9 %."""
    m, _ = _model(name, 1, perturb)
    x = ref_nets.make_input(0, 1, 64, 128)
    with torch.no_grad():
        y = m(x.cuda())
    ref = torch.from_numpy(load_net_outputs()[f"{name}.{tag}"])
    assert y.shape == ref.shape and rel_err(y, ref) < TOL_FP32
    assert (y.float().cpu().max(1)[1] == ref.max(1)[1]).float().mean().item() >= ARGMAX_AGREE_FP32


def test_config4_state_dict_with_aux_heads_has_the_reference_keys():
    """use_aux=True (the class default, models/bisenetv2.py:19, ddrnet.py:19): aux heads exist under the reference's names."""
    from rtseg_b200 import nets
    keys = set(nets.BiSeNetv2(19).state_dict())
    assert {"semantic_branch.seg_head2.0.0.weight", "semantic_branch.seg_head5.1.weight", "bga_layer.conv_last.0.weight"} <= keys
    keys = set(nets.DDRNet(19).state_dict())
    assert {"aux_head.0.0.weight", "aux_head.1.weight", "conv5.dappm.conv_last.0.weight"} <= keys


@pytest.mark.parametrize("name", ["bisenetv2", "ddrnet"])
def test_config4_nets_at_512x1024_fp32_bf16_and_fused_head(name):
    """Batch 2 at 512 x 1024 (config 4 runs at 1024 x 2048; the oracle on the host finishes this size in seconds):
    fp32 <= 1e-4 with >= 99.9 % class agreement; bf16 against the bf16-storage emulation of the oracle; the fused
    upsample + argmax head equals the oracle head on the same low-resolution logits; uint8 HWC input accepted."""
    from rtseg_b200 import ops
    m, sd = _model(name, 2)
    x = ref_nets.make_input(9, 2, 512, 1024)
    with torch.no_grad():
        ref = ref_nets.FORWARD[name](sd, x)
        ref64 = ref_nets.FORWARD[name]({k: (v.double() if v.dtype.is_floating_point else v) for k, v in sd.items()}, x.double())
        y32 = m(x.cuda())
        labels32 = m.predict(x.cuda())
        ref_blocks.STORAGE = torch.bfloat16
        try:
            emu = ref_nets.FORWARD[name](sd, x)
        finally:
            ref_blocks.STORAGE = None
        m.set_act_dtype(torch.bfloat16)
        y16 = m(x.cuda())
        low = m.lowres_logits(x.cuda())
        labels16 = m.predict(x.cuda())
    assert ops.tc_timeout_flag() == 0
    want = ref.max(1)[1]
    # BiSeNetv2's aggregation layer feeds un-normalised semantic features (|x| ~ 1e4 with random weights) through a
    # sigmoid: wherever they cross zero an fp32 rounding difference of 1e-6 relative becomes 2.5e-4 absolute.  The fp32
    # oracle itself is that far from its own fp64 evaluation, so the bound is taken against fp64 and scaled by the
    # oracle's own fp32 distance (1e-4 wherever the net is well conditioned: DDRNet, and both nets on the golden input)
    own = rel_err(ref, ref64)
    from tests.helpers import rel_err_robust
    assert tuple(y32.shape) == (2, 19, 512, 1024) and rel_err(y32, ref64) < max(TOL_FP32, 6.0 * own), (rel_err(y32, ref64), own)
    assert rel_err_robust(y32, ref64, 1e-5)[0] < max(TOL_FP32, 3.0 * own)   # all but 1e-5 of the logits
    assert (labels32.cpu().long() == want).float().mean().item() >= ARGMAX_AGREE_FP32
    err_fmt = rel_err(emu, ref)
    agree16 = (labels16.cpu().long() == want).float().mean().item()
    print(f"{name} 2x3x512x1024: fp32 {rel_err(y32, ref64):.2e} from the fp64 oracle (fp32 oracle itself {own:.2e}); bf16 vs fp32 oracle {rel_err(y16, ref):.2e}, vs emulation "
          f"{rel_err(y16, emu):.2e} (format {err_fmt:.2e}); bf16 argmax agreement with fp32 oracle {agree16:.5f}")
    assert rel_err(y16, emu) < max(TOL_BF16, 2.0 * err_fmt) and rel_err(y16, ref) < max(TOL_BF16, 2.0 * err_fmt)
    head = ref_head.upsample_argmax(low.float().cpu(), (512, 1024), True)
    assert (labels16.cpu().long() == head).float().mean().item() >= ARGMAX_AGREE_FP32
