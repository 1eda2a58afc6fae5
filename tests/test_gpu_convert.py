"""GPU test of the block-level drop-in: a model in the REFERENCE's module layout (tests/ref_layout.py —
the reference's own forwards with torch.cat / MaxPool2d / `+=` / BatchNorm2d / bare Conv2d / F.interpolate glue) goes
through `rtseg_b200.convert` and must then (i) match the oracle, (ii) keep its state dict, and (iii) launch NO
cuDNN / ATen convolution, batch-norm, pooling, cat or interpolation kernel — checked with torch.profiler."""
import pytest
import torch

from oracle import ref_nets
from tests import ref_layout
from tests.helpers import ARGMAX_AGREE_FP32, TOL_FP32, load_specs, rel_err

pytestmark = pytest.mark.gpu

FORBIDDEN = ("cudnn", "conv", "batch_norm", "batchnorm", "max_pool", "cat", "upsample", "interpolate", "implicit_gemm",
             "winograd", "gemm", "sgemm", "cutlass")


def _kernels(fn):
    """Names of the CUDA kernels fn() launches (torch.profiler, CUDA activity)."""
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        out = fn()
        torch.cuda.synchronize()
    names = [e.name for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    return out, names


def _build(name, seed=3):
    sd = ref_nets.make_state(load_specs()[name]["spec"], seed, True)
    m = ref_layout.LAYOUTS[name](num_class=19)
    m.load_state_dict(sd, strict=True)          # the restated layout carries the reference's state-dict keys
    return m.eval().cuda(), sd


@pytest.mark.parametrize("name", ["erfnet", "fastscnn"])
def test_convert_swaps_blocks_and_removes_every_aten_kernel(name):
    import rtseg_b200
    from rtseg_b200 import nets
    m, sd = _build(name)
    x = ref_nets.make_input(6, 2, 128, 256)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False      # the ATen glue convolutions would otherwise run in TF32 (1e-3 off)
    try:
        with torch.no_grad():
            ref = ref_nets.FORWARD[name](sd, x)
            y_before, k_before = _kernels(lambda: m(x.cuda()))
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    aten_before = [k for k in k_before if any(f in k.lower() for f in FORBIDDEN) and "rtseg" not in k]
    assert aten_before, "the unconverted reference layout was expected to launch ATen / cuDNN glue kernels"
    assert rel_err(y_before, ref) < TOL_FP32     # injected leaves + ATen glue already match the oracle

    keys = list(m.state_dict().keys())
    ptrs = {k: v.data_ptr() for k, v in m.state_dict().items()}
    out = rtseg_b200.convert(m)
    assert out is m and isinstance(m, nets.MODEL_HUB[name]) and m.num_class == 19
    assert list(m.state_dict().keys()) == keys and {k: v.data_ptr() for k, v in m.state_dict().items()} == ptrs
    swapped = m._rtseg_converted
    if name == "erfnet":
        assert swapped["NonBt1DBlock"] == 17 and swapped["InitialBlock"] == 3 and swapped["Conv2d"] == 17
    else:
        assert swapped["InvertedResidual"] == 8 and swapped["FeatureFusionModule"] == 1 and swapped["LearningToDownsample"] == 1

    with torch.no_grad():
        y_after, k_after = _kernels(lambda: m(x.cuda()))
        labels, k_pred = _kernels(lambda: m.predict(x.cuda()))
    leftovers = [k for k in k_after + k_pred if any(f in k.lower() for f in FORBIDDEN) and "rtseg" not in k]
    assert not leftovers, f"ATen / cuDNN kernels still launched after convert(): {sorted(set(leftovers))}"
    assert all("rtseg" in k for k in k_pred if "memcpy" not in k.lower() and "memset" not in k.lower()), sorted(set(k_pred))
    assert rel_err(y_after, ref) < TOL_FP32
    want = ref.max(1)[1]
    assert (y_after.float().cpu().max(1)[1] == want).float().mean().item() >= ARGMAX_AGREE_FP32
    assert (labels.cpu().long() == want).float().mean().item() >= ARGMAX_AGREE_FP32
    assert rtseg_b200.convert(m)._rtseg_converted == {}      # idempotent


def test_fused_head_mode_defers_the_final_interpolate():
    """Unconverted Fast-SCNN layout under fused_head(): `preds.max(dim=1)[1]` (reference core/seg_trainer.py:172) runs
    the fused upsample+argmax kernel — no full-resolution logits tensor, no ATen upsample — and equals the oracle;
    any other use of the deferred tensor yields the real logits."""
    import rtseg_b200
    m, sd = _build("fastscnn", 5)
    x = ref_nets.make_input(7, 1, 256, 512)
    with torch.no_grad():
        ref = ref_nets.fastscnn_forward(sd, x)

        def run():
            with rtseg_b200.fused_head():
                preds = m(x.cuda())
                assert isinstance(preds, rtseg_b200.DeferredLogits) and tuple(preds.shape) == (1, 19, 256, 512)
                return preds, preds.max(dim=1)[1]
        (preds, labels), names = _kernels(run)
        assert any("upsample_argmax" in k for k in names)
        # the feature-fusion resize is also a bilinear interpolate of a CUDA tensor: it materialises through the rtseg
        # kernel; no ATen upsample kernel may run at all
        assert not [k for k in names if "upsample" in k.lower() and "rtseg" not in k]
        assert labels.dtype == torch.int64 and tuple(labels.shape) == (1, 256, 512)
        assert (labels.cpu() == ref.max(1)[1]).float().mean().item() >= ARGMAX_AGREE_FP32
        vals, idx = torch.max(preds, 1)
        assert torch.equal(idx, labels)
        full = preds + 0                      # any other op: the logits themselves
        assert not isinstance(full, rtseg_b200.DeferredLogits) and rel_err(full, ref) < TOL_FP32
        assert rel_err(vals, ref.max(1)[0]) < TOL_FP32
        assert torch.equal(rtseg_b200.predict_labels(m, x.cuda()), labels)


@pytest.mark.parametrize("name,hw", [("enet", (64, 96)), ("bisenetv2", (64, 128)), ("ddrnet", (64, 128))])
def test_convert_of_the_round2_nets_runs_the_fused_forward(name, hw):
    """A net in the reference's attribute layout (same-named stand-in classes, bare nn.Conv2d) after convert(): the same
    class map and logits as the fused net built directly, bit for bit, and no cuDNN / ATen convolution, pooling, cat or
    interpolation kernel in the profile."""
    import rtseg_b200
    from rtseg_b200 import nets
    from tests.test_host_logic import _defused
    sd = ref_nets.make_state(load_specs()[name]["spec"], 3, True)
    kw = {} if name == "enet" else {"use_aux": False}
    fused = nets.MODEL_HUB[name](num_class=19, **kw)
    fused.load_state_dict(sd, strict=True)
    m = rtseg_b200.convert(_defused(fused)).eval().cuda()
    fused = fused.eval().cuda()
    x = ref_nets.make_input(4, 2, *hw).cuda()
    with torch.no_grad():
        want = fused(x)
        m(x)
        got, names = _kernels(lambda: m(x))
    assert torch.equal(got, want)
    bad = [n for n in names if any(f in n.lower() for f in FORBIDDEN) and "rtseg" not in n.lower()]
    assert not bad, bad[:5]
    with torch.no_grad():
        assert torch.equal(m.predict(x), fused.predict(x))
