"""GPU parity tests at network level: ERFNet / Fast-SCNN against the CPU oracle and the reference's
golden vectors; fused-head, engine and full-size (1024x2048) checks."""
import numpy as np
import pytest
import torch

from oracle import ref_head, ref_nets
from tests.helpers import NETS, NETS_BF16, ARGMAX_AGREE_FP32, TOL_BF16, TOL_FP32, load_net_outputs, load_specs, rel_err

pytestmark = pytest.mark.gpu


def _model(name, seed=1, perturb=True):
    from rtseg_b200 import nets
    spec = load_specs()[name]["spec"]
    sd = ref_nets.make_state(spec, seed, perturb)
    m = nets.MODEL_HUB[name](num_class=19)
    m.load_state_dict(sd, strict=True)
    return m.eval().cuda(), sd


@pytest.mark.parametrize("name", NETS)
@pytest.mark.parametrize("tag,perturb", [("perturbed", True), ("identity_bn", False)])
def test_fp32_logits_match_reference_golden_vectors(name, tag, perturb):
    """The reference's own outputs (tests/golden, generated from /root/reference) — fp32 <= 1e-4."""
    m, _ = _model(name, 1, perturb)
    x = ref_nets.make_input(0, 1, 32, 64)
    with torch.no_grad():
        y = m(x.cuda())
    ref = torch.from_numpy(load_net_outputs()[f"{name}.{tag}"])
    assert y.shape == ref.shape
    assert rel_err(y, ref) < TOL_FP32
    assert (y.float().cpu().max(1)[1] == ref.max(1)[1]).float().mean().item() >= ARGMAX_AGREE_FP32


def _emulated_bf16(name, sd, x):
    """Oracle with bf16 STORAGE rounding at the kernels' own rounding points, fp32 arithmetic."""
    from oracle import ref_blocks
    ref_blocks.STORAGE = torch.bfloat16
    try:
        with torch.no_grad():
            return ref_nets.FORWARD[name](sd, x)
    finally:
        ref_blocks.STORAGE = None


@pytest.mark.parametrize("name", NETS)
def test_fp32_and_bf16_vs_oracle_medium(name):
    """Seeded, BN-perturbed weights (a wrong BN fold cannot hide).

    fp32 path: <= 1e-4 against the fp32 oracle.  bf16 path: these weights amplify rounding (the oracle
    itself, with nothing but bf16 storage rounding inserted, is 1.5-4e-2 away from fp32), so the kernel
    is held to the storage-emulating oracle instead and the distance to fp32 is reported."""
    from rtseg_b200 import ops
    m, sd = _model(name, 2)
    x = ref_nets.make_input(3, 2, 128, 256)
    with torch.no_grad():
        ref = ref_nets.FORWARD[name](sd, x)
        y32 = m(x.cuda())
    assert rel_err(y32, ref) < TOL_FP32
    agree32 = (y32.float().cpu().max(1)[1] == ref.max(1)[1]).float().mean().item()
    assert agree32 >= ARGMAX_AGREE_FP32
    m.set_act_dtype(torch.bfloat16)  # bf16 activations, fp32 master weights
    with torch.no_grad():
        y16 = m(x.cuda())
        labels = m.predict(x.cuda())
    assert ops.tc_timeout_flag() == 0
    assert y16.dtype == torch.bfloat16
    emu = _emulated_bf16(name, sd, x)
    err_emu, err16, err_fmt = rel_err(y16, emu), rel_err(y16, ref), rel_err(emu, ref)
    agree16 = (y16.float().cpu().max(1)[1] == ref.max(1)[1]).float().mean().item()
    print(f"{name}: fp32 argmax agreement {agree32:.5f}; bf16 vs storage-emulating oracle {err_emu:.3e}; "
          f"bf16 vs fp32 {err16:.3e} (format alone: {err_fmt:.3e}), argmax agreement {agree16:.5f}")
    # Two bf16 evaluations that differ only in accumulation order decorrelate after ~70 layers (one
    # flipped rounding propagates), so kernel-vs-emulation sits near sqrt(2) x the format's own error.
    assert err_emu < 2.0 * err_fmt
    assert err16 < max(TOL_BF16, 1.5 * err_fmt)
    if name == "erfnet":  # fused argmax epilogue == argmax of the kernel's own bf16 logits
        own = y16.float().max(1)[1].cpu()
        assert (labels.cpu().long() == own).float().mean().item() == 1.0
    else:  # fused x8 upsample + argmax == oracle head applied to the kernel's own low-res logits
        with torch.no_grad():
            low = m.lowres_logits(x.cuda())
        want = ref_head.upsample_argmax(low.float().cpu(), (128, 256), getattr(m, "HEAD_ALIGN_CORNERS", True))
        assert (labels.cpu().long() == want).float().mean().item() >= ARGMAX_AGREE_FP32


@pytest.mark.parametrize("name", NETS_BF16)
def test_bf16_default_init_within_2e_2(name):
    """The contractual recipe (BASELINE north_star): random-init weights as the reference creates them
    (torch.manual_seed(1), default initialisers), bf16 logits within 2e-2 of fp32."""
    from rtseg_b200 import nets
    torch.manual_seed(1)
    m = nets.MODEL_HUB[name](num_class=19).eval()
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    x = ref_nets.make_input(5, 1, 256, 512)
    with torch.no_grad():
        ref = ref_nets.FORWARD[name](sd, x)
        y16 = m.cuda().set_act_dtype(torch.bfloat16)(x.cuda())
    err = rel_err(y16, ref)
    agree = (y16.float().cpu().max(1)[1] == ref.max(1)[1]).float().mean().item()
    print(f"{name} default init: bf16 rel err {err:.3e}, argmax agreement {agree:.5f}, "
          f"{ref.max(1)[1].unique().numel()} classes present")
    assert err < TOL_BF16


# Argmax agreement the bf16 path commits to at full size.  Against the storage-emulating oracle (same rounding
# points, fp32 arithmetic) only accumulation order differs; against the fp32 oracle the bf16 format itself moves
# near-tied pixels.  BN-perturbed weights amplify rounding far more than real checkpoints do.
AGREE_BF16_VS_EMULATION = 0.97
AGREE_BF16_DEFAULT_INIT = 0.985


@pytest.mark.parametrize("name", NETS_BF16)
def test_bf16_full_size_2x1024x2048_vs_storage_emulating_oracle(name):
    """The benchmark's own regime — bf16 activations, 1024x2048, more than one image, every persistent kernel in its
    multi-tile steady state — against the CPU oracle: logits and the fused-head class map."""
    from rtseg_b200 import ops
    m, sd = _model(name, 4)
    m.set_act_dtype(torch.bfloat16)
    x = ref_nets.make_input(5, 2, 1024, 2048)
    with torch.no_grad():
        ref = ref_nets.FORWARD[name](sd, x)
        emu = _emulated_bf16(name, sd, x)
        y16 = m(x.cuda())
        labels = m.predict(x.cuda())
    assert ops.tc_timeout_flag() == 0
    assert tuple(y16.shape) == (2, 19, 1024, 2048) and y16.dtype == torch.bfloat16
    y16 = y16.float().cpu()
    err_emu, err16, err_fmt = rel_err(y16, emu), rel_err(y16, ref), rel_err(emu, ref)
    want = emu.max(1)[1]
    agree_logits = (y16.max(1)[1] == want).float().mean().item()
    agree_head = (labels.cpu().long() == want).float().mean().item()
    agree_fp32 = (labels.cpu().long() == ref.max(1)[1]).float().mean().item()
    print(f"{name} 2x1024x2048 bf16: vs storage-emulating oracle {err_emu:.3e}, vs fp32 {err16:.3e} (format alone "
          f"{err_fmt:.3e}); argmax agreement with the emulation: logits {agree_logits:.5f}, fused head {agree_head:.5f}; "
          f"fused head vs fp32 oracle {agree_fp32:.5f}; {want.unique().numel()} classes present")
    assert err_emu < 2.0 * err_fmt
    assert err16 < max(TOL_BF16, 1.5 * err_fmt)
    assert agree_logits >= AGREE_BF16_VS_EMULATION and agree_head >= AGREE_BF16_VS_EMULATION
    # both images, not just the first: a tile walk that loses its way shows up in the last image
    per_image = (labels.cpu().long() == want).float().mean(dim=(1, 2))
    assert per_image.min().item() >= AGREE_BF16_VS_EMULATION - 0.01


@pytest.mark.parametrize("name", NETS_BF16)
def test_bf16_default_init_full_size_within_2e_2(name):
    """The contractual recipe at the contractual size: reference-style random init (torch.manual_seed(1), default
    initialisers), one 1024x2048 image, bf16 logits within 2e-2 of the fp32 oracle and the class map agreeing."""
    from rtseg_b200 import nets, ops
    torch.manual_seed(1)
    m = nets.MODEL_HUB[name](num_class=19).eval()
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    x = ref_nets.make_input(5, 1, 1024, 2048)
    with torch.no_grad():
        ref = ref_nets.FORWARD[name](sd, x)
        m = m.cuda().set_act_dtype(torch.bfloat16)
        y16 = m(x.cuda())
        labels = m.predict(x.cuda())
    assert ops.tc_timeout_flag() == 0
    err = rel_err(y16, ref)
    want = ref.max(1)[1]
    agree = (y16.float().cpu().max(1)[1] == want).float().mean().item()
    agree_head = (labels.cpu().long() == want).float().mean().item()
    print(f"{name} default init 1x1024x2048: bf16 rel err {err:.3e}, argmax agreement logits {agree:.5f} / fused head "
          f"{agree_head:.5f}, {want.unique().numel()} classes present")
    assert err < TOL_BF16
    assert agree >= AGREE_BF16_DEFAULT_INIT and agree_head >= AGREE_BF16_DEFAULT_INIT


def test_bfloat16_module_conversion_also_works():
    """model.bfloat16() additionally rounds parameters and BN statistics; it must still run and stay close."""
    m, sd = _model("fastscnn", 2)
    x = ref_nets.make_input(3, 1, 64, 128)
    with torch.no_grad():
        ref = ref_nets.fastscnn_forward(sd, x)
        y = m.bfloat16()(x.cuda().bfloat16())
    assert y.dtype == torch.bfloat16 and rel_err(y, ref) < 1e-1


@pytest.mark.parametrize("name", NETS)
def test_full_size_1024x2048_fp32(name):
    """BASELINE size: one 1024x2048 image, fp32 path vs the CPU oracle (a few seconds of CPU)."""
    m, sd = _model(name, 4)
    x = ref_nets.make_input(5, 1, 1024, 2048)
    with torch.no_grad():
        y = m(x.cuda())
        ref = ref_nets.FORWARD[name](sd, x)
    assert tuple(y.shape) == (1, 19, 1024, 2048)
    if name == "enet":   # pooling indices: a near-tied window may flip under 1e-6 input noise and move one value by a pixel
        from tests.helpers import rel_err_robust
        # (each flipped index spreads through the receptive field of the layers behind it: a few 50 x 50 patches)
        bound, outliers = rel_err_robust(y, ref, 1e-3)
        print(f"enet 1x1024x2048 fp32: error bound met by all but 1e-3 of the elements {bound:.2e}; {outliers} of {ref.numel()} "
              f"elements above 1e-3")
        assert bound < TOL_FP32 and outliers < 1e-3 * ref.numel()
    else:
        assert rel_err(y, ref) < TOL_FP32
    pred, want = y.float().cpu().max(1)[1], ref.max(1)[1]
    assert (pred == want).float().mean().item() >= ARGMAX_AGREE_FP32
    # confusion matrix / mIoU from the same labels: bit-exact (int64) and identical IoU
    from rtseg_b200 import metrics, ops
    g = torch.Generator().manual_seed(9)
    target = torch.randint(0, 19, (1, 1024, 2048), generator=g)
    target[torch.rand((1, 1024, 2048), generator=g) < 0.1] = 255
    met = metrics.SegMetrics(19, 255).to("cuda")
    met.update(ops.argmax_channels(y, torch.uint8), target.cuda())
    cm = ref_head.confusion_matrix(pred.numpy(), target.numpy(), 19, 255)
    assert np.array_equal(met.confmat.cpu().numpy(), cm)
    assert np.array_equal(met.compute().cpu().numpy(), ref_head.iou_from_confmat(cm))


def test_fastscnn_fused_head_with_confmat_full_size():
    m, _ = _model("fastscnn", 6)
    m.set_act_dtype(torch.bfloat16)
    x = ref_nets.make_input(7, 2, 1024, 2048).cuda()
    g = torch.Generator().manual_seed(10)
    target = torch.randint(0, 19, (2, 1024, 2048), generator=g)
    target[torch.rand((2, 1024, 2048), generator=g) < 0.1] = 255
    cm = torch.zeros((19, 19), dtype=torch.int64, device="cuda")
    with torch.no_grad():
        labels = m.predict(x, target=target.cuda(), confmat=cm)
        low = m.lowres_logits(x)
    want_labels = ref_head.upsample_argmax(low.float().cpu(), (1024, 2048), True)
    assert (labels.cpu().long() == want_labels).float().mean().item() >= ARGMAX_AGREE_FP32
    assert np.array_equal(cm.cpu().numpy(), ref_head.confusion_matrix(labels.cpu().numpy(), target.numpy(), 19, 255))
    assert cm.sum().item() == (target != 255).sum().item()  # every valid pixel counted exactly once


@pytest.mark.parametrize("name", NETS)
@pytest.mark.parametrize("use_graph", [False, True])
def test_inference_engine_host_io(name, use_graph):
    import rtseg_b200
    m, _ = _model(name, 8)
    m.set_act_dtype(torch.bfloat16)
    eng = rtseg_b200.InferenceEngine(m, batch=2, height=128, width=256, use_graph=use_graph)
    assert eng.launches_per_step > 10
    xs = [ref_nets.make_input(20 + i, 2, 128, 256).pin_memory() for i in range(3)]
    with torch.no_grad():
        want = [m.predict(x.cuda()).cpu() for x in xs]
    # pipelined: submit two, collect in order
    s0 = eng.submit(xs[0])
    s1 = eng.submit(xs[1])
    out0 = eng.collect(s0).clone()
    out1 = eng.collect(s1).clone()
    out2 = eng(xs[2]).clone()
    for got, ref in zip((out0, out1, out2), want):
        assert got.dtype == torch.uint8 and torch.equal(got, ref)
    assert eng.h2d_bytes() == 2 * 3 * 128 * 256 * 4 and eng.d2h_bytes() == 2 * 128 * 256


@pytest.mark.parametrize("name", ["dabnet", "fastscnn"])
def test_two_engine_slots_replayed_concurrently_stay_deterministic(name):
    """bench.py's end-to-end phase runs two CUDA-graph slots on two streams at once (conv_tc, dw_tma and stem_tc
    CTAs of different launches share SMs).  Replay both slots a few hundred times without waiting in between: every
    class map must equal the slot's first result, the barrier watchdog must stay silent and no launch may fault.
    DABNet and Fast-SCNN are the nets whose residual convs run with a ring shorter than one unit (conv_tc.cu)."""
    import rtseg_b200
    from rtseg_b200 import ops
    m, _ = _model(name, 8)
    m.set_act_dtype(torch.bfloat16)
    eng = rtseg_b200.InferenceEngine(m, batch=4, height=512, width=1024, use_graph=True, slots=2)
    for i, s in enumerate(eng.slots):
        s.x.copy_(ref_nets.make_input(30 + i, 4, 512, 1024))
    torch.cuda.synchronize()
    first = []
    for i in range(2):
        out = eng.run_device(i)           # enqueued on the slot's own stream
        torch.cuda.synchronize()
        first.append(out.clone())
    torch.cuda.synchronize()
    mismatches = 0
    for it in range(150):
        a, b = eng.run_device(0), eng.run_device(1)       # both streams busy at once
        if it % 10 == 9:
            torch.cuda.synchronize()
            mismatches += int(not torch.equal(a, first[0])) + int(not torch.equal(b, first[1]))
    torch.cuda.synchronize()
    assert ops.tc_timeout_flag() == 0, "a kernel hit its barrier watchdog under concurrent replay"
    assert mismatches == 0


def test_modules_refuse_training_mode_and_cpu():
    from rtseg_b200 import modules
    m = modules.DWConvBNAct(16, 32, 3).cuda()   # depthwise channel multipliers have no gradient kernels: no silent fallback
    with pytest.raises(NotImplementedError, match="depthwise"):
        m(torch.randn(1, 16, 8, 8, device="cuda"))
    m = modules.DWConvBNAct(16, 16, 3).cuda()   # ... and a slice / residual call has no training form
    with pytest.raises(NotImplementedError, match="training-mode"):
        m(torch.randn(1, 32, 8, 8, device="cuda").contiguous(memory_format=torch.channels_last), x_coff=16)
    m = modules.ConvBNAct(16, 16, 3).cuda()
    with pytest.raises(TypeError):
        m.eval()(torch.randn(1, 16, 8, 8, device="cuda").half())
