"""GPU parity tests in the regime bench.py times: persistent kernels in their multi-tile steady state.

The kernel tests elsewhere are small (a CTA sees one or two tiles).  Here every case is sized so that
  * `conv_tc_kernel` — one persistent CTA per SM — walks at least 8 unit tiles per CTA (asserted through
    rtseg_conv_tc_plan_info), i.e. the mbarrier ring wraps several times, both TMEM accumulator stages and both
    staging buffers are reused, and the incremental (image, row, column) tile walk crosses image boundaries;
  * `dw_tma_kernel` is above its 65,536-output-pixel gate (asserted through rtseg_dw_tma_launch_count);
and the result is compared with the CPU oracle (oracle/ref_blocks.py, the restatement of reference
models/modules.py:46-108 and models/erfnet.py:62-82).  Each launch is also repeated and required to be bit-identical
run to run: a barrier-phase race shows up as a run that differs.

Tolerance: bf16 <= 2e-2 of the largest reference magnitude (BASELINE.json north_star).
"""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from oracle import ref_blocks as B
from tests.helpers import TOL_BF16, rel_err, seeded

pytestmark = pytest.mark.gpu

MIN_UNITS_PER_CTA = 8
REPEATS = 6


def _randomise(module, seed=0):
    from tests.test_gpu_kernels import _randomise as r
    return r(module, seed)


def _sd(module, prefix="m."):
    return {prefix + k: v.detach().clone().float().cpu() for k, v in module.state_dict().items()}


def _plan(conv, n, h, w, x_ct=None, y_ct=None, res=False):
    """The plan rtseg_conv_tc would use for this nn.Conv2d on an [n, cin, h, w] bf16 input."""
    from rtseg_b200 import _lib
    kh, kw = conv.kernel_size
    d = _lib.ConvDesc(n, h, w, conv.in_channels, conv.out_channels, kh, kw, conv.stride[0], conv.dilation[0],
                      conv.padding[0], conv.padding[1], 1, _lib.BF16, 1, 0.0, 0.0, 1 if res else 0, 0,
                      x_ct or conv.in_channels, 0, y_ct or conv.out_channels, 0, conv.out_channels if res else 0, 0)
    info = (C.c_int32 * 12)()
    _lib.check(_lib.load().rtseg_conv_tc_plan_info(C.byref(d), info), "rtseg_conv_tc_plan_info")
    keys = ("kb", "bn", "n_tiles", "ms", "stages", "out_bufs", "share", "units", "grid_x", "smem", "tmem_cols", "res_boxes")
    return dict(zip(keys, list(info)))


def _batch_for(conv, h, w, res=False, start=4, limit=64):
    """Smallest batch >= start for which every persistent CTA walks MIN_UNITS_PER_CTA unit tiles."""
    n = start
    while n <= limit:
        pl = _plan(conv, n, h, w, res=res)
        if pl["units"] >= MIN_UNITS_PER_CTA * pl["grid_x"]:
            return n, pl
        n += 1
    raise AssertionError(f"no batch <= {limit} reaches {MIN_UNITS_PER_CTA} units per CTA: {pl}")


def _repeat_identical(fn):
    """fn() launches the kernel and returns its output tensor; all REPEATS outputs must be bit-identical."""
    from rtseg_b200 import ops
    first = fn().clone()
    for i in range(1, REPEATS):
        again = fn()
        assert torch.equal(first, again), f"run {i} differs from run 0: the launch is not deterministic"
    assert ops.tc_timeout_flag() == 0, "a tensor-core kernel hit its barrier watchdog"
    return first


# ERFNet's dense asymmetric convs at their own resolutions (reference models/erfnet.py:15-33, 62-82):
# (channels, h, w) of the stage, kernel, dilation, residual (the block's fourth conv: + x -> BN -> act)
ERFNET_CONVS = []
for _c, _h, _w, _dils in ((16, 512, 1024, (1,)), (64, 256, 512, (1,)), (128, 128, 256, (2, 4, 8, 16))):
    ERFNET_CONVS += [(_c, _h, _w, (3, 1), 1, False), (_c, _h, _w, (1, 3), 1, False)]
    for _d in _dils:
        if _d != 1:
            ERFNET_CONVS.append((_c, _h, _w, (3, 1), _d, False))
        ERFNET_CONVS.append((_c, _h, _w, (1, 3), _d, True))


@pytest.mark.parametrize("c,h,w,k,dil,res", ERFNET_CONVS,
                         ids=[f"c{c}-{k[0]}x{k[1]}-d{d}{'-res' if r else ''}" for c, h, w, k, d, r in ERFNET_CONVS])
def test_conv_tc_steady_state_erfnet_layers(c, h, w, k, dil, res):
    from rtseg_b200 import _lib, modules, nets, ops
    if res:   # the block tail: Conv2d (no BN of its own) + x -> BatchNorm -> ReLU in the epilogue
        blk = _randomise(nets.NonBt1DBlock(c, dilation=dil), 31).cuda()
        conv = blk.conv[3]
    else:
        m = _randomise(modules.ConvBNAct(c, c, k, dilation=dil, act_type="relu"), 32).cuda()
        conv = m[0]
    n, pl = _batch_for(conv, h, w, res=res)
    assert pl["units"] >= MIN_UNITS_PER_CTA * pl["grid_x"]
    print(f"plan c={c} {k} d={dil} res={res}: n={n} {pl} -> {pl['units'] / pl['grid_x']:.1f} units per CTA")
    x = seeded((n, c, h, w), 33)
    xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
    xq = x.bfloat16().float()      # the oracle sees the same bf16-rounded input the kernel reads
    if res:
        r = seeded((n, c, h, w), 34)
        rb = ops.to_nhwc(r.to("cuda", torch.bfloat16))
        sd = _sd(blk)
        before = _lib.launch_count()
        with torch.no_grad():
            t = F.conv2d(xq, sd["m.conv.3.weight"], None, 1, (0, dil), dil)
            ref = F.relu(B.batch_norm(t + r.bfloat16().float(), sd, "m.bn_act.0."))
            out = _repeat_identical(lambda: conv(xb, res=rb, res_mode=ops.RES_PRE, bn=blk.bn_act[0], act=blk.bn_act[1].activation))
    else:
        before = _lib.launch_count()
        with torch.no_grad():
            ref = B.conv_bn_act(xq, _sd(m), "m.", k, 1, dil, "relu")
            out = _repeat_identical(lambda: m(xb))
    assert _lib.launch_count() - before == REPEATS
    assert out.dtype == torch.bfloat16 and tuple(out.shape) == tuple(ref.shape)
    err = rel_err(out, ref)
    assert err < TOL_BF16, f"steady-state conv_tc off by {err:.3e}"
    # a stale ring stage or accumulator corrupts whole tiles: beyond the max-abs bound, no 128-pixel tile may be off
    diff = (out.float().cpu() - ref).abs().amax(dim=1)          # [n, h, w]
    assert (diff > 0.05 * ref.abs().max()).sum().item() == 0


# Layers whose ring is SHORTER than one unit's boxes (K blocks + residual boxes): Fast-SCNN's inverted-residual
# projections (reference models/fastscnn.py:105-121) and DABNet's layer8 main conv (models/dabnet.py:24).
SHORT_RING = [
    # cin, cout, kernel, stride, res_mode, h, w, ring shorter than a unit?
    (384, 64, 1, 1, "post", 128, 256, False),   # 6 K blocks + 1 residual box in an 8-stage ring: fits
    (576, 96, 1, 1, "post", 64, 128, True),     # 9 + 3 boxes, 5 stages
    (768, 128, 1, 1, "post", 64, 128, True),    # 12 + 1 boxes, 6 stages
    (128, 128, 3, 2, "pre", 256, 512, True),    # 18 + 1 boxes, 3 stages
]


@pytest.mark.parametrize("cin,cout,k,stride,mode,h,w,short", SHORT_RING)
def test_conv_tc_steady_state_residual_ring_shorter_than_a_unit(cin, cout, k, stride, mode, h, w, short):
    from rtseg_b200 import modules, ops
    m = _randomise(modules.ConvBNAct(cin, cout, k, stride, act_type="none" if mode == "post" else "prelu"), 35).cuda()
    n, pl = _batch_for(m[0], h, w, res=True)
    kb_per_unit = (m[0].in_channels // pl["kb"]) * (1 if pl["share"] else k * k)
    print(f"plan {cin}->{cout} k={k} s={stride}: n={n} {pl}; boxes per unit {kb_per_unit} + {pl['res_boxes']}, ring {pl['stages']}")
    assert pl["res_boxes"] > 0, "the residual does not travel through the ring"
    if short:
        assert pl["stages"] < kb_per_unit + pl["res_boxes"], "case no longer exercises a ring shorter than a unit"
    ho, wo = h // stride, w // stride
    x, r = seeded((n, cin, h, w), 36), seeded((n, cout, ho, wo), 37)
    xb, rb = ops.to_nhwc(x.to("cuda", torch.bfloat16)), ops.to_nhwc(r.to("cuda", torch.bfloat16))
    sd = _sd(m)
    with torch.no_grad():
        conv = F.conv2d(x.bfloat16().float(), sd["m.0.weight"], None, stride, (k - 1) // 2)
        rq = r.bfloat16().float()
        if mode == "pre":
            ref = F.prelu(B.batch_norm(conv + rq, sd, "m.1."), sd["m.2.activation.weight"])
        else:
            ref = B.batch_norm(conv, sd, "m.1.") + rq
        out = _repeat_identical(lambda: m(xb, res=rb, res_mode=ops.RES_PRE if mode == "pre" else ops.RES_POST))
    assert rel_err(out, ref) < TOL_BF16


def test_conv_tc_steady_state_downsampler_and_deconv():
    """ERFNet's stride-2 conv branches (parity views) and transposed convs (four accumulators) at batch 4."""
    from rtseg_b200 import modules, ops
    for cin, cout, h, w in ((16, 48, 512, 1024), (64, 64, 256, 512)):
        m = _randomise(modules.ConvBNAct(cin, cout, 3, 2, act_type="relu"), 38).cuda()
        n, pl = _batch_for(m[0], h, w)
        x = seeded((n, cin, h, w), 39)
        xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
        with torch.no_grad():
            ref = B.conv_bn_act(x.bfloat16().float(), _sd(m), "m.", 3, 2, 1, "relu")
            out = _repeat_identical(lambda: m(xb))
        assert rel_err(out, ref) < TOL_BF16, (cin, cout)
    for cin, cout, h, w in ((128, 64, 128, 256), (64, 16, 256, 512)):
        m = _randomise(modules.DeConvBNAct(cin, cout, act_type="relu"), 40).cuda()
        x = seeded((4, cin, h, w), 41)
        xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
        with torch.no_grad():
            ref = B.deconv_bn_act(x.bfloat16().float(), _sd(m), "m.", 2, "relu")
            out = _repeat_identical(lambda: m(xb))
        assert rel_err(out, ref) < TOL_BF16, (cin, cout)


def test_labels_head_steady_state_full_resolution():
    """ERFNet's layer23 (transposed conv 16 -> 19 with the argmax inside the kernel) on 4 x 512 x 1024 inputs: labels
    equal the argmax of the oracle's logits wherever the oracle's top-2 margin exceeds bf16 resolution."""
    from rtseg_b200 import modules, ops
    m = _randomise(modules.DeConvBNAct(16, 19, act_type="relu"), 42).cuda()
    x = seeded((4, 16, 512, 1024), 43)
    xb = ops.to_nhwc(x.to("cuda", torch.bfloat16))
    labels = torch.empty((4, 1024, 2048), dtype=torch.uint8, device="cuda")

    def run():
        m(xb, labels=labels, want_logits=False)
        return labels
    with torch.no_grad():
        ref = B.deconv_bn_act(x.bfloat16().float(), _sd(m), "m.", 2, "relu")
        got = _repeat_identical(run).cpu().long()
    top2 = ref.topk(2, dim=1)[0]
    clear = (top2[:, 0] - top2[:, 1]) > 2e-2 * ref.abs().max()
    assert clear.float().mean().item() > 0.5
    assert (got[clear] == ref.max(1)[1][clear]).all()


# ------------------------------------------------------------------------------ dw_tma ----
DW_TMA_CASES = [
    # c, stride, n, h, w, act, (x_ctotal, x_coff, y_ctotal, y_coff)
    (32, 2, 2, 512, 1024, "relu", None),          # Fast-SCNN DS1.dw at its own size (CB 32)
    (48, 2, 4, 256, 512, "relu", None),           # DS2.dw (CB 16)
    (384, 2, 8, 128, 256, "relu", None),          # first inverted residual (CB 64, stride 2)
    (128, 1, 2, 128, 256, "relu", None),          # classifier / fusion depthwise (CB 64, stride 1)
    (96, 1, 3, 131, 259, "relu6", None),          # CB 32, ragged H and W
    (48, 1, 5, 120, 136, "prelu", None),          # CB 16, slope epilogue
    (64, 2, 3, 301, 517, "hardswish", None),      # odd sizes under stride 2, general epilogue
    (32, 1, 2, 256, 192, "relu", (96, 64, 80, 16)),   # channel-slice input and output
    (64, 2, 2, 384, 512, "relu", (136, 72, 64, 0)),
]


@pytest.mark.parametrize("c,stride,n,h,w,act,slices", DW_TMA_CASES)
def test_dw_tma_above_the_size_gate(c, stride, n, h, w, act, slices):
    from rtseg_b200 import _lib, modules, ops
    lib = _lib.load()
    m = _randomise(modules.DWConvBNAct(c, c, 3, stride, act_type=act), 44).cuda()
    x = seeded((n, c, h, w), 45)
    ho, wo = (h - 1) // stride + 1, (w - 1) // stride + 1
    assert n * ho * wo >= 65536
    with torch.no_grad():
        ref = B.conv_bn_act(x.bfloat16().float(), _sd(m), "m.", 3, stride, 1, act, groups=c)
    kw = {}
    if slices is None:
        xin = ops.to_nhwc(x.to("cuda", torch.bfloat16))
    else:
        x_ct, x_co, y_ct, y_co = slices
        xin = torch.full((n, x_ct, h, w), float("nan"), dtype=torch.bfloat16, device="cuda").contiguous(memory_format=torch.channels_last)
        xin[:, x_co:x_co + c] = x.to("cuda", torch.bfloat16)
        out_buf = torch.zeros((n, y_ct, ho, wo), dtype=torch.bfloat16, device="cuda").contiguous(memory_format=torch.channels_last)
        kw = dict(x_coff=x_co, out=out_buf, y_coff=y_co)
    before = lib.rtseg_dw_tma_launch_count()
    with torch.no_grad():
        out = _repeat_identical(lambda: m(xin, **kw))
    assert lib.rtseg_dw_tma_launch_count() - before == REPEATS, "the layer did not take dw_tma_kernel"
    if slices is not None:
        x_ct, x_co, y_ct, y_co = slices
        mask = torch.ones(y_ct, dtype=torch.bool)
        mask[y_co:y_co + c] = False
        assert (out[:, mask.cuda()] == 0).all(), "the kernel wrote outside its channel slice"
        out = out[:, y_co:y_co + c]
    assert tuple(out.shape) == tuple(ref.shape)
    assert rel_err(out, ref) < TOL_BF16
