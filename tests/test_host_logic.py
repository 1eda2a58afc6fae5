"""CPU: host-side logic — C-ABI surface, state-dict layout, BN folding, error behaviour, sharding."""
import copy
import ctypes
import os
import re
import types

import numpy as np
import pytest
import torch
import torch.nn as nn

import rtseg_b200
from oracle import ref_blocks, ref_head, ref_nets
from rtseg_b200 import _lib, metrics, modules, nets, ops, parallel
from tests.helpers import NETS, load_specs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "rtseg.h")).read()
    declared = set(re.findall(r"\b(rtseg_[a-z0-9_]+)\s*\(", header))
    declared -= {"rtseg_status", "rtseg_dtype", "rtseg_act", "rtseg_res_mode"}
    assert len(declared) >= 25
    lib = _lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"librtseg.so does not export {name}"
    assert declared == set(_lib.PROTOTYPES), "ctypes prototypes and include/rtseg.h disagree"
    assert lib.rtseg_version() == 100
    assert lib.rtseg_ohem_workspace_bytes() % 256 == 0


def test_conv_desc_matches_header_layout():
    header = open(os.path.join(ROOT, "include", "rtseg.h")).read()
    body = header[header.index("typedef struct rtseg_conv_desc {"):header.index("} rtseg_conv_desc;")]
    fields = re.findall(r"\b(?:int32_t|float)\s+([a-z0-9_, ]+);", body)
    names = [n.strip() for grp in fields for n in grp.split(",")]
    assert names == [f[0] for f in _lib.ConvDesc._fields_]
    assert ctypes.sizeof(_lib.ConvDesc) == 4 * len(names)


def test_out_hw_and_validation_without_gpu():
    lib = _lib.load()
    d = _lib.ConvDesc(1, 128, 256, 64, 64, 3, 1, 1, 16, 16, 0, 1, 1, 1, 0, 0, 0, 0, 64, 0, 64, 0, 0, 0)
    ho, wo = ctypes.c_int32(), ctypes.c_int32()
    assert lib.rtseg_out_hw(ctypes.byref(d), ctypes.byref(ho), ctypes.byref(wo)) == 0
    assert (ho.value, wo.value) == (128, 256)
    d.kh, d.pad_h, d.dil, d.stride = 3, 1, 1, 2
    d.kw, d.pad_w = 3, 1
    lib.rtseg_out_hw(ctypes.byref(d), ctypes.byref(ho), ctypes.byref(wo))
    assert (ho.value, wo.value) == (64, 128)
    # tensor-core eligibility is pure host logic
    assert lib.rtseg_conv_tc_supported(ctypes.byref(d)) == 1
    d.cin = 13
    d.x_ctotal = 13
    assert lib.rtseg_conv_tc_supported(ctypes.byref(d)) == 0
    # invalid descriptor -> negative status and a message, no crash
    d.cin = 0
    assert lib.rtseg_conv_dense(ctypes.byref(d), None, None, None, None, None, None, None, None) < 0
    assert b"extent" in lib.rtseg_last_error()


@pytest.mark.parametrize("name", NETS)
def test_state_dict_layout_equals_reference(name):
    spec = load_specs()[name]
    model = nets.MODEL_HUB[name](num_class=19)
    ours = [(k, tuple(v.shape), str(v.dtype).replace("torch.", "")) for k, v in model.state_dict().items()]
    assert ours == spec["spec"]
    assert sum(p.numel() for p in model.parameters()) == spec["n_params"]
    # a reference-shaped state dict loads strictly, caches never leak into it
    sd = ref_nets.make_state(spec["spec"], 1)
    model.load_state_dict(sd, strict=True)
    assert not any("rtseg" in k for k in model.state_dict())


def test_fold_matches_unfused_arithmetic():
    torch.manual_seed(0)
    m = modules.PWConvBNAct(8, 5, act_type="prelu")
    with torch.no_grad():
        m[1].running_mean.normal_()
        m[1].running_var.uniform_(0.5, 1.5)
        m[1].weight.normal_()
        m[1].bias.normal_()
    m.eval()
    f, post = ops.fold_conv(m[0], m[1], m[2].activation)
    assert post is None and f.kind == "dense" and f.act == _lib.ACT_CODES["prelu"]
    x = torch.randn(2, 8, 4, 4)
    ref = m[1](m[0](x))
    w = f.w_f32.reshape(8, 5)  # [1][1][cin][cout]
    y = torch.einsum("nchw,co->nohw", x, w) * f.scale.view(1, -1, 1, 1) + f.shift.view(1, -1, 1, 1)
    assert torch.allclose(y, ref, atol=1e-5)
    assert f.slope.shape == (5,) and torch.allclose(f.slope, torch.full((5,), 0.25))
    # K padded to 16 with zeros (the tensor-core kernel's TMA map stops at the last real channel)
    assert f.w_tc.shape == (1, 16, 16) and (f.w_tc[:, :, 8:] == 0).all() and (f.w_tc[:, 5:] == 0).all()
    f3, _ = ops.fold_conv(modules.ConvBNAct(3, 8, 3).eval()[0])
    assert f3.w_tc is None  # image stems stay on the fp32 stem / CUDA-core kernels
    m2 = modules.ConvBNAct(32, 19, (3, 1), dilation=4).eval()
    f2, _ = ops.fold_conv(m2[0], m2[1], m2[2].activation)
    assert f2.w_tc.shape == (3, 32, 32) and f2.w_tc.dtype == torch.bfloat16
    assert (f2.w_tc[:, 19:] == 0).all() and (f2.pad_h, f2.pad_w, f2.dil) == (4, 0, 4)


def test_fold_cache_invalidation_and_deepcopy():
    m = modules.ConvBNAct(4, 4, 3).eval()
    a = m._folded()
    assert m._folded() is a
    with torch.no_grad():
        m[0].weight.add_(1.0)
    b = m._folded()
    assert b is not a
    m.load_state_dict(copy.deepcopy(m.state_dict()))
    assert m._folded() is not b
    c = copy.deepcopy(m)  # EMA path, reference utils/model_ema.py:20
    assert "_rtseg_cache" not in c.__dict__ and list(c.state_dict()) == list(m.state_dict())


def test_error_behaviour_mirrors_reference():
    with pytest.raises(NotImplementedError):
        modules.Activation("swishish")  # reference models/modules.py:125-126
    with pytest.raises(TypeError):
        modules.ConvBNAct(3, 3, kernel_size=3.0)
    with pytest.raises(TypeError):  # kwargs reach the activation only (App. B.2: regseg's groups= crashes)
        modules.ConvBNAct(3, 3, 3, groups=2)
    m = modules.ConvBNAct(3, 8).eval()
    with pytest.raises(RuntimeError, match="no CPU implementation"):
        m(torch.randn(1, 3, 8, 8))
    with pytest.raises(AssertionError):
        nets.InitialBlock(16, 16, "relu")  # reference models/enet.py:41
    cfg = types.SimpleNamespace(model="nope", num_class=19, use_aux=False, use_detail_head=False)
    with pytest.raises(NotImplementedError):
        nets.get_model(cfg)


def test_activation_hub_has_the_reference_names():
    assert set(modules.Activation._HUB) == {
        "relu", "relu6", "leakyrelu", "prelu", "celu", "elu", "hardswish", "hardtanh", "gelu", "glu", "selu", "silu",
        "sigmoid", "softmax", "tanh", "none"}
    assert isinstance(modules.Activation("PReLU").activation, nn.PReLU)
    assert ops.act_spec(nn.LeakyReLU(0.2))[:2] == (_lib.ACT_CODES["leakyrelu"], pytest.approx(0.2))
    assert ops.act_spec(nn.GLU())[4] is not None


def test_iou_host_function_equals_oracle():
    rng = np.random.default_rng(1)
    cm = rng.integers(0, 1000, size=(19, 19)).astype(np.int64)
    cm[5] = 0
    cm[:, 5] = 0
    ours = metrics.iou_from_confmat(torch.from_numpy(cm)).numpy()
    np.testing.assert_array_equal(ours, ref_head.iou_from_confmat(cm))
    assert ours[5] == 0.0


def test_ohem_threshold_value():
    assert rtseg_b200.OhemCELoss(0.7).thresh == pytest.approx(0.35667494, abs=1e-7)


def test_shard_range_partitions_batch():
    for total in (1, 7, 16, 32):
        for world in (1, 2, 4, 8):
            spans = [parallel.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


# ------------------------------------------------------------------ convert() / fused_head() host logic (no GPU) ----
def test_convert_reclasses_reference_layout_in_place():
    import rtseg_b200
    from rtseg_b200 import modules, nets
    from tests import ref_layout
    for name, counts in (("erfnet", {"NonBt1DBlock": 17, "InitialBlock": 3, "Conv2d": 17, "ERFNet": 1}),
                         ("fastscnn", {"InvertedResidual": 8, "FeatureFusionModule": 1, "LearningToDownsample": 1, "FastSCNN": 1})):
        m = ref_layout.LAYOUTS[name](num_class=19)
        before = {k: v.data_ptr() for k, v in m.state_dict().items()}
        assert rtseg_b200.convert(m) is m
        assert isinstance(m, nets.MODEL_HUB[name]) and m.num_class == 19 and hasattr(m, "predict")
        assert {k: v.data_ptr() for k, v in m.state_dict().items()} == before      # same tensors, same keys, same order
        assert m._rtseg_converted == counts
        assert rtseg_b200.convert(m)._rtseg_converted == {}                         # idempotent
    assert isinstance(m.feature_fusion.higher_res_conv, modules.Conv2d)


def _defused(model):
    """A copy of a fused net whose block / net classes are same-named stand-ins from outside rtseg_b200 (no forward of
    ours), and whose bare convolutions are plain nn.Conv2d: the attribute layout the reference's model files produce."""
    import copy
    import sys
    import torch.nn as nn
    import rtseg_b200
    from rtseg_b200 import modules
    conv = sys.modules[rtseg_b200.convert.__module__]
    leaves = (modules.ConvBNAct, modules.DWConvBNAct, modules.PWConvBNAct, modules.DSConvBNAct, modules.DeConvBNAct, modules.Activation)
    m = copy.deepcopy(model)
    for mod in m.modules():
        cls = type(mod)
        if cls is modules.Conv2d:
            mod.__class__ = nn.Conv2d
        elif cls.__module__.startswith("rtseg_b200") and not isinstance(mod, leaves) and (cls.__name__ in conv._BLOCKS or cls.__name__ in conv._NETS):
            mod.__class__ = type(cls.__name__, (nn.Sequential if isinstance(mod, nn.Sequential) else nn.Module,), {})
    return m


def test_convert_reclasses_the_round2_nets_from_their_attribute_layout():
    """ENet / BiSeNetv2 / DDRNet (reference models/enet.py, bisenetv2.py, ddrnet.py): every block and the net itself come
    back as the fused classes, same tensors, same keys."""
    import rtseg_b200
    from rtseg_b200 import nets
    for name in ("enet", "bisenetv2", "ddrnet"):
        fused = nets.MODEL_HUB[name](num_class=19)
        m = _defused(fused)
        assert not type(m).__module__.startswith("rtseg_b200")   # (pure containers — Stage2, BottleNeck1 ... — keep their class)
        before = {k: v.data_ptr() for k, v in m.state_dict().items()}
        rtseg_b200.convert(m)
        assert [type(x) for x in m.modules()] == [type(x) for x in fused.modules()]
        assert {k: v.data_ptr() for k, v in m.state_dict().items()} == before and m.num_class == 19
        assert rtseg_b200.convert(m)._rtseg_converted == {}


def test_convert_on_the_real_reference_models():
    """Build container only: the reference's OWN BiSeNetv2 / DDRNet / ENet objects (its classes, its leaf modules) are
    re-classed in place; what stays a reference class is a pure container whose forward only calls its children."""
    import pytest
    import rtseg_b200
    from oracle import ref_import
    from rtseg_b200 import nets
    if not ref_import.available():
        pytest.skip("needs /root/reference")
    for name in ("ddrnet", "bisenetv2", "enet"):
        m = ref_import.reference_model(name, 19)
        keys = list(m.state_dict().keys())
        rtseg_b200.convert(m)
        assert type(m) is nets.MODEL_HUB[name] and list(m.state_dict().keys()) == keys
        left = {type(x).__name__ for x in m.modules() if not type(x).__module__.startswith(("torch", "rtseg_b200"))}
        assert left <= {"Stage2", "Stage3", "BottleNeck1", "BottleNeck23", "BottleNeck45"}, left


def test_convert_leaves_lookalike_classes_alone():
    """A class that only shares a NAME with a known block (other model files of the reference define their own
    InitialBlock / InvertedResidual variants) must keep its forward."""
    import torch.nn as nn
    import rtseg_b200

    class NonBt1DBlock(nn.Module):   # same name, different structure
        def __init__(self):
            super().__init__()
            self.conv = nn.Sequential(nn.Conv2d(4, 4, 3, padding=1), nn.ReLU())

        def forward(self, x):
            return self.conv(x)

    m = nn.Sequential(NonBt1DBlock())
    rtseg_b200.convert(m)
    assert type(m[0]) is NonBt1DBlock and type(m[0].conv[0]) is nn.Conv2d and m._rtseg_converted == {}


def test_deferred_logits_routes_max_to_the_fused_head(monkeypatch):
    """fused_head() logic with the two kernels stubbed by their torch definitions (CPU): max / argmax over dim 1 never
    materialise the logits, every other use does, and the values equal plain interpolate + max."""
    import torch.nn.functional as F
    import sys
    import rtseg_b200
    cv = sys.modules[rtseg_b200.convert.__module__]      # the module (the package re-exports the function under the same name)
    calls = {"argmax": 0, "bilinear": 0}

    def fake_argmax(low, size, align, out_dtype=torch.int64, **kw):
        calls["argmax"] += 1
        return F.interpolate(low, size, mode="bilinear", align_corners=align).max(1)[1].to(out_dtype)

    def fake_bilinear(low, size, align, **kw):
        calls["bilinear"] += 1
        return F.interpolate(low, size, mode="bilinear", align_corners=align)
    monkeypatch.setattr(cv.ops, "upsample_argmax", fake_argmax)
    monkeypatch.setattr(cv.ops, "upsample_bilinear", fake_bilinear)
    low = torch.randn(2, 5, 4, 6)
    want = F.interpolate(low, (16, 24), mode="bilinear", align_corners=True)
    d = cv.DeferredLogits(low, (16, 24), True)
    assert tuple(d.shape) == (2, 5, 16, 24) and d.size()[2:] == (16, 24) and d.dim() == 4 and d.dtype == torch.float32
    assert torch.equal(d.max(dim=1)[1], want.max(1)[1]) and torch.equal(torch.max(d, 1)[1], want.max(1)[1])
    assert torch.equal(d.argmax(dim=1), want.argmax(1)) and torch.equal(d.max(1, keepdim=True)[1], want.max(1, keepdim=True)[1])
    assert calls == {"argmax": 4, "bilinear": 0}
    values, indices = d.max(dim=1)
    assert torch.equal(values, want.max(1)[0]) and calls["bilinear"] == 1           # values need the logits
    assert torch.equal(d + 1, want + 1) and torch.equal(torch.softmax(d, 1), torch.softmax(want, 1))
    assert calls["bilinear"] == 1                                                    # materialised once, then cached
    # the mode only defers CUDA tensors: on the CPU F.interpolate stays itself
    with cv.fused_head():
        assert not isinstance(F.interpolate(low, (16, 24), mode="bilinear", align_corners=True), cv.DeferredLogits)


def test_pyramid_branch_weights_are_stacked_once_and_follow_the_parameters():
    """PyramidPoolingModule (reference models/modules.py:134-158): the one-launch branch kernel reads the four conv1x1
    weights as ONE fp32 [4, hid, c] tensor; the stack is cached on parameter versions (like the folded convolutions) and
    never enters state_dict."""
    import copy
    import torch
    from rtseg_b200 import modules
    m = modules.PyramidPoolingModule(16, 16, "relu")
    convs = [m.stage1[1], m.stage2[1], m.stage3[1], m.stage4[1]]
    w, b = m._stacked_branch_weights(convs)
    assert tuple(w.shape) == (4, 4, 16) and w.dtype == torch.float32 and w.is_contiguous() and b is None
    for i, cv in enumerate(convs):
        assert torch.equal(w[i], cv.weight.detach().reshape(4, 16))
    assert m._stacked_branch_weights(convs)[0] is w                      # cached
    with torch.no_grad():
        convs[2].weight.mul_(2.0)                                         # in-place update bumps the version
    w2, _ = m._stacked_branch_weights(convs)
    assert w2 is not w and torch.equal(w2[2], convs[2].weight.detach().reshape(4, 16))
    assert all("_branch" not in k for k in m.state_dict())
    sd = {k: (torch.randn_like(v) if v.is_floating_point() else v.clone()) for k, v in m.state_dict().items()}
    m.load_state_dict(sd)
    assert torch.equal(m._stacked_branch_weights(convs)[0][0], sd["stage1.1.weight"].reshape(4, 16))
    copy.deepcopy(m)                                                       # the cache does not break module copies
