"""CPU, build container only: the restatement and the drop-in classes against the REAL reference.

Skipped wherever /root/reference is absent (the GPU box).  The committed golden vectors
(tests/test_oracle_golden.py) carry the same evidence there.
"""
import os
import subprocess
import sys
import textwrap

import pytest
import torch

from oracle import ref_blocks as B
from oracle import ref_import, ref_nets
from tests.helpers import rel_err

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not ref_import.available(), reason="/root/reference not present")


def _randomise(module, seed=0):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, v in module.state_dict().items():
            if k.endswith("running_var"):
                v.copy_(0.5 + torch.rand(v.shape, generator=g))
            elif k.endswith("num_batches_tracked"):
                pass
            else:
                v.copy_(torch.randn(v.shape, generator=g) * 0.3 + (1.0 if k.endswith("1.weight") else 0.0))
    return module.eval()


def _prefixed(module, prefix):
    return {prefix + k: v for k, v in module.state_dict().items()}


CASES = [
    ("ConvBNAct", dict(in_channels=6, out_channels=9, kernel_size=(3, 1), dilation=4, act_type="prelu"),
     lambda x, sd: B.conv_bn_act(x, sd, "m.", (3, 1), 1, 4, "prelu")),
    ("ConvBNAct", dict(in_channels=6, out_channels=4, kernel_size=3, stride=2, bias=True, act_type="leakyrelu",
                       negative_slope=0.2),
     lambda x, sd: B.conv_bn_act(x, sd, "m.", 3, 2, 1, "leakyrelu", negative_slope=0.2)),
    ("DWConvBNAct", dict(in_channels=6, out_channels=12, kernel_size=(1, 3), dilation=2, act_type="relu6"),
     lambda x, sd: B.dw_conv_bn_act(x, sd, "m.", (1, 3), 1, 2, "relu6")),
    ("PWConvBNAct", dict(in_channels=6, out_channels=5, act_type="hardswish"),
     lambda x, sd: B.pw_conv_bn_act(x, sd, "m.", "hardswish")),
    ("DSConvBNAct", dict(in_channels=6, out_channels=8, kernel_size=3, stride=2, act_type="selu"),
     lambda x, sd: B.ds_conv_bn_act(x, sd, "m.", 3, 2, 1, "selu")),
    ("DeConvBNAct", dict(in_channels=6, out_channels=5, act_type="elu"),
     lambda x, sd: B.deconv_bn_act(x, sd, "m.", 2, "elu")),
    ("PyramidPoolingModule", dict(in_channels=8, out_channels=6, act_type="relu", bias=True),
     lambda x, sd: B.pyramid_pooling(x, sd, "m.", "relu")),
]


@pytest.mark.parametrize("cls,kwargs,fn", CASES)
def test_block_restatement_matches_reference_module(cls, kwargs, fn):
    ref_modules = __import__("importlib").import_module("models.modules") if ref_import.reference_models() else None
    m = _randomise(getattr(ref_modules, cls)(**kwargs))
    x = torch.randn(2, kwargs["in_channels"], 12, 20, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        ref = m(x)
        out = fn(x, _prefixed(m, "m."))
    assert out.shape == ref.shape and rel_err(out, ref) < 1e-6


def test_initial_block_and_nonbt1d_match_reference():
    models = ref_import.reference_models()
    enet = __import__("importlib").import_module("models.enet")
    erf = __import__("importlib").import_module("models.erfnet")
    x = torch.randn(1, 8, 16, 24, generator=torch.Generator().manual_seed(4))
    ib = _randomise(enet.InitialBlock(8, 20, "prelu"))
    nb = _randomise(erf.NonBt1DBlock(8, dilation=3, act_type="relu"), 1)
    with torch.no_grad():
        assert rel_err(B.initial_block(x, _prefixed(ib, "m."), "m.", "prelu"), ib(x)) < 1e-6
        assert rel_err(ref_nets.non_bt_1d(x, _prefixed(nb, "m."), "m.", 3, "relu"), nb(x.clone())) < 1e-6
    assert models is not None


def test_drop_in_classes_keep_reference_state_dict_keys():
    import rtseg_b200.modules as ours
    theirs = __import__("importlib").import_module("models.modules")
    builds = {
        "ConvBNAct": dict(in_channels=4, out_channels=6, kernel_size=(1, 3), dilation=2, bias=True, act_type="prelu"),
        "DWConvBNAct": dict(in_channels=4, out_channels=8, kernel_size=3, stride=2),
        "PWConvBNAct": dict(in_channels=4, out_channels=6),
        "DSConvBNAct": dict(in_channels=4, out_channels=6, kernel_size=3),
        "DeConvBNAct": dict(in_channels=4, out_channels=6),
        "PyramidPoolingModule": dict(in_channels=8, out_channels=6, act_type="relu"),
        "SegHead": dict(in_channels=4, num_class=19, act_type="relu"),
        "Activation": dict(act_type="prelu"),
    }
    ref_import.reference_models()
    for name, kw in builds.items():
        a, b = getattr(ours, name)(**kw), getattr(theirs, name)(**kw)
        sa = [(k, tuple(v.shape)) for k, v in a.state_dict().items()]
        sb = [(k, tuple(v.shape)) for k, v in b.state_dict().items()]
        assert sa == sb, name
    for fn in ("conv1x1", "conv3x3"):
        a, b = getattr(ours, fn)(4, 6, 2, True), getattr(theirs, fn)(4, 6, 2, True)
        assert (a.kernel_size, a.stride, a.padding) == (b.kernel_size, b.stride, b.padding)
        assert list(a.state_dict()) == list(b.state_dict())
    x = torch.arange(2 * 8 * 3 * 3, dtype=torch.float32).reshape(2, 8, 3, 3)
    assert torch.equal(ours.channel_shuffle(x, 2), theirs.channel_shuffle(x, 2))
    assert set(n for n in dir(theirs) if not n.startswith("_") and n not in ("torch", "nn", "F")) <= set(dir(ours))


def test_reference_model_files_build_on_injected_modules():
    """Unchanged reference models/*.py pick up the rtseg classes via sys.modules (needs a fresh interpreter)."""
    code = textwrap.dedent("""
        import sys, types
        sys.path.insert(0, %r)
        from rtseg_b200 import inject, modules as ours
        models = inject.install('/root/reference')
        import torch
        for name in ['erfnet', 'fastscnn', 'dabnet', 'lednet', 'enet', 'bisenetv2', 'ddrnet']:
            cfg = types.SimpleNamespace(model=name, num_class=19, use_aux=False, use_detail_head=False)
            m = models.get_model(cfg)
            fused = [x for x in m.modules() if type(x).__module__ == ours.__name__]
            assert len(fused) > 5, name
            print(name, sum(p.numel() for p in m.parameters()), len(fused))
        import models.erfnet as e
        assert e.ConvBNAct is ours.ConvBNAct
    """ % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    counts = dict(line.split()[:2] for line in r.stdout.strip().splitlines())
    assert counts["erfnet"] == "2065458" and counts["fastscnn"] == "1023769" and counts["dabnet"] == "750269"
    assert counts["lednet"] == "1461383" and counts["enet"] == "365100" and counts["bisenetv2"] == "2272688"
