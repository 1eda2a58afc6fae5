"""CPU: the oracle restatement against vectors produced by the real reference (tests/golden)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import ref_head, ref_nets, ref_ohem
from tests.helpers import NETS, GOLDEN, load_net_outputs, load_specs, rel_err


@pytest.mark.parametrize("name", NETS)
@pytest.mark.parametrize("tag,perturb", [("perturbed", True), ("identity_bn", False)])
def test_net_restatement_matches_reference_vectors(name, tag, perturb):
    specs, golden = load_specs(), load_net_outputs()
    sd = ref_nets.make_state(specs[name]["spec"], seed=1, perturb_bn=perturb)
    x = ref_nets.make_input(0, 1, 32, 64)
    with torch.no_grad():
        y = ref_nets.FORWARD[name](sd, x)
    ref = torch.from_numpy(golden[f"{name}.{tag}"])
    assert y.shape == ref.shape
    assert rel_err(y, ref) < 1e-5
    assert (y.argmax(1) == ref.argmax(1)).float().mean().item() >= 0.9999


def test_param_counts_match_readme():
    # README.md:148,153 "my" column; SURVEY §4 known answers
    specs = load_specs()
    assert specs["erfnet"]["n_params"] == 2065458
    assert specs["fastscnn"]["n_params"] == 1023769
    assert specs["dabnet"]["n_params"] == 750269
    assert specs["lednet"]["n_params"] == 1461383


def test_ohem_restatement_matches_reference_losses():
    with open(os.path.join(GOLDEN, "ohem_cases.json")) as f:
        cases = json.load(f)
    branches = []
    for c in cases:
        g = torch.Generator().manual_seed(100 + c["case"])
        logits = torch.randn((2, 19, 24, 40), generator=g)
        labels = torch.randint(0, 19, (2, 24, 40), generator=g)
        if c["margin"]:
            logits.scatter_add_(1, labels.unsqueeze(1), torch.full((2, 1, 24, 40), c["margin"]))
        labels[torch.rand((2, 24, 40), generator=g) < c["ignore_frac"]] = 255
        logits.requires_grad_(True)
        loss, branch = ref_ohem.ohem_ce_loss(logits, labels, 0.7, 255)
        loss.backward()
        branches.append(branch)
        assert abs(float(loss) - c["loss"]) <= 1e-6 * max(1.0, abs(c["loss"]))
        assert abs(float(logits.grad.abs().sum()) - c["grad_abs_sum"]) <= 1e-5 * c["grad_abs_sum"]
    assert set(branches) == {0, 1}, "golden cases must exercise both OHEM branches"


def test_ohem_all_ignored_is_nan():
    logits = torch.randn(1, 19, 4, 4)
    labels = torch.full((1, 4, 4), 255)
    loss, _ = ref_ohem.ohem_ce_loss(logits, labels)
    assert torch.isnan(loss)  # empty selection -> NaN


def test_confusion_matrix_restatement():
    rng = np.random.default_rng(0)
    pred = rng.integers(0, 19, size=5000)
    target = rng.integers(0, 19, size=5000)
    target[rng.random(5000) < 0.1] = 255
    cm = ref_head.confusion_matrix(pred, target, 19, 255)
    assert cm.dtype == np.int64 and cm.shape == (19, 19)
    assert (cm == ref_head.confusion_matrix_loops(pred, target, 19, 255)).all()
    assert cm.sum() == (target != 255).sum()
    # empty / all ignored
    assert ref_head.confusion_matrix(np.zeros(0, int), np.zeros(0, int), 19).sum() == 0
    assert ref_head.confusion_matrix(pred, np.full(5000, 255), 19).sum() == 0


def test_iou_known_answer():
    cm = np.array([[3, 1, 0], [2, 4, 0], [0, 0, 0]], dtype=np.int64)
    iou = ref_head.iou_from_confmat(cm)
    np.testing.assert_allclose(iou, [3 / 6, 4 / 7, 0.0], rtol=1e-6)
    assert abs(ref_head.miou(cm) - (3 / 6 + 4 / 7) / 3) < 1e-6  # absent class counts as 0


def test_upsample_argmax_first_max_on_ties():
    x = torch.zeros(1, 3, 2, 2)  # all ties -> index 0 everywhere
    assert (ref_head.upsample_argmax(x, (8, 8)) == 0).all()


def test_training_restatement_matches_reference_training_step():
    """oracle/ref_blocks.TRAINING (batch-statistics BatchNorm) + torch autograd vs the REAL reference's ERFNet.train()
    step recorded in tests/golden/train_step.json: loss, every gradient's norm, the updated running statistics."""
    from oracle import ref_blocks as B
    with open(os.path.join(GOLDEN, "train_step.json")) as f:
        gold = json.load(f)
    spec = load_specs()["erfnet"]["spec"]
    sd = ref_nets.make_state(spec, 3, True)
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.dtype.is_floating_point and "running" not in k}
    x = ref_nets.make_input(4, 2, 64, 128)
    g = torch.Generator().manual_seed(5)
    labels = torch.randint(0, 19, (2, 64, 128), generator=g)
    labels[torch.rand((2, 64, 128), generator=g) < 0.1] = 255
    B.TRAINING = True
    try:
        logits = ref_nets.erfnet_forward(sd, x)
    finally:
        B.TRAINING = False
    loss = ref_ohem.ohem_ce_loss(logits, labels, 0.7, 255)[0]
    loss.backward()
    assert abs(float(loss.detach()) - gold["loss"]) <= 1e-6 * max(1.0, abs(gold["loss"]))
    assert set(params) == set(gold["grad_norms"])
    for k, want in gold["grad_norms"].items():
        got = float(params[k].grad.double().norm())
        # same ATen kernels, same order: agreement is at rounding level; biases in front of a BatchNorm have ~0 gradients
        assert abs(got - want) <= 1e-4 * max(want, 1e-3), k
    for k, want in gold["running_norms"].items():
        assert abs(float(sd[k].double().norm()) - want) <= 1e-6 * max(want, 1.0), k


def test_colorize_blend_restatement_matches_pillow_and_reference_colormap():
    """ref_head.colorize_blend == colormap LUT + PIL.Image.blend (what reference core/seg_trainer.py:172-191 calls)."""
    from PIL import Image
    rng = np.random.default_rng(11)
    cmap = [(128, 64, 128), (244, 35, 232), (70, 70, 70), (102, 102, 156), (190, 153, 153), (153, 153, 153), (250, 170, 30),
            (220, 220, 0), (107, 142, 35), (152, 251, 152), (70, 130, 180), (220, 20, 60), (255, 0, 0), (0, 0, 142), (0, 0, 70),
            (0, 60, 100), (0, 80, 100), (0, 0, 230), (119, 11, 32)]     # utils/utils.py:59-66
    labels = rng.integers(0, 19, size=(37, 53)).astype(np.uint8)
    image = rng.integers(0, 256, size=(37, 53, 3)).astype(np.uint8)
    pred = ref_head.colorize_blend(labels, cmap)
    assert pred.dtype == np.uint8 and np.array_equal(pred, np.asarray(cmap, dtype=np.uint8)[labels])
    for alpha in (0.0, 0.3, 0.4, 0.5, 0.77, 1.0):
        want = np.asarray(Image.blend(Image.fromarray(image), Image.fromarray(pred), alpha))
        assert np.array_equal(ref_head.colorize_blend(labels, cmap, image, alpha), want), alpha


def test_uint8_normalisation_restatement_matches_the_textbook_formula():
    """oracle/ref_input.normalize_u8 (albumentations' float32 arithmetic, parity unpinned: the library is absent)
    agrees with (x/255 - mean)/std evaluated in float64 to float32 rounding, for all 256 byte values."""
    import numpy as np
    from oracle import ref_input
    img = np.arange(256, dtype=np.uint8).reshape(1, 16, 16, 1).repeat(3, axis=3)
    got = ref_input.normalize_u8(img).double()
    mean = torch.tensor(ref_input.CITYSCAPES_MEAN, dtype=torch.float64).view(1, 3, 1, 1)
    std = torch.tensor(ref_input.CITYSCAPES_STD, dtype=torch.float64).view(1, 3, 1, 1)
    want = (torch.from_numpy(img).permute(0, 3, 1, 2).double() / 255.0 - mean) / std
    assert got.shape == (1, 3, 16, 16) and got.dtype == torch.float64
    assert (got - want).abs().max().item() < 2e-6


def test_sgd_restatement_matches_torch_optim_sgd():
    """oracle/ref_optim.sgd_step against torch.optim.SGD (the reference's optimizer, utils/optimizer.py:8-12): three steps."""
    from oracle import ref_optim
    g = torch.Generator().manual_seed(0)
    ps = [torch.randn(s, generator=g) for s in ((7, 3), (5,), (2, 2, 3, 1))]
    mine = [p.clone() for p in ps]
    theirs = [torch.nn.Parameter(p.clone()) for p in ps]
    opt = torch.optim.SGD(theirs, lr=0.08, momentum=0.9, weight_decay=1e-4)
    bufs = [None] * len(ps)
    for _ in range(3):
        grads = [torch.randn(p.shape, generator=g) for p in ps]
        for t, gr in zip(theirs, grads):
            t.grad = gr.clone()
        opt.step()
        ref_optim.sgd_step(mine, grads, bufs, 0.08, 0.9, 1e-4)
    for a, b in zip(mine, theirs):
        assert torch.allclose(a, b.detach(), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name,seed,in_seed,label_seed,hw", [("fastscnn", 13, 14, 15, (128, 256)), ("dabnet", 23, 24, 25, (64, 128))])
def test_training_restatement_matches_reference_step_of_the_depthwise_nets(name, seed, in_seed, label_seed, hw):
    """Row T: the oracle's Fast-SCNN / DABNet in TRAINING mode + torch autograd vs the REAL reference's .train() step
    (tests/golden/train_step_dw.json, written by make_golden_train.py): depthwise convolutions, PReLU slopes,
    pyramid pooling and the bilinear resizes all carry gradients here."""
    from oracle import ref_blocks as B
    with open(os.path.join(GOLDEN, "train_step_dw.json")) as f:
        gold = json.load(f)[name]
    sd = ref_nets.make_state(load_specs()[name]["spec"], seed, True)
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.dtype.is_floating_point and "running" not in k}
    x = ref_nets.make_input(in_seed, 2, *hw)
    g = torch.Generator().manual_seed(label_seed)
    labels = torch.randint(0, 19, (2, *hw), generator=g)
    labels[torch.rand((2, *hw), generator=g) < 0.1] = 255
    B.TRAINING = True
    try:
        logits = ref_nets.FORWARD[name](sd, x)
    finally:
        B.TRAINING = False
    loss = ref_ohem.ohem_ce_loss(logits, labels, 0.7, 255)[0]
    loss.backward()
    assert abs(float(loss.detach()) - gold["loss"]) <= 1e-6 * max(1.0, abs(gold["loss"]))
    assert set(params) == set(gold["grad_norms"])
    for k, want in gold["grad_norms"].items():
        got = float(params[k].grad.double().norm())
        assert abs(got - want) <= 1e-4 * max(want, 1e-3), k
    for k, want in gold["running_norms"].items():
        assert abs(float(sd[k].double().norm()) - want) <= 1e-6 * max(want, 1.0), k


@pytest.mark.parametrize("name", ["bisenetv2", "ddrnet"])
@pytest.mark.parametrize("tag,perturb", [("perturbed", True), ("identity_bn", False)])
def test_config4_restatement_matches_reference_vectors(name, tag, perturb):
    """oracle/ref_nets_dual.py vs the REAL reference's BiSeNetv2 / DDRNet logits (tests/golden/cfg4_outputs.npz)."""
    specs, golden = load_specs(), load_net_outputs()
    sd = ref_nets.make_state(specs[name]["spec"], seed=1, perturb_bn=perturb)
    x = ref_nets.make_input(0, 1, 64, 128)
    with torch.no_grad():
        y = ref_nets.FORWARD[name](sd, x)
    ref = torch.from_numpy(golden[f"{name}.{tag}"])
    assert y.shape == ref.shape
    assert rel_err(y, ref) < 1e-5
    assert (y.argmax(1) == ref.argmax(1)).float().mean().item() >= 0.9999
