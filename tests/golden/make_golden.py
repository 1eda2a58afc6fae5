"""Generates tests/golden/*.json|*.npz by running the REAL reference (read-only /root/reference).

Run in the build container only:   python tests/golden/make_golden.py
The GPU box has no reference; it uses the committed files.  Everything is seeded, CPU, fp32.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import ref_import, ref_nets  # noqa: E402

NETS = ["erfnet", "fastscnn", "dabnet", "lednet"]
H, W = 32, 64  # multiple of 32 so every stage of both nets has a valid size


def spec_of(model):
    return [[k, list(v.shape), str(v.dtype).replace("torch.", "")] for k, v in model.state_dict().items()]


def main():
    assert ref_import.available(), "needs /root/reference"
    torch.set_num_threads(8)
    specs, arrays = {}, {}
    for name in NETS:
        torch.manual_seed(1)  # the reference's seed, configs/base_config.py:70
        model = ref_import.reference_model(name, 19)
        spec = spec_of(model)
        n_params = sum(p.numel() for p in model.parameters())
        specs[name] = {"num_class": 19, "n_params": n_params, "spec": spec}
        x = ref_nets.make_input(0, 1, H, W)
        for tag, perturb in (("perturbed", True), ("identity_bn", False)):
            sd = ref_nets.make_state([tuple(s) for s in spec], seed=1, perturb_bn=perturb)
            model.load_state_dict(sd, strict=True)
            with torch.no_grad():
                y = model(x)
            assert y.shape == (1, 19, H, W)
            arrays[f"{name}.{tag}"] = y.numpy().astype(np.float32)
            print(name, tag, "params", n_params, "out mean %.5f std %.5f classes %d" % (
                y.mean().item(), y.std().item(), y.argmax(1).unique().numel()))
    with open(os.path.join(HERE, "state_specs.json"), "w") as f:
        json.dump(specs, f)
    np.savez_compressed(os.path.join(HERE, "net_outputs.npz"), **arrays)

    # OHEM: the reference module calls .cuda() in __init__ (core/loss.py:9); make that a no-op on this CPU box.
    ref_loss = ref_import.reference_loss_module()
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        crit = ref_loss.OhemCELoss(0.7, 255)
    finally:
        torch.Tensor.cuda = orig_cuda
    cases = []
    for case, (margin, ignore_frac) in enumerate([(0.0, 0.1), (12.0, 0.1), (4.0, 0.5), (12.0, 0.0)]):
        g = torch.Generator().manual_seed(100 + case)
        logits = torch.randn((2, 19, 24, 40), generator=g)
        labels = torch.randint(0, 19, (2, 24, 40), generator=g)
        if margin:
            logits.scatter_add_(1, labels.unsqueeze(1), torch.full((2, 1, 24, 40), margin))
        mask = torch.rand((2, 24, 40), generator=g) < ignore_frac
        labels[mask] = 255
        logits.requires_grad_(True)
        loss = crit(logits, labels)
        loss.backward()
        cases.append({"case": case, "margin": margin, "ignore_frac": ignore_frac, "loss": float(loss),
                      "grad_abs_sum": float(logits.grad.abs().sum()), "grad_sq_sum": float((logits.grad ** 2).sum())})
        print("ohem", cases[-1])
    with open(os.path.join(HERE, "ohem_cases.json"), "w") as f:
        json.dump(cases, f, indent=1)

    # Training step of the REAL reference (ERFNet in .train(): batch-statistics BatchNorm; its own OhemCELoss; autograd):
    # loss, every parameter gradient's L2 norm and the updated running statistics' norms pin oracle/ref_blocks.TRAINING.
    model = ref_import.reference_model("erfnet", 19)
    spec = [tuple(s) for s in specs["erfnet"]["spec"]]
    model.load_state_dict(ref_nets.make_state(spec, seed=3, perturb_bn=True), strict=True)
    model.train()
    x = ref_nets.make_input(4, 2, 64, 128)
    g = torch.Generator().manual_seed(5)
    labels = torch.randint(0, 19, (2, 64, 128), generator=g)
    labels[torch.rand((2, 64, 128), generator=g) < 0.1] = 255
    loss = crit(model(x), labels)
    loss.backward()
    train = {"loss": float(loss.detach()),
             "grad_norms": {k: float(p.grad.double().norm()) for k, p in model.named_parameters()},
             "running_norms": {k: float(v.double().norm()) for k, v in model.state_dict().items() if "running" in k}}
    print("train step: loss", train["loss"], "params with grads", len(train["grad_norms"]))
    with open(os.path.join(HERE, "train_step.json"), "w") as f:
        json.dump(train, f)


if __name__ == "__main__":
    main()
