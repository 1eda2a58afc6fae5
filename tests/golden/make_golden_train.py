"""Training steps of the REAL reference for the depthwise nets (read-only /root/reference), recorded for the oracle.

Run in the build container only:   python tests/golden/make_golden_train.py
Writes tests/golden/train_step_dw.json: for Fast-SCNN and DABNet in .train() (batch-statistics BatchNorm, the
reference's own OhemCELoss, torch autograd) the loss, every parameter gradient's L2 norm and the norms of the updated
running statistics.  tests/test_oracle_golden.py replays the same step on oracle/ref_nets with ref_blocks.TRAINING.
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import ref_import, ref_nets  # noqa: E402

CASES = {"fastscnn": dict(seed=13, in_seed=14, label_seed=15, hw=(128, 256)),
         "dabnet": dict(seed=23, in_seed=24, label_seed=25, hw=(64, 128))}


def step_inputs(case):
    h, w = case["hw"]
    x = ref_nets.make_input(case["in_seed"], 2, h, w)
    g = torch.Generator().manual_seed(case["label_seed"])
    labels = torch.randint(0, 19, (2, h, w), generator=g)
    labels[torch.rand((2, h, w), generator=g) < 0.1] = 255
    return x, labels


def main():
    assert ref_import.available(), "needs /root/reference"
    torch.set_num_threads(8)
    with open(os.path.join(HERE, "state_specs.json")) as f:
        specs = json.load(f)
    ref_loss = ref_import.reference_loss_module()
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self   # core/loss.py:9 calls .cuda() in __init__
    try:
        crit = ref_loss.OhemCELoss(0.7, 255)
    finally:
        torch.Tensor.cuda = orig_cuda
    out = {}
    for name, case in CASES.items():
        model = ref_import.reference_model(name, 19)
        spec = [tuple(s) for s in specs[name]["spec"]]
        model.load_state_dict(ref_nets.make_state(spec, seed=case["seed"], perturb_bn=True), strict=True)
        model.train()
        x, labels = step_inputs(case)
        loss = crit(model(x), labels)
        loss.backward()
        out[name] = {"loss": float(loss.detach()),
                     "grad_norms": {k: float(p.grad.double().norm()) for k, p in model.named_parameters()},
                     "running_norms": {k: float(v.double().norm()) for k, v in model.state_dict().items() if "running" in k}}
        print(name, "train step: loss", out[name]["loss"], "params with grads", len(out[name]["grad_norms"]))
    with open(os.path.join(HERE, "train_step_dw.json"), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()
