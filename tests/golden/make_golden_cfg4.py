"""Golden vectors of the REAL reference (read-only /root/reference) for the config-4 nets: BiSeNetv2 and DDRNet.

Run in the build container only:   python tests/golden/make_golden_cfg4.py
Writes tests/golden/cfg4_outputs.npz (full-resolution logits of each net for a perturbed-BatchNorm and an
identity-BatchNorm seeded state on a seeded 1 x 3 x 64 x 128 input) and adds the two state-dict specs to
tests/golden/state_specs.json.  Both nets are built as get_model does with use_aux = False (models/__init__.py).
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

NETS = ["bisenetv2", "ddrnet"]
H, W = 64, 128   # DDRNet's pyramid ends at 1/64 resolution


def main():
    assert ref_import.available(), "needs /root/reference"
    torch.set_num_threads(8)
    with open(os.path.join(HERE, "state_specs.json")) as f:
        specs = json.load(f)
    arrays = {}
    for name in NETS:
        torch.manual_seed(1)
        model = ref_import.reference_model(name, 19)
        spec = [[k, list(v.shape), str(v.dtype).replace("torch.", "")] for k, v in model.state_dict().items()]
        specs[name] = {"num_class": 19, "n_params": sum(p.numel() for p in model.parameters()), "spec": spec}
        x = ref_nets.make_input(0, 1, H, W)
        for tag, perturb in (("perturbed", True), ("identity_bn", False)):
            model.load_state_dict(ref_nets.make_state([tuple(s) for s in spec], seed=1, perturb_bn=perturb), strict=True)
            with torch.no_grad():
                y = model(x)
            assert y.shape == (1, 19, H, W)
            arrays[f"{name}.{tag}"] = y.numpy().astype(np.float32)
            print(name, tag, "params", specs[name]["n_params"], "out mean %.5f std %.5f classes %d" % (
                y.mean().item(), y.std().item(), y.argmax(1).unique().numel()))
    with open(os.path.join(HERE, "state_specs.json"), "w") as f:
        json.dump(specs, f)
    np.savez_compressed(os.path.join(HERE, "cfg4_outputs.npz"), **arrays)


if __name__ == "__main__":
    main()
