"""Adds ENet (BASELINE configs[0]) to the golden fixtures by running the REAL reference (read-only /root/reference).

Run in the build container only:   python tests/golden/make_golden_enet.py
Writes tests/golden/enet_outputs.npz and merges the `enet` entry into state_specs.json; the other fixtures are left
untouched (their generator is make_golden.py).  Seeded, CPU, fp32.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))

from oracle import ref_import, ref_nets  # noqa: E402


def main():
    assert ref_import.available(), "needs /root/reference"
    torch.manual_seed(1)
    model = ref_import.reference_model("enet", 19).eval()
    spec = [[k, list(v.shape), str(v.dtype).replace("torch.", "")] for k, v in model.state_dict().items()]
    n_params = sum(p.numel() for p in model.parameters())
    arrays = {}
    for (h, w), seed in (((32, 64), 0), ((64, 96), 7)):
        x = ref_nets.make_input(seed, 1, h, w)
        for tag, perturb in (("perturbed", True), ("identity_bn", False)):
            sd = ref_nets.make_state([tuple(s) for s in spec], seed=1, perturb_bn=perturb)
            model.load_state_dict(sd, strict=True)
            with torch.no_grad():
                y = model(x)
            # the 32 x 64 vectors carry the same keys as the other nets' (seed-0 input of make_golden.py)
            arrays[f"enet.{tag}" if (h, w) == (32, 64) else f"enet.{tag}.{h}x{w}"] = y.numpy().astype(np.float32)
            print("enet", tag, (h, w), "params", n_params, "mean %.5f std %.5f classes %d" % (
                y.mean().item(), y.std().item(), y.argmax(1).unique().numel()))
    np.savez_compressed(os.path.join(HERE, "enet_outputs.npz"), **arrays)
    path = os.path.join(HERE, "state_specs.json")
    specs = json.load(open(path))
    specs["enet"] = {"num_class": 19, "n_params": n_params, "spec": spec}
    json.dump(specs, open(path, "w"))


if __name__ == "__main__":
    main()
