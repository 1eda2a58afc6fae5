"""Imports the REAL reference (read-only, /root/reference) when it is present — build container only.

Used by tests/golden/make_golden.py to generate the committed vectors and by the CPU test-suite to
check the restatement live.  The GPU box has no /root/reference; nothing that runs there calls this.
Recipe: SURVEY App. C (stub the missing `segmentation_models_pytorch`, never import utils/core/datasets).
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("RTSEG_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "modules.py"))


def reference_models():
    """The reference's `models` package, untouched."""
    if "models" in sys.modules and getattr(sys.modules["models"], "__file__", "").startswith(REFERENCE_ROOT):
        return sys.modules["models"]
    if "segmentation_models_pytorch" not in sys.modules:
        stub = types.ModuleType("segmentation_models_pytorch")
        for name in ("DeepLabV3", "DeepLabV3Plus", "FPN", "Linknet", "MAnet", "PAN", "PSPNet", "Unet", "UnetPlusPlus"):
            setattr(stub, name, object)
        sys.modules["segmentation_models_pytorch"] = stub
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    return importlib.import_module("models")


def reference_model(name: str, num_class: int = 19):
    cfg = types.SimpleNamespace(model=name, num_class=num_class, use_aux=False, use_detail_head=False)
    return reference_models().get_model(cfg).eval()


def reference_loss_module():
    spec = importlib.util.spec_from_file_location("ref_loss", os.path.join(REFERENCE_ROOT, "core", "loss.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
