"""Run the reference's own, unmodified model files on the rtseg building blocks.

The reference has no plugin layer; its seam is `from .modules import ConvBNAct, ...` at the top of
every models/*.py.  Seeding `sys.modules['models.modules']` with our module object *before*
`import models` makes those imports bind to the fused classes.  The reference
module itself is never modified.  `segmentation_models_pytorch` (imported unconditionally at
models/__init__.py:2 but absent from this image) is satisfied with an attribute stub.
"""
from __future__ import annotations

import importlib
import sys
import types

_SMP_NAMES = ("DeepLabV3", "DeepLabV3Plus", "FPN", "Linknet", "MAnet", "PAN", "PSPNet", "Unet", "UnetPlusPlus")


def install(reference_root: str, use_rtseg_modules: bool = True):
    """Returns the reference `models` package, built on rtseg modules if requested."""
    if "models" in sys.modules:
        raise RuntimeError("`models` is already imported; install() must run first")
    if "segmentation_models_pytorch" not in sys.modules:
        try:
            importlib.import_module("segmentation_models_pytorch")
        except ImportError:
            stub = types.ModuleType("segmentation_models_pytorch")
            for name in _SMP_NAMES:
                setattr(stub, name, object)
            sys.modules["segmentation_models_pytorch"] = stub
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    if use_rtseg_modules:
        from . import modules as rt_modules
        sys.modules["models.modules"] = rt_modules
    return importlib.import_module("models")
