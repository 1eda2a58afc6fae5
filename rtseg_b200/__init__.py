"""Import shim: `rtseg_b200` -> ../realtime-semantic-segmentation-pytorch_b200/ (a hyphenated
directory name cannot be imported directly)."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "realtime-semantic-segmentation-pytorch_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
