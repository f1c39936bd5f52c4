"""Import alias for the framework package.

The source tree lives in ``layer-wise-aaai20_b200/`` (the directory name the
project layout prescribes).  A hyphen cannot appear in a Python identifier, so
this stub package re-points its ``__path__`` at that directory and executes its
``__init__`` -- ``import layerwise_b200.ops`` therefore loads
``layer-wise-aaai20_b200/ops``.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
_SRC = _os.path.join(_ROOT, "layer-wise-aaai20_b200")
__path__ = [_SRC]
with open(_os.path.join(_SRC, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_SRC, "__init__.py"), "exec"))
del _f
