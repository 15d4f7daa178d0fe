"""Importable alias of the framework package.

The source tree lives in ``distributed-training-and-deepspeed_b200/`` (the name the project brief
asks for); a hyphenated directory cannot be imported, so this shim exposes it as ``dtds_b200``:
its ``__path__`` points at the real directory and the real ``__init__`` is executed in this
module's namespace.  ``import dtds_b200.ops`` therefore loads
``distributed-training-and-deepspeed_b200/ops/__init__.py``.
"""
import os as _os

_REAL = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "distributed-training-and-deepspeed_b200")
__path__ = [_REAL]
__file__ = _os.path.join(_REAL, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
