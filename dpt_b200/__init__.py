"""Importable alias for the package that lives in ``distributed-pytorch-training_b200/``.

The project directory keeps the upstream repository's name (hyphens included), which Python
cannot import directly; this shim makes ``import dpt_b200`` resolve to that directory.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "distributed-pytorch-training_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f, _os
