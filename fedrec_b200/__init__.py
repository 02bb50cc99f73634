"""Importable alias for the ``fedrec-with-pytorchdistributed_b200`` source tree.

The framework lives in a directory whose name carries a hyphen (it mirrors the
upstream project name), which Python cannot import directly.  This package
re-points its ``__path__`` at that directory so ``import fedrec_b200.models``
resolves to ``fedrec-with-pytorchdistributed_b200/models`` with no copies.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
_SRC = _os.path.join(_ROOT, "fedrec-with-pytorchdistributed_b200")
__path__ = [_SRC]
__file__ = _os.path.join(_SRC, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
