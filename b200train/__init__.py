"""Import shim: the framework lives in ``lambda-labs_distributed-training-guide_b200/`` (a directory
name that is not a Python identifier); ``import b200train`` resolves to that tree."""
import os as _os

_real = _os.path.join(
    _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
    "lambda-labs_distributed-training-guide_b200",
)
__path__.insert(0, _real)
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _f
