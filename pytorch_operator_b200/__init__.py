"""Import alias: the package directory is ``pytorch-operator_b200/`` (not a Python
identifier), so ``import pytorch_operator_b200`` resolves here and is redirected there."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "pytorch-operator_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
