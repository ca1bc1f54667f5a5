"""Import alias for the framework package.

The framework lives in ``sagemaker-distributed-training-pytorch-kr_b200/`` (the name the project
brief fixes); hyphens are not importable, so ``import b200dist`` maps onto that directory.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "sagemaker-distributed-training-pytorch-kr_b200")
__path__.insert(0, _real)  # type: ignore[name-defined]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
