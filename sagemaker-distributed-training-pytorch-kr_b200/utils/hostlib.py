"""ctypes loader for ``lib/libb200dist_host.so`` (CPU-side C++ runtime: reducer core, dataset index
builders, pipeline schedule generator).  Built by ``b200dist.build_native``; built on demand if absent."""
from __future__ import annotations

import ctypes
import os

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_PATH = os.path.join(_PKG, "lib", "libb200dist_host.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            from .. import build_native  # type: ignore
            build_native.build(verbose=False)
        _lib = ctypes.CDLL(_PATH)
        _lib.b2h_reducer_create.restype = ctypes.c_void_p
    return _lib
