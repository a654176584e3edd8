"""qpdo_b200 -- B200-native QPDO Newton hot path behind the reference's qpdo.h C API.

The product is the shared library qpdo_b200/lib/libqpdo_b200.so (host C + sm_100a CUDA, built by
build.sh / __graft_entry__.build()).  This package is only the ctypes mirror of the reference's
user-facing wrapper class (interfaces/mex/qpdo.m) over that C ABI; there is no Python or CPU
compute path, and loading fails loudly when the library is missing.
"""
import os
from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libqpdo_b200.so")
_lib = None


def lib():
    """dlopen the CUDA library (once)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("qpdo_b200: %s is missing -- run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)" % LIB_PATH)
        _lib = capi.load_library(LIB_PATH)
        capi.declare_device_layer(_lib)
    return _lib


class QPDO(capi.Solver):
    """setup / warm_start / update_q / update_bounds / solve / delete, like interfaces/mex/qpdo.m."""

    def __init__(self):
        capi.Solver.__init__(self, lib())

    @property
    def device(self):
        """opaque qpdo_dev* (QPDOCholmod.LD) for the diagnostics entry points of qpdo_cuda.h"""
        return self._work.contents.chol.contents.LD
