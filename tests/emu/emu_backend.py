"""numpy 'device' + the host-emulation build of the kernels (tests only; see csrc/rm_emu.h)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import build_emu  # noqa: E402

from radixmesh_b200 import capi  # noqa: E402


class NumpyBackend:
    name = "emu"

    def empty(self, n, dtype):
        return np.zeros(int(n), dtype=dtype)

    def to_device(self, a):
        return np.ascontiguousarray(a).copy()

    def to_host(self, t):
        return np.asarray(t)

    def stream(self):
        return None


_lib = None


def emu_lib():
    global _lib
    if _lib is None:
        _lib = capi.load_library(build_emu.build())
    return _lib


def make_index(**kw):
    return capi.DeviceIndex(backend=NumpyBackend(), library=emu_lib(), **kw)
