"""Tiny device-buffer helper over b2_dev_* so host code (tests, bench) can drive the op-level entry
points with numpy arrays and no other CUDA binding."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import check


class DevBuf:
    def __init__(self, nbytes: int):
        self.lib = _lib.load()
        self.nbytes = int(nbytes)
        self.p = C.c_void_p()
        check(self.lib.b2_dev_alloc(self.nbytes, C.byref(self.p)))

    @classmethod
    def from_numpy(cls, a: np.ndarray) -> "DevBuf":
        a = np.ascontiguousarray(a)
        b = cls(max(a.nbytes, 1))
        if a.nbytes:
            check(b.lib.b2_dev_upload(b.p, a.ctypes.data, a.nbytes))
        return b

    def to_numpy(self, dtype, shape) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        assert out.nbytes <= self.nbytes
        if out.nbytes:
            check(self.lib.b2_dev_download(out.ctypes.data, self.p, out.nbytes))
        return out

    @property
    def ptr(self) -> C.c_void_p:
        return self.p

    def free(self):
        if self.p:
            self.lib.b2_dev_free(self.p)
            self.p = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
