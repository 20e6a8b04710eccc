"""ctypes front-end for oracle B (oracle/ekf_struct.c) and the compiled reference turtlelib.

TEST INFRASTRUCTURE ONLY - see oracle/ekf_literal.py.  Imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libekf_oracle.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libturtlelib_ref.so")
_lib = None
_ref = None


def build(force: bool = False) -> None:
    """(Re)build the oracle libraries with oracle/Makefile."""
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "ekf_struct.c")):
        subprocess.check_call(["make", "-C", _HERE, "libekf_oracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/turtlelib/src") and (force or not os.path.exists(_REF_PATH)):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        L.ekfo_create.restype = C.c_void_p
        L.ekfo_create.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_double]
        L.ekfo_destroy.argtypes = [C.c_void_p]
        L.ekfo_state.restype = dp
        L.ekfo_state.argtypes = [C.c_void_p]
        L.ekfo_sigma.restype = dp
        L.ekfo_sigma.argtypes = [C.c_void_p]
        L.ekfo_n.argtypes = [C.c_void_p]
        L.ekfo_count.restype = C.c_uint
        L.ekfo_count.argtypes = [C.c_void_p]
        L.ekfo_set_count.argtypes = [C.c_void_p, C.c_uint]
        L.ekfo_set_prev.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.ekfo_get_prev.argtypes = [C.c_void_p, dp]
        L.ekfo_predict.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.ekfo_update_known.argtypes = [C.c_void_p, C.c_int, ip, dp, C.POINTER(C.c_ubyte)]
        L.ekfo_update_unknown.argtypes = [C.c_void_p, C.c_int, dp, ip]
        L.ekfo_update_unknown_ex.argtypes = [C.c_void_p, C.c_int, dp, ip, dp, dp]
        L.ekfo_normalize_angle.restype = C.c_double
        L.ekfo_normalize_angle.argtypes = [C.c_double]
        _lib = L
    return _lib


def ref_lib():
    """The reference's own turtlelib object code, or None when it was never built."""
    global _ref
    if _ref is None:
        if not os.path.exists(_REF_PATH):
            build()
        if not os.path.exists(_REF_PATH):
            return None
        R = C.CDLL(_REF_PATH)
        dp = C.POINTER(C.c_double)
        R.ref_normalize_angle.restype = C.c_double
        R.ref_normalize_angle.argtypes = [C.c_double]
        for name in ("ref_compose", "ref_apply_point"):
            getattr(R, name).argtypes = [dp, dp, dp]
        R.ref_inverse.argtypes = [dp, dp]
        R.ref_integrate_twist.argtypes = [dp, dp]
        R.ref_fkin_sequence.argtypes = [C.c_double, C.c_double, C.c_int, dp, dp]
        _ref = R
    return _ref


def _dp(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class StructEkf:
    """Oracle B with the same method names as the C ABI / the GPU host mirror."""

    def __init__(self, n_landmarks=50, q_noise=1.0e-2, r_noise=1.0e-2, mah_threshold=2.0, prior=10e6):
        self._L = lib()
        self._h = self._L.ekfo_create(n_landmarks, q_noise, r_noise, mah_threshold, prior)
        if not self._h:
            raise MemoryError("ekfo_create failed")
        self.n_landmarks = n_landmarks
        self.n = 3 + 2 * n_landmarks
        self.state = np.ctypeslib.as_array(self._L.ekfo_state(self._h), shape=(self.n,))
        # column-major n x n  ->  expose as sigma[i, j]
        self.sigma = np.ctypeslib.as_array(self._L.ekfo_sigma(self._h), shape=(self.n, self.n)).T

    def close(self):
        if self._h:
            self.state = self.sigma = None
            self._L.ekfo_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def count(self) -> int:
        return int(self._L.ekfo_count(self._h))

    @count.setter
    def count(self, c: int):
        self._L.ekfo_set_count(self._h, int(c))

    def set_prev(self, x, y):
        self._L.ekfo_set_prev(self._h, float(x), float(y))

    def get_prev(self):
        out = np.zeros(2)
        self._L.ekfo_get_prev(self._h, _dp(out))
        return out

    def predict(self, theta, x, y):
        self._L.ekfo_predict(self._h, float(theta), float(x), float(y))

    def update_known(self, ids, xy, action=None):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1)
        act = None
        if action is not None:
            action = np.ascontiguousarray(action, dtype=np.uint8)
            act = action.ctypes.data_as(C.POINTER(C.c_ubyte))
        rc = self._L.ekfo_update_known(self._h, len(ids), ids.ctypes.data_as(C.POINTER(C.c_int)), _dp(xy), act)
        if rc:
            raise IndexError(f"oracle update_known failed rc={rc}")

    def update_unknown(self, xy):
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1)
        m = len(xy) // 2
        assoc = np.zeros(m, dtype=np.int32)
        rc = self._L.ekfo_update_unknown(self._h, m, _dp(xy), assoc.ctypes.data_as(C.POINTER(C.c_int)))
        if rc:
            raise IndexError(f"oracle update_unknown failed rc={rc}")
        return assoc

    def update_unknown_margins(self, xy):
        """update_unknown that also returns (best, second) existing-landmark distance per measurement"""
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1)
        m = len(xy) // 2
        assoc = np.zeros(m, dtype=np.int32)
        best, second = np.zeros(m), np.zeros(m)
        rc = self._L.ekfo_update_unknown_ex(self._h, m, _dp(xy), assoc.ctypes.data_as(C.POINTER(C.c_int)),
                                            _dp(best), _dp(second))
        if rc:
            raise IndexError(f"oracle update_unknown failed rc={rc}")
        return assoc, best, second
