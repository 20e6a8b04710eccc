"""ctypes front-end for oracle/_ref/libnusim_ref.so: the REFERENCE's own simulator source
(/root/reference/nusim/src/nusim.cpp, compiled unmodified by oracle/Makefile against the stub
headers in oracle/stubs/), driven through oracle/nusim_ref_shim.cpp.

TEST INFRASTRUCTURE ONLY.  Pins oracle/nusim_oracle.py (tests/test_nusim.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libnusim_ref.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH) and os.path.isdir("/root/reference/nusim/src"):
            subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
        if not os.path.exists(_PATH):
            return None
        L = C.CDLL(_PATH)
        dp, ip, fp, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p
        L.nusim_ref_create.restype = vp
        L.nusim_ref_create.argtypes = [C.c_int, dp] + [C.c_double] * 8 + [C.c_uint]
        L.nusim_ref_destroy.argtypes = [vp]
        L.nusim_ref_reseed.argtypes = [vp, C.c_uint]
        L.nusim_ref_normal_draws.argtypes = [C.c_uint, C.c_double, C.c_int, dp]
        L.nusim_ref_set_pose.argtypes = [vp, C.c_double, C.c_double, C.c_double]
        L.nusim_ref_get_pose.argtypes = [vp, dp]
        L.nusim_ref_sensor_tick.argtypes = [vp, fp, dp, ip, ip]
        L.nusim_ref_lidar.argtypes = [vp, fp]
        L.nusim_ref_walls.argtypes = [vp, dp]
        L.nusim_ref_drive.argtypes = [vp, C.c_int, C.c_int, C.c_int, dp]
        _lib = L
    return _lib


def available() -> bool:
    return lib() is not None


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def normal_draws(seed: int, stddev: float, count: int) -> np.ndarray:
    """what std::normal_distribution<>{0, stddev} yields from std::mt19937{seed} (libstdc++)"""
    out = np.zeros(count)
    lib().nusim_ref_normal_draws(seed, stddev, count, _dp(out))
    return out


class RefNuSim:
    """One `NuSim` node object of the reference (nusim.cpp:65-755)."""

    def __init__(self, obstacles_xyr, arena_x, arena_y, max_range=5.0, sensor_std=0.0, input_noise=0.0,
                 slip_fraction=0.0, collision_radius=0.11, rate=200.0, seed=0):
        self._L = lib()
        if self._L is None:
            raise RuntimeError("oracle/_ref/libnusim_ref.so is not built (needs /root/reference)")
        obs = np.ascontiguousarray(obstacles_xyr, dtype=np.float64).reshape(-1, 3)
        self.n_obst = len(obs)
        self._h = self._L.nusim_ref_create(self.n_obst, _dp(obs), arena_x, arena_y, max_range, sensor_std,
                                           input_noise, slip_fraction, collision_radius, rate, seed)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.nusim_ref_destroy(self._h)
            self._h = None

    def reseed(self, seed):
        self._L.nusim_ref_reseed(self._h, seed)

    def set_pose(self, theta, x, y):
        self._L.nusim_ref_set_pose(self._h, float(x), float(y), float(theta))

    def pose(self):
        out = np.zeros(3)
        self._L.nusim_ref_get_pose(self._h, _dp(out))
        return out[2], out[0], out[1]

    def sensor_tick(self):
        """-> (ranges float32[360], xy[n,2], ids[n], action[n]) of one sensor_timer_callback"""
        ranges = np.zeros(360, dtype=np.float32)
        xy = np.zeros((self.n_obst, 2))
        ids = np.zeros(self.n_obst, dtype=np.int32)
        act = np.zeros(self.n_obst, dtype=np.int32)
        k = self._L.nusim_ref_sensor_tick(self._h, ranges.ctypes.data_as(C.POINTER(C.c_float)), _dp(xy),
                                          ids.ctypes.data_as(C.POINTER(C.c_int)), act.ctypes.data_as(C.POINTER(C.c_int)))
        assert k == self.n_obst
        return ranges, xy, ids, act

    def lidar(self):
        ranges = np.zeros(360, dtype=np.float32)
        self._L.nusim_ref_lidar(self._h, ranges.ctypes.data_as(C.POINTER(C.c_float)))
        return ranges

    def walls(self):
        out = np.zeros((4, 5))
        self._L.nusim_ref_walls(self._h, _dp(out))
        return out

    def drive(self, left_cmd, right_cmd, steps):
        out = np.zeros(7)
        self._L.nusim_ref_drive(self._h, int(left_cmd), int(right_cmd), int(steps), _dp(out))
        return out
