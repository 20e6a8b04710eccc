"""ctypes front-end for oracle/_ref/libslam_ref.so: the REFERENCE's own EKF source
(/root/reference/nuslam/src/slam.cpp, compiled unmodified by oracle/Makefile against the
stub headers in oracle/stubs/), driven through oracle/slam_ref_shim.cpp.

TEST INFRASTRUCTURE ONLY - see oracle/ekf_literal.py.  `RefSlam` exposes the same surface
as `ekf_literal.LiteralEkf` (set_odometry, fake_sensor_cb, sensor_cb, state, sigma,
counter_obstacles, composed_pose) so tests and tests/golden/make_golden.py can swap one for
the other.  The library is built in this container only (the reference tree is absent on
the GPU box, where the prebuilt file travels with the snapshot).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libslam_ref.so")
_lib = None


def available() -> bool:
    return lib() is not None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH) and os.path.isdir("/root/reference/nuslam/src"):
            subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
        if not os.path.exists(_PATH):
            return None
        L = C.CDLL(_PATH)
        dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
        L.slam_ref_create.restype = vp
        L.slam_ref_create.argtypes = [C.c_int, C.c_char_p, C.c_double, C.c_double]
        L.slam_ref_destroy.argtypes = [vp]
        L.slam_ref_error.restype = C.c_char_p
        L.slam_ref_error.argtypes = [vp]
        L.slam_ref_n.argtypes = [vp]
        L.slam_ref_count.argtypes = [vp]
        L.slam_ref_threshold.restype = C.c_double
        L.slam_ref_threshold.argtypes = [vp]
        L.slam_ref_set_noise.argtypes = [vp, C.c_double, C.c_double, C.c_double]
        L.slam_ref_set_odom.argtypes = [vp, C.c_double, C.c_double, C.c_double]
        L.slam_ref_joint_state.argtypes = [vp, C.c_double, C.c_double]
        for name in ("slam_ref_get_odom", "slam_ref_get_map_odom", "slam_ref_composed_pose", "slam_ref_get_state",
                     "slam_ref_get_sigma", "slam_ref_last_tf", "slam_ref_set_sigma"):
            getattr(L, name).argtypes = [vp, dp]
            getattr(L, name).restype = None
        L.slam_ref_fake_sensor.argtypes = [vp, C.c_int, ip, dp, C.POINTER(C.c_ubyte)]
        L.slam_ref_sensor.argtypes = [vp, C.c_int, dp, ip, dp, ip]
        L.slam_ref_sensor_ex.argtypes = [vp, C.c_int, dp, ip, dp, ip, dp, dp]
        L.slam_ref_set_state.argtypes = [vp, dp, C.c_int]
        L.slam_ref_set_previous.argtypes = [vp, C.c_double, C.c_double, C.c_double]
        L.slam_ref_set_map_odom.argtypes = [vp, C.c_double, C.c_double, C.c_double]
        L.slam_ref_published_markers.argtypes = [vp, dp, ip, C.c_int]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


class RefOutOfRange(IndexError):
    """The reference's own bounds check fired (Armadillo element access / vector::at)."""


class RefSlam:
    """One `Slam` node object of the reference (slam.cpp:45-677)."""

    def __init__(self, n_landmarks: int = 50, sensor_source: str = "assoc", wheel_radius: float = 0.033,
                 track_width: float = 0.16, q_noise=None, r_noise=None, mah_threshold=None):
        self._L = lib()
        if self._L is None:
            raise RuntimeError("oracle/_ref/libslam_ref.so is not built (needs /root/reference)")
        self._h = self._L.slam_ref_create(n_landmarks, sensor_source.encode(), wheel_radius, track_width)
        self.n_landmarks = n_landmarks
        self.n = self._L.slam_ref_n(self._h)
        self.second_branch_taken = 0
        self.last_min = []
        self.last_best = self.last_second = np.zeros(0)
        if q_noise is not None or r_noise is not None or mah_threshold is not None:
            self._L.slam_ref_set_noise(self._h, 1e-2 if q_noise is None else q_noise,
                                       1e-2 if r_noise is None else r_noise,
                                       2.0 if mah_threshold is None else mah_threshold)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.slam_ref_destroy(self._h)
            self._h = None

    def _raise(self, rc):
        msg = self._L.slam_ref_error(self._h).decode()
        if rc == 1:
            raise RefOutOfRange(msg)
        raise RuntimeError(msg)

    # inputs -----------------------------------------------------------------------------
    def set_odometry(self, x, y, theta):
        self._L.slam_ref_set_odom(self._h, float(x), float(y), float(theta))

    def joint_state(self, left, right):
        """Slam::jointStateCallback (slam.cpp:599-634) with wheel positions (rad)."""
        rc = self._L.slam_ref_joint_state(self._h, float(left), float(right))
        if rc:
            self._raise(rc)

    def fake_sensor_cb(self, markers):
        """markers: iterable of (id, x, y, action) - slam.cpp:180-316."""
        markers = list(markers)
        m = len(markers)
        ids = np.array([int(a[0]) for a in markers], dtype=np.int32)
        xy = np.array([[a[1], a[2]] for a in markers], dtype=np.float64).reshape(-1)
        act = np.array([int(a[3]) for a in markers], dtype=np.uint8)
        rc = self._L.slam_ref_fake_sensor(self._h, m, _ip(ids), _dp(xy), act.ctypes.data_as(C.POINTER(C.c_ubyte)))
        if rc:
            self._raise(rc)

    def sensor_cb(self, markers):
        """markers: iterable of (x, y) - slam.cpp:318-530.  Returns the association list."""
        xy = np.array([[a[0], a[1]] for a in markers], dtype=np.float64).reshape(-1)
        m = xy.size // 2
        assoc = np.full(max(m, 1), -1, dtype=np.int32)
        dist = np.zeros(max(m, 1))
        second = C.c_int(0)
        best, second_d = np.zeros(max(m, 1)), np.zeros(max(m, 1))
        rc = self._L.slam_ref_sensor_ex(self._h, m, _dp(xy), _ip(assoc), _dp(dist), C.byref(second), _dp(best),
                                        _dp(second_d))
        self.second_branch_taken += second.value
        self.last_min = dist[:m].tolist()
        self.last_best = best[:m].copy()          # smallest distance over the existing landmarks, per measurement
        self.last_second = second_d[:m].copy()    # second smallest
        if rc:
            self._raise(rc)
        return assoc[:m].tolist()

    # outputs ----------------------------------------------------------------------------
    def _vec3(self, fn):
        out = np.zeros(3)
        fn(self._h, _dp(out))
        return out

    @property
    def state(self):
        out = np.zeros(self.n)
        self._L.slam_ref_get_state(self._h, _dp(out))
        return out

    @property
    def sigma(self):
        out = np.zeros((self.n, self.n))
        self._L.slam_ref_get_sigma(self._h, _dp(out))
        return out

    @property
    def counter_obstacles(self):
        return self._L.slam_ref_count(self._h)

    @property
    def t_map_odom(self):
        return self._vec3(self._L.slam_ref_get_map_odom)

    @property
    def t_odom_robot(self):
        return self._vec3(self._L.slam_ref_get_odom)

    def composed_pose(self):
        x, y, th = self._vec3(self._L.slam_ref_composed_pose)
        return th, x, y

    def last_tf(self):
        out = np.zeros(4)
        self._L.slam_ref_last_tf(self._h, _dp(out))
        return out

    def published_markers(self):
        cap = self.n_landmarks + 64
        xy = np.zeros(2 * cap)
        ids = np.zeros(cap, dtype=np.int32)
        k = self._L.slam_ref_published_markers(self._h, _dp(xy), _ip(ids), cap)
        return ids[:k].copy(), xy[:2 * k].reshape(-1, 2).copy()

    # fixtures ---------------------------------------------------------------------------
    def set_state(self, state, count):
        s = np.ascontiguousarray(state, dtype=np.float64)
        assert s.size == self.n
        self._L.slam_ref_set_state(self._h, _dp(s), int(count))

    def set_sigma(self, sigma):
        s = np.ascontiguousarray(sigma, dtype=np.float64)
        assert s.shape == (self.n, self.n)
        self._L.slam_ref_set_sigma(self._h, _dp(s))

    def set_previous(self, x, y, theta=0.0):
        self._L.slam_ref_set_previous(self._h, float(x), float(y), float(theta))

    def set_map_odom(self, x, y, theta):
        self._L.slam_ref_set_map_odom(self._h, float(x), float(y), float(theta))
