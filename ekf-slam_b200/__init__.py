"""ekf_slam_b200 - Python host mirror of the C ABI (include/ekf_slam_b200.h).

The product is the sm_100a shared library `libekf_slam_b200.so` next to this file; this module
only binds it with ctypes so that tests and bench.py can drive it.  There is NO fallback: if the
library is missing or no B200 is present, construction raises.

The directory is called `ekf-slam_b200` (repo convention); import it as `ekf_slam_b200` through
the loader at the repo root (`import ekf_slam_b200`), see /ekf_slam_b200.py.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# EKF_SLAM_B200_LIB selects another build of the same library for A/B measurements (tools/near_tie_probe.py:
# the FMA-free association variant).  It must be a library built from this tree; there is no other backend.
LIB_PATH = os.environ.get("EKF_SLAM_B200_LIB") or os.path.join(_HERE, "libekf_slam_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "ekf_slam_b200.h")

EKF_OK, EKF_ERR_INVALID, EKF_ERR_CAPACITY, EKF_ERR_EMPTY_SCAN, EKF_ERR_BAD_ID, EKF_ERR_CUDA, \
    EKF_ERR_NO_DEVICE, EKF_ERR_NOMEM = range(8)

MARKER_ADD, MARKER_DELETE = 0, 2


class EkfError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str = ""):
        self.status = status
        super().__init__(f"{where}: status {status} ({status_string(status)}) {detail}".strip())


class EkfConfig(C.Structure):
    _fields_ = [("n_landmarks", C.c_int32), ("q_noise", C.c_double), ("r_noise", C.c_double),
                ("mah_threshold", C.c_double), ("prior", C.c_double), ("device", C.c_int32),
                ("max_scan", C.c_int32), ("shard_rank", C.c_int32), ("shard_count", C.c_int32)]


_lib = None


def build(force: bool = False) -> str:
    """Compile the library in-tree with nvcc for sm_100a (ekf-slam_b200/Makefile)."""
    if force:
        subprocess.check_call(["make", "-C", _HERE, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return LIB_PATH


def load() -> C.CDLL:
    """Load the C-ABI library; raises if it was never built (no silent fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
    u8p = C.POINTER(C.c_uint8)
    i32, i64p = C.c_int32, C.POINTER(C.c_int64)
    sigs = {
        "ekf_default_config": (None, [C.POINTER(EkfConfig)]),
        "ekf_status_string": (C.c_char_p, [C.c_int]),
        "ekf_version": (C.c_char_p, []),
        "ekf_bag_open": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(vp)]),
        "ekf_bag_close": (C.c_int, [vp]),
        "ekf_bag_last_error": (C.c_char_p, []),
        "ekf_bag_counts": (C.c_int, [vp, i64p, i64p, i64p, i64p, i64p]),
        "ekf_bag_event": (C.c_int, [vp, C.c_int64, ip, ip, C.POINTER(C.c_uint64)]),
        "ekf_bag_joint_state": (C.c_int, [vp, i32, i64p, ip, i32, dp]),
        "ekf_bag_detections": (C.c_int, [vp, i32, ip, i32, dp, ip, ip]),
        "ekf_selftest_division": (C.c_int, [i32, C.c_uint64, i32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
        "ekf_selftest_atan2": (C.c_int, [i32, C.c_uint64, i32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
        "ekf_create": (C.c_int, [C.POINTER(EkfConfig), C.POINTER(vp)]),
        "ekf_destroy": (C.c_int, [vp]),
        "ekf_set_stream": (C.c_int, [vp, vp]),
        "ekf_set_overlap": (C.c_int, [vp, i32]),
        "ekf_synchronize": (C.c_int, [vp]),
        "ekf_last_status": (C.c_int, [vp]),
        "ekf_last_cuda_error": (C.c_char_p, [vp]),
        "ekf_predict": (C.c_int, [vp, C.c_double, C.c_double, C.c_double]),
        "ekf_update_known": (C.c_int, [vp, i32, ip, dp, u8p]),
        "ekf_update_unknown": (C.c_int, [vp, i32, dp, ip, ip]),
        "ekf_scan_dev": (C.c_int, [vp, vp, i32, vp, vp, vp, vp]),
        "ekf_scan": (C.c_int, [vp, dp, i32, dp, ip, u8p, ip, ip, dp]),
        "ekf_state_dim": (C.c_int, [vp, ip]),
        "ekf_get_pose": (C.c_int, [vp, dp]),
        "ekf_get_state": (C.c_int, [vp, dp, i32]),
        "ekf_get_count": (C.c_int, [vp, ip]),
        "ekf_get_landmarks": (C.c_int, [vp, dp, i32, ip]),
        "ekf_get_sigma_block": (C.c_int, [vp, i32, i32, i32, i32, dp]),
        "ekf_set_state": (C.c_int, [vp, dp, i32]),
        "ekf_set_count": (C.c_int, [vp, i32]),
        "ekf_set_previous_xy": (C.c_int, [vp, C.c_double, C.c_double]),
        "ekf_get_previous_xy": (C.c_int, [vp, dp]),
        "ekf_set_sigma_block": (C.c_int, [vp, i32, i32, i32, i32, dp]),
        "ekf_fill_procedural": (C.c_int, [vp, C.c_uint64, i32, i32]),
        "ekf_mahalanobis": (C.c_int, [vp, C.c_double, C.c_double, dp, i32, C.POINTER(i32)]),
        "ekf_mahalanobis_noalloc": (C.c_int, [vp, C.c_double, C.c_double, dp, i32, C.POINTER(i32)]),
        "ekf_sigma_storage": (C.c_int, [vp, i64p, i64p]),
        "ekf_launch_count": (C.c_int, [vp, i64p]),
        "ekf_read_timing": (C.c_int, [vp, C.POINTER(C.c_float), C.POINTER(C.c_float), i32, ip]),
        "ekf_enable_timing": (C.c_int, [vp, i32]),
        "ekf_read_timing_shard": (C.c_int, [vp, C.POINTER(C.c_float), C.POINTER(C.c_float), i32, ip]),
        "ekf_shard_info": (C.c_int, [vp, ip, ip, i64p]),
        "ekf_shard_sync_begin": (C.c_int, [vp, C.POINTER(vp), i64p]),
        "ekf_shard_sync_finish": (C.c_int, [vp]),
        "ekf_shard_scan_begin": (C.c_int, [vp, vp, i32, vp, vp, i32, C.POINTER(vp), i64p]),
        "ekf_shard_peer_export": (C.c_int, [vp, vp, C.POINTER(vp), i64p]),
        "ekf_shard_peer_attach": (C.c_int, [vp, i32, vp, vp]),
        "ekf_shard_scan_finish": (C.c_int, [vp, vp, vp, vp, vp]),
        "ekf_shard_scan_finish_ex": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, C.c_int64, vp]),
        "ekf_landmarks_create": (C.c_int, [i32, i32, i32, i32, C.POINTER(vp)]),
        "ekf_landmarks_destroy": (C.c_int, [vp]),
        "ekf_landmarks_detect": (C.c_int, [vp, i32, C.POINTER(C.c_float), C.c_float, C.c_float, C.c_double, dp, ip, ip]),
        "ekf_landmarks_detect_dev": (C.c_int, [vp, i32, vp, C.c_float, C.c_float, C.c_double, vp, vp, vp, vp]),
        "ekf_nusim_create": (C.c_int, [i32, dp, C.c_double, C.c_double, i32, C.POINTER(vp)]),
        "ekf_nusim_destroy": (C.c_int, [vp]),
        "ekf_nusim_wall_lines": (C.c_int, [vp, dp]),
        "ekf_nusim_fake_sensor": (C.c_int, [vp, i32, dp, C.c_double, C.c_double, C.c_uint64, dp, dp, ip]),
        "ekf_nusim_fake_sensor_dev": (C.c_int, [vp, i32, vp, C.c_double, C.c_double, C.c_uint64, vp, vp, vp, vp]),
        "ekf_nusim_lidar": (C.c_int, [vp, i32, dp, C.c_double, C.c_double, C.c_uint64, dp, C.POINTER(C.c_float)]),
        "ekf_nusim_lidar_dev": (C.c_int, [vp, i32, vp, C.c_double, C.c_double, C.c_uint64, vp, vp, vp]),
        "ekf_batch_create": (C.c_int, [C.POINTER(EkfConfig), i32, C.POINTER(vp)]),
        "ekf_batch_destroy": (C.c_int, [vp]),
        "ekf_batch_set_stream": (C.c_int, [vp, vp]),
        "ekf_batch_synchronize": (C.c_int, [vp]),
        "ekf_batch_step": (C.c_int, [vp, dp, i32, dp, ip, ip, ip, ip, dp]),
        "ekf_batch_step_dev": (C.c_int, [vp, vp, i32, vp, vp, vp, vp, vp, vp]),
        "ekf_batch_last_status": (C.c_int, [vp]),
        "ekf_batch_launch_count": (C.c_int, [vp, i64p]),
        "ekf_batch_get_state": (C.c_int, [vp, i32, dp, i32]),
        "ekf_batch_get_sigma": (C.c_int, [vp, i32, dp]),
        "ekf_batch_get_count": (C.c_int, [vp, i32, ip]),
        "ekf_batch_set_state": (C.c_int, [vp, i32, dp, i32]),
        "ekf_batch_set_sigma": (C.c_int, [vp, i32, dp]),
        "ekf_batch_set_count": (C.c_int, [vp, i32, i32]),
        "ekf_batch_set_previous_xy": (C.c_int, [vp, i32, C.c_double, C.c_double]),
        "ekf_batch_step_bytes": (C.c_int, [vp, i32, i64p]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def status_string(status: int) -> str:
    return load().ekf_status_string(int(status)).decode()


def selftest_division(count: int = 1 << 22, seed: int = 1, device: int = -1):
    """(mismatches, outside_guard) of the kernels' shared-reciprocal division against IEEE `/` on the device."""
    mis, slow = C.c_uint64(0), C.c_uint64(0)
    st = load().ekf_selftest_division(device, seed, count, C.byref(mis), C.byref(slow))
    if st != 0:
        raise RuntimeError(f"ekf_selftest_division: {status_string(st)}")
    return int(mis.value), int(slow.value)


def selftest_atan2(count: int = 1 << 22, seed: int = 1, device: int = -1):
    """(mismatches, outside_guard) of the kernels' branch-free atan2 against the CUDA library's on the device."""
    mis, slow = C.c_uint64(0), C.c_uint64(0)
    st = load().ekf_selftest_atan2(device, seed, count, C.byref(mis), C.byref(slow))
    if st != 0:
        raise RuntimeError(f"ekf_selftest_atan2: {status_string(st)}")
    return int(mis.value), int(slow.value)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class EkfSlam:
    """One EKF-SLAM filter on one B200.  Method names follow the C ABI; argument meaning and
    error behaviour follow the reference callbacks (nuslam/src/slam.cpp:180-316, 318-530)."""

    def __init__(self, n_landmarks: int = 50, q_noise: float = 1.0e-2, r_noise: float = 1.0e-2,
                 mah_threshold: float = 2.0, prior: float = 10e6, device: int = -1, max_scan: int = 0,
                 shard_rank: int = 0, shard_count: int = 1):
        self._L = load()
        cfg = EkfConfig()
        self._L.ekf_default_config(C.byref(cfg))
        cfg.n_landmarks, cfg.q_noise, cfg.r_noise = n_landmarks, q_noise, r_noise
        cfg.mah_threshold, cfg.prior, cfg.device, cfg.max_scan = mah_threshold, prior, device, max_scan
        cfg.shard_rank, cfg.shard_count = shard_rank, shard_count
        self.shard_rank, self.shard_count = shard_rank, shard_count
        h = C.c_void_p()
        rc = self._L.ekf_create(C.byref(cfg), C.byref(h))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_create")
        self._h = h
        self.n_landmarks = n_landmarks
        self.n = 3 + 2 * n_landmarks

    # -- lifetime ---------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._L.ekf_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc: int, where: str, allow=()):
        if rc != EKF_OK and rc not in allow:
            detail = self._L.ekf_last_cuda_error(self._h).decode() if rc == EKF_ERR_CUDA else ""
            raise EkfError(rc, where, detail)
        return rc

    def set_stream(self, cuda_stream_ptr: int):
        self._check(self._L.ekf_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "ekf_set_stream")

    def set_overlap(self, on: bool = True) -> bool:
        """Overlap the covariance pass of a scan with the sequential side of the next one (ekf_set_overlap).
        Returns False (and stays on the classic path) when the second covariance buffer does not fit."""
        rc = self._L.ekf_set_overlap(self._h, 1 if on else 0)
        if rc == 7:          # EKF_ERR_NOMEM
            return False
        self._check(rc, "ekf_set_overlap")
        return bool(on)

    def synchronize(self):
        self._check(self._L.ekf_synchronize(self._h), "ekf_synchronize")

    def last_status(self) -> int:
        return self._L.ekf_last_status(self._h)

    # -- the scan update --------------------------------------------------------------------------
    def predict(self, theta: float, x: float, y: float):
        self._check(self._L.ekf_predict(self._h, float(theta), float(x), float(y)), "ekf_predict")

    def update_known(self, ids, xy, action=None):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1)
        act = None
        if action is not None:
            action = np.ascontiguousarray(action, dtype=np.uint8)
            act = action.ctypes.data_as(C.POINTER(C.c_uint8))
        self._check(self._L.ekf_update_known(self._h, len(ids), _ip(ids), _dp(xy), act), "ekf_update_known")

    def update_unknown(self, xy, allow_capacity_error: bool = False):
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1)
        m = len(xy) // 2
        assoc = np.full(m, -7, dtype=np.int32)
        cnt = C.c_int32(0)
        rc = self._L.ekf_update_unknown(self._h, m, _dp(xy), _ip(assoc), C.byref(cnt))
        self._check(rc, "ekf_update_unknown", allow=(EKF_ERR_CAPACITY,) if allow_capacity_error else ())
        return assoc

    def scan(self, pose, xy, ids=None, action=None):
        """predict + update in one call; returns (assoc, count, pose_after)."""
        pose = np.ascontiguousarray(pose, dtype=np.float64)
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1)
        m = len(xy) // 2
        idp = actp = None
        if ids is not None:
            ids = np.ascontiguousarray(ids, dtype=np.int32)
            idp = _ip(ids)
        if action is not None:
            action = np.ascontiguousarray(action, dtype=np.uint8)
            actp = action.ctypes.data_as(C.POINTER(C.c_uint8))
        assoc = np.full(m, -7, dtype=np.int32)
        cnt = C.c_int32(0)
        out_pose = np.zeros(3)
        self._check(self._L.ekf_scan(self._h, _dp(pose), m, _dp(xy), idp, actp, _ip(assoc), C.byref(cnt),
                                     _dp(out_pose)), "ekf_scan")
        return assoc, int(cnt.value), out_pose

    def scan_dev(self, pose_ptr: int, m: int, xy_ptr: int, ids_ptr: int = 0, action_ptr: int = 0,
                 assoc_ptr: int = 0):
        """asynchronous scan from raw device pointers (ints)."""
        vp = C.c_void_p
        self._check(self._L.ekf_scan_dev(self._h, vp(pose_ptr), m, vp(xy_ptr), vp(ids_ptr or None),
                                         vp(action_ptr or None), vp(assoc_ptr or None)), "ekf_scan_dev")

    # -- accessors ----------------------------------------------------------------------------------
    @property
    def count(self) -> int:
        c = C.c_int32(0)
        self._check(self._L.ekf_get_count(self._h, C.byref(c)), "ekf_get_count")
        return int(c.value)

    @count.setter
    def count(self, value: int):
        self._check(self._L.ekf_set_count(self._h, int(value)), "ekf_set_count")

    def pose(self) -> np.ndarray:
        out = np.zeros(3)
        self._check(self._L.ekf_get_pose(self._h, _dp(out)), "ekf_get_pose")
        return out

    def state(self, n: int | None = None) -> np.ndarray:
        n = self.n if n is None else n
        out = np.zeros(n)
        self._check(self._L.ekf_get_state(self._h, _dp(out), n), "ekf_get_state")
        return out

    def set_state(self, values):
        values = np.ascontiguousarray(values, dtype=np.float64)
        self._check(self._L.ekf_set_state(self._h, _dp(values), len(values)), "ekf_set_state")

    def landmarks(self) -> np.ndarray:
        c = self.count
        out = np.zeros((max(c, 1), 2))
        cc = C.c_int32(0)
        self._check(self._L.ekf_get_landmarks(self._h, _dp(out), c, C.byref(cc)), "ekf_get_landmarks")
        return out[:c]

    def sigma_block(self, i0: int, i1: int, j0: int, j1: int) -> np.ndarray:
        out = np.zeros((i1 - i0, j1 - j0))
        self._check(self._L.ekf_get_sigma_block(self._h, i0, i1, j0, j1, _dp(out)), "ekf_get_sigma_block")
        return out

    def sigma(self) -> np.ndarray:
        return self.sigma_block(0, self.n, 0, self.n)

    def set_sigma_block(self, i0: int, j0: int, block):
        block = np.ascontiguousarray(block, dtype=np.float64)
        i1, j1 = i0 + block.shape[0], j0 + block.shape[1]
        self._check(self._L.ekf_set_sigma_block(self._h, i0, i1, j0, j1, _dp(block)), "ekf_set_sigma_block")

    def set_sigma(self, full):
        self.set_sigma_block(0, 0, full)

    def set_previous_xy(self, x: float, y: float):
        self._check(self._L.ekf_set_previous_xy(self._h, float(x), float(y)), "ekf_set_previous_xy")

    def previous_xy(self) -> np.ndarray:
        out = np.zeros(2)
        self._check(self._L.ekf_get_previous_xy(self._h, _dp(out)), "ekf_get_previous_xy")
        return out

    def mahalanobis(self, x: float, y: float) -> np.ndarray:
        """distances of the body-frame sighting (x, y) to every mapped landmark (slam.cpp:358-404), read-only"""
        out = np.zeros(max(self.n_landmarks, 1))
        c = C.c_int32(0)
        self._check(self._L.ekf_mahalanobis(self._h, float(x), float(y), _dp(out), self.n_landmarks, C.byref(c)),
                    "ekf_mahalanobis")
        return out[:c.value].copy()

    def fill_procedural(self, seed: int, count: int, grid_w: int):
        self._check(self._L.ekf_fill_procedural(self._h, C.c_uint64(seed), count, grid_w), "ekf_fill_procedural")

    # -- row-sharded large map (see ekf_slam_b200.shard for the drivers) --------------------------
    def shard_info(self):
        """(row_begin, row_end, stored_doubles): internal padded row range this handle stores"""
        a, b, c = C.c_int32(0), C.c_int32(0), C.c_int64(0)
        self._check(self._L.ekf_shard_info(self._h, C.byref(a), C.byref(b), C.byref(c)), "ekf_shard_info")
        return int(a.value), int(b.value), int(c.value)

    def shard_sync_begin(self):
        ptr, cnt = C.c_void_p(), C.c_int64(0)
        self._check(self._L.ekf_shard_sync_begin(self._h, C.byref(ptr), C.byref(cnt)), "ekf_shard_sync_begin")
        return int(ptr.value or 0), int(cnt.value)

    def shard_sync_finish(self):
        self._check(self._L.ekf_shard_sync_finish(self._h), "ekf_shard_sync_finish")

    def shard_scan_begin(self, pose_ptr: int, m: int, xy_ptr: int, ids_ptr: int = 0, last_round: bool = True):
        ptr, cnt = C.c_void_p(), C.c_int64(0)
        vp = C.c_void_p
        self._check(self._L.ekf_shard_scan_begin(self._h, vp(pose_ptr or None), m, vp(xy_ptr), vp(ids_ptr or None),
                                                 1 if last_round else 0, C.byref(ptr), C.byref(cnt)),
                    "ekf_shard_scan_begin")
        return int(ptr.value or 0), int(cnt.value)

    def shard_peer_export(self):
        """(64-byte CUDA IPC handle, device pointer, bytes) of this rank's peer-visible exchange region"""
        buf = C.create_string_buffer(64)
        base, nbytes = C.c_void_p(0), C.c_int64(0)
        self._check(self._L.ekf_shard_peer_export(self._h, buf, C.byref(base), C.byref(nbytes)),
                    "ekf_shard_peer_export")
        return bytes(buf.raw), int(base.value), int(nbytes.value)

    def shard_peer_attach(self, peer_rank: int, ipc_handle: bytes = None, base_ptr: int = 0):
        """another rank's region: its IPC handle (other process) or its device pointer (same process)"""
        buf = C.create_string_buffer(ipc_handle, 64) if ipc_handle is not None else None
        self._check(self._L.ekf_shard_peer_attach(self._h, peer_rank, buf, C.c_void_p(base_ptr or None)),
                    "ekf_shard_peer_attach")

    def shard_scan_finish(self, assoc_ptr: int = 0, done_ptr: int = 0, pose_ptr: int = 0, count_ptr: int = 0,
                          results_dev: int = 0, results_host: int = 0, results_bytes: int = 0, results_event: int = 0):
        """results_*: device block -> pinned host copy and event, enqueued before the covariance pass"""
        vp = C.c_void_p
        self._check(self._L.ekf_shard_scan_finish_ex(self._h, vp(assoc_ptr or None), vp(done_ptr or None),
                                                     vp(pose_ptr or None), vp(count_ptr or None),
                                                     vp(results_dev or None), vp(results_host or None),
                                                     results_bytes, vp(results_event or None)),
                    "ekf_shard_scan_finish_ex")

    def storage(self):
        a, b = C.c_int64(0), C.c_int64(0)
        self._check(self._L.ekf_sigma_storage(self._h, C.byref(a), C.byref(b)), "ekf_sigma_storage")
        return int(a.value), int(b.value)

    def launch_count(self) -> int:
        a = C.c_int64(0)
        self._check(self._L.ekf_launch_count(self._h, C.byref(a)), "ekf_launch_count")
        return int(a.value)

    def enable_timing(self, ring_scans: int):
        self._check(self._L.ekf_enable_timing(self._h, int(ring_scans)), "ekf_enable_timing")

    def read_timing_shard(self, max_scans: int):
        """(assoc_ms[], push_ms[]) of the begin phase of the most recent row-sharded scans (synchronises)."""
        a = np.zeros(max_scans, dtype=np.float32)
        b = np.zeros(max_scans, dtype=np.float32)
        n = C.c_int32(0)
        fp = C.POINTER(C.c_float)
        self._check(self._L.ekf_read_timing_shard(self._h, a.ctypes.data_as(fp), b.ctypes.data_as(fp), max_scans,
                                                  C.byref(n)), "ekf_read_timing_shard")
        return a[:n.value].astype(np.float64), b[:n.value].astype(np.float64)

    def read_timing(self, max_scans: int):
        """(seq_ms[], downdate_ms[]) of the most recent scans, oldest first (synchronises)."""
        a = np.zeros(max_scans, dtype=np.float32)
        b = np.zeros(max_scans, dtype=np.float32)
        n = C.c_int32(0)
        fp = C.POINTER(C.c_float)
        self._check(self._L.ekf_read_timing(self._h, a.ctypes.data_as(fp), b.ctypes.data_as(fp), max_scans,
                                            C.byref(n)), "ekf_read_timing")
        return a[:n.value].astype(np.float64), b[:n.value].astype(np.float64)


class EkfBatch:
    """B independent small filters on one B200, one CTA each (BASELINE config 3).  Mirrors the
    ekf_batch_* entry points; per filter the semantics are those of EkfSlam.scan."""

    def __init__(self, n_filters: int, n_landmarks: int = 20, q_noise: float = 1.0e-2, r_noise: float = 1.0e-2,
                 mah_threshold: float = 2.0, prior: float = 10e6, device: int = -1):
        self._L = load()
        cfg = EkfConfig()
        self._L.ekf_default_config(C.byref(cfg))
        cfg.n_landmarks, cfg.q_noise, cfg.r_noise = n_landmarks, q_noise, r_noise
        cfg.mah_threshold, cfg.prior, cfg.device = mah_threshold, prior, device
        h = C.c_void_p()
        rc = self._L.ekf_batch_create(C.byref(cfg), n_filters, C.byref(h))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_batch_create")
        self._h = h
        self.n_filters, self.n_landmarks, self.n = n_filters, n_landmarks, 3 + 2 * n_landmarks

    def close(self):
        if getattr(self, "_h", None):
            self._L.ekf_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc, where, allow=()):
        if rc != EKF_OK and rc not in allow:
            raise EkfError(rc, where)
        return rc

    def set_stream(self, cuda_stream_ptr: int):
        self._check(self._L.ekf_batch_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "ekf_batch_set_stream")

    def synchronize(self):
        self._check(self._L.ekf_batch_synchronize(self._h), "ekf_batch_synchronize")

    def last_status(self) -> int:
        return self._L.ekf_batch_last_status(self._h)

    def step(self, poses, xy, ids=None, m_count=None, allow=(), out=None):
        """poses [B,3] (theta, x, y); xy [B,m,2]; ids [B,m] or None; m_count [B] or None.
        Returns (assoc [B,m], count [B], pose_after [B,3]).  Page-locked inputs / `out` arrays (e.g. numpy views of
        torch pin_memory tensors) are copied from / into directly, pageable ones through the handle's staging."""
        B = self.n_filters
        poses = np.ascontiguousarray(poses, dtype=np.float64).reshape(B, 3)
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(B, -1, 2)
        m = xy.shape[1]
        idp = mcp = None
        if ids is not None:
            ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(B, m)
            idp = _ip(ids)
        if m_count is not None:
            m_count = np.ascontiguousarray(m_count, dtype=np.int32).reshape(B)
            mcp = _ip(m_count)
        if out is None:
            assoc = np.full((B, m), -7, dtype=np.int32)
            cnt = np.zeros(B, dtype=np.int32)
            pose = np.zeros((B, 3))
        else:       # caller-owned (ideally page-locked: copied into directly) int32 [B,m], int32 [B], float64 [B,3]
            assoc, cnt, pose = out
            assert assoc.dtype == np.int32 and assoc.shape == (B, m) and assoc.flags.c_contiguous
            assert cnt.dtype == np.int32 and cnt.shape == (B,) and pose.dtype == np.float64 and pose.shape == (B, 3)
        self._check(self._L.ekf_batch_step(self._h, _dp(poses), m, _dp(xy), idp, mcp, _ip(assoc), _ip(cnt),
                                           _dp(pose)), "ekf_batch_step", allow)
        return assoc, cnt, pose

    def step_dev(self, poses_ptr: int, m: int, xy_ptr: int, ids_ptr: int = 0, m_count_ptr: int = 0,
                 assoc_ptr: int = 0, count_ptr: int = 0, pose_ptr: int = 0):
        vp = C.c_void_p
        self._check(self._L.ekf_batch_step_dev(self._h, vp(poses_ptr), m, vp(xy_ptr), vp(ids_ptr or None),
                                               vp(m_count_ptr or None), vp(assoc_ptr or None),
                                               vp(count_ptr or None), vp(pose_ptr or None)), "ekf_batch_step_dev")

    def state(self, index: int) -> np.ndarray:
        out = np.zeros(self.n)
        self._check(self._L.ekf_batch_get_state(self._h, index, _dp(out), self.n), "ekf_batch_get_state")
        return out

    def sigma(self, index: int) -> np.ndarray:
        out = np.zeros((self.n, self.n))
        self._check(self._L.ekf_batch_get_sigma(self._h, index, _dp(out)), "ekf_batch_get_sigma")
        return out

    def count(self, index: int) -> int:
        c = C.c_int32(0)
        self._check(self._L.ekf_batch_get_count(self._h, index, C.byref(c)), "ekf_batch_get_count")
        return int(c.value)

    def set_state(self, index: int, values):
        values = np.ascontiguousarray(values, dtype=np.float64)
        self._check(self._L.ekf_batch_set_state(self._h, index, _dp(values), len(values)), "ekf_batch_set_state")

    def set_sigma(self, index: int, full):
        full = np.ascontiguousarray(full, dtype=np.float64)
        self._check(self._L.ekf_batch_set_sigma(self._h, index, _dp(full)), "ekf_batch_set_sigma")

    def set_count(self, index: int, count: int):
        self._check(self._L.ekf_batch_set_count(self._h, index, int(count)), "ekf_batch_set_count")

    def set_previous_xy(self, index: int, x: float, y: float):
        self._check(self._L.ekf_batch_set_previous_xy(self._h, index, float(x), float(y)),
                    "ekf_batch_set_previous_xy")

    def step_bytes(self, m: int) -> int:
        a = C.c_int64(0)
        self._check(self._L.ekf_batch_step_bytes(self._h, m, C.byref(a)), "ekf_batch_step_bytes")
        return int(a.value)

    def launch_count(self) -> int:
        a = C.c_int64(0)
        self._check(self._L.ekf_batch_launch_count(self._h, C.byref(a)), "ekf_batch_launch_count")
        return int(a.value)


class LandmarkExtractor:
    """Batched LaserScan -> circle detections (ekf_landmarks_*; nuslam/src/landmarks.cpp:58-156)."""

    def __init__(self, max_scans: int, n_ranges: int = 360, max_out: int = 16, device: int = -1):
        self._L = load()
        h = C.c_void_p()
        rc = self._L.ekf_landmarks_create(max_scans, n_ranges, max_out, device, C.byref(h))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_landmarks_create")
        self._h, self.max_scans, self.n_ranges, self.max_out = h, max_scans, n_ranges, max_out

    def close(self):
        if getattr(self, "_h", None):
            self._L.ekf_landmarks_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def detect(self, ranges, angle_min: float, angle_increment: float, cluster_threshold: float = 0.2):
        """ranges [n_scans, n_ranges] float32 -> (list of [k,3] arrays, status [n_scans])"""
        ranges = np.ascontiguousarray(ranges, dtype=np.float32).reshape(-1, self.n_ranges)
        ns = ranges.shape[0]
        out = np.zeros((ns, self.max_out, 3))
        cnt = np.zeros(ns, dtype=np.int32)
        st = np.zeros(ns, dtype=np.int32)
        rc = self._L.ekf_landmarks_detect(self._h, ns, ranges.ctypes.data_as(C.POINTER(C.c_float)),
                                          C.c_float(angle_min), C.c_float(angle_increment), cluster_threshold,
                                          _dp(out), _ip(cnt), _ip(st))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_landmarks_detect")
        return [out[s, :min(int(cnt[s]), self.max_out)].copy() for s in range(ns)], st


class NuSim:
    """nusim's sensor models on the device for many robots at once (ekf_nusim_*): the fake landmark sensor
    (nusim/src/nusim.cpp:310-349) and the 360-beam lidar (nusim.cpp:559-710)."""

    def __init__(self, obstacles_xyr, arena_x: float, arena_y: float, device: int = -1):
        self._L = load()
        obs = np.ascontiguousarray(obstacles_xyr, dtype=np.float64).reshape(-1, 3)
        h = C.c_void_p()
        rc = self._L.ekf_nusim_create(obs.shape[0], _dp(obs), arena_x, arena_y, device, C.byref(h))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_nusim_create")
        self._h, self.n_obstacles = h, obs.shape[0]

    def close(self):
        if getattr(self, "_h", None):
            self._L.ekf_nusim_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def wall_lines(self):
        out = np.zeros((4, 4))
        rc = self._L.ekf_nusim_wall_lines(self._h, _dp(out))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_nusim_wall_lines")
        return out

    def fake_sensor(self, poses, max_range: float, noise_std: float = 0.0, seed: int = 0, noise=None):
        """poses [n, 3] (theta, x, y) -> (xy [n, n_obstacles, 2], ids [n, n_obstacles])"""
        poses = np.ascontiguousarray(poses, dtype=np.float64).reshape(-1, 3)
        n = poses.shape[0]
        xy = np.zeros((n, self.n_obstacles, 2))
        ids = np.zeros((n, self.n_obstacles), dtype=np.int32)
        nz = None if noise is None else np.ascontiguousarray(noise, dtype=np.float64).reshape(n, self.n_obstacles, 2)
        rc = self._L.ekf_nusim_fake_sensor(self._h, n, _dp(poses), max_range, noise_std, seed,
                                           None if nz is None else _dp(nz), _dp(xy), _ip(ids))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_nusim_fake_sensor")
        return xy, ids

    def lidar(self, poses, max_range: float, noise_std: float = 0.0, seed: int = 0, noise=None):
        """poses [n, 3] (theta, x, y) -> ranges [n, 360] float32"""
        poses = np.ascontiguousarray(poses, dtype=np.float64).reshape(-1, 3)
        n = poses.shape[0]
        out = np.zeros((n, 360), dtype=np.float32)
        nz = None if noise is None else np.ascontiguousarray(noise, dtype=np.float64).reshape(n, 360)
        rc = self._L.ekf_nusim_lidar(self._h, n, _dp(poses), max_range, noise_std, seed,
                                     None if nz is None else _dp(nz), out.ctypes.data_as(C.POINTER(C.c_float)))
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_nusim_lidar")
        return out

    def fake_sensor_dev(self, n, poses_ptr, max_range, noise_std, seed, xy_ptr, ids_ptr, stream=0, noise_ptr=0):
        rc = self._L.ekf_nusim_fake_sensor_dev(self._h, n, poses_ptr, max_range, noise_std, seed, noise_ptr or None,
                                               xy_ptr, ids_ptr, stream or None)
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_nusim_fake_sensor_dev")

    def lidar_dev(self, n, poses_ptr, max_range, noise_std, seed, ranges_ptr, stream=0, noise_ptr=0):
        rc = self._L.ekf_nusim_lidar_dev(self._h, n, poses_ptr, max_range, noise_std, seed, noise_ptr or None,
                                         ranges_ptr, stream or None)
        if rc != EKF_OK:
            raise EkfError(rc, "ekf_nusim_lidar_dev")
