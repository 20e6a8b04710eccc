"""ekf_mahalanobis_noalloc: the association sweep for callers that must not allocate (no counterpart in the reference).
Its staging lives in a file-scope static __device__ array instead of the
cudaMalloc'ed per-handle buffer of ekf_mahalanobis; same kernel, so the distances must be bit-identical (the tests'
tolerance for this quantity is 1e-12 relative; 0 is what is asserted).  Called through the C ABI, twice in a row - the
second call finds the first call's contents in the shared array.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

from _harness import ekf_slam_b200

pytestmark = pytest.mark.gpu

N, COUNT, SEED = 400, 390, 7


def _filter():
    f = ekf_slam_b200.EkfSlam(n_landmarks=N)
    f.fill_procedural(SEED, COUNT, int(np.ceil(np.sqrt(COUNT))))
    return f


def _noalloc(f, x, y, cap):
    L = ekf_slam_b200.load()
    out = np.full(max(cap, 1), np.nan)
    cnt = C.c_int32(-1)
    rc = L.ekf_mahalanobis_noalloc(f._h, float(x), float(y), out.ctypes.data_as(C.POINTER(C.c_double)), cap, C.byref(cnt))
    return rc, out, cnt.value


def test_noalloc_sweep_equals_the_allocating_entry_point_twice_in_a_row():
    f = _filter()
    st = f.state()
    rng = np.random.default_rng(3)
    # the pose the sweep linearises at is the filter's own; sightings near two different landmarks
    for call, k in enumerate((17, 301)):
        th, px, py = st[:3]
        dx, dy = st[3 + 2 * k] - px, st[4 + 2 * k] - py
        c, s = np.cos(th), np.sin(th)
        x, y = c * dx + s * dy + rng.normal(0, 1e-3), -s * dx + c * dy + rng.normal(0, 1e-3)
        want = f.mahalanobis(x, y)                       # existing entry point (cudaMalloc'ed staging)
        rc, got, cnt = _noalloc(f, x, y, N)
        assert rc == 0 and cnt == COUNT == len(want)
        assert np.isfinite(want).all() and int(np.argmin(want)) == k
        assert np.array_equal(got[:COUNT], want), call   # same kernel: bit-identical
        assert np.isnan(got[COUNT:]).all()               # nothing written past the count
    # a smaller cap returns the first `cap` distances only (and the second call's data, not the first's)
    rc, got, cnt = _noalloc(f, x, y, 10)
    assert rc == 0 and cnt == COUNT and np.array_equal(got[:10], want[:10])
    assert f.last_status() == 0
    f.close()


def test_noalloc_rejects_sizes_beyond_the_static_array():
    L = ekf_slam_b200.load()
    f = _filter()
    big = 65536 + 1                                     # EKF_NOALLOC_MAX_LANDMARKS + 1
    out = np.zeros(big)
    cnt = C.c_int32(0)
    rc = L.ekf_mahalanobis_noalloc(f._h, 1.0, 0.5, out.ctypes.data_as(C.POINTER(C.c_double)), big, C.byref(cnt))
    assert rc == ekf_slam_b200.EKF_ERR_INVALID
    assert b"EKF_NOALLOC_MAX_LANDMARKS" in L.ekf_last_cuda_error(f._h)
    # the handle is still usable, on both entry points
    rc, got, cnt = _noalloc(f, 1.0, 0.5, N)
    assert rc == 0 and np.array_equal(got[:COUNT], f.mahalanobis(1.0, 0.5))
    f.close()
