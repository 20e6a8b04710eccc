"""Landmark extraction.

CPU: oracle/landmarks_oracle.py against the reference's OWN known answers for fitCircle
(turtlelib/tests/circle_tests.cpp:8-34, tolerance 1e-4 as in the reference) and against hand-checkable
clustering cases including the two quirks of getClusters (nuslam/src/landmarks.cpp:80-85, :93-103).
GPU: the CUDA kernel through the C ABI against the oracle on synthetic lidar scans (nusim-style ray casts):
same number of detections in the same order, (c_x, c_y, r) within 1e-7 absolute."""
import math
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import landmarks_oracle as lo  # noqa: E402

WORLD = [(-0.5, -0.7), (0.8, -0.8), (0.4, 0.8), (-0.6, 0.65)]     # nusim/config/basic_world.yaml:8-9


def test_fit_circle_reference_vector_1():
    cx, cy, r = lo.fit_circle([(1, 7), (2, 6), (5, 8), (7, 7), (9, 5), (3, 7)])     # circle_tests.cpp:8-20
    assert abs(cx - 4.615482) < 1e-4 and abs(cy - 2.807354) < 1e-4 and abs(r - 4.8275) < 1e-4


def test_fit_circle_reference_vector_2():
    cx, cy, r = lo.fit_circle([(-1, 0), (-0.3, -0.06), (0.3, 0.1), (1, 0)])          # circle_tests.cpp:22-34
    assert abs(cx - 0.4908357) < 1e-4 and abs(cy + 22.15212) < 1e-4 and abs(r - 22.17979) < 1e-4


def test_get_clusters_drops_the_gap_reading_and_merges_the_wrap():
    # 10 readings on a line with spacing 0.1 except one jump of 1.0 between index 4 and 5
    pts = [(0.1 * i + (1.0 if i >= 5 else 0.0), 0.0) for i in range(10)]
    cl = lo.get_clusters(pts, thr=0.2)
    assert [len(c) for c in cl] == [5, 4]          # reading 5 opened the gap and is dropped
    assert cl[1][0] == pts[6]
    # a ring: last reading close to the first -> the tail is appended to cluster 0
    ring = [(math.cos(a), math.sin(a)) for a in np.linspace(0, 2 * math.pi, 40, endpoint=False)]
    ring[20] = (3.0, 3.0)                           # two gaps: before and after the outlier
    cl = lo.get_clusters(ring, thr=0.2)
    assert len(cl) == 2 and len(cl[0]) == 20 + 18 and len(cl[1]) == 0
    with pytest.raises(IndexError):                 # no gap at all: the reference throws at clusters.at(0)
        lo.get_clusters([(0.01 * i, 0.0) for i in range(10)], thr=0.2)


def test_detect_finds_the_basic_world_cylinders():
    rng = np.random.default_rng(0)
    r, amin, inc = lo.synth_scan(WORLD, 0.05, noise=0.002, rng=rng)
    det = lo.detect(r, amin, inc)
    assert len(det) == 4
    for cx, cy, rad in det:
        d = min(math.hypot(cx + 0.032 - wx, cy - wy) for wx, wy in WORLD)
        assert d < 0.02 and abs(rad - 0.05) < 0.02


def _kernel_fit(cluster):
    """The arithmetic of the CUDA kernel's fit (csrc/ekf_landmarks.cuh, fit_circle), restated with loops: modified
    Gram-Schmidt of Z = [z x y 1] evaluated point by point (Z is never stored), SVD of the 4 x 4 factor R in place of
    the SVD of Z, then the reference's remaining steps on (V, s)."""
    c = np.array(cluster, dtype=np.float64)
    n = len(c)
    mx, my = sum(p[0] for p in c) / n, sum(p[1] for p in c) / n
    rows = []
    for x, y in c:
        xi, yi = x - mx, y - my
        rows.append([xi * xi + yi * yi, xi, yi, 1.0])
    zm = sum(r[0] for r in rows) / n
    R = np.zeros((4, 4))
    inv = [0.0] * 4
    for col in range(4):                                   # pass `col`: recompute every point's row up to this column
        acc = np.zeros(4)
        for r in rows:
            v = list(r)
            for k in range(col):                           # what the earlier columns left of the row
                q = v[k] * inv[k]
                for j in range(k + 1, 4):
                    v[j] -= R[k, j] * q
            for j in range(col, 4):
                acc[j] += v[col] * v[j]
        R[col, col] = math.sqrt(acc[col])
        inv[col] = 1.0 / R[col, col] if R[col, col] > 0.0 else 0.0
        for j in range(col + 1, 4):
            R[col, j] = acc[j] * inv[col]
    _, s, Vt = np.linalg.svd(R)
    V = Vt.T
    if s.min() < 10.0e-12:
        A = V[:, 3]
    else:
        H_inv = np.eye(4)
        H_inv[0, 0] = 0.0; H_inv[3, 0] = 0.5; H_inv[0, 3] = 0.5; H_inv[3, 3] = -2.0 * zm
        Y = V @ np.diag(s) @ V.T
        Q = Y @ H_inv @ Y
        w, E = np.linalg.eigh((Q + Q.T) / 2.0)
        idx, best = 0, 10.0e6
        for i in range(4):
            if 0.0 < w[i] < best:
                best, idx = w[i], i
        A = np.linalg.solve(Y, E[:, idx])
    a, b = -A[1] / 2.0 / A[0], -A[2] / 2.0 / A[0]
    return a + mx, b + my, math.sqrt((A[1] * A[1] + A[2] * A[2] - 4.0 * A[0] * A[3]) / 4.0 / A[0] / A[0])


def test_triangular_factor_route_of_the_kernel_equals_the_reference_route():
    """The kernel takes the SVD of R (Z = QR by on-the-fly modified Gram-Schmidt) instead of the SVD of Z
    (landmark_detection.cpp:79-83): same singular values and right singular vectors.  On the reference's two vectors
    and on every cluster of 24 synthetic scans that reaches the fit: |difference| <= 1e-12 (measured 1e-15), far inside
    the 1e-7 the GPU test holds the kernel to."""
    for pts in ([(1, 7), (2, 6), (5, 8), (7, 7), (9, 5), (3, 7)], [(-1, 0), (-0.3, -0.06), (0.3, 0.1), (1, 0)]):
        assert np.allclose(_kernel_fit(pts), lo.fit_circle(pts), rtol=1e-13, atol=0)
    ranges, amin, inc = _scans(24, seed=3)
    n_fit, worst = 0, 0.0
    for s in range(ranges.shape[0]):
        for cl in lo.get_clusters(lo.scan_points(ranges[s], amin, inc)):
            if 3 < len(cl) < 40 and lo.check_circle(cl):
                worst = max(worst, float(np.abs(np.array(_kernel_fit(cl)) - np.array(lo.fit_circle(cl))).max()))
                n_fit += 1
    assert n_fit > 60 and worst <= 1e-12, (n_fit, worst)


def _scans(n_scans, seed):
    rng = np.random.default_rng(seed)
    out, amin, inc = [], None, None
    for s in range(n_scans):
        pose = (rng.uniform(-0.3, 0.3), rng.uniform(-0.3, 0.3), rng.uniform(-math.pi, math.pi))
        extra = [tuple(rng.uniform(-1.8, 1.8, 2)) for _ in range(rng.integers(0, 4))]
        cyl = [c for c in WORLD + extra if math.hypot(c[0] - pose[0], c[1] - pose[1]) > 0.25]
        r, amin, inc = lo.synth_scan(cyl, 0.05, pose=pose, noise=0.0015, rng=rng)
        out.append(r)
    return np.stack(out), float(amin), float(inc)


@pytest.mark.gpu
def test_cuda_extraction_matches_oracle():
    import ekf_slam_b200
    ranges, amin, inc = _scans(48, seed=3)
    ex = ekf_slam_b200.LandmarkExtractor(max_scans=64, n_ranges=ranges.shape[1], max_out=16)
    det, st = ex.detect(ranges, amin, inc)
    assert not st.any()
    n_det = 0
    for s in range(ranges.shape[0]):
        want = np.array(lo.detect(ranges[s], amin, inc)).reshape(-1, 3)
        assert det[s].shape == want.shape, s
        assert np.allclose(det[s], want, rtol=0, atol=1e-7), (s, np.abs(det[s] - want).max())
        n_det += len(want)
    assert n_det > 100
    ex.close()


@pytest.mark.gpu
def test_cuda_fit_on_the_reference_vectors_and_no_gap_status():
    """the reference's two circle-fit vectors pushed through the kernel as tiny 'scans' is not possible
    (they are not range data), so the kernel's fit is pinned through the oracle above; here: statuses."""
    import ekf_slam_b200
    ex = ekf_slam_b200.LandmarkExtractor(max_scans=2, n_ranges=360, max_out=4)
    flat = np.full((2, 360), 1.0, dtype=np.float32)      # a perfect circle of readings: no gap anywhere
    r, amin, inc = lo.synth_scan(WORLD, 0.05)
    flat[1] = r
    det, st = ex.detect(flat, float(amin), float(inc))
    assert st[0] == ekf_slam_b200.EKF_ERR_EMPTY_SCAN and len(det[0]) == 0
    assert st[1] == 0 and len(det[1]) == len(lo.detect(r, amin, inc))
    ex.close()
