"""Oracle for the landmark-extraction step: LaserScan -> clusters -> circle
test -> circle fit -> (c_x, c_y, r) detections, the step right before the EKF hot path.

TEST INFRASTRUCTURE ONLY (same rule as oracle/ekf_literal.py).

Restates, with numpy in place of Armadillo (svd / eigh / solve are the same LAPACK drivers):
  nuslam/src/landmarks.cpp:58-106   getClusters (including its two quirks: the reading that opens a new
                                    cluster is dropped, :80-85; the last cluster is merged into the first
                                    when the scan wraps, :93-103)
  nuslam/src/landmarks.cpp:109-156  laserCallback: size filter 4..39, circle test, fit, r < 0.2 and
                                    range < 2.0 filter
  turtlelib/src/landmark_detection.cpp:5-48    checkCircle (inscribed-angle mean / stddev test)
  turtlelib/src/landmark_detection.cpp:50-135  fitCircle (hyper-accurate algebraic fit)
PINNED: fit_circle reproduces the reference's own known answers turtlelib/tests/circle_tests.cpp:8-34
(tests/test_landmarks.py); the clustering / circle test have no reference vectors.
"""
from __future__ import annotations

import math

import numpy as np

PI = 3.14159265358979323846
LIDAR_OFFSET_X = 0.032            # landmarks.cpp:68
CLUSTER_THRESHOLD = 0.2           # landmarks.cpp:195


def normalize_angle(rad):
    d = math.fmod(rad + PI, 2 * PI)
    return d + PI if d <= 0.0 else d - PI


def scan_points(ranges, angle_min, angle_increment):
    """landmarks.cpp:66-69; ranges / angles are float32 in the message, promoted to double"""
    pts = []
    inc, amin = float(np.float32(angle_increment)), float(np.float32(angle_min))
    for i, r in enumerate(ranges):
        r = float(np.float32(r))
        ang = normalize_angle(float(i) * inc) + amin
        pts.append((r * math.cos(ang) - LIDAR_OFFSET_X, r * math.sin(ang)))
    return pts


def get_clusters(pts, thr=CLUSTER_THRESHOLD):
    """landmarks.cpp:58-106.  Raises IndexError where the reference's clusters.at(0) throws."""
    clusters, current, prev = [], [], (0.0, 0.0)
    for i, p in enumerate(pts):
        if i == 0:
            current.append(p)
        elif math.sqrt((p[0] - prev[0]) ** 2 + (p[1] - prev[1]) ** 2) <= thr:
            current.append(p)
        else:
            clusters.append(current)
            current = []                      # the reading that opened the gap is NOT kept (:80-85)
        prev = p
    first = clusters[0][0]                    # .at(0).at(0): throws on no gap / empty first cluster
    if math.sqrt((first[0] - prev[0]) ** 2 + (first[1] - prev[1]) ** 2) <= thr:
        clusters[0] = clusters[0] + current
    else:
        clusters.append(current)
    return clusters


def check_circle(cluster):
    """landmark_detection.cpp:5-48 (arma::stddev is the N-1 normalised one)"""
    s, e = cluster[0], cluster[-1]
    ang = []
    for j in range(1, len(cluster) - 1):
        p = cluster[j]
        a = math.sqrt((s[0] - p[0]) ** 2 + (s[1] - p[1]) ** 2)
        b = math.sqrt((e[0] - p[0]) ** 2 + (e[1] - p[1]) ** 2)
        c = math.sqrt((s[0] - e[0]) ** 2 + (s[1] - e[1]) ** 2)
        v = (c * c - a * a - b * b) / (-2.0 * a * b)
        ang.append(math.acos(v) if -1.0 <= v <= 1.0 else float("nan"))
    ang = np.array(ang)
    stdev = float(np.std(ang, ddof=1)) if len(ang) > 1 else 0.0
    mn = float(np.mean(ang))
    return bool(stdev < 0.2 and 1.3 < mn < 2.6)


def fit_circle(cluster):
    """landmark_detection.cpp:50-135"""
    c = np.array(cluster, dtype=np.float64)
    n = c.shape[0]
    means = c.mean(axis=0)
    c = c - means
    z = c[:, 0] ** 2 + c[:, 1] ** 2
    z_mean = z.mean()
    Z = np.stack([z, c[:, 0], c[:, 1], np.ones(n)], axis=1)
    H_inv = np.eye(4)
    H_inv[0, 0] = 0.0; H_inv[3, 0] = 0.5; H_inv[0, 3] = 0.5; H_inv[3, 3] = -2.0 * z_mean
    U, s, Vt = np.linalg.svd(Z)
    V = Vt.T
    if s.min() < 10.0e-12:
        A = V[:, 3]
    else:
        Y = V @ np.diag(s) @ V.T
        Q = Y @ H_inv @ Y
        eigval, eigvec = np.linalg.eigh((Q + Q.T) / 2.0)
        idx, min_val = 0, 10.0e6
        for i in range(4):
            if eigval[i] < min_val and eigval[i] > 0.0:
                min_val, idx = eigval[i], i
        A = np.linalg.solve(Y, eigvec[:, idx])
    a = -A[1] / 2.0 / A[0]
    b = -A[2] / 2.0 / A[0]
    R = math.sqrt((A[1] * A[1] + A[2] * A[2] - 4.0 * A[0] * A[3]) / 4.0 / A[0] / A[0])
    return a + means[0], b + means[1], R


def detect(ranges, angle_min, angle_increment, thr=CLUSTER_THRESHOLD):
    """landmarks.cpp:109-156: list of (c_x, c_y, r) in cluster order"""
    out = []
    for cl in get_clusters(scan_points(ranges, angle_min, angle_increment), thr):
        if 3 < len(cl) < 40 and check_circle(cl):
            cx, cy, r = fit_circle(cl)
            if r < 0.2 and math.sqrt(cx * cx + cy * cy) < 2.0:
                out.append((cx, cy, r))
    return out


def synth_scan(cylinders, radius, pose=(0.0, 0.0, 0.0), n=360, max_range=3.5, noise=0.0, rng=None, wall=2.5):
    """Input side of nusim's lidar (nusim/src/nusim.cpp:559-710): n beams from the lidar frame, nearest hit
    on cylinders or on a square arena of half-width `wall`; float32 ranges like sensor_msgs/LaserScan."""
    x0, y0, th0 = pose
    inc = 2 * math.pi / n
    out = np.zeros(n, dtype=np.float32)
    for i in range(n):
        ang = th0 + i * inc
        dx, dy = math.cos(ang), math.sin(ang)
        best = max_range
        for (cx, cy) in cylinders:
            fx, fy = x0 - cx, y0 - cy
            b = 2 * (fx * dx + fy * dy)
            cc = fx * fx + fy * fy - radius * radius
            disc = b * b - 4 * cc
            if disc >= 0:
                t = (-b - math.sqrt(disc)) / 2
                if 0 < t < best:
                    best = t
        for (wx, wy) in ((wall, None), (-wall, None), (None, wall), (None, -wall)):
            if wx is not None and abs(dx) > 1e-12:
                t = (wx - x0) / dx
                if 0 < t < best and abs(y0 + t * dy) <= wall:
                    best = t
            if wy is not None and abs(dy) > 1e-12:
                t = (wy - y0) / dy
                if 0 < t < best and abs(x0 + t * dx) <= wall:
                    best = t
        if noise and rng is not None:
            best += rng.normal(0.0, noise)
        out[i] = best
    return out, np.float32(0.0), np.float32(inc)
