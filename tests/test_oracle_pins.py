"""Pins the oracle's boundary arithmetic: normalize_angle and SE(2) helpers of oracle A
(oracle/ekf_literal.py) and oracle B (oracle/ekf_struct.c) against the reference's Catch2 vectors
(turtlelib/tests/test_geometry2d.cpp:8-17, test_se2d.cpp, test_diff_drive.cpp) and against the
reference's own turtlelib object code compiled into oracle/_ref."""
import ctypes as C
import math
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ekf_literal as lit  # noqa: E402
import struct_oracle as so  # noqa: E402

PI = lit.PI
VECTORS = [(0, 0), (PI, PI), (-PI, PI), (-PI / 4, -PI / 4), (3 * PI / 2, -PI / 2), (-3 * PI / 2, PI / 2),
           (-5 * PI / 2, -PI / 2), (5 * PI / 2, PI / 2)]


def test_normalize_angle_reference_vectors():
    L = so.lib()
    for x, want in VECTORS:
        assert abs(lit.normalize_angle(x) - want) < 1e-8
        assert abs(L.ekfo_normalize_angle(float(x)) - want) < 1e-8


def test_normalize_angle_bitwise_vs_reference_object_code():
    ref = so.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    L = so.lib()
    rng = np.random.default_rng(0)
    for v in [0.0, PI, -PI, 3 * PI, -3 * PI, 1e-310, 7.0, -7.0] + list(rng.uniform(-100, 100, 5000)):
        r = ref.ref_normalize_angle(float(v))
        assert lit.normalize_angle(float(v)) == r
        assert L.ekfo_normalize_angle(float(v)) == r


def test_se2_and_odometry_bitwise_vs_reference_object_code():
    ref = so.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    dp = C.POINTER(C.c_double)
    rng = np.random.default_rng(1)
    for _ in range(300):
        a, b = rng.uniform(-3, 3, 3), rng.uniform(-3, 3, 3)
        out = np.zeros(3)
        ref.ref_compose(a.ctypes.data_as(dp), b.ctypes.data_as(dp), out.ctypes.data_as(dp))
        t = lit.Transform2D(*a) * lit.Transform2D(*b)
        assert (t.x, t.y, t.rot) == tuple(out)
        ref.ref_inverse(a.ctypes.data_as(dp), out.ctypes.data_as(dp))
        t = lit.Transform2D(*a).inv()
        assert (t.x, t.y, t.rot) == tuple(out)
    wheels = np.cumsum(rng.uniform(-0.3, 0.5, (200, 2)), axis=0)
    out = np.zeros(600)
    flat = wheels.reshape(-1).copy()
    ref.ref_fkin_sequence(0.16, 0.033, 200, flat.ctypes.data_as(dp), out.ctypes.data_as(dp))
    dd = lit.DiffDrive(0.16, 0.033)
    for i, (l, r) in enumerate(wheels):
        t = dd.fkin(float(l), float(r))
        assert (t.x, t.y, t.rot) == tuple(out[3 * i:3 * i + 3])


def test_diff_drive_reference_vectors():
    dd = lit.DiffDrive(0.1, 0.2)
    t = dd.fkin((0.4 * PI / 4.0) / (2.0 * PI * 0.2) * 2.0 * PI, (0.2 * PI / 4.0) / (2.0 * PI * 0.2) * 2.0 * PI)
    assert abs(t.x - 0.15) < 1e-8 and abs(t.y + 0.15) < 1e-8 and abs(t.rot + PI / 2) < 1e-8
    t = lit.integrate_twist(1.0, 1.0, -1.0)
    assert abs(t.x - (-math.cos(1.0) + math.sin(1.0) + 1)) < 1e-3
