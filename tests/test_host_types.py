"""Host-side boundary types (ekf-slam_b200/host/turtlelib/*.hpp): the reference's own Catch2
vectors, and bit-for-bit agreement with the reference's turtlelib object code (oracle/_ref)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
HOST_INC = os.path.join(ROOT, "ekf-slam_b200", "host")


def test_reference_vectors(tmp_path):
    exe = tmp_path / "tt"
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", "-Wextra", "-pedantic",
                           "-I" + HOST_INC, os.path.join(ROOT, "tests", "cpp", "test_turtlelib.cpp"), "-o", str(exe)])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout


def test_bitwise_against_reference_object_code(tmp_path):
    import struct_oracle as so
    ref = so.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    lib = tmp_path / "libhost.so"
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-I" + HOST_INC,
                           os.path.join(ROOT, "tests", "cpp", "host_shim.cpp"), "-o", str(lib)])
    mine = C.CDLL(str(lib))
    dp = C.POINTER(C.c_double)
    for L in (mine,):
        L.ref_normalize_angle.restype = C.c_double
        L.ref_normalize_angle.argtypes = [C.c_double]
    rng = np.random.default_rng(3)
    specials = [0.0, np.pi, -np.pi, 3 * np.pi, -3 * np.pi, 2 * np.pi, -2 * np.pi, 1e-300, -1e-300, 1e6, -1e6]
    for v in specials + list(rng.uniform(-50, 50, 2000)):
        a, b = mine.ref_normalize_angle(float(v)), ref.ref_normalize_angle(float(v))
        assert a == b, (v, a, b)

    def call(L, name, *arrs, out_n=3):
        out = np.zeros(out_n)
        getattr(L, name)(*[x.ctypes.data_as(dp) for x in arrs], out.ctypes.data_as(dp))
        return out
    for _ in range(500):
        a, b = rng.uniform(-3, 3, 3), rng.uniform(-3, 3, 3)
        assert np.array_equal(call(mine, "ref_compose", a, b), call(ref, "ref_compose", a, b))
        assert np.array_equal(call(mine, "ref_inverse", a), call(ref, "ref_inverse", a))
        assert np.array_equal(call(mine, "ref_apply_point", a, b[:2].copy(), out_n=2),
                              call(ref, "ref_apply_point", a, b[:2].copy(), out_n=2))
        tw = rng.uniform(-2, 2, 3)
        if _ % 7 == 0:
            tw[0] = 0.0
        assert np.array_equal(call(mine, "ref_integrate_twist", tw), call(ref, "ref_integrate_twist", tw))
    wheels = np.cumsum(rng.uniform(-0.3, 0.5, (400, 2)), axis=0).reshape(-1).copy()
    o1, o2 = np.zeros(1200), np.zeros(1200)
    for L, o in ((mine, o1), (ref, o2)):
        L.ref_fkin_sequence.argtypes = [C.c_double, C.c_double, C.c_int, dp, dp]
        L.ref_fkin_sequence(0.16, 0.033, 400, wheels.ctypes.data_as(dp), o.ctypes.data_as(dp))
    assert np.array_equal(o1, o2)
