"""The C-ABI library loads and exports every function include/ekf_slam_b200.h declares
(no compute calls here: there is no GPU on the CPU test box)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    txt = open(os.path.join(ROOT, "include", "ekf_slam_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ekf_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(ekf_mod):
    if not os.path.exists(ekf_mod.LIB_PATH):
        ekf_mod.build()
    lib = ctypes.CDLL(ekf_mod.LIB_PATH)
    names = declared_functions()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_mirror_binds_every_symbol(ekf_mod):
    if not os.path.exists(ekf_mod.LIB_PATH):
        ekf_mod.build()
    L = ekf_mod.load()
    for n in declared_functions():
        assert getattr(L, n).argtypes is not None, n
    assert ekf_mod.status_string(2) == "landmark capacity exceeded"


def test_no_device_is_an_error_not_a_fallback(ekf_mod):
    import torch
    if torch.cuda.is_available():
        return
    try:
        ekf_mod.EkfSlam(n_landmarks=4)
    except ekf_mod.EkfError as e:
        assert e.status in (ekf_mod.EKF_ERR_NO_DEVICE, ekf_mod.EKF_ERR_CUDA)
    else:
        raise AssertionError("constructing a filter without a GPU must fail")
