"""bench.py's contract lines that can be checked without a GPU: the reference arm (the CPU oracle timed on the host
cores) prints ONE JSON line with the driver's keys, and the product arm refuses to run without a CUDA device."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          env=env, timeout=600)


def test_reference_arm_line():
    out = _run("--impl", "reference", "--landmarks", "400", "--steps", "2", "--warmup", "1")
    assert out.returncode == 0, out.stderr
    lines = [l for l in out.stdout.strip().split("\n") if l.strip()]
    assert len(lines) == 1                                   # exactly one JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "ekf_scan_updates_per_s_N10k" and d["unit"] == "scan-updates/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1
    assert d["value"] > 0 and abs(d["ms_per_step"] * d["value"] - 1000.0) < 1e-6 * 1000.0
    assert d["vs_baseline"] is None and d["dtype"] == "f64" and d["data"] == "synthetic"
    assert d["config"]["landmarks"] == 400 and "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0
    # every step is a WHOLE scan (all m measurements): `steps` means steps, nothing is projected
    assert "whole scan" in cb["sample"] and cb["association_equals_expected_ids"] is True
    assert d["scaling"] == "strong"
    # the reference's own compiled slam.cpp is timed at its native size beside it
    native = cb["others"][0]
    assert native["kind"] == "reference" and native["value"] > 0 and native["cores"] == 1


def test_both_arms_describe_the_workload_identically():
    """the driver compares the two arms' `config` objects: both come from the one function"""
    sys.path.insert(0, ROOT)
    import importlib
    bench = importlib.import_module("bench")
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"config": config_of(N, m)') == 2
    c = bench.config_of(10000, 20)
    assert set(c) == {"workload", "landmarks", "m", "map", "l2"} and "model" not in c


def test_reference_arm_under_torchrun_only_rank0_prints():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = _run("--impl", "reference", "--gpus", "2", "--landmarks", "400", "--steps", "1", "--warmup", "0", env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_product_arm_has_no_cpu_path():
    import torch
    if torch.cuda.is_available():
        return
    out = _run("--steps", "1", "--warmup", "0", "--landmarks", "100")
    assert out.returncode != 0
    assert "no CUDA device" in (out.stderr + out.stdout)
