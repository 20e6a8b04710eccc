#!/usr/bin/env python
"""CPU timing of the LITERAL dense algorithm of the reference (SURVEY 8(d) plan (i)):

  * the reference's own object code (oracle/_ref/libslam_ref.so = slam.cpp compiled unmodified; its matrix products are
    the stub Armadillo's plain loops, 1 core) at N = 50 (native) and N = 200;
  * oracle A (oracle/ekf_literal.py: the same dense expressions on numpy / OpenBLAS dgemm, all cores) at N = 50, 200,
    1,000 and 2,000 - every measurement pays the dense (I - K H) Sigma product (2 n^3 flop) and one dense H Sigma H^T
    per candidate landmark, as in slam.cpp:361-416,485-486;
  * the projection to N = 10,000 from the measured dgemm rate (clearly a projection: 1.6e13 flop per measurement).

    python tools/cpu_literal_timing.py > profiles/r2_cpu_literal_timing.json       (no GPU needed)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import ekf_literal as lit  # noqa: E402
import slam_ref  # noqa: E402
from procedural_map import ProceduralMap  # noqa: E402
import bench  # noqa: E402


def scan_inputs(N, m, seed=7):
    # capacity N + 1 with N mapped landmarks: the reference appends a tentative landmark before it associates
    # (slam.cpp:351-356), so a completely full map cannot take a sighting at all
    pm = ProceduralMap(seed, N + 1, N, int(np.ceil(np.sqrt(N))))
    poses, xy, ids = bench.make_workload(pm.state(), N, m, 2, seed + 100)
    return pm, poses, xy, ids


def time_oracle_a(N, m):
    pm, poses, xy, ids = scan_inputs(N, m)
    f = lit.LiteralEkf(n_landmarks=N + 1)
    f.state[:] = pm.state(); f.sigma[:] = pm.sigma0_rows(0, pm.n); f.counter_obstacles = N
    f.filter_previous_configuration = lit.Transform2D(poses[0][1], poses[0][2], 0.0)
    f.set_odometry(poses[0][1], poses[0][2], poses[0][0])
    t0 = time.perf_counter()
    a = f.sensor_cb([(x, y) for (x, y) in xy[0]])
    dt = time.perf_counter() - t0
    return dt, list(a) == list(ids[0])


def time_reference(N, m):
    if not slam_ref.available():
        return None, None
    pm, poses, xy, ids = scan_inputs(N, m)
    f = slam_ref.RefSlam(n_landmarks=N + 1)
    f.set_state(pm.state(), N); f.set_sigma(pm.sigma0_rows(0, pm.n))
    f.set_previous(poses[0][1], poses[0][2]); f.set_odometry(poses[0][1], poses[0][2], poses[0][0])
    t0 = time.perf_counter()
    a = f.sensor_cb([(x, y) for (x, y) in xy[0]])
    dt = time.perf_counter() - t0
    return dt, list(a) == list(ids[0])


def main():
    out = {"cores": os.cpu_count(), "where": "development container (no GPU); the GPU box's host has 16 cores", "rows": []}
    n = 4000
    A = np.random.default_rng(0).normal(size=(n, n)); B = A.T.copy()
    t0 = time.perf_counter(); A @ B; dg = time.perf_counter() - t0
    out["dgemm_gflops"] = 2.0 * n ** 3 / dg / 1e9
    for N, m in ((50, 4), (200, 10), (1000, 4), (2000, 2)):
        dt, ok = time_oracle_a(N, m)
        out["rows"].append({"impl": "oracle A (numpy / OpenBLAS, all cores)", "N": N, "n": 5 + 2 * N, "measurements": m,
                            "seconds_per_measurement": dt / m, "scans_per_s_at_m20": 1.0 / (dt / m * 20),
                            "association_as_expected": ok})
    for N, m in ((50, 4), (200, 2)):
        dt, ok = time_reference(N, m)
        if dt is not None:
            out["rows"].append({"impl": "reference slam.cpp object code (stub Armadillo: plain loops, 1 core)", "N": N,
                                "n": 5 + 2 * N, "measurements": m, "seconds_per_measurement": dt / m,
                                "scans_per_s_at_m20": 1.0 / (dt / m * 20), "association_as_expected": ok})
    n10 = 20003
    flop = 2.0 * n10 ** 3 + 10000 * 4.0 * n10 * n10          # (I - K H) Sigma + one dense H Sigma H^T per candidate
    out["projection_N10k"] = {"flop_per_measurement": flop, "seconds_per_measurement_at_measured_dgemm_rate":
                              flop / (out["dgemm_gflops"] * 1e9), "label": "PROJECTED, not measured"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
