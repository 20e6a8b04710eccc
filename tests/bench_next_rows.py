#!/usr/bin/env python
"""Throughput of the "next" rows either side of the filter, device-resident, next to their
CPU oracles on a bounded sample (which is why this script lives under tests/: only tests, smoke() and bench.py's
cpu_baseline may execute oracle/).  One JSON object on stdout:

  N2 lidar        nusim.cpp:559-710   one thread per (robot, beam)        ekf_nusim_lidar_dev
  N1 landmarks    landmarks.cpp:58-156 warp per scan + thread per cluster   ekf_landmarks_detect_dev
  N2 fake sensor  nusim.cpp:310-349   one thread per (robot, obstacle)    ekf_nusim_fake_sensor_dev
  chain           lidar -> landmarks without leaving the device (the Monte-Carlo producer loop)

Algorithmic bytes per scan: lidar 24 (pose) + 1,440 (360 float ranges); landmarks 1,440 + 16 x 24 + 8 (circles, count,
status); fake sensor 24 + n_obstacles x (16 + 4).  Timing: CUDA events on the launching stream, 3 warm-up + `--reps`
timed launches."""
import argparse
import ctypes as C
import json
import math
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ekf_slam_b200 as ekf  # noqa: E402
import landmarks_oracle as lo  # noqa: E402
import nusim_oracle as no  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scans", type=int, default=65536)
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--cpu-scans", type=int, default=48)
a = ap.parse_args()

ARENA_X, ARENA_Y, MAX_RANGE = 10.0, 5.0, 5.0                      # basic_world.yaml:5-6, nusim.cpp:90
OBST = np.array([[-0.5, -0.7, 0.05], [0.8, -0.8, 0.05], [0.4, 0.8, 0.05], [-0.6, 0.65, 0.05]])
n = a.scans
rng = np.random.default_rng(3)
poses = np.stack([rng.uniform(-math.pi, math.pi, n), rng.uniform(-1.2, 1.2, n), rng.uniform(-1.0, 1.0, n)], axis=1)
peak = 6650.0
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pk):
    peak = float(json.load(open(pk))["hbm_gbs"])

sim = ekf.NuSim(OBST, ARENA_X, ARENA_Y)
lm = ekf.LandmarkExtractor(n, 360, 16)
st = torch.cuda.Stream()
torch.cuda.set_stream(st)
d_poses = torch.from_numpy(poses).cuda()
d_ranges = torch.zeros((n, 360), dtype=torch.float32, device="cuda")
d_out = torch.zeros((n, 16, 3), dtype=torch.float64, device="cuda")
d_cnt = torch.zeros(n, dtype=torch.int32, device="cuda")
d_st = torch.zeros(n, dtype=torch.int32, device="cuda")
d_xy = torch.zeros((n, len(OBST), 2), dtype=torch.float64, device="cuda")
d_ids = torch.zeros((n, len(OBST)), dtype=torch.int32, device="cuda")


def lidar():
    sim.lidar_dev(n, d_poses.data_ptr(), MAX_RANGE, 0.001, 7, d_ranges.data_ptr(), stream=st.cuda_stream)


def detect():
    rc = lm._L.ekf_landmarks_detect_dev(lm._h, n, d_ranges.data_ptr(), C.c_float(0.0), C.c_float(2.0 * math.pi / 360.0),
                                        0.2, d_out.data_ptr(), d_cnt.data_ptr(), d_st.data_ptr(), st.cuda_stream)
    assert rc == 0


def fake():
    sim.fake_sensor_dev(n, d_poses.data_ptr(), 2.0, 0.001, 9, d_xy.data_ptr(), d_ids.data_ptr(), stream=st.cuda_stream)


def chain():
    lidar(); detect()


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(a.reps):
        fn()
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / a.reps


rows = {}
for name, fn, byt in (("lidar", lidar, 24 + 1440), ("landmarks", detect, 1440 + 16 * 24 + 8),
                      ("fake_sensor", fake, 24 + len(OBST) * 20), ("lidar_to_landmarks", chain, 24 + 2 * 1440 + 16 * 24 + 8)):
    ms = timed(fn)
    rows[name] = {"ms_per_launch": ms, "scans_per_s": n / ms * 1e3, "algorithmic_bytes_per_scan": byt,
                  "GBps": n * byt / ms * 1e-6, "hbm_frac": n * byt / ms * 1e-6 / peak}
found = int(d_cnt.sum().item())

# CPU oracles (numpy restatements of the reference's node code) on a bounded sample
k = a.cpu_scans
t0 = time.perf_counter()
cpu_ranges = [no.lidar(poses[r], OBST, ARENA_X, ARENA_Y, MAX_RANGE) for r in range(k)]
t1 = time.perf_counter()
cpu_found = [lo.detect(r, np.float32(0.0), np.float32(2.0 * math.pi / 360.0)) for r in cpu_ranges]
t2 = time.perf_counter()
rows["lidar"]["cpu_oracle_scans_per_s"] = k / (t1 - t0)
rows["landmarks"]["cpu_oracle_scans_per_s"] = k / (t2 - t1)
rows["lidar_to_landmarks"]["cpu_oracle_scans_per_s"] = k / (t2 - t0)
print(json.dumps({"scans_per_launch": n, "reps": a.reps, "hbm_peak_gbs": peak, "circles_found_last_launch": found,
                  "cpu_sample_scans": k, "cpu_cores": 1, "rows": rows}))
