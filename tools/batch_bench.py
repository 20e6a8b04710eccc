#!/usr/bin/env python
"""Quick device-resident timing of the batched mode (development aid; bench.py reports the number)."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import ekf_slam_b200 as ekf
from ekf_slam_b200 import synth

ap = argparse.ArgumentParser()
ap.add_argument("--filters", type=int, default=65536)
ap.add_argument("--landmarks", type=int, default=20)
ap.add_argument("--m", type=int, default=20)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--warmup", type=int, default=4)
a = ap.parse_args()
B, N, m = a.filters, a.landmarks, a.m
g = ekf.EkfBatch(B, n_landmarks=N)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); g.set_stream(stream.cuda_stream)
fields = synth.batch_fields(B, N, seed=1)
rng = np.random.default_rng(2)
T = a.warmup + a.steps
d_pose, d_xy, ids = [], [], []
for s in range(T):
    p = synth.batch_pose(B, s)
    xy, order = synth.batch_scan(fields, p, rng, m=N if s < 2 else m)   # first scans map everything
    d_pose.append(torch.from_numpy(p).cuda()); d_xy.append(torch.from_numpy(xy).cuda()); ids.append(order)
d_assoc = torch.zeros((B, N), dtype=torch.int32, device="cuda")
d_cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
def step(s):
    g.step_dev(d_pose[s].data_ptr(), d_xy[s].shape[1], d_xy[s].data_ptr(), 0, 0, d_assoc.data_ptr(), d_cnt.data_ptr(), 0)
for s in range(a.warmup):
    step(s)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for s in range(a.warmup, T):
    step(s)
e1.record(stream)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
byt = g.step_bytes(m)
print(json.dumps({"filters": B, "N": N, "m": m, "ms_per_step": ms, "filter_steps_per_s": B / ms * 1e3,
                  "GBps": B * byt / ms * 1e-6, "status": g.last_status(), "cnt_min": int(d_cnt.min()), "cnt_max": int(d_cnt.max())}))
