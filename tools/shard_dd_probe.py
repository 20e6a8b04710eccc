#!/usr/bin/env python
"""Per-shard covariance-pass time of a row-sharded map, all R shards emulated on ONE GPU
(LocalShards): which shard of the equal-area split is the slow one?"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as ekf  # noqa: E402
from ekf_slam_b200 import shard  # noqa: E402
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--landmarks", type=int, default=60000)
ap.add_argument("--shards", type=int, default=8)
ap.add_argument("--m", type=int, default=20)
ap.add_argument("--steps", type=int, default=6)
a = ap.parse_args()
N, R, m, K = a.landmarks, a.shards, a.m, a.steps
side = int(np.ceil(np.sqrt(N)))
drv = shard.LocalShards(N, R, 0, max_scan=m)
drv.fill_procedural(7, N, side)
drv.sync()
poses, xy, ids = bench.make_workload(drv.state(), N, m, K + 2, 107)
drv.set_previous_xy(poses[0][1], poses[0][2])
d_pose, d_xy = torch.from_numpy(poses).cuda(), torch.from_numpy(xy).cuda()
d_assoc = torch.full((K + 2, m), -9, dtype=torch.int32, device="cuda")
for s in range(2):
    drv.scan_dev(d_pose[s], d_xy[s], d_assoc[s], check=True)
for h in drv.handles():
    h.enable_timing(K)
for s in range(2, 2 + K):
    drv.scan_dev(d_pose[s], d_xy[s], d_assoc[s], check=True)
torch.cuda.synchronize()
for r, h in enumerate(drv.handles()):
    seq_ms, dd_ms = h.read_timing(K)
    stored, n_pad = h.storage()
    gb = stored * 8 * 2 / 1e9
    print(f"shard {r}: rows stored {stored * 8 / 1e9:7.2f} GB  seq {np.mean(seq_ms):.3f} ms  pass {np.mean(dd_ms):.3f} ms "
          f"-> {gb / np.mean(dd_ms) * 1e3 / 1e3:.2f} TB/s")
