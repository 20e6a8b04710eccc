#!/usr/bin/env python
"""Per-kernel durations of the row-sharded scan on ONE GPU: R handles emulate R ranks (ekf_slam_b200.shard.LocalShards,
peer-memory exchange with plain pointers), so `ncu --metrics gpu__time_duration.sum` can list every launch of a scan
(ncu cannot wrap a multi-rank command).  Each handle's kernels do exactly the work one rank does.

    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/X.csv \
        python tools/shard_launch_probe.py --shards 8 --scans 6
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
sys.argv0 = sys.argv[:]
ap = argparse.ArgumentParser()
ap.add_argument("--shards", type=int, default=8)
ap.add_argument("--landmarks", type=int, default=10000)
ap.add_argument("--m", type=int, default=20)
ap.add_argument("--scans", type=int, default=6)
ap.add_argument("--no-peer", action="store_true")
a = ap.parse_args()
sys.argv = sys.argv[:1]
import torch  # noqa: E402
import bench  # noqa: E402
from ekf_slam_b200 import shard  # noqa: E402

N, m = a.landmarks, a.m
g = shard.LocalShards(N, a.shards, peer=not a.no_peer, max_scan=m)
g.fill_procedural(7, N, int(np.ceil(np.sqrt(N))))
g.sync()
poses, xy, ids = bench.make_workload(g.state(), N, m, a.scans, 107)
g.set_previous_xy(poses[0][1], poses[0][2])
ok = True
for s in range(a.scans):
    assoc, cnt, _, rounds = g.scan(poses[s], xy[s])
    ok = ok and list(assoc) == list(ids[s]) and rounds == 1
torch.cuda.synchronize()
print("association ok:", ok)
g.close()
