#!/usr/bin/env python
"""Probe: can the latency-bound sequential phase of one filter hide under the covariance pass of
another?  Two independent N-landmark filters on two streams, scans issued alternately, against one
filter alone.  (Feasibility check for pipelining seq(s+1) under downdate(s) inside ONE filter.)"""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as ekf  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--landmarks", type=int, default=10000)
    ap.add_argument("--m", type=int, default=20)
    ap.add_argument("--steps", type=int, default=200)
    a = ap.parse_args()
    N, m, K = a.landmarks, a.m, a.steps
    side = int(np.ceil(np.sqrt(N)))
    hs, streams, inp = [], [], []
    for i in range(2):
        st = torch.cuda.Stream()
        f = ekf.EkfSlam(n_landmarks=N, device=0, max_scan=m)
        f.set_stream(st.cuda_stream)
        f.fill_procedural(7, N, side)
        poses, xy, ids = bench.make_workload(f.state(), N, m, K + 8, 107)
        f.set_previous_xy(poses[0][1], poses[0][2])
        d_pose, d_xy = torch.from_numpy(poses).cuda(), torch.from_numpy(xy).cuda()
        d_assoc = torch.full((K + 8, m), -9, dtype=torch.int32, device="cuda")
        hs.append(f); streams.append(st); inp.append((d_pose, d_xy, d_assoc))

    def scan(i, s):
        d_pose, d_xy, d_assoc = inp[i]
        hs[i].scan_dev(d_pose[s].data_ptr(), m, d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr())

    for s in range(4):
        scan(0, s); scan(1, s)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(4, 4 + K // 2):
        scan(0, s)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    for s in range(4 + K // 2, 4 + K):
        scan(0, s); scan(1, s)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    alone = (t1 - t0) / (K // 2) * 1e3
    both = (t2 - t1) / (K - K // 2) / 2 * 1e3
    print(f"one filter alone: {alone:.4f} ms/scan; two filters interleaved: {both:.4f} ms/scan each "
          f"(ratio {alone / both:.3f})")


if __name__ == "__main__":
    main()
