#!/usr/bin/env python
"""Phase timing of the sequential kernel (N = 10,000, m = 20).  Needs the library built with
`make -C ekf-slam_b200 EXTRA=-DEKF_SEQ_PROF` (thread 0 of the first / last CTA stamps clock64() at the
phase boundaries of every measurement behind the association output); rebuild without it afterwards."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as ekf  # noqa: E402
import bench  # noqa: E402

N, m, K = 10000, 20, 40
side = int(np.ceil(np.sqrt(N)))
st = torch.cuda.Stream()
torch.cuda.set_stream(st)
f = ekf.EkfSlam(n_landmarks=N, device=0, max_scan=m)
f.set_stream(st.cuda_stream)
f.fill_procedural(7, N, side)
poses, xy, ids = bench.make_workload(f.state(), N, m, K, 107)
f.set_previous_xy(poses[0][1], poses[0][2])
d_pose, d_xy = torch.from_numpy(poses).cuda(), torch.from_numpy(xy).cuda()
WORDS = 32 + 2 * (2 * 24 * 8)                     # int32 words: assoc (padded to 32) + 2 CTAs x 24 x 8 stamps
d_assoc = torch.zeros((K, WORDS), dtype=torch.int32, device="cuda")
for s in range(K):
    f.scan_dev(d_pose[s].data_ptr(), m, d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr())
torch.cuda.synchronize()
raw = d_assoc.cpu().numpy()
full = raw[:, 32:].copy().view(np.int64).reshape(K, 2, 24, 8)[:, :, :m, :]
stamps = full[:, :, :, :7]
names = ["A: association math", "A: arg-min + slot publish", "B: grid barrier", "C: decision + record + W row",
         "D: columns, gain, state, panels", "end sync"]
d = np.diff(stamps[8:], axis=3).astype(np.float64)            # skip warm-up scans
per_t = (stamps[8:, :, 1:, 0] - stamps[8:, :, :-1, 0]).astype(np.float64)
ghz = 1.965
for c, lab in enumerate(["first CTA", "last CTA"]):
    print(lab, "cycles per measurement (mean over scans, t) -> us at %.3f GHz" % ghz)
    for i, nm in enumerate(names):
        print(f"  {nm:34s} {d[:, c, :, i].mean():8.0f}  {d[:, c, :, i].mean() / ghz / 1e3:6.2f} us")
    print(f"  {'whole measurement':34s} {per_t[:, c].mean():8.0f}  {per_t[:, c].mean() / ghz / 1e3:6.2f} us")
    c0 = (full[8:, c, :, 1] - full[8:, c, :, 3]).astype(np.float64)
    print("  C0 (slot loads arrive) by t:", " ".join(f"{c0[:, t].mean():.0f}" for t in range(m)))
    c1 = (full[8:, c, :, 7] - full[8:, c, :, 3]).astype(np.float64)
    c2 = (full[8:, c, :, 4] - full[8:, c, :, 7]).astype(np.float64)
    print("  C1 (slot read + reduce) by t:", " ".join(f"{c1[:, t].mean():.0f}" for t in range(m)))
    print("  C2 (record / W row + sync) by t:", " ".join(f"{c2[:, t].mean():.0f}" for t in range(m)))
    print("  phase D by t:", " ".join(f"{d[:, c, t, 4].mean():.0f}" for t in range(m)))
    print("  barrier by t:", " ".join(f"{d[:, c, t, 2].mean():.0f}" for t in range(m)))
