"""End-to-end cost of ekf_batch_step per route: pageable arrays (staged) vs page-locked arrays (copied directly)."""
import os, sys, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as ekf
from ekf_slam_b200 import synth
B, N, m = 65536, 20, 20
g = ekf.EkfBatch(B, n_landmarks=N)
fields = synth.batch_fields(B, N, seed=1)
rng = np.random.default_rng(2)
P, X = [], []
for s in range(8):
    p = synth.batch_pose(B, s); xy, _ = synth.batch_scan(fields, p, rng, m=N if s < 2 else m)
    P.append(p); X.append(xy)
for s in range(3): g.step(P[s], X[s])
t0 = time.perf_counter()
for s in range(3, 8): g.step(P[s], X[s])
print("pageable ms/step", (time.perf_counter() - t0) / 5 * 1e3)
tp = [torch.from_numpy(p).pin_memory() for p in P]; tx = [torch.from_numpy(x).pin_memory() for x in X]
ta = torch.empty((B, m), dtype=torch.int32).pin_memory(); tc = torch.empty((B,), dtype=torch.int32).pin_memory(); tq = torch.empty((B, 3), dtype=torch.float64).pin_memory()
out = (ta.numpy(), tc.numpy(), tq.numpy())
for rep in range(2):
    t0 = time.perf_counter()
    for s in range(3, 8): g.step(tp[s].numpy(), tx[s].numpy(), out=out)
    print("pinned ms/step", (time.perf_counter() - t0) / 5 * 1e3)
t0 = time.perf_counter()
for s in range(3, 8): g.step(tp[s].numpy(), tx[s].numpy())
print("pinned in, pageable out ms/step", (time.perf_counter() - t0) / 5 * 1e3)
