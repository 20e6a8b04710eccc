#!/usr/bin/env python
"""Soak test of the sequential kernel's hand-written grid barrier and hot-winner area.

    python tools/seq_soak.py [--scans 100000] [--out gpurun_out/seq_soak.json]

compute-sanitizer is closed on this GPU pool, so the evidence that the release/acquire barrier (csrc/ekf_seq.cuh
grid_barrier) and the hot-winner area are race-free is behavioural: the SAME random scans go through

  * a filter whose sequential kernel runs as ONE CTA-wide grid slice per SM in the default shape (the shape under
    test: up to 160 CTAs, hot-winner area on), and
  * a twin with another grid shape / hot area off (EKF_SEQ_THREADS, EKF_SEQ_BLOCKS, EKF_SEQ_HOT - read at create),

and after every chunk the two must agree BIT FOR BIT: every association index, the whole state, the counters and
sampled covariance blocks.  Any lost update, stale slot or early barrier exit shows up as a different bit.  Maps:
procedural fields of several sizes and seeds, scans with re-sightings and brand-new landmarks.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]


def workload(state0, count, m, n_scans, rng, grid_w, n_new):
    lm = state0[3:3 + 2 * count].reshape(-1, 2)
    cx = cy = grid_w / 2.0
    rad = max(1.5, grid_w / 5.0)
    poses = np.zeros((n_scans, 3)); xy = np.zeros((n_scans, m, 2))
    new_at = set(rng.choice(n_scans, size=min(n_new, n_scans), replace=False).tolist())
    for s in range(n_scans):
        ang = rng.uniform(0, 2 * np.pi)
        x, y = cx + rad * np.cos(ang) * rng.uniform(0.2, 1.0), cy + rad * np.sin(ang) * rng.uniform(0.2, 1.0)
        th = rng.uniform(-np.pi, np.pi)
        d = np.hypot(lm[:, 0] - x, lm[:, 1] - y)
        near = np.argpartition(d, m)[:m]
        rng.shuffle(near)
        c, sn = np.cos(th), np.sin(th)
        dx, dy = lm[near, 0] - x, lm[near, 1] - y
        xy[s] = np.stack([c * dx + sn * dy, -sn * dx + c * dy], axis=1) + rng.normal(0, 1e-3, (m, 2))
        if s in new_at:                       # one sighting of nothing mapped: far outside the field
            far = np.array([grid_w + 5.0 + 3.0 * len(new_at) * rng.uniform(), -7.0 - 40.0 * rng.uniform()])
            dd = far - np.array([x, y])
            xy[s, 0] = (c * dd[0] + sn * dd[1], -sn * dd[0] + c * dd[1])
        poses[s] = (th, x, y)
    return poses, xy


def make(ekf, N, env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return ekf.EkfSlam(n_landmarks=N, max_scan=24)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scans", type=int, default=100000)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "seq_soak.json"))
    a = ap.parse_args()
    import torch
    import ekf_slam_b200 as ekf
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    # (N, m, share of the scan budget, twin's environment)
    plans = [
        (10000, 20, 0.10, {"EKF_SEQ_HOT": 0}),
        (10000, 20, 0.05, {"EKF_SEQ_THREADS": 256}),
        (5000, 20, 0.10, {"EKF_SEQ_BLOCKS": 37, "EKF_SEQ_HOT": 0}),
        (2000, 24, 0.15, {"EKF_SEQ_BLOCKS": 1}),
        (1000, 20, 0.20, {"EKF_SEQ_BLOCKS": 3, "EKF_SEQ_HOT": 0}),
        (300, 7, 0.20, {"EKF_SEQ_THREADS": 256, "EKF_SEQ_BLOCKS": 2}),
        (60, 5, 0.20, {"EKF_SEQ_BLOCKS": 1, "EKF_SEQ_HOT": 0}),
    ]
    report, total, t_start = [], 0, time.time()
    for pi, (N, m, share, env) in enumerate(plans):
        S = max(200, int(a.scans * share))
        chunk = 500
        seed = 11 + pi
        rng = np.random.default_rng(seed)
        count = N - min(64, N // 4)
        grid_w = int(np.ceil(np.sqrt(count)))
        A, B = make(ekf, N, {}), make(ekf, N, env)
        for f in (A, B):
            f.set_stream(stream.cuda_stream)
            f.fill_procedural(seed, count, grid_w)
        state0 = A.state()
        done, mismatches, new_seen = 0, 0, 0
        while done < S:
            k = min(chunk, S - done)
            poses, xy = workload(state0, count, m, k, rng, grid_w, n_new=2 if A.count < N - 2 else 0)
            if done == 0:
                for f in (A, B):
                    f.set_previous_xy(poses[0][1], poses[0][2])
            d_pose, d_xy = torch.from_numpy(poses).cuda(), torch.from_numpy(xy).cuda()
            outs = []
            for f in (A, B):
                d_assoc = torch.full((k, m), -9, dtype=torch.int32, device="cuda")
                for s in range(k):
                    f.scan_dev(d_pose[s].data_ptr(), m, d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr())
                torch.cuda.synchronize()
                outs.append(d_assoc.cpu().numpy())
            same = np.array_equal(outs[0], outs[1]) and A.count == B.count
            sa, sb = A.state(), B.state()
            same = same and np.array_equal(sa, sb) and np.isfinite(sa).all()
            n = A.n
            for i0 in (0, min(n - 40, 3 + 2 * int(outs[0][-1, 0])), max(0, n - 40)):
                same = same and np.array_equal(A.sigma_block(i0, i0 + 40, 0, min(n, 600)), B.sigma_block(i0, i0 + 40, 0, min(n, 600)))
            same = same and A.last_status() == 0 and B.last_status() == 0
            mismatches += 0 if same else 1
            new_seen = A.count - count
            done += k
        total += done
        report.append({"N": N, "m": m, "scans": done, "twin_env": env, "chunks_with_a_difference": mismatches,
                       "new_landmarks_created": int(new_seen), "final_count": int(A.count)})
        print(report[-1], flush=True)
        A.close(); B.close()
    out = {"total_scans_per_filter": total, "filters_compared": 2 * len(plans), "seconds": time.time() - t_start,
           "all_bitwise_equal": all(r["chunks_with_a_difference"] == 0 for r in report), "plans": report,
           "lib": os.path.basename(ekf.LIB_PATH)}
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps({k: v for k, v in out.items() if k != "plans"}))
    sys.exit(0 if out["all_bitwise_equal"] else 1)


if __name__ == "__main__":
    main()
