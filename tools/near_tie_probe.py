#!/usr/bin/env python
"""Near-tie association evidence, as numbers.

    python tools/near_tie_probe.py [--out gpurun_out/near_tie.json]

Runs, for the default build and for the FMA-free association build (`make -C ekf-slam_b200 nofma`), each in its own
process (EKF_SLAM_B200_LIB):
  * the threshold sweep of tests/test_near_tie.py: sightings whose REFERENCE distance is threshold * (1 +- 10^-k),
    k = 3..15 -> the smallest margin at which the GPU decision equals the reference's, the first at which it differs;
  * the distance sweep: max relative difference between ekf_mahalanobis and the reference's distances;
  * the sequential kernel's time at N = 10,000, m = 20 (what the FMA-free arithmetic costs).
The checker is oracle/_ref/libslam_ref.so (the reference's slam.cpp compiled unmodified).
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
    import numpy as np
    import test_near_tie as nt
    import ekf_slam_b200 as ekf
    state, S, count = nt.fixture_filter()
    p = nt.Probe(state, S, count)
    rows = nt.sweep(p, p.gpu)
    per_k = {}
    for (k, side, ra, ga, rd, gd) in rows:
        e = per_k.setdefault(k, {"agree": True, "max_rel_d_diff": 0.0})
        e["agree"] = e["agree"] and (ra == ga)
        e["max_rel_d_diff"] = max(e["max_rel_d_diff"], abs(gd - rd) / rd)
    fd = nt.first_disagreement(rows)
    rng = np.random.default_rng(3)
    worst, n_bit_equal, n_tot = 0.0, 0, 0
    for k in range(count):
        for _ in range(20):
            xy = nt.body_frame(state, k, rng.normal(0.0, 0.02, 2))
            _, rb, rs = p.reference(xy)
            _, gb, gs = p.gpu(xy)
            worst = max(worst, abs(gb - rb) / rb, abs(gs - rs) / rs)
            n_bit_equal += int(gb == rb) + int(gs == rs)
            n_tot += 2
    # cost: sequential kernel at N = 10,000, m = 20
    import torch
    sys.argv = ["bench.py"]
    import bench
    N, m, K = 10000, 20, 60
    f = ekf.EkfSlam(n_landmarks=N, max_scan=m)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    f.set_stream(stream.cuda_stream)
    f.fill_procedural(7, N, 100)
    poses, xy, ids = bench.make_workload(f.state(), N, m, K + 5, 107)
    f.set_previous_xy(poses[0][1], poses[0][2])
    d_pose, d_xy = torch.from_numpy(poses).cuda(), torch.from_numpy(xy).cuda()
    d_assoc = torch.zeros((K + 5, m), dtype=torch.int32, device="cuda")
    for s in range(5):
        f.scan_dev(d_pose[s].data_ptr(), m, d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr())
    f.enable_timing(K)
    for s in range(5, 5 + K):
        f.scan_dev(d_pose[s].data_ptr(), m, d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr())
    torch.cuda.synchronize()
    seq_ms, dd_ms = f.read_timing(K)
    ok = bool((d_assoc[5:5 + K].cpu().numpy() == ids[5:5 + K]).all())
    f.close()
    print(json.dumps({"lib": os.path.basename(ekf.LIB_PATH), "first_disagreeing_margin_exp": fd,
                      "per_margin_exp": per_k, "distance_sweep_max_rel_diff": worst,
                      "distance_sweep_bit_equal": [n_bit_equal, n_tot],
                      "seq_kernel_ms_N10k_m20": float(np.mean(seq_ms)), "pass_ms": float(np.mean(dd_ms)),
                      "assoc_ok_N10k": ok}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "near_tie.json"))
    ap.add_argument("--child", action="store_true")
    a = ap.parse_args()
    if a.child:
        return child()
    out = {}
    for name, lib in (("default_fma", "libekf_slam_b200.so"), ("no_fma", "libekf_slam_b200_nofma.so")):
        path = os.path.join(ROOT, "ekf-slam_b200", lib)
        if not os.path.exists(path):
            out[name] = {"unavailable": f"{lib} not built"}
            continue
        env = dict(os.environ, EKF_SLAM_B200_LIB=path)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env, capture_output=True, text=True)
        if r.returncode != 0:
            out[name] = {"failed": r.stderr[-2000:]}
        else:
            out[name] = json.loads(r.stdout.strip().splitlines()[-1])
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
