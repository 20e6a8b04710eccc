#!/usr/bin/env python
"""Generates the golden input/output vectors of the EKF hot path FROM THE REFERENCE'S OWN SOURCE:
oracle/_ref/libslam_ref.so is /root/reference/nuslam/src/slam.cpp compiled unmodified (oracle/Makefile,
oracle/slam_ref_shim.cpp, stub ROS / Armadillo headers under oracle/stubs/); this script drives its
Slam::fake_sensor_cb (slam.cpp:180-316) and Slam::sensor_cb (:318-530) and records what they leave in
the node's members.

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz (needs /root/reference)

nuslam ships no test vectors of its own (nuslam/CMakeLists.txt:37-47 is lint only), so these are the
pins: oracle A, oracle B, the numpy models and the CUDA path are all held to them.  The reference tree
does not exist on the GPU box, which is why the vectors are committed.

Each file holds: the inputs (per scan: composed odometry (x, y, theta) as the node shell passes it,
body-frame sightings, ids, action flags, ragged via `m`), and the outputs (association per sighting,
final state, final Sigma, counter_obstacles, and the pose after every scan).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]

import slam_ref  # noqa: E402
from ekf_slam_b200 import synth  # noqa: E402

CASES = {
    # BASELINE config 1: nusim circle drive around 4 landmarks, known association (fake sensor, with DELETE flags)
    "config1_known": dict(known=True, N=50, kw=dict(n_scans=60, seed=0, max_range=1.2)),
    # BASELINE config 2 stand-in: detections only, shuffled, Mahalanobis association
    "config2_unknown": dict(known=False, N=50, kw=dict(n_scans=60, seed=1, drop_out_of_range=True, shuffle=True)),
    # config 3 shape: one filter of the batch, N = 20 landmarks
    "config3_n43_unknown": dict(known=False, N=20, field=(20, 21), kw=dict(n_scans=25, seed=5, drop_out_of_range=True,
                                                                            shuffle=True, max_range=1.8)),
}


def generate(name, known, N, kw, field=None):
    if field is not None:
        kw = dict(kw, landmarks=synth.scattered_field(field[0], seed=field[1], extent=4.0, min_sep=0.5))
    scans = synth.circle_drive(**kw)
    mmax = max(len(s.xy) for s in scans)
    S = len(scans)
    odom = np.zeros((S, 3)); m = np.zeros(S, dtype=np.int32)
    xy = np.zeros((S, mmax, 2)); ids = np.full((S, mmax), -1, dtype=np.int32); act = np.zeros((S, mmax), dtype=np.uint8)
    assoc = np.full((S, mmax), -1, dtype=np.int32); poses = np.zeros((S, 3))
    f = slam_ref.RefSlam(n_landmarks=N, sensor_source="sim" if known else "assoc")
    for i, s in enumerate(scans):
        k = len(s.xy)
        odom[i] = s.odom; m[i] = k
        xy[i, :k] = s.xy; ids[i, :k] = s.ids; act[i, :k] = s.action
        f.set_odometry(*s.odom)
        if known:
            f.fake_sensor_cb([(int(a), x, y, int(c)) for a, (x, y), c in zip(s.ids, s.xy, s.action)])
            assoc[i, :k] = s.ids
        elif k:
            assoc[i, :k] = f.sensor_cb([(x, y) for (x, y) in s.xy])
        poses[i] = f.state[:3]
    np.savez_compressed(os.path.join(HERE, name + ".npz"), known=known, n_landmarks=N, odom=odom, m=m, xy=xy, ids=ids,
                        action=act, assoc=assoc, poses=poses, state=f.state.copy(), sigma=f.sigma.copy(),
                        count=f.counter_obstacles, t_map_odom=f.t_map_odom, source="reference slam.cpp via oracle/_ref")
    assert f.second_branch_taken == 0   # slam.cpp:423-432 is unreachable
    print(name, "scans", S, "max m", mmax, "count", f.counter_obstacles)


if __name__ == "__main__":
    for name, c in CASES.items():
        generate(name, c["known"], c["N"], c["kw"], c.get("field"))
