"""The C++ host path: nuslam::SlamCore (ekf-slam_b200/host/nuslam/slam_core.hpp, the node-shell adapter
of SURVEY.md 8(f) N3) linked against the C-ABI library, fed with wheel angles and marker lists like
the reference's callbacks, against oracle A driven by the same trace."""
import os
import subprocess

import numpy as np
import pytest

from _harness import ekf_slam_b200, run_literal, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    exe = str(tmp_path / "slam_core_replay")
    pkg = os.path.join(ROOT, "ekf-slam_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-pedantic", "-ffp-contract=off",
                           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(pkg, "host"),
                           os.path.join(ROOT, "tests", "cpp", "slam_core_replay.cpp"), "-o", exe,
                           "-L" + pkg, "-lekf_slam_b200", "-Wl,-rpath," + pkg])
    return exe


@pytest.mark.parametrize("known", [True, False])
def test_slam_core_replay_matches_oracle_a(tmp_path, known):
    if not os.path.exists(ekf_slam_b200.LIB_PATH):
        ekf_slam_b200.build()
    exe = _build(tmp_path)
    scans = synth.circle_drive(n_scans=60, seed=21) if known else \
        synth.circle_drive(n_scans=60, seed=22, drop_out_of_range=True, shuffle=True)
    scans = [s for s in scans if len(s.xy)]
    lines = [f"{int(known)} 50 {len(scans)}"]
    for s in scans:
        lines.append(f"{s.wheels[0]!r} {s.wheels[1]!r} {len(s.xy)}")
        for i, (x, y), a in zip(s.ids, s.xy, s.action):
            lines.append(f"{int(i)} {float(x)!r} {float(y)!r} {int(a)}")
    out = subprocess.run([exe], input="\n".join(lines) + "\n", capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    rows = out.stdout.strip().split("\n")
    aL, sL, SL, cL = run_literal(scans, known)
    last = rows[len(scans) - 1].split()
    pose = np.array([float(v) for v in last[:3]])
    assert np.allclose(pose, sL[:3], rtol=0, atol=1e-9)
    if not known:
        assert int(last[3]) == cL
        got = [[int(v) for v in r.split()[4:]] for r in rows[:len(scans)]]
        assert got == [list(map(int, a)) for a in aL]
    assert rows[-1].startswith("MAPODOM")


def test_replay_bag_program_matches_oracle_a(tmp_path):
    """tools/replay_bag.cpp: MCAP file -> rosbag::McapReader -> nuslam::SlamCore (C++ end to end, SURVEY 8f N3 + N4)
    against oracle A fed the same wheel angles and sightings."""
    import _mcap_writer as mw
    pkg = os.path.join(ROOT, "ekf-slam_b200")
    exe = str(tmp_path / "replay_bag")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-pedantic", "-ffp-contract=off",
                           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(pkg, "host"),
                           os.path.join(ROOT, "tools", "replay_bag.cpp"), "-o", exe,
                           "-L" + pkg, "-lekf_slam_b200", "-Wl,-rpath," + pkg, "-ldl"])
    scans = [s for s in synth.circle_drive(n_scans=60, seed=22, drop_out_of_range=True, shuffle=True) if len(s.xy)]
    channels = {1: ("/blue/joint_states", "sensor_msgs/msg/JointState"),
                2: ("/green/detected_obstacles", "visualization_msgs/msg/MarkerArray")}
    msgs, t = [], 1710718557269449803
    for s in scans:
        t += 200_000_000
        msgs.append((1, t, mw.encode_joint_state(t, ["wheel_left_joint", "wheel_right_joint"], s.wheels)))
        msgs.append((2, t + 1000, mw.encode_marker_array(t, [{"id": int(i), "x": float(x), "y": float(y)}
                                                             for i, (x, y) in zip(s.ids, s.xy)])))
    path = str(tmp_path / "bag_0.mcap")
    mw.write_mcap(path, channels, msgs, chunk_messages=16)
    out = subprocess.run([exe, path], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    rows = out.stdout.strip().split("\n")
    aL, sL, _, cL = run_literal(scans, False)
    assert len(rows) == len(scans) + 1 and rows[-1].startswith("MAPODOM")
    assert [[int(v) for v in r.split()[5:]] for r in rows[:-1]] == [list(map(int, a)) for a in aL]
    last = rows[-2].split()
    assert np.allclose([float(v) for v in last[1:4]], sL[:3], rtol=0, atol=1e-9) and int(last[4]) == cL


def _build_shard_replay(tmp_path, nccl=False):
    exe = str(tmp_path / "shard_replay")
    pkg = os.path.join(ROOT, "ekf-slam_b200")
    cmd = ["nvcc", "-x", "cu", "-std=c++17", "-O2", "-gencode", "arch=compute_100a,code=sm_100a",
           "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "shard_replay.cpp"), "-o", exe,
           "-L" + pkg, "-lekf_slam_b200", "-Xlinker", "-rpath=" + pkg]
    if nccl:
        cmd += ["-DSHARD_REPLAY_NCCL", "-lnccl"]
    subprocess.check_call(cmd)
    return exe


@pytest.mark.parametrize("shards,extra", [(3, []), (2, ["--allreduce"]), (4, ["--m", "24"])])
def test_cpp_multi_gpu_host_loop_on_the_c_abi(tmp_path, shards, extra):
    """tools/shard_replay.cpp (INTEGRATION.md section 10 as a program): R row-shard handles driven from C++ on
    the C ABI - peer-memory exchange or summed buffers - against the unsharded filter on the same scans.  On a
    one-GPU box the R handles share the device; with R <= devices it is one shard per GPU."""
    import json
    if not os.path.exists(ekf_slam_b200.LIB_PATH):
        ekf_slam_b200.build()
    exe = _build_shard_replay(tmp_path)
    out = subprocess.run([exe, "--shards", str(shards), "--landmarks", "1500", "--scans", "6"] + extra,
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["ok"] and d["association_as_expected"] and d["association_equals_unsharded"] and d["replicas_bitwise_equal"]
    assert d["state_rel_vs_unsharded"] <= 1e-12 and d["sigma_rel_vs_unsharded"] <= 1e-12
    assert d["rounds"] == 6
