"""Row-sharded large map: R shards of one filter as R handles on ONE GPU (ekf_slam_b200.shard.LocalShards;
the exchange of the multi-rank protocol becomes a plain sum of the R buffers), against oracle B and
against the unsharded GPU filter.  Bars as in tests/test_gpu_parity.py."""
import os

import numpy as np
import pytest

from _harness import ekf_slam_b200, rel_fro

pytestmark = pytest.mark.gpu


def _scan_inputs(state, count, rng, m, pose):
    lm = state[3:3 + 2 * count].reshape(-1, 2)
    th, x, y = pose
    d = np.hypot(lm[:, 0] - x, lm[:, 1] - y)
    near = np.argsort(d)[:m]
    rng.shuffle(near)
    c, s = np.cos(th), np.sin(th)
    dx, dy = lm[near, 0] - x, lm[near, 1] - y
    xy = np.stack([c * dx + s * dy, -s * dx + c * dy], axis=1) + rng.normal(0, 1e-3, (m, 2))
    return near.astype(np.int32), xy


def _pair(N, count, shards, seed=7, peer=False):
    import struct_oracle as so
    from ekf_slam_b200 import shard
    g = shard.LocalShards(N, shards, peer=peer)
    grid_w = int(np.ceil(np.sqrt(count)))
    g.fill_procedural(seed, count, grid_w)
    g.sync()
    o = so.StructEkf(n_landmarks=N)
    o.state[:] = g.state()
    for i0 in range(0, g.n, 2048):
        i1 = min(g.n, i0 + 2048)
        o.sigma[i0:i1, :] = g.sigma_block(i0, i1, 0, g.n)
    o.count = count
    return g, o, grid_w


def _drive(g, o, count, grid_w, m, scans, new_in_last=0, expect_rounds=None, resight=False):
    rng = np.random.default_rng(5)
    pose = np.array([0.3, grid_w / 2.0, grid_w / 2.0])
    g.set_previous_xy(pose[1], pose[2]); o.set_prev(pose[1], pose[2])
    for s in range(scans):
        pose = pose + np.array([0.05, 0.1, 0.07])
        near, xy = _scan_inputs(o.state, count, rng, m, pose)
        if s == scans - 1 and new_in_last:
            xy[:new_in_last] += 0.45
            if resight:        # the new landmarks are sighted a second time later in the SAME scan
                xy[m - new_in_last:] = xy[:new_in_last] + rng.normal(0, 1e-3, (new_in_last, 2))
        aG, cG, pose_after, rounds = g.scan(pose, xy)
        o.predict(*pose)
        aO = o.update_unknown(xy)
        assert list(aG) == list(aO), s
        assert cG == o.count
        if expect_rounds is not None:
            assert rounds == expect_rounds, rounds
    assert rel_fro(g.state(), o.state) <= 1e-9
    SG = g.sigma()
    assert np.array_equal(SG, SG.T)
    assert rel_fro(SG, o.sigma) <= 1e-9
    k = 3 + 2 * count
    assert rel_fro(SG[:k, :k], o.sigma[:k, :k]) <= 1e-8


@pytest.mark.parametrize("shards", [1, 2, 3])
def test_sharded_map_against_oracle_b(shards):
    g, o, grid_w = _pair(300, 290, shards)
    _drive(g, o, 290, grid_w, 20, 4, new_in_last=4, expect_rounds=1)
    g.close(); o.close()


@pytest.mark.parametrize("shards", [2, 3])
def test_peer_memory_exchange_against_oracle_b(shards, monkeypatch):
    """the per-round exchange as one kernel over peer memory (every shard stores its columns into every
    shard's buffer and raises arrival flags) instead of the caller's all-reduce; includes the stop / resume
    path, which re-pushes into the other parity buffer"""
    g, o, grid_w = _pair(300, 290, shards, peer=True)
    _drive(g, o, 290, grid_w, 20, 5, new_in_last=4, expect_rounds=1)
    monkeypatch.setenv("EKF_SHARD_DROP_GUESS", "5")
    _drive(g, o, 294, grid_w, 20, 2)
    assert g.hs[0].last_status() == 0
    g.close(); o.close()


def test_landmark_created_and_resighted_in_one_scan_needs_no_extra_round():
    """a landmark born in this scan has an all-zero Sigma_0 column: its second sighting in the same scan
    is handled from the panels alone (no exchange, no stop)"""
    g, o, grid_w = _pair(300, 290, 3)
    _drive(g, o, 290, grid_w, 20, 2, new_in_last=3, expect_rounds=1, resight=True)
    g.close(); o.close()


def test_sharded_scan_longer_than_one_round():
    g, o, grid_w = _pair(400, 380, 4)
    _drive(g, o, 380, grid_w, 30, 2, new_in_last=3, expect_rounds=2)
    g.close(); o.close()


def test_failed_speculation_resumes_with_identical_results():
    """fault injection: the columns of sighting 5 of every round are not fetched; each round stops there
    and the driver continues with the rest (20 sightings -> rounds of 5, 5, 5, 5) - same numbers as ever."""
    os.environ["EKF_SHARD_DROP_GUESS"] = "5"
    try:
        g, o, grid_w = _pair(300, 290, 3)
        _drive(g, o, 290, grid_w, 20, 2, expect_rounds=4)
        g.close(); o.close()
    finally:
        del os.environ["EKF_SHARD_DROP_GUESS"]


def test_sharded_matches_unsharded_at_N10k():
    """BASELINE config 4 at full size over 4 row shards vs the single-handle filter: same associations,
    same state and Sigma to 1e-12 (the two differ only in the rounding of the pose-panel / diagonal
    entries, which the sharded path takes from the sequentially updated side state)."""
    from ekf_slam_b200 import shard
    N, m = 10000, 20
    grid_w = 100
    u = ekf_slam_b200.EkfSlam(n_landmarks=N)
    u.fill_procedural(7, N, grid_w)
    g = shard.LocalShards(N, 4)
    g.fill_procedural(7, N, grid_w)
    g.sync()
    rng = np.random.default_rng(9)
    pose = np.array([0.7, 41.3, 57.9])
    u.set_previous_xy(pose[1] - 0.1, pose[2] + 0.05); g.set_previous_xy(pose[1] - 0.1, pose[2] + 0.05)
    state0 = u.state()
    for s in range(2):
        pose = pose + np.array([0.02, 0.3, -0.2])
        near, xy = _scan_inputs(state0, N, rng, m, pose)
        aU, cU, pU = u.scan(pose, xy)
        aG, cG, pG, rounds = g.scan(pose, xy)
        assert list(aU) == list(aG) == list(near) and cU == cG == N and rounds == 1
    assert rel_fro(g.state(), u.state()) <= 1e-12
    worst = 0.0
    for i0 in range(0, u.n, 4096):
        i1 = min(u.n, i0 + 4096)
        a, b = g.sigma_block(i0, i1, 0, u.n), u.sigma_block(i0, i1, 0, u.n)
        worst = max(worst, float(np.linalg.norm(a - b) / np.linalg.norm(b)))
    assert worst <= 1e-12, worst
    # every shard stores its equal-area part, together exactly the triangle
    stored = [h.shard_info() for h in g.hs]
    assert [s[0] for s in stored] == shard.shard_bounds(N, 4)[:-1]
    assert sum(s[2] for s in stored) == u.storage()[0]
    u.close(); g.close()


def test_sharded_handle_refuses_the_single_handle_entry_points():
    m = ekf_slam_b200
    with m.EkfSlam(n_landmarks=300, shard_rank=1, shard_count=2) as h:
        with pytest.raises(m.EkfError) as e:
            h.predict(0.0, 0.0, 0.0)
        assert e.value.status == m.EKF_ERR_INVALID
        L = m.load()
        import ctypes as C
        ptr, cnt = C.c_void_p(), C.c_int64(0)
        # scan before the side state was synchronised
        assert L.ekf_shard_scan_begin(h._h, None, 1, C.c_void_p(8), None, 1, C.byref(ptr), C.byref(cnt)) == m.EKF_ERR_INVALID


@pytest.mark.parametrize("threads", [64, 256])
def test_device_resident_pipeline_both_sequential_shapes(threads, monkeypatch):
    """scan_dev back to back without host round trips (what bench.py times), peer-memory exchange: the 64-thread
    sequential shape runs concurrently with the push kernel and polls its columns (flag-in-data), the 256-thread shape
    of large maps cannot share an SM with the push CTAs and is ordered after it.  Same numbers as the host loop."""
    import torch
    from ekf_slam_b200 import shard
    monkeypatch.setenv("EKF_SEQ_THREADS", str(threads))
    N, count, m, scans = 300, 290, 20, 6
    g, o, grid_w = _pair(N, count, 3, peer=True)
    rng = np.random.default_rng(5)
    pose = np.array([0.3, grid_w / 2.0, grid_w / 2.0])
    g.set_previous_xy(pose[1], pose[2]); o.set_prev(pose[1], pose[2])
    poses, xys, want = [], [], []
    for s in range(scans):
        pose = pose + np.array([0.05, 0.1, 0.07])
        near, xy = _scan_inputs(o.state, count, rng, m, pose)
        o.predict(*pose)
        want.append(list(o.update_unknown(xy)))
        poses.append(pose.copy()); xys.append(xy)
    with torch.cuda.stream(g.stream):
        d_pose = torch.from_numpy(np.array(poses)).cuda()
        d_xy = torch.from_numpy(np.array(xys)).cuda()
        d_assoc = torch.full((scans, m), -9, dtype=torch.int32, device="cuda")
        d_done = torch.zeros(scans, dtype=torch.int32, device="cuda")
    for s in range(scans):
        g.scan_dev(d_pose[s], d_xy[s], d_assoc[s], check=False, done_t=d_done[s:s + 1])
    torch.cuda.synchronize()
    assert d_done.cpu().tolist() == [m] * scans
    assert d_assoc.cpu().numpy().tolist() == want
    assert g.hs[0].last_status() == 0
    assert rel_fro(g.state(), o.state) <= 1e-9
    assert rel_fro(g.sigma(), o.sigma) <= 1e-9
    g.close(); o.close()
