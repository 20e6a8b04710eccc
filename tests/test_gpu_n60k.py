"""BASELINE config 5 (N = 60,000 landmarks, n = 120,003; the stored triangle is 57.7 GB, the full matrix 115 GB)
against an ORACLE: tests/_streamed_model.py evaluates the reference's scan semantics with O(n m) memory over the
host re-implementation of the procedural map (oracle/procedural_map.py), so nothing has to hold Sigma (SURVEY.md 7.2
item 7, 8(d)).  tests/test_streamed_model.py pins that model to oracle B at sizes oracle B can hold.

Checked on one handle and on the row-sharded path (2 shards as 2 handles on one GPU): association identical, the
whole state <= 1e-9, 72 sampled 128-row block rows of Sigma_new (full width) <= 1e-9 norm-wise each, and the observed
block (pose + the 20 sighted landmarks) <= 1e-6 (the reference's own 1e7-prior cancellation floor)."""
import numpy as np
import pytest

from _harness import ekf_slam_b200, rel_fro
from procedural_map import ProceduralMap
from _streamed_model import StreamedModel

pytestmark = pytest.mark.gpu

N, COUNT, M, SEED = 60000, 59990, 20, 7
GRID_W = 245
TOL = 1e-9


def _need_memory(gb):
    import torch
    free, _ = torch.cuda.mem_get_info()
    if free < gb * 1e9:
        pytest.skip(f"needs {gb} GB of free HBM")


def _inputs(state, rng, pose):
    lm = state[3:3 + 2 * COUNT].reshape(-1, 2)
    th, x, y = pose
    near = np.argsort(np.hypot(lm[:, 0] - x, lm[:, 1] - y))[:M]
    rng.shuffle(near)
    c, s = np.cos(th), np.sin(th)
    dx, dy = lm[near, 0] - x, lm[near, 1] - y
    xy = np.stack([c * dx + s * dy, -s * dx + c * dy], axis=1) + rng.normal(0, 1e-3, (M, 2))
    # four sightings of nothing mapped (the field ends at y = 244): new landmarks
    fresh = np.array([[214.0, 246.0], [216.5, 246.3], [219.0, 246.1], [221.5, 246.4]])
    d = fresh - np.array([x, y])
    xy[:4] = np.stack([c * d[:, 0] + s * d[:, 1], -s * d[:, 0] + c * d[:, 1]], axis=1)
    return near, xy


def _check_against_streamed_oracle(g, scan):
    pm = ProceduralMap(SEED, N, COUNT, GRID_W)
    n = pm.n
    rng = np.random.default_rng(23)
    # the device's generator and the host's produce the same Sigma_0 (sampled block rows, before the scan)
    for i0 in (0, 3 + 2 * 101, 3 + 2 * 30011, 3 + 2 * 59900, n - 64):
        dev = g.sigma_block(i0, i0 + 64, 0, n)
        host = pm.sigma0_rows(i0, i0 + 64)
        assert np.abs(dev - host).max() <= 1e-18 + 1e-13 * np.abs(host).max()
    state0 = g.state()
    assert np.abs(state0 - pm.state()).max() <= 1e-12
    pose = np.array([0.9, 215.3, 242.6])
    prev = (pose[1] - 0.1, pose[2] + 0.05)
    g.set_previous_xy(*prev)
    near, xy = _inputs(state0, rng, pose)
    sm = StreamedModel(N, pm.sigma0_rows, state0, COUNT, prev)
    a_o = sm.scan(pose, xy)
    a_g, cnt = scan(pose, xy)
    assert list(a_g) == list(a_o)                       # association identical
    assert list(a_o[:4]) == [COUNT, COUNT + 1, COUNT + 2, COUNT + 3] and list(a_o[4:]) == list(near[4:])
    assert cnt == sm.cnt == COUNT + 4
    # no decision of this scan is a near-tie: the margins the oracle saw
    margin = np.minimum(np.abs(sm.best_d - 2.0) / 2.0, (sm.second_d - sm.best_d) / np.maximum(sm.best_d, 1e-300))
    assert margin.min() > 0.05
    assert rel_fro(g.state(), sm.mu) <= TOL             # the whole state
    # 72 block rows of 128 rows, full width: spread over the matrix + the rows this scan touched most
    starts = list(np.linspace(0, n - 128, 56).astype(int))
    starts += [0, n - 128] + [max(0, min(n - 128, 3 + 2 * int(k) - 40)) for k in list(near[4:18])]
    worst = 0.0
    for i0 in starts[:72]:
        dev = g.sigma_block(int(i0), int(i0) + 128, 0, n)
        ref = sm.sigma_new_rows(int(i0), int(i0) + 128)
        worst = max(worst, float(np.linalg.norm(dev - ref) / np.linalg.norm(ref)))
    assert worst <= TOL, worst
    # the observed block: pose + every sighted landmark (old and new)
    idx = [0, 1, 2] + [3 + 2 * int(k) + e for k in a_o for e in (0, 1)]
    rows = np.stack([sm.sigma_new_rows(i, i + 1)[0, idx] for i in idx])
    dev = np.stack([g.sigma_block(i, i + 1, 0, n)[0, idx] for i in idx])
    assert rel_fro(dev, rows) <= 1e-6
    return worst, rel_fro(dev, rows)


def test_n60k_single_handle_against_streamed_oracle():
    _need_memory(70)
    g = ekf_slam_b200.EkfSlam(n_landmarks=N)
    g.fill_procedural(SEED, COUNT, GRID_W)

    def scan(pose, xy):
        a, cnt, _ = g.scan(pose, xy)
        return a, cnt
    worst, active = _check_against_streamed_oracle(g, scan)
    print(f"N=60k single handle: worst block-row rel {worst:.2e}, observed block rel {active:.2e}")
    assert g.last_status() == 0
    g.close()


def test_n60k_row_sharded_against_streamed_oracle():
    _need_memory(75)
    from ekf_slam_b200 import shard
    g = shard.LocalShards(N, 2, peer=True)
    g.fill_procedural(SEED, COUNT, GRID_W)
    g.sync()

    def scan(pose, xy):
        a, cnt, _, rounds = g.scan(pose, xy)
        assert rounds == 1
        return a, cnt
    worst, active = _check_against_streamed_oracle(g, scan)
    print(f"N=60k 2 row shards: worst block-row rel {worst:.2e}, observed block rel {active:.2e}")
    g.close()
