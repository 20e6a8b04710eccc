"""The streamed block-wise oracle (tests/_streamed_model.py, used for BASELINE config 5 where no host can hold
Sigma) against oracle B entry for entry at sizes oracle B can hold, on the host generator's procedural map."""
import numpy as np

from _harness import rel_fro
from procedural_map import ProceduralMap, hash_sym
from _streamed_model import StreamedModel

import struct_oracle as so


def _inputs(state, count, rng, m, pose):
    lm = state[3:3 + 2 * count].reshape(-1, 2)
    th, x, y = pose
    near = np.argsort(np.hypot(lm[:, 0] - x, lm[:, 1] - y))[:m]
    rng.shuffle(near)
    c, s = np.cos(th), np.sin(th)
    dx, dy = lm[near, 0] - x, lm[near, 1] - y
    return near, np.stack([c * dx + s * dy, -s * dx + c * dy], axis=1) + rng.normal(0, 1e-3, (m, 2))


def test_generator_is_symmetric_spd_and_blockwise_consistent():
    pm = ProceduralMap(7, 40, 36, 6)
    S = pm.sigma0_rows(0, pm.n)
    assert np.array_equal(S, S.T)
    k = pm.n_active
    assert np.all(np.linalg.eigvalsh(S[:k, :k]) > 0)
    assert np.array_equal(np.diag(S)[k:], np.full(pm.n - k, 10e6)) and not S[k:, :k].any()
    assert np.array_equal(pm.sigma0_rows(10, 30, 5, 50), S[10:30, 5:50])
    assert np.array_equal(pm.sigma0_cols(11, 13), S[:, 11:13])
    assert -1.0 <= hash_sym(7, np.arange(1000), 3).min() and hash_sym(7, np.arange(1000), 3).max() < 1.0


def test_streamed_model_equals_oracle_b():
    N, count, m = 600, 590, 20
    grid_w = int(np.ceil(np.sqrt(count)))
    pm = ProceduralMap(7, N, count, grid_w)
    o = so.StructEkf(n_landmarks=N)
    o.state[:] = pm.state(); o.sigma[:] = pm.sigma0_rows(0, pm.n); o.count = count
    rng = np.random.default_rng(5)
    pose = np.array([0.3, grid_w / 2.0, grid_w / 2.0])
    o.set_prev(pose[1] - 0.1, pose[2] + 0.05)
    near, xy = _inputs(np.array(o.state), count, rng, m, pose)
    xy[:3] += 0.45                                     # three sightings of nothing mapped -> new landmarks
    sm = StreamedModel(N, pm.sigma0_rows, pm.state(), count, (pose[1] - 0.1, pose[2] + 0.05))
    a_s = sm.scan(pose, xy)
    o.predict(*pose)
    a_o, best, second = o.update_unknown_margins(xy)
    assert list(a_s) == list(a_o) and sm.cnt == o.count == count + 3
    # runner-up distances may involve the landmarks created in this very scan, whose covariance carries the
    # 1e7-prior cancellation noise (~1e-8 relative)
    assert np.allclose(sm.best_d, best, rtol=1e-9) and np.allclose(sm.second_d, second, rtol=1e-6)
    assert rel_fro(sm.mu, o.state) <= 1e-12
    for i0 in range(0, pm.n, 257):
        i1 = min(pm.n, i0 + 257)
        ref = o.sigma[i0:i1, :]
        assert np.linalg.norm(sm.sigma_new_rows(i0, i1) - ref) <= 1e-11 * np.linalg.norm(ref)
    o.close()


def test_streamed_model_two_scans_chain_through_an_accessor():
    """second scan: the accessor is the first scan's sigma_new_rows - what a caller does to chain scans"""
    N, count, m = 200, 190, 12
    grid_w = 14
    pm = ProceduralMap(3, N, count, grid_w)
    o = so.StructEkf(n_landmarks=N)
    o.state[:] = pm.state(); o.sigma[:] = pm.sigma0_rows(0, pm.n); o.count = count
    rng = np.random.default_rng(8)
    pose = np.array([0.2, 6.0, 6.5])
    o.set_prev(pose[1], pose[2])
    sm = StreamedModel(N, pm.sigma0_rows, pm.state(), count, (pose[1], pose[2]))
    for s in range(2):
        pose = pose + np.array([0.05, 0.1, 0.07])
        _, xy = _inputs(np.array(o.state), count, rng, m, pose)
        a_s = sm.scan(pose, xy)
        o.predict(*pose)
        assert list(a_s) == list(o.update_unknown(xy))
        full = sm.sigma_new_rows(0, pm.n)
        sm = StreamedModel(N, lambda i0, i1, j0=0, j1=None, F=full: F[i0:i1, j0:(F.shape[1] if j1 is None else j1)],
                           sm.mu, sm.cnt, sm.prev)
    assert rel_fro(sm.mu, o.state) <= 1e-12
    assert rel_fro(full, np.array(o.sigma)) <= 1e-11
    o.close()
