"""Overlapped scans (ekf_set_overlap): the covariance pass of scan s runs out of place on a second stream while the
sequential side of scan s+1 already runs.  Same bars as tests/test_gpu_parity.py against the oracle, and the
overlapped filter is compared with the classic path of the same library."""
import numpy as np
import pytest

from _harness import ekf_slam_b200, rel_fro, run_gpu, run_literal, run_struct, synth
from test_gpu_parity import TOL, _check, _procedural_pair, _scan_inputs

pytestmark = pytest.mark.gpu


def _same_filter(a, b, tol=TOL):   # both paths carry the 1e7-prior cancellation noise of first sightings
    assert a[0] == b[0] and a[3] == b[3]
    assert rel_fro(a[1], b[1]) <= tol and rel_fro(a[2], b[2]) <= tol
    assert np.array_equal(a[2], a[2].T)


def test_config2_unknown_association_overlapped():
    scans = synth.circle_drive(n_scans=120, seed=1, drop_out_of_range=True, shuffle=True)
    scans = [s for s in scans if len(s.xy)]
    gpu = run_gpu(scans, False, fused=True, overlap=True)
    _check(gpu, run_literal(scans, False), 4)
    _same_filter(gpu, run_gpu(scans, False, fused=True))


def test_config1_known_association_overlapped():
    scans = synth.circle_drive(n_scans=80, seed=0, max_range=10.0)       # every landmark in range: no DELETE
    gpu = run_gpu(scans, True, fused=True, overlap=True, no_action=True)
    _check(gpu, run_literal(scans, True), 4)
    _same_filter(gpu, run_gpu(scans, True, fused=True))


def test_new_landmarks_partial_visibility_and_accessors_in_the_middle():
    """Landmarks appear over time; every 7th scan the test reads Sigma / state (drains the pipeline) and every
    11th it runs a classic-path call (predict-less update) through the same handle."""
    field = synth.scattered_field(12, seed=5, extent=3.0, min_sep=0.5)
    scans = synth.circle_drive(n_scans=60, seed=2, landmarks=field, drop_out_of_range=True, shuffle=True, max_range=1.6)
    scans = [s for s in scans if len(s.xy)]
    seen = []

    def probe(f, i):
        if i % 7 == 3:
            S = f.sigma()
            assert np.array_equal(S, S.T) and np.isfinite(S).all()
            seen.append(f.state()[:3].copy())
        if i % 5 == 1:
            f.synchronize()

    gpu = run_gpu(scans, False, fused=True, overlap=True, probe=probe)
    ref = run_literal(scans, False)
    _check(gpu, ref, ref[3])
    assert len(seen) >= 5
    _same_filter(gpu, run_gpu(scans, False, fused=True))


def test_scan_longer_than_one_round_overlapped():
    field = synth.scattered_field(30, seed=8, extent=3.2, min_sep=0.45)
    scans = synth.circle_drive(n_scans=6, seed=4, landmarks=field, max_range=10.0, shuffle=True)
    gpu = run_gpu(scans, False, n_landmarks=40, fused=True, overlap=True)
    _check(gpu, run_struct(scans, False, n_landmarks=40), 30)


@pytest.mark.parametrize("drop", [0, 2])
def test_failed_speculation_stops_and_resumes(monkeypatch, drop):
    """drop = 2: the round stops before sighting 2 and another round takes the rest; drop = 0: a round that cannot get
    past its first sighting hands the rest of the scan to the classic path."""
    monkeypatch.setenv("EKF_SHARD_DROP_GUESS", str(drop))
    scans = synth.circle_drive(n_scans=40, seed=6, drop_out_of_range=True, shuffle=True)
    scans = [s for s in scans if len(s.xy)]
    gpu = run_gpu(scans, False, fused=True, overlap=True)
    monkeypatch.delenv("EKF_SHARD_DROP_GUESS")
    _check(gpu, run_literal(scans, False), 4)


def test_switching_modes_on_a_live_filter():
    scans = synth.circle_drive(n_scans=60, seed=12, drop_out_of_range=True, shuffle=True)
    scans = [s for s in scans if len(s.xy)]

    def probe(f, i):
        if i % 9 == 4:
            f.set_overlap(False)
        if i % 9 == 7:
            assert f.set_overlap(True)

    gpu = run_gpu(scans, False, fused=True, overlap=True, probe=probe)
    _check(gpu, run_literal(scans, False), 4)


@pytest.mark.parametrize("N,count,m,scans", [(1000, 990, 20, 5)])
def test_procedural_map_device_resident_overlapped(N, count, m, scans):
    """Larger map (threads own several landmarks at 148 CTAs x 64), several scans in flight back to back."""
    g, o, grid_w = _procedural_pair(N, count)
    assert g.set_overlap(True)
    rng = np.random.default_rng(5)
    pose = np.array([0.3, grid_w / 2.0, grid_w / 2.0])
    g.set_previous_xy(pose[1], pose[2]); o.set_prev(pose[1], pose[2])
    for s in range(scans):
        pose = pose + np.array([0.05, 0.1, 0.07])
        near, xy = _scan_inputs(o.state, count, grid_w, rng, m, pose)
        if s == scans - 2:
            xy[:4] += 0.45                                   # four sightings of nothing mapped -> new landmarks
        aG, cG, _ = g.scan(pose, xy)
        o.predict(*pose)
        aO = o.update_unknown(xy)
        assert list(aG) == list(aO)
        assert cG == o.count
    assert rel_fro(g.state(), o.state) <= TOL
    SG = g.sigma()
    assert rel_fro(SG, o.sigma) <= TOL
    g.close(); o.close()
