"""Pins the oracles to the REFERENCE'S OWN EKF SOURCE.

oracle/_ref/libslam_ref.so is /root/reference/nuslam/src/slam.cpp compiled unmodified (oracle/Makefile,
oracle/slam_ref_shim.cpp, stub headers under oracle/stubs/): Slam::fake_sensor_cb (slam.cpp:180-316),
Slam::sensor_cb (:318-530), Slam::jointStateCallback (:599-634) and the constructor's filter set-up
(:127-139) run as the reference wrote them.  Every oracle and model used by the GPU parity tests is
compared with it here: association identical, state <= 1e-9, Sigma <= 1e-9 norm-wise (whole matrix) and
the observed block at the reference's own 1e7-prior cancellation floor (<= 1e-6).

The library is built in this container (the reference tree does not exist on the GPU box; the prebuilt
file travels there).  Tests skip only when it is genuinely absent.
"""
import math
import os
import sys

import numpy as np
import pytest

from _harness import rel_fro, run_literal, run_model, run_struct, synth

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import ekf_literal as lit  # noqa: E402
import slam_ref  # noqa: E402

pytestmark = pytest.mark.skipif(not slam_ref.available(), reason="oracle/_ref/libslam_ref.so not built")

STATE_TOL = 1e-9
SIGMA_TOL = 1e-9
ACTIVE_TOL = 1e-6


def run_reference(scans, known, n_landmarks=50, prior=None, **kw):
    f = slam_ref.RefSlam(n_landmarks=n_landmarks, sensor_source="sim" if known else "assoc", **kw)
    if prior is not None:       # the reference hard-codes 10e6 (slam.cpp:128-132): another prior is a fixture injection
        f.set_sigma(np.diag([0.0] * 3 + [float(prior)] * (2 * n_landmarks)))
    assoc = []
    for s in scans:
        f.set_odometry(*s.odom)
        if known:
            f.fake_sensor_cb([(int(i), x, y, int(a)) for i, (x, y), a in zip(s.ids, s.xy, s.action)])
            assoc.append(list(s.ids))
        elif len(s.xy):
            assoc.append(f.sensor_cb([(x, y) for (x, y) in s.xy]))
        else:
            assoc.append([])
    assert f.second_branch_taken == 0       # slam.cpp:423-432 never runs
    return assoc, f.state, f.sigma, f.counter_obstacles, f


def _active(state):
    idx = [0, 1, 2] + [int(i) for i in np.nonzero(state[3:])[0] + 3]
    return np.ix_(idx, idx)


def _same(ref, got, known, state_tol=STATE_TOL):
    ra, rs, rS, rc = ref[:4]
    ga, gs, gS, gc = got[:4]
    if not known:
        assert [list(a) for a in ra] == [list(a) for a in ga]
        assert rc == gc
    assert rel_fro(gs, rs) <= state_tol
    assert rel_fro(gS, rS) <= SIGMA_TOL
    blk = _active(rs)
    assert rel_fro(gS[blk], rS[blk]) <= ACTIVE_TOL


TRACES = {
    "config1_known": dict(known=True, N=50, kw=dict(n_scans=80, seed=0, max_range=1.2)),
    "config1_known_all_in_range": dict(known=True, N=50, kw=dict(n_scans=40, seed=3)),
    "config2_unknown": dict(known=False, N=50, kw=dict(n_scans=80, seed=1, drop_out_of_range=True, shuffle=True)),
    "config2_unknown_noisy": dict(known=False, N=50, kw=dict(n_scans=40, seed=7, drop_out_of_range=True, shuffle=True,
                                                              sensor_std=0.01, slip=0.3)),
    "config3_n43": dict(known=False, N=20, field=(20, 21), kw=dict(n_scans=25, seed=5, drop_out_of_range=True,
                                                                     shuffle=True, max_range=1.8)),
    "twelve_in_fifty": dict(known=False, N=50, field=(12, 4), kw=dict(n_scans=30, seed=9, drop_out_of_range=True,
                                                                       shuffle=True, max_range=1.5)),
}


def _scans(c):
    kw = dict(c["kw"])
    if "field" in c:
        kw["landmarks"] = synth.scattered_field(c["field"][0], seed=c["field"][1], extent=4.0, min_sep=0.5)
    return synth.circle_drive(**kw)


@pytest.mark.parametrize("name", sorted(TRACES))
def test_oracle_a_equals_reference_source(name):
    c = TRACES[name]
    scans = [s for s in _scans(c) if c["known"] or len(s.xy)]
    ref = run_reference(scans, c["known"], c["N"])
    _same(ref, run_literal(scans, c["known"], c["N"]), c["known"])


@pytest.mark.parametrize("name", sorted(TRACES))
def test_oracle_b_and_gpu_model_equal_reference_source(name):
    c = TRACES[name]
    scans = [s for s in _scans(c) if c["known"] or len(s.xy)]
    ref = run_reference(scans, c["known"], c["N"])
    _same(ref, run_struct(scans, c["known"], c["N"]), c["known"])
    _same(ref, run_model(scans, c["known"], c["N"]), c["known"])


def test_reference_constructor_matches_documented_defaults():
    """slam.cpp:127-139,665-671: Sigma0 = diag(0,0,0, 1e7 I), state 0, threshold 2.0, N = 50."""
    f = slam_ref.RefSlam()
    assert f.n == 103 and f.counter_obstacles == 0
    S = f.sigma
    assert np.array_equal(S, np.diag([0.0] * 3 + [10e6] * 100))
    assert not f.state.any()
    assert slam_ref.lib().slam_ref_threshold(f._h) == 2.0
    a = lit.LiteralEkf()
    assert np.array_equal(a.sigma, S) and a.mah_threshold == 2.0


def test_capacity_hook_changes_only_the_size():
    """the shim's construction hook sets n_landmarks (slam.cpp:665) before the matrices are sized"""
    for N in (4, 20, 120):
        f = slam_ref.RefSlam(n_landmarks=N)
        n = 2 * N + 3
        assert f.n == n and f.sigma.shape == (n, n) and f.state.shape == (n,)
        assert np.array_equal(f.sigma, np.diag([0.0] * 3 + [10e6] * (2 * N)))


def test_larger_capacity_n243_against_reference_source():
    """N = 120 (n = 243): a map larger than the reference ships, still its own code path"""
    field = synth.scattered_field(40, seed=2, extent=5.0, min_sep=0.45)
    scans = [s for s in synth.circle_drive(n_scans=12, seed=2, landmarks=field, drop_out_of_range=True, shuffle=True,
                                           max_range=2.2) if len(s.xy)]
    ref = run_reference(scans, False, 120)
    _same(ref, run_literal(scans, False, 120), False)
    _same(ref, run_struct(scans, False, 120), False)
    _same(ref, run_model(scans, False, 120), False)


def test_non_default_noise_and_threshold():
    scans = [s for s in synth.circle_drive(n_scans=30, seed=11, drop_out_of_range=True, shuffle=True, sensor_std=0.02)
             if len(s.xy)]
    kw = dict(q_noise=3e-3, r_noise=5e-2, mah_threshold=3.5)
    ref = run_reference(scans, False, 50, **kw)
    _same(ref, run_literal(scans, False, 50, **kw), False)
    _same(ref, run_struct(scans, False, 50, **kw), False)


def test_wheel_odometry_through_joint_state_callback():
    """Slam::jointStateCallback (slam.cpp:599-634) -> DiffDrive::FKin: the reference's t_odom_robot from
    the encoder angles equals oracle A's DiffDrive bit for bit, and the filter run on it matches."""
    scans = [s for s in synth.circle_drive(n_scans=25, seed=4, drop_out_of_range=True, shuffle=True) if len(s.xy)]
    f = slam_ref.RefSlam(sensor_source="assoc")
    a = lit.LiteralEkf()
    dd = lit.DiffDrive(synth.TRACK_WIDTH, synth.WHEEL_RADIUS)
    for s in scans:
        f.joint_state(*s.wheels)
        t = dd.fkin(*s.wheels)
        assert tuple(f.t_odom_robot) == (t.x, t.y, t.rot)
        a.set_odometry(t.x, t.y, t.rot)
        assert f.sensor_cb([(x, y) for (x, y) in s.xy]) == a.sensor_cb([(x, y) for (x, y) in s.xy])
    assert rel_fro(a.state, f.state) <= STATE_TOL
    assert rel_fro(a.sigma, f.sigma) <= SIGMA_TOL
    tmo = np.array([a.t_map_odom.x, a.t_map_odom.y, a.t_map_odom.rot])
    assert np.abs(tmo - f.t_map_odom).max() <= 1e-9


def test_published_tf_and_markers_are_the_filter_outputs():
    """slam.cpp:281-289,512-528: the tf carries t_map_odom, green/obstacles carries the map estimate"""
    scans = [s for s in synth.circle_drive(n_scans=10, seed=6, drop_out_of_range=True) if len(s.xy)]
    _, state, _, count, f = run_reference(scans, False)
    x, y, th = f.t_map_odom
    assert np.allclose(f.last_tf(), [x, y, math.sin(th / 2.0), math.cos(th / 2.0)], rtol=0, atol=0)
    ids, xy = f.published_markers()
    assert list(ids) == list(range(count))
    assert np.array_equal(xy.reshape(-1), state[3:3 + 2 * count])


def test_capacity_overflow_is_the_reference_bounds_error():
    """51st landmark: state(3 + 2*50) is out of bounds (slam.cpp:351 with :665) -> std::out_of_range from the
    element access; oracle A raises IndexError at the same measurement; the C ABI's EKF_ERR_CAPACITY mirrors it."""
    N = 6
    field = synth.scattered_field(10, seed=8, extent=3.0, min_sep=0.6)
    scans = [s for s in synth.circle_drive(n_scans=6, seed=8, landmarks=field, drop_out_of_range=True, max_range=3.0)
             if len(s.xy)]
    f = slam_ref.RefSlam(n_landmarks=N)
    a = lit.LiteralEkf(n_landmarks=N)
    f.set_odometry(*scans[0].odom)
    a.set_odometry(*scans[0].odom)
    xy = [(x, y) for (x, y) in scans[0].xy]
    assert len(xy) > N
    with pytest.raises(slam_ref.RefOutOfRange):
        f.sensor_cb(xy)
    with pytest.raises(IndexError):
        a.sensor_cb(xy)
    # both stopped at the 7th sighting with the first six fused
    assert f.counter_obstacles == N == a.counter_obstacles
    assert rel_fro(a.state, f.state) <= STATE_TOL


def test_empty_scan_throws_after_predicting():
    """an empty MarkerArray reaches msg.markers.at(0) (slam.cpp:498): std::out_of_range, after the prediction
    and the t_map_odom update ran but before filter_previous_configuration moved (:508)"""
    f = slam_ref.RefSlam()
    f.set_odometry(0.3, -0.1, 0.2)
    before = f.sigma
    with pytest.raises(slam_ref.RefOutOfRange):
        f.sensor_cb([])
    after = f.sigma
    assert after[0, 0] == before[0, 0] + 1e-2 and after[1, 1] == 1e-2


def test_exact_ties_go_to_the_lowest_index():
    """two landmarks that are mirror images about the robot's heading axis with a sighting on the axis have
    bit-identical Mahalanobis distances when their covariances mirror too: index_min (slam.cpp:418) takes the
    first.  Built by fixture injection (set_state / set_sigma) on the reference, oracle A and oracle B."""
    import struct_oracle as so
    N = 6
    n = 2 * N + 3
    state = np.zeros(n)
    state[3:7] = [1.0, 0.25, 1.0, -0.25]            # landmark 0 at (1, .25), landmark 1 at (1, -.25)
    sig = np.zeros((n, n))
    sig[np.arange(3, n), np.arange(3, n)] = 10e6
    sig[0, 0] = 1e-3; sig[1, 1] = 2e-3; sig[2, 2] = 2e-3
    for k in (0, 1):
        sig[3 + 2 * k, 3 + 2 * k] = 0.5; sig[4 + 2 * k, 4 + 2 * k] = 0.5
    thr = 50.0                                       # wide gate so the tie is inside it
    f = slam_ref.RefSlam(n_landmarks=N, mah_threshold=thr)
    f.set_state(state, 2); f.set_sigma(sig)
    a = lit.LiteralEkf(n_landmarks=N, mah_threshold=thr)
    a.state[:] = state; a.sigma[:] = sig; a.counter_obstacles = 2
    b = so.StructEkf(n_landmarks=N, mah_threshold=thr)
    b.state[:] = state; b.sigma[:] = sig; b.count = 2
    got_ref = f.sensor_cb([(1.0, 0.0)])
    got_a = a.sensor_cb([(1.0, 0.0)])
    b.predict(0.0, 0.0, 0.0)
    got_b = list(b.update_unknown(np.array([[1.0, 0.0]])))
    assert got_ref == [0] and got_a == [0] and got_b == [0]
    b.close()


def test_landmark_on_top_of_the_robot_nan_never_wins():
    """d = 0 -> H has inf/NaN entries -> NaN distance (slam.cpp:384-401); index_min's strict '<' never
    selects it, so the sighting opens a new landmark.  Reference, oracle A and oracle B agree."""
    import struct_oracle as so
    N = 6
    n = 2 * N + 3
    state = np.zeros(n)
    state[1] = 0.5; state[2] = 0.5
    state[3:5] = [0.5, 0.5]                          # landmark 0 exactly at the robot position
    sig = np.diag([1e-2] * 3 + [0.1, 0.1] + [10e6] * (n - 5))
    f = slam_ref.RefSlam(n_landmarks=N)
    f.set_state(state, 1); f.set_sigma(sig)
    f.set_odometry(0.5, 0.5, 0.0); f.set_previous(0.5, 0.5)
    a = lit.LiteralEkf(n_landmarks=N)
    a.state[:] = state; a.sigma[:] = sig; a.counter_obstacles = 1
    a.set_odometry(0.5, 0.5, 0.0); a.filter_previous_configuration = lit.Transform2D(0.5, 0.5, 0.0)
    with np.errstate(all="ignore"):
        got_a = a.sensor_cb([(0.7, 0.1)])
    got_ref = f.sensor_cb([(0.7, 0.1)])
    b = so.StructEkf(n_landmarks=N)
    b.state[:] = state; b.sigma[:] = sig; b.count = 1; b.set_prev(0.5, 0.5)
    b.predict(0.0, 0.5, 0.5)
    got_b = list(b.update_unknown(np.array([[0.7, 0.1]])))
    assert got_ref == [1] and got_a == [1] and got_b == [1]
    assert f.counter_obstacles == 2 == a.counter_obstacles == b.count
    b.close()


def _elementwise(ref, got, known, tol):
    """max |difference| over the observed block, relative to that block's largest entry; and the same for the state"""
    ra, rs, rS, rc = ref[:4]
    ga, gs, gS, gc = got[:4]
    if not known:
        assert [list(a) for a in ra] == [list(a) for a in ga] and rc == gc
    blk = _active(rs)
    e_sig = np.abs(gS[blk] - rS[blk]).max() / np.abs(rS[blk]).max()
    e_state = np.abs(gs - rs).max() / np.abs(rs).max()
    assert e_sig <= tol and e_state <= tol, (e_sig, e_state)
    return e_sig, e_state


NO_CANCEL_TOL = 1e-11


@pytest.mark.parametrize("name", ["config2_unknown", "config1_known", "config3_n43"])
def test_small_prior_regime_is_tight_elementwise(name):
    """With a prior of 1e2 instead of 1e7 a landmark's first sighting cancels nothing, so two correct evaluation
    orders agree far below the 1e-6 the observed block is held to above: every oracle against the reference's object
    code, ELEMENT-WISE on the observed block, <= 1e-11 (measured 2e-13; state 1e-15).  This is the check that the 1e-8 seen at the
    reference's own prior is its cancellation and not a defect of the restatements."""
    c = TRACES[name]
    scans = [s for s in _scans(c) if c["known"] or len(s.xy)]
    ref = run_reference(scans, c["known"], c["N"], prior=1e2)
    _elementwise(ref, run_literal(scans, c["known"], c["N"], prior=1e2), c["known"], NO_CANCEL_TOL)
    _elementwise(ref, run_struct(scans, c["known"], c["N"], prior=1e2), c["known"], NO_CANCEL_TOL)
    _elementwise(ref, run_model(scans, c["known"], c["N"], prior=1e2), c["known"], NO_CANCEL_TOL)


# ------------------------------------------------------------------------------------------------
# the CUDA path against the reference's object code (the prebuilt oracle/_ref travels to the GPU box)
@pytest.mark.gpu
@pytest.mark.parametrize("fused", [False, True], ids=["split_calls", "ekf_scan"])
@pytest.mark.parametrize("name", sorted(TRACES))
def test_cuda_equals_reference_source(name, fused):
    from _harness import run_gpu
    c = TRACES[name]
    scans = [s for s in _scans(c) if c["known"] or len(s.xy)]
    ref = run_reference(scans, c["known"], c["N"])
    _same(ref, run_gpu(scans, c["known"], c["N"], fused=fused), c["known"])


@pytest.mark.gpu
def test_cuda_larger_capacity_n243_equals_reference_source():
    from _harness import run_gpu
    field = synth.scattered_field(40, seed=2, extent=5.0, min_sep=0.45)
    scans = [s for s in synth.circle_drive(n_scans=12, seed=2, landmarks=field, drop_out_of_range=True, shuffle=True,
                                           max_range=2.2) if len(s.xy)]
    ref = run_reference(scans, False, 120)
    _same(ref, run_gpu(scans, False, 120, fused=True), False)


@pytest.mark.gpu
def test_cuda_non_default_noise_equals_reference_source():
    from _harness import run_gpu
    scans = [s for s in synth.circle_drive(n_scans=30, seed=11, drop_out_of_range=True, shuffle=True, sensor_std=0.02)
             if len(s.xy)]
    kw = dict(q_noise=3e-3, r_noise=5e-2, mah_threshold=3.5)
    ref = run_reference(scans, False, 50, **kw)
    _same(ref, run_gpu(scans, False, 50, fused=True, **kw), False)


@pytest.mark.gpu
@pytest.mark.parametrize("fused", [False, True], ids=["split_calls", "ekf_scan"])
@pytest.mark.parametrize("name", ["config2_unknown", "config1_known", "config3_n43"])
def test_cuda_small_prior_regime_is_tight_elementwise(name, fused):
    """the CUDA path against the reference's object code where nothing cancels (prior 1e2): element-wise <= 1e-11 on
    the observed block and on the state - the deferred rank-2m form included"""
    from _harness import run_gpu
    c = TRACES[name]
    scans = [s for s in _scans(c) if c["known"] or len(s.xy)]
    ref = run_reference(scans, c["known"], c["N"], prior=1e2)
    _elementwise(ref, run_gpu(scans, c["known"], c["N"], fused=fused, prior=1e2), c["known"], NO_CANCEL_TOL)


# ------------------------------------------------------------------------------------------------
# seeded random configurations: capacity, field, drive, noise, gate and mode all vary with the seed
def _random_case(seed):
    rng = np.random.default_rng(1000 + seed)
    known = bool(seed % 3 == 0)
    n_field = int(rng.integers(3, 15))
    N = int(n_field + rng.integers(2, 12))      # (an exactly full map makes the reference write past the end: pinned above)
    field = synth.scattered_field(n_field, seed=200 + seed, extent=float(rng.uniform(2.5, 4.5)), min_sep=0.55)
    kw = dict(n_scans=int(rng.integers(12, 36)), seed=300 + seed, landmarks=field, radius=float(rng.uniform(0.2, 0.5)),
              slip=float(rng.uniform(0.0, 0.25)), sensor_std=float(10.0 ** rng.uniform(-3.3, -2.2)),
              max_range=float(rng.uniform(1.2, 2.6)), theta0=float(rng.uniform(-3.0, 3.0)))
    if not known:
        kw.update(drop_out_of_range=True, shuffle=True)
    filt = dict(q_noise=float(10.0 ** rng.uniform(-3.0, -1.7)), r_noise=float(10.0 ** rng.uniform(-2.5, -1.5)),
                mah_threshold=float(rng.uniform(1.5, 4.0)))
    scans = [s for s in synth.circle_drive(**kw) if known or len(s.xy)]
    return scans, known, N, filt


FUZZ_SEEDS = list(range(12))
FUZZ_STATE_TOL = 1e-8     # the 1e7-prior cancellation reaches the state through the gain: 1.2e-9 on seed 9 (SURVEY 7.2 item 1
                          # measured 4e-9 on its own probe); with a prior of 1e2 the same set-ups agree to 1e-11 element-wise


def _fuzz_check(ref, got, known, prior):
    if prior is None:
        _same(ref, got, known, state_tol=FUZZ_STATE_TOL)
    else:
        _elementwise(ref, got, known, NO_CANCEL_TOL)


@pytest.mark.parametrize("prior", [None, 1e2], ids=["prior_1e7", "prior_1e2"])
@pytest.mark.parametrize("seed", FUZZ_SEEDS)
def test_random_configurations_against_reference_source(seed, prior):
    """oracle A, oracle B and the deferred model against the reference's object code on seeded random set-ups (map
    capacity 5..25, 3..14 landmarks in the field, noise, slip, gate, process / sensor covariance, known or unknown
    association).  At the reference's prior: association identical, Sigma 1e-9 norm-wise, observed block 1e-6, state
    1e-8; with a prior of 1e2 (nothing cancels) the same set-ups: element-wise 1e-11."""
    scans, known, N, filt = _random_case(seed)
    pk = {} if prior is None else dict(prior=prior)
    ref = run_reference(scans, known, N, prior=prior, **filt)
    _fuzz_check(ref, run_literal(scans, known, N, **filt, **pk), known, prior)
    _fuzz_check(ref, run_struct(scans, known, N, **filt, **pk), known, prior)
    kw = dict(q=filt["q_noise"], r=filt["r_noise"], thr=filt["mah_threshold"])
    _fuzz_check(ref, run_model(scans, known, N, **kw, **pk), known, prior)


@pytest.mark.gpu
@pytest.mark.parametrize("prior", [None, 1e2], ids=["prior_1e7", "prior_1e2"])
@pytest.mark.parametrize("seed", FUZZ_SEEDS)
def test_cuda_random_configurations_against_reference_source(seed, prior):
    from _harness import run_gpu
    scans, known, N, filt = _random_case(seed)
    pk = {} if prior is None else dict(prior=prior)
    ref = run_reference(scans, known, N, prior=prior, **filt)
    _fuzz_check(ref, run_gpu(scans, known, N, fused=bool(seed % 2), **filt, **pk), known, prior)
