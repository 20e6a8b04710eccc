"""Near-tie evidence for "association indices bit-exact" (SURVEY.md 7.2 item 4).

Reference semantics at stake: the tentative landmark sits LAST with distance == threshold and index_min takes
the FIRST minimum (slam.cpp:406-408, 418-421): a sighting matches landmark k* = argmin d_k iff d_k* <= threshold,
lowest index on ties.  A decision can differ between two correct implementations only when the best distance is
within their arithmetic difference of the threshold, or of the runner-up.  The tests below start every
implementation from the SAME injected filter (state, symmetric Sigma, count), so what is compared is one
prediction + one association sweep:

  (i)   threshold sweep: sightings placed so that the REFERENCE's own best distance is threshold * (1 +- 10^-k),
        k = 3 .. 15 (bisection on the reference's distance, read from the index_min tap of oracle/stubs);
  (ii)  exact ties: mirror-image landmarks whose distances are bit-identical -> index 0 must win;
  (iii) twin landmarks 1e-9 apart: the distances differ in the last digits, the smaller must win;
  (iv)  the distances themselves: GPU sweep (ekf_mahalanobis) vs the reference's, relative difference.

Checker = oracle/_ref/libslam_ref.so (the reference's slam.cpp, compiled unmodified) and oracle B.
"""
import math
import os
import sys

import numpy as np
import pytest

from _harness import ekf_slam_b200, synth

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import slam_ref  # noqa: E402
import struct_oracle as so  # noqa: E402

pytestmark = pytest.mark.skipif(not slam_ref.available(), reason="oracle/_ref/libslam_ref.so not built")

N = 12
THR = 2.0
AGREE_DOWN_TO_K = 12        # decisions must agree while |d - thr| / thr >= 1e-12


def fixture_filter():
    """a 9-landmark map after 25 scans of the reference itself, Sigma symmetrised (lower triangle mirrored:
    what a triangle-storing implementation holds), pose = the filter's own pose"""
    field = synth.scattered_field(9, seed=13, extent=3.0, min_sep=0.6)
    scans = [s for s in synth.circle_drive(n_scans=25, seed=13, landmarks=field, drop_out_of_range=True, shuffle=True,
                                           max_range=2.5) if len(s.xy)]
    f = slam_ref.RefSlam(n_landmarks=N)
    for s in scans:
        f.set_odometry(*s.odom)
        f.sensor_cb([(x, y) for (x, y) in s.xy])
    S = f.sigma
    S = np.tril(S) + np.tril(S, -1).T
    return f.state, S, f.counter_obstacles


class Probe:
    """one prediction + one sighting from the injected filter, on each implementation"""

    def __init__(self, state, S, count, thr=THR, n_landmarks=N):
        self.state, self.S, self.count, self.thr, self.N = state.copy(), S.copy(), count, thr, n_landmarks
        self.th, self.x, self.y = state[0], state[1], state[2]

    def reference(self, xy):
        f = slam_ref.RefSlam(n_landmarks=self.N, mah_threshold=self.thr)
        f.set_state(self.state, self.count); f.set_sigma(self.S)
        f.set_previous(self.x, self.y); f.set_map_odom(0.0, 0.0, 0.0); f.set_odometry(self.x, self.y, self.th)
        a = f.sensor_cb([tuple(xy)])
        return a[0], float(f.last_best[0]), float(f.last_second[0])

    def oracle_b(self, xy):
        b = so.StructEkf(n_landmarks=self.N, mah_threshold=self.thr)
        b.state[:] = self.state; b.sigma[:] = self.S; b.count = self.count; b.set_prev(self.x, self.y)
        b.predict(self.th, self.x, self.y)
        a, best, second = b.update_unknown_margins(np.array([xy]))
        b.close()
        return int(a[0]), float(best[0]), float(second[0])

    def gpu(self, xy):
        f = ekf_slam_b200.EkfSlam(n_landmarks=self.N, mah_threshold=self.thr)
        f.set_state(self.state); f.set_sigma(self.S); f.count = self.count; f.set_previous_xy(self.x, self.y)
        f.predict(self.th, self.x, self.y)
        d = f.mahalanobis(*xy)
        a = f.update_unknown(np.array([xy]))
        f.close()
        order = np.sort(d)
        return int(a[0]), float(order[0]), float(order[1]) if len(order) > 1 else math.inf

    def gpu_batched(self, xy):
        g = ekf_slam_b200.EkfBatch(2, n_landmarks=self.N, mah_threshold=self.thr)
        for i in range(2):
            g.set_state(i, self.state); g.set_sigma(i, self.S); g.set_count(i, self.count)
            g.set_previous_xy(i, self.x, self.y)
        a, _, _ = g.step(np.tile([self.th, self.x, self.y], (2, 1)), np.tile(np.array([xy]), (2, 1, 1)))
        g.close()
        assert a[0, 0] == a[1, 0]
        return int(a[0, 0])


def body_frame(state, k, off=(0.0, 0.0)):
    """body-frame coordinates of landmark k's estimate, displaced by `off` (metres, body frame)"""
    th, x, y = state[:3]
    dx, dy = state[3 + 2 * k] - x, state[4 + 2 * k] - y
    c, s = math.cos(th), math.sin(th)
    return (c * dx + s * dy + off[0], -s * dx + c * dy + off[1])


def sighting_at_distance(probe, k, target, direction=(1.0, 0.3)):
    """bisection on the displacement so that the REFERENCE's best distance for landmark k is `target`"""
    lo, hi = 0.0, 0.05
    u = np.array(direction) / np.hypot(*direction)
    while probe.reference(body_frame(probe.state, k, hi * u))[1] < target:
        hi *= 2.0
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        if mid == lo or mid == hi:
            break
        if probe.reference(body_frame(probe.state, k, mid * u))[1] < target:
            lo = mid
        else:
            hi = mid
    return body_frame(probe.state, k, lo * u), body_frame(probe.state, k, hi * u)


def sweep(probe, impl, ks=range(3, 16)):
    """-> list of (k, side, reference decision, impl decision, reference d, impl d)"""
    rows = []
    j = 2
    for k in ks:
        for side in (-1.0, +1.0):
            target = probe.thr * (1.0 + side * 10.0 ** (-k))
            below, above = sighting_at_distance(probe, j, target)
            for xy in (below, above):
                ra, rd, _ = probe.reference(xy)
                ga, gd, _ = impl(xy)
                rows.append((k, side, ra, ga, rd, gd))
    return rows


def first_disagreement(rows):
    bad = [k for (k, _, ra, ga, _, _) in rows if ra != ga]
    return min(bad) if bad else None


# ------------------------------------------------------------------------------------------------ CPU
def test_threshold_sweep_oracle_b_vs_reference():
    state, S, count = fixture_filter()
    p = Probe(state, S, count)
    rows = sweep(p, p.oracle_b)
    # the sweep really reaches the margins it claims
    for (k, side, _, _, rd, _) in rows:
        assert abs(rd / p.thr - 1.0) <= 1.5 * 10.0 ** (-k) + 1e-14     # d(displacement) is granular at ~3e-15
    fd = first_disagreement(rows)
    assert fd is None or fd > AGREE_DOWN_TO_K, rows
    rel = max(abs(gd - rd) / rd for (_, _, _, _, rd, gd) in rows)
    assert rel <= 1e-12
    print(f"oracle B vs reference: first disagreeing margin 1e-{fd}, max rel distance difference {rel:.2e}")


def test_reference_threshold_rule_is_inclusive():
    """d == threshold exactly still matches (the tentative entry is LAST and index_min takes the first minimum).
    Built with a threshold equal to the reference's own distance for the sighting."""
    state, S, count = fixture_filter()
    xy = body_frame(state, 1, (0.01, -0.02))
    d = Probe(state, S, count).reference(xy)[1]
    p = Probe(state, S, count, thr=d)
    assert p.reference(xy)[0] == 1 and p.oracle_b(xy)[0] == 1
    p2 = Probe(state, S, count, thr=np.nextafter(d, 0.0))
    assert p2.reference(xy)[0] == count and p2.oracle_b(xy)[0] == count      # one ulp tighter: new landmark


def mirror_fixture():
    n = 2 * N + 3
    state = np.zeros(n)
    state[3:7] = [1.0, 0.25, 1.0, -0.25]
    S = np.zeros((n, n))
    S[np.arange(3, n), np.arange(3, n)] = 10e6
    S[0, 0] = 1e-3; S[1, 1] = 2e-3; S[2, 2] = 2e-3
    for k in (0, 1):
        S[3 + 2 * k, 3 + 2 * k] = 0.5; S[4 + 2 * k, 4 + 2 * k] = 0.5
    return state, S, 2


def test_exact_tie_reference_and_oracle_b():
    state, S, count = mirror_fixture()
    p = Probe(state, S, count, thr=50.0)
    ra, rb, rs = p.reference((1.0, 0.0))
    assert rb == rs          # the two distances are bit-identical in the reference
    assert ra == 0 and p.oracle_b((1.0, 0.0))[0] == 0


def twin_fixture():
    state, S, count = fixture_filter()
    # landmark `count` becomes a twin of landmark 3, 1e-9 m away, with the same covariance rows
    k, t = 3, count
    state = state.copy(); S = S.copy()
    state[3 + 2 * t] = state[3 + 2 * k] + 1e-9
    state[4 + 2 * t] = state[4 + 2 * k]
    a, b = 3 + 2 * k, 3 + 2 * t
    S[b:b + 2, :] = S[a:a + 2, :]; S[:, b:b + 2] = S[:, a:a + 2]
    S[b:b + 2, b:b + 2] = S[a:a + 2, a:a + 2]
    S[b:b + 2, a:a + 2] = 0.5 * S[a:a + 2, a:a + 2]; S[a:a + 2, b:b + 2] = S[b:b + 2, a:a + 2].T
    return state, S, count + 1, k, t


def test_twin_landmarks_reference_and_oracle_b():
    state, S, count, k, t = twin_fixture()
    p = Probe(state, S, count, thr=1e3)
    for off in ((0.004, 0.0), (-0.004, 0.0), (0.0, 0.003), (0.002, -0.002)):
        xy = body_frame(state, k, off)
        ra, rb, rs = p.reference(xy)
        assert ra in (k, t) and rs > rb and (rs - rb) / rb < 1e-5       # a near-tie, resolved by the last digits
        assert p.oracle_b(xy)[0] == ra


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_gpu_threshold_sweep_vs_reference():
    state, S, count = fixture_filter()
    p = Probe(state, S, count)
    rows = sweep(p, p.gpu)
    fd = first_disagreement(rows)
    rel = max(abs(gd - rd) / rd for (_, _, _, _, rd, gd) in rows)
    print(f"GPU vs reference: first disagreeing margin 1e-{fd}, max rel distance difference {rel:.2e}")
    assert fd is None or fd > AGREE_DOWN_TO_K, rows
    assert rel <= 1e-12
    # the batched kernel takes the same decisions
    for (k, side) in ((6, -1.0), (6, 1.0), (10, -1.0), (10, 1.0)):
        below, above = sighting_at_distance(p, 2, p.thr * (1.0 + side * 10.0 ** (-k)))
        for xy in (below, above):
            assert p.gpu_batched(xy) == p.reference(xy)[0]


@pytest.mark.gpu
def test_gpu_exact_tie_lowest_index_wins():
    state, S, count = mirror_fixture()
    p = Probe(state, S, count, thr=50.0)
    ga, gb, gs = p.gpu((1.0, 0.0))
    assert gb == gs          # bit-identical distances on the device too
    assert ga == 0 and p.gpu_batched((1.0, 0.0)) == 0
    # the same pair stored in the other order: still the LOWER index
    state2 = state.copy(); state2[3:7] = [1.0, -0.25, 1.0, 0.25]
    p2 = Probe(state2, S, count, thr=50.0)
    assert p2.reference((1.0, 0.0))[0] == 0 and p2.gpu((1.0, 0.0))[0] == 0 and p2.gpu_batched((1.0, 0.0)) == 0


@pytest.mark.gpu
def test_gpu_twin_landmarks_vs_reference():
    state, S, count, k, t = twin_fixture()
    p = Probe(state, S, count, thr=1e3)
    for off in ((0.004, 0.0), (-0.004, 0.0), (0.0, 0.003), (0.002, -0.002), (0.0, 0.0)):
        xy = body_frame(state, k, off)
        ra = p.reference(xy)[0]
        assert p.gpu(xy)[0] == ra and p.gpu_batched(xy) == ra


@pytest.mark.gpu
def test_gpu_distance_sweep_matches_reference_distances():
    """ekf_mahalanobis (the kernels' association arithmetic) against the reference's own distances on random
    sightings around every landmark: relative difference at the 1e-13 level, i.e. a decision can only differ
    from the reference's when its margin is below that"""
    state, S, count = fixture_filter()
    p = Probe(state, S, count)
    rng = np.random.default_rng(3)
    worst = 0.0
    for k in range(count):
        for _ in range(6):
            xy = body_frame(state, k, rng.normal(0.0, 0.02, 2))
            ra, rb, rs = p.reference(xy)
            ga, gb, gs = p.gpu(xy)
            assert ga == ra
            worst = max(worst, abs(gb - rb) / rb, abs(gs - rs) / rs)
    print(f"max relative difference of best / second distance, GPU vs reference: {worst:.2e}")
    assert worst <= 1e-12
