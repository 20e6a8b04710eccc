"""nusim's sensor models: the oracle's known answers and its SE(2) pin on the CPU; the
device generators (ekf_nusim_*) against the oracle on the GPU, alone and chained into the landmark
extraction and the batched filters.

Bars: fake-sensor body-frame coordinates within 1e-12 (absolute, metres), ids exact; lidar ranges (float32)
within 1e-6 m of the oracle's, and bit-identical on all but a handful of beams."""
import ctypes as C
import math

import numpy as np
import pytest

from _harness import ekf_slam_b200

import nusim_oracle as no
import landmarks_oracle as lo

# nusim/config/basic_world.yaml:5-10
ARENA_X, ARENA_Y = 10.0, 5.0
OBST = np.array([[-0.5, -0.7, 0.05], [0.8, -0.8, 0.05], [0.4, 0.8, 0.05], [-0.6, 0.65, 0.05]])
MAX_RANGE = 5.0    # nusim.cpp:90


# ---------------------------------------------------------------- CPU: the oracle itself
def test_oracle_se2_matches_reference_turtlelib_bit_for_bit():
    import struct_oracle as so
    ref = so.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    dp = C.POINTER(C.c_double)
    ref.ref_compose.argtypes = [dp, dp, dp]
    ref.ref_inverse.argtypes = [dp, dp]
    rng = np.random.default_rng(3)
    for _ in range(300):
        a, b = rng.normal(0, 3, 3), rng.normal(0, 3, 3)       # (x, y, rot) as the shim takes them
        out = np.zeros(3)
        ref.ref_compose(a.ctypes.data_as(dp), b.ctypes.data_as(dp), out.ctypes.data_as(dp))
        m = no.T2D(a[0], a[1], a[2]) * no.T2D(b[0], b[1], b[2])
        assert (m.x, m.y, m.rot) == tuple(out)
        ref.ref_inverse(a.ctypes.data_as(dp), out.ctypes.data_as(dp))
        i = no.T2D(a[0], a[1], a[2]).inv()
        assert (i.x, i.y, i.rot) == tuple(out)


def test_oracle_known_answers():
    # a cylinder straight ahead: the scanner sits 0.032 m behind the body origin (nusim.cpp:578)
    r = no.lidar((0.0, 0.0, 0.0), [(1.0, 0.0, 0.05)], ARENA_X, ARENA_Y, MAX_RANGE)
    assert abs(r[0] - (1.032 - 0.05)) < 1e-6
    # sideways: the left wall's line sits a whole wall thickness inside the marker centre: y = 2.55 - 0.1
    assert abs(r[90] - 2.45) < 1e-3
    # backwards: x = -(10 + 0.1)/2 + 0.1, seen from the scanner at x = -0.032
    assert abs(r[180] - (4.95 - 0.032)) < 1e-3
    # nothing within max_range and no wall closer: the initial value max_range stays (nusim.cpp:588)
    r2 = no.lidar((0.0, 0.0, 0.0), [(1.0, 0.0, 0.05)], 100.0, 100.0, 3.0)
    assert np.all(r2[5:355] == np.float32(3.0))
    # clamp to range_min (nusim.cpp:702-704): the scanner inside a big cylinder's near surface
    r3 = no.lidar((0.0, 0.0, 0.0), [(0.05, 0.0, 0.05)], ARENA_X, ARENA_Y, MAX_RANGE)
    assert r3[0] == np.float32(0.11)
    # fake sensor: robot at (1, 1) facing +y sees (1, 2) one metre straight ahead
    xy, ids = no.fake_sensor((math.pi / 2, 1.0, 1.0), [(1.0, 2.0), (9.0, 1.0)], 5.0)
    assert np.allclose(xy[0], (1.0, 0.0), atol=1e-15) and list(ids) == [0, -2]


def test_oracle_lidar_feeds_landmark_extraction():
    pose = (0.4, 0.1, -0.05)
    r = no.lidar(pose, OBST, ARENA_X, ARENA_Y, MAX_RANGE)
    det = lo.detect(r, 0.0, 2.0 * math.pi / 360.0)
    T = no.T2D(pose[1], pose[2], pose[0]).inv()
    truth = np.array([[(T * no.T2D(o[0], o[1])).x, (T * no.T2D(o[0], o[1])).y] for o in OBST])
    assert len(det) == 4
    for d in det:
        assert np.min(np.hypot(truth[:, 0] - d[0], truth[:, 1] - d[1])) < 0.02 and abs(d[2] - 0.05) < 0.02


# ---------------------------------------------------------------- CPU: oracle == the reference's nusim.cpp
def _ref_or_skip():
    import nusim_ref
    if not nusim_ref.available():
        pytest.skip("oracle/_ref/libnusim_ref.so not built (reference tree absent)")
    return nusim_ref


def test_oracle_equals_reference_nusim_source_noise_free():
    """NuSim::sensor_timer_callback / publish_lidar_data run from the reference's own object code
    (nusim.cpp:310-349, 559-710): every lidar range, sighting coordinate, id and ADD/DELETE flag is
    bit-identical to the restatement."""
    nr = _ref_or_skip()
    rng = np.random.default_rng(0)
    for world, arena, rmax in [(OBST, (ARENA_X, ARENA_Y), MAX_RANGE), (OBST, (ARENA_X, ARENA_Y), 1.0),
                               (np.array([[1.0, 0.0, 0.3], [0.05, 0.0, 0.05], [-2.0, 1.0, 0.1]]), (6.0, 6.0), 3.0)]:
        f = nr.RefNuSim(world, arena[0], arena[1], max_range=rmax)
        walls = f.walls()
        for (px, py, sx, ang), w in zip(no.wall_markers(*arena), walls):
            assert (px, py, sx) == tuple(w[:3])
            assert ang == 2.0 * math.acos(max(-1.0, min(1.0, w[4])))
        for _ in range(40):
            pose = (rng.uniform(-math.pi, math.pi), rng.uniform(-1.5, 1.5), rng.uniform(-1.2, 1.2))
            f.set_pose(*pose)
            ranges, xy, ids, act = f.sensor_tick()
            assert np.array_equal(ranges, no.lidar(pose, world, arena[0], arena[1], rmax))
            xyo, ido = no.fake_sensor(pose, world[:, :2], rmax)
            assert np.array_equal(xy, xyo)
            assert list(ids) == list(range(len(world)))
            assert [int(a) == 2 for a in act] == [int(i) < 0 for i in ido]      # Marker::DELETE beyond max_range


def test_oracle_equals_reference_nusim_source_with_noise():
    """the same with sensor noise on: the reference's draws (libstdc++ normal_distribution over mt19937) fed
    to the restatement in the reference's draw order reproduce its output bit for bit"""
    nr = _ref_or_skip()
    std = 0.001                                   # nusim.cpp:88 default
    f = nr.RefNuSim(OBST, ARENA_X, ARENA_Y, max_range=1.2, sensor_std=std, seed=5)
    rng = np.random.default_rng(1)
    ticks = 6
    draws = nr.normal_draws(5, std, ticks * (360 + 2 * len(OBST)))
    for t in range(ticks):
        pose = (rng.uniform(-math.pi, math.pi), rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 0.5))
        f.set_pose(*pose)
        d = draws[t * (360 + 2 * len(OBST)):(t + 1) * (360 + 2 * len(OBST))]
        ranges, xy, ids, act = f.sensor_tick()
        assert np.array_equal(ranges, no.lidar(pose, OBST, ARENA_X, ARENA_Y, 1.2, noise=d[:360]))
        yx = d[360:].reshape(-1, 2)               # Y is drawn before X
        xyo, ido = no.fake_sensor(pose, OBST[:, :2], 1.2, noise=yx[:, ::-1])
        assert np.array_equal(xy, xyo)
        assert [int(a) == 2 for a in act] == [int(i) < 0 for i in ido]


# ---------------------------------------------------------------- GPU: the device generators
def _poses(n, seed):
    rng = np.random.default_rng(seed)
    return np.stack([rng.uniform(-math.pi, math.pi, n), rng.uniform(-1.5, 1.5, n), rng.uniform(-1.2, 1.2, n)], axis=1)


@pytest.mark.gpu
def test_wall_lines_and_fake_sensor_against_oracle():
    sim = ekf_slam_b200.NuSim(OBST, ARENA_X, ARENA_Y)
    assert np.allclose(sim.wall_lines(), np.array(no.wall_lines(ARENA_X, ARENA_Y)), rtol=0, atol=1e-15)
    poses = _poses(257, 11)
    noise = np.random.default_rng(12).normal(0, 1e-3, (257, 4, 2))
    for nz in (None, noise):
        xy, ids = sim.fake_sensor(poses, 1.3, noise=nz)
        for r in range(257):
            xo, io = no.fake_sensor(poses[r], OBST[:, :2], 1.3, None if nz is None else nz[r])
            assert np.max(np.abs(xy[r] - xo)) <= 1e-12
            assert list(ids[r]) == list(io)
    assert (ids < 0).any() and (ids >= 0).any()               # both sides of the range gate were exercised
    sim.close()


@pytest.mark.gpu
def test_lidar_against_oracle():
    sim = ekf_slam_b200.NuSim(OBST, ARENA_X, ARENA_Y)
    poses = _poses(48, 21)
    noise = np.random.default_rng(22).normal(0, 1e-3, (48, 360))
    for nz in (None, noise):
        got = sim.lidar(poses, MAX_RANGE, noise=nz)
        differ = 0
        for r in range(48):
            want = no.lidar(poses[r], OBST, ARENA_X, ARENA_Y, MAX_RANGE, None if nz is None else nz[r])
            assert np.max(np.abs(got[r].astype(np.float64) - want.astype(np.float64))) <= 1e-6, r
            differ += int(np.sum(got[r] != want))
        assert differ <= 48 * 360 // 100                      # fused multiply-adds move a last float bit at most
    hits = got < np.float32(2.0)
    assert hits.any() and (~hits).any()
    sim.close()


@pytest.mark.gpu
def test_internal_noise_generator():
    sim = ekf_slam_b200.NuSim(OBST, 100.0, 100.0)
    poses = np.tile(np.array([[0.3, 20.0, 20.0]]), (2048, 1))      # nothing in range: every beam = max_range + noise
    a = sim.lidar(poses, 3.0, noise_std=0.01, seed=5).astype(np.float64) - 3.0
    b = sim.lidar(poses, 3.0, noise_std=0.01, seed=5).astype(np.float64) - 3.0
    c = sim.lidar(poses, 3.0, noise_std=0.01, seed=6).astype(np.float64) - 3.0
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert abs(a.mean()) < 5e-5 and abs(a.std() - 0.01) < 1e-4
    assert abs(np.corrcoef(a[:-1].ravel(), a[1:].ravel())[0, 1]) < 5e-3      # robots draw independently
    kurt = np.mean((a / a.std()) ** 4)
    assert abs(kurt - 3.0) < 0.05
    xy0, _ = sim.fake_sensor(poses[:512], 50.0)
    xy1, _ = sim.fake_sensor(poses[:512], 50.0, noise_std=0.001, seed=9)
    d = xy1 - xy0
    assert abs(d.std() - 0.001) < 5e-5 and abs(d.mean()) < 1e-4
    sim.close()


@pytest.mark.gpu
def test_device_pipeline_lidar_to_landmarks():
    """lidar -> landmark extraction without leaving the device, against the two oracles chained"""
    import torch
    n = 24
    sim = ekf_slam_b200.NuSim(OBST, ARENA_X, ARENA_Y)
    lm = ekf_slam_b200.LandmarkExtractor(n, 360, 16)
    poses = _poses(n, 31) * np.array([1.0, 0.3, 0.3])
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        d_poses = torch.from_numpy(poses).cuda()
        d_ranges = torch.zeros((n, 360), dtype=torch.float32, device="cuda")
        d_out = torch.zeros((n, 16, 3), dtype=torch.float64, device="cuda")
        d_cnt = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_st = torch.zeros(n, dtype=torch.int32, device="cuda")
        sim.lidar_dev(n, d_poses.data_ptr(), MAX_RANGE, 0.0, 0, d_ranges.data_ptr(), stream=st.cuda_stream)
        rc = lm._L.ekf_landmarks_detect_dev(lm._h, n, d_ranges.data_ptr(), C.c_float(0.0),
                                            C.c_float(2.0 * math.pi / 360.0), 0.2, d_out.data_ptr(),
                                            d_cnt.data_ptr(), d_st.data_ptr(), st.cuda_stream)
        assert rc == 0
    st.synchronize()
    out, cnt = d_out.cpu().numpy(), d_cnt.cpu().numpy()
    total = 0
    for r in range(n):
        want = lo.detect(no.lidar(poses[r], OBST, ARENA_X, ARENA_Y, MAX_RANGE), np.float32(0.0),
                         np.float32(2.0 * math.pi / 360.0))
        assert cnt[r] == len(want), r
        for k in range(len(want)):
            assert np.max(np.abs(out[r, k] - np.array(want[k]))) <= 1e-6
        total += len(want)
    assert total >= 2 * n
    sim.close(); lm.close()


@pytest.mark.gpu
def test_device_monte_carlo_loop_fake_sensor_into_batched_filters():
    """B robots on a circle: fake sensor on the device -> ekf_batch_step_dev (known association) with no host
    round trip; filter 0..3 are replayed through oracle B from the same device-generated sightings"""
    import torch
    import struct_oracle as so
    B, steps = 64, 12
    sim = ekf_slam_b200.NuSim(OBST, ARENA_X, ARENA_Y)
    g = ekf_slam_b200.EkfBatch(B, n_landmarks=4)
    rng = np.random.default_rng(41)
    phase = rng.uniform(0, 2 * math.pi, B)
    st = torch.cuda.Stream()
    g.set_stream(st.cuda_stream)
    oracles = [so.StructEkf(n_landmarks=4) for _ in range(4)]
    with torch.cuda.stream(st):
        d_xy = torch.zeros((B, 4, 2), dtype=torch.float64, device="cuda")
        d_ids = torch.zeros((B, 4), dtype=torch.int32, device="cuda")
        for s in range(steps):
            ang = phase + 0.15 * s
            poses = np.stack([ang + math.pi / 2, 0.3 * np.cos(ang), 0.3 * np.sin(ang)], axis=1)
            d_poses = torch.from_numpy(poses).cuda()
            sim.fake_sensor_dev(B, d_poses.data_ptr(), 1.2, 0.001, 100 + s, d_xy.data_ptr(), d_ids.data_ptr(),
                                stream=st.cuda_stream)
            g.step_dev(d_poses.data_ptr(), 4, d_xy.data_ptr(), d_ids.data_ptr(), 0, 0, 0, 0)
            xy, ids = d_xy.cpu().numpy(), d_ids.cpu().numpy()
            for f, o in enumerate(oracles):
                o.predict(*poses[f])
                keep = ids[f] >= 0
                if keep.any():
                    o.update_known(ids[f][keep], xy[f][keep])
                else:
                    o.set_prev(o.state[1], o.state[2])
    st.synchronize()
    assert g.last_status() == 0
    for f, o in enumerate(oracles):
        assert np.linalg.norm(g.state(f) - o.state) / np.linalg.norm(o.state) <= 1e-9
        # every landmark of this 4-landmark map has been seen: all of Sigma is the "observed block", whose
        # bar is the 1e7-prior cancellation floor (tests/test_gpu_parity.py), not 1e-9
        assert np.linalg.norm(g.sigma(f) - o.sigma) / np.linalg.norm(o.sigma) <= 1e-6
        o.close()
    # the maps converge on the world's cylinders wherever a cylinder was ever in range
    stt = g.state(5)
    for j in range(4):
        if stt[3 + 2 * j] != 0.0:
            assert np.hypot(stt[3 + 2 * j] - OBST[j, 0], stt[4 + 2 * j] - OBST[j, 1]) < 0.05
    sim.close(); g.close()
