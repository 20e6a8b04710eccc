"""N4: the rosbag2 / MCAP reader behind ekf_bag_* against an independent writer (tests/_mcap_writer.py).
Topic / type names and message counts follow the reference's recording
(rosbags/rosbag2_2024_03_17-18_35_57/metadata.yaml:116-121,158-163); its payload is missing from the checkout, so
these round trips are the only pin there is ("parity unpinned" for the real bag)."""
import os
import struct

import numpy as np
import pytest

import _mcap_writer as mw
from _harness import ekf_slam_b200, synth

bag_mod = pytest.importorskip("ekf_slam_b200.bag")

JOINTS, DETS, SCAN, TF = 1, 2, 3, 4
CHANNELS = {JOINTS: ("/blue/joint_states", "sensor_msgs/msg/JointState"),
            DETS: ("/green/detected_obstacles", "visualization_msgs/msg/MarkerArray"),
            SCAN: ("/scan", "sensor_msgs/msg/LaserScan"),
            TF: ("/tf", "tf2_msgs/msg/TFMessage")}
T0 = 1710718557269449803          # metadata.yaml: starting_time


def _trace(n_scans=24, joint_per_scan=22, seed=3):
    """The bag's shape in small: ~22 joint states between detection arrays (metadata.yaml: 8,419 / 380)."""
    scans = synth.circle_drive(n_scans=n_scans, seed=seed, drop_out_of_range=True, shuffle=True)
    rng = np.random.default_rng(seed)
    events, t, prev = [], T0, (0.0, 0.0)
    for k, s in enumerate(scans):
        for j in range(1, joint_per_scan + 1):
            f = j / joint_per_scan
            left, right = prev[0] + f * (s.wheels[0] - prev[0]), prev[1] + f * (s.wheels[1] - prev[1])
            t += 10_000_000
            events.append(("joint", t, (left, right)))
        prev = s.wheels
        t += 1_234_567
        markers = []
        for q, (x, y) in enumerate(s.xy):
            m = {"id": int(q), "x": float(x), "y": float(y), "action": 0, "ns": "detected" if q % 2 else "o"}
            if q == 1:      # every variable-length field of Marker non-empty once: alignment after each
                m.update(points=[(0.1, 0.2, 0.3)] * (k % 3), colors=[(1.0, 0.0, 0.0, 1.0)] * (k % 2), text="cyl %d" % k,
                         mesh_resource="package://x/y.dae", texture_data=bytes(range(k % 5)), uv=[(0.5, 0.25)] * (k % 4),
                         mesh_filename="m", mesh_data=b"\x01\x02\x03"[: k % 4], texture_format="png", texture_resource="t")
            markers.append(m)
        events.append(("det", t, markers))
        if k % 5 == 0:
            events.append(("junk", t + 5, rng.integers(0, 255, size=37, dtype=np.uint8).tobytes()))
    return scans, events


def _messages(events):
    out = []
    for kind, t, payload in events:
        if kind == "joint":
            out.append((JOINTS, t, mw.encode_joint_state(t - 777, ["wheel_left_joint", "wheel_right_joint"], payload,
                                                         velocity=(0.5, 0.25), frame_id="")))
        elif kind == "det":
            out.append((DETS, t, mw.encode_marker_array(t, payload)))
        else:
            out.append((SCAN if t % 2 else TF, t, payload))
    return out


def _check(reader, events):
    wanted = [e for e in events if e[0] != "junk"]
    assert reader.n_events == len(wanted)
    assert reader.messages_total == len(events)
    assert reader.n_joint_states == sum(e[0] == "joint" for e in events)
    assert reader.n_detection_arrays == sum(e[0] == "det" for e in events)
    for (kind, index, log_time), (wk, wt, payload) in zip(reader.events(), sorted(wanted, key=lambda e: e[1])):
        assert log_time == wt
        if wk == "joint":
            assert kind == bag_mod.JOINT_STATE
            stamp, pos = reader.joint_state(index)
            assert stamp == wt - 777
            assert pos.tobytes() == np.array(payload, dtype=np.float64).tobytes()      # bit-exact doubles
        else:
            assert kind == bag_mod.DETECTIONS
            xy, ids, act = reader.detections(index)
            assert xy.tobytes() == np.array([(m["x"], m["y"]) for m in payload], dtype=np.float64).reshape(-1, 2).tobytes()
            assert list(ids) == [m["id"] for m in payload]
            assert list(act) == [m["action"] for m in payload]


@pytest.mark.parametrize("chunk,compression", [(0, ""), (16, ""), (1, ""), (64, "zstd"), (64, "lz4")])
def test_round_trip(tmp_path, chunk, compression):
    if not mw.compression_available(compression):
        pytest.skip(f"lib for {compression} not on this box")
    _, events = _trace()
    path = tmp_path / "bag_0.mcap"
    mw.write_mcap(path, CHANNELS, _messages(events), chunk_messages=chunk, compression=compression)
    r = bag_mod.BagReader(path)
    _check(r, events)
    assert r.chunks == (0 if chunk == 0 else -(-len(events) // chunk))
    r.close()


def test_file_order_need_not_be_time_order(tmp_path):
    """rosbag2 writes per-topic caches: messages of different topics interleave out of time order in the file."""
    _, events = _trace(n_scans=8)
    shuffled = [events[i] for i in np.random.default_rng(0).permutation(len(events))]
    path = tmp_path / "bag_0.mcap"
    mw.write_mcap(path, CHANNELS, _messages(shuffled), chunk_messages=7, channels_in_chunks=False)
    _check(bag_mod.BagReader(path), events)


def test_other_topic_names(tmp_path):
    """nusim/fake_sensor (known association, slam.cpp:116-119) instead of the detector's topic."""
    _, events = _trace(n_scans=4)
    ch = dict(CHANNELS)
    ch[DETS] = ("/nusim/fake_sensor", CHANNELS[DETS][1])
    path = tmp_path / "bag_0.mcap"
    mw.write_mcap(path, ch, _messages(events))
    assert bag_mod.BagReader(path).n_detection_arrays == 0
    _check(bag_mod.BagReader(path, detections_topic="/nusim/fake_sensor"), events)


def test_empty_detection_array_and_empty_bag(tmp_path):
    path = tmp_path / "a.mcap"
    mw.write_mcap(path, CHANNELS, [(DETS, T0, mw.encode_marker_array(T0, []))])
    r = bag_mod.BagReader(path)
    xy, ids, act = r.detections(0)
    assert xy.shape == (0, 2) and len(ids) == 0
    with pytest.raises(IndexError):
        r.detections(1)
    with pytest.raises(IndexError):
        r.event(1)
    mw.write_mcap(path, CHANNELS, [])
    assert bag_mod.BagReader(path).n_events == 0


def test_errors_are_reported_not_thrown_across_the_abi(tmp_path):
    _, events = _trace(n_scans=3)
    good = tmp_path / "good.mcap"
    mw.write_mcap(good, CHANNELS, _messages(events), chunk_messages=8)
    raw = open(good, "rb").read()
    with pytest.raises(bag_mod.BagError, match="cannot open"):
        bag_mod.BagReader(tmp_path / "missing.mcap")
    bad = tmp_path / "bad.mcap"
    bad.write_bytes(b"not an mcap file at all")
    with pytest.raises(bag_mod.BagError, match="magic"):
        bag_mod.BagReader(bad)
    bad.write_bytes(raw[: len(raw) // 2])                       # cut inside a chunk
    with pytest.raises(bag_mod.BagError, match="past"):
        bag_mod.BagReader(bad)
    be = bytearray(mw.encode_joint_state(5, ["a"], (1.0, 2.0)))
    be[1] = 0x00                                                # CDR_BE encapsulation
    mw.write_mcap(bad, CHANNELS, [(JOINTS, T0, bytes(be))])
    with pytest.raises(bag_mod.BagError, match="little-endian"):
        bag_mod.BagReader(bad)
    mw.write_mcap(bad, CHANNELS, [(DETS, T0, mw.encode_marker_array(T0, [{"id": 1, "x": 1.0, "y": 2.0}])[:-9])])
    with pytest.raises(bag_mod.BagError, match="past the end"):
        bag_mod.BagReader(bad)
    L = ekf_slam_b200.load()
    assert L.ekf_bag_open(None, None, None, None) == 1          # EKF_ERR_INVALID
    assert ekf_slam_b200.status_string(8).startswith("bag file")


@pytest.mark.gpu
def test_replay_from_the_bag_equals_direct_replay(tmp_path):
    """Config 2 through the file: joint states -> DiffDrive::FKin, detection arrays -> ekf_scan (unknown
    association), against the same wheel angles / sightings handed over directly."""
    _, events = _trace(n_scans=40)
    path = tmp_path / "bag_0.mcap"
    mw.write_mcap(path, CHANNELS, _messages(events), chunk_messages=32,
                  compression="zstd" if mw.compression_available("zstd") else "")

    def replay(stream):
        f = ekf_slam_b200.EkfSlam(n_landmarks=50)
        host, drive, assoc = synth.MapHost(), synth._Drive(), []
        for kind, payload in stream:
            if kind == "joint":
                drive.fkin(float(payload[0]), float(payload[1]))
            else:
                x, y, th = host.composed(drive.cfg)
                a, _, pose = f.scan((th, x, y), payload, None, None)
                assoc.append(list(a))
                host.after_update(tuple(pose), drive.cfg)
        out = assoc, f.state(), f.sigma()
        f.close()
        return out

    r = bag_mod.BagReader(path)
    from_bag = replay(("joint", r.joint_state(i)[1]) if k == bag_mod.JOINT_STATE else ("det", r.detections(i)[0])
                      for k, i, _ in r.events())
    direct = replay((k, np.array(p) if k == "joint" else np.array([(m["x"], m["y"]) for m in p]).reshape(-1, 2))
                    for k, _, p in events if k != "junk")
    assert from_bag[0] == direct[0] and len(direct[0]) == 40
    assert np.array_equal(from_bag[1], direct[1]) and np.array_equal(from_bag[2], direct[2])


def test_corrupted_files_are_rejected_or_read_never_crash(tmp_path):
    """Bit flips, overwritten words and truncations of a good file (chunked + zstd when available, and unchunked):
    the reader either returns a consistent bag or reports EKF_ERR_IO / EKF_ERR_NOMEM - it never faults."""
    _, events = _trace(n_scans=6)
    comp = "zstd" if mw.compression_available("zstd") else ""
    good = tmp_path / "good.mcap"
    mw.write_mcap(good, CHANNELS, _messages(events), chunk_messages=16, compression=comp)
    raw_a = open(good, "rb").read()
    mw.write_mcap(good, CHANNELS, _messages(events), chunk_messages=0)
    raw_b = open(good, "rb").read()
    rng = np.random.default_rng(0)
    read = rejected = 0
    path = tmp_path / "fuzzed.mcap"
    for it in range(300):
        b = bytearray(raw_a if it % 2 else raw_b)
        if it % 3 == 0:
            for _ in range(int(rng.integers(1, 6))):
                b[int(rng.integers(8, len(b)))] = int(rng.integers(0, 256))
        elif it % 3 == 1:
            b = b[: int(rng.integers(9, len(b)))]
        else:
            p = int(rng.integers(8, len(b) - 8))
            b[p:p + 8] = rng.integers(0, 256, 8, dtype=np.uint8).tobytes()
        path.write_bytes(bytes(b))
        try:
            r = bag_mod.BagReader(path)
        except bag_mod.BagError:
            rejected += 1
            continue
        for e in range(r.n_events):
            kind, index, _ = r.event(e)
            (r.joint_state if kind == bag_mod.JOINT_STATE else r.detections)(index)
        r.close()
        read += 1
    assert read + rejected == 300 and rejected > 50
