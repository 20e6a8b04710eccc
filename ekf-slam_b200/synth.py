"""Seeded synthetic inputs for the EKF hot path (BASELINE.json configs).

Host-side input generation only - restates the *input side* of the reference's simulator:
  * fake landmark sensor: nusim/src/nusim.cpp:310-349 (relative x,y in the body frame, one
    Gaussian draw per axis with stddev `sensor_variance` = 0.001 (:88,202), ADD inside
    `sensor_max_range`, DELETE outside (:331-336));
  * wheel odometry: turtlelib DiffDrive::FKin fed with encoder angles, slip U(+-slip) on the
    simulated wheels (nusim.cpp:200,224-225);
  * world: nusim/config/basic_world.yaml:8-10, robot: nuturtle_description/config/diff_params.yaml.
Nothing here touches the filter; tests and bench.py feed these traces to both the GPU path and
the CPU oracle.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

WHEEL_RADIUS = 0.033   # diff_params.yaml:3
TRACK_WIDTH = 0.160    # diff_params.yaml:4
BASIC_WORLD = np.array([[-0.5, -0.7], [0.8, -0.8], [0.4, 0.8], [-0.6, 0.65]])  # basic_world.yaml:8-9
MARKER_ADD, MARKER_DELETE = 0, 2


@dataclass
class Scan:
    odom: tuple            # t_odom_robot as (x, y, theta): what jointStateCallback leaves (slam.cpp:633)
    xy: np.ndarray         # (m, 2) body-frame sightings
    ids: np.ndarray        # (m,) landmark ids (known association) - ignored by the unknown path
    action: np.ndarray     # (m,) MARKER_ADD / MARKER_DELETE
    truth: tuple = (0.0, 0.0, 0.0)
    wheels: tuple = (0.0, 0.0)   # encoder angles (left, right) that produced `odom` (joint_states)


def _compose(a, b):
    """SE(2) compose of (x, y, th) tuples - turtlelib/src/se2d.cpp:69-77."""
    c, s = math.cos(a[2]), math.sin(a[2])
    return (c * b[0] - s * b[1] + a[0], s * b[0] + c * b[1] + a[1], a[2] + b[2])


def _inverse(a):
    c, s = math.cos(a[2]), math.sin(a[2])
    return (-a[0] * c - a[1] * s, -a[1] * c + a[0] * s, -a[2])


def _integrate_twist(w, vx, vy):
    """turtlelib/src/se2d.cpp:127-138."""
    if w == 0.0:
        return (vx, vy, 0.0)
    tsb = (vy / w, -vx / w, 0.0)
    return _compose(_compose(_inverse(tsb), (0.0, 0.0, w)), tsb)


class _Drive:
    """DiffDrive::FKin - turtlelib/src/diff_drive.cpp:10-28."""

    def __init__(self, pose=(0.0, 0.0, 0.0)):
        self.pl = self.pr = 0.0
        self.cfg = pose

    def fkin(self, left, right):
        dl, dr = left - self.pl, right - self.pr
        tw = (WHEEL_RADIUS / TRACK_WIDTH * (-dl + dr), WHEEL_RADIUS / 2.0 * (dl + dr), 0.0)
        self.cfg = _compose(self.cfg, _integrate_twist(*tw))
        self.pl, self.pr = left, right
        return self.cfg


def circle_drive(n_scans=300, seed=0, landmarks=BASIC_WORLD, radius=0.3, speed=0.1, scan_hz=5.0,
                 odom_hz=100.0, slip=0.1, sensor_std=0.001, max_range=2.0, theta0=0.0,
                 drop_out_of_range=False, shuffle=False):
    """Config 1 / 2 trace: the robot drives a circle (what the `circle` node commands) among
    `landmarks`; returns a list of Scan.  With drop_out_of_range the out-of-range sightings are
    removed (what the `landmarks` node publishes: only detections, nuslam/src/landmarks.cpp:145)
    instead of being flagged DELETE (what nusim's fake sensor does)."""
    rng = np.random.default_rng(seed)
    true_drive = _Drive((0.0, 0.0, theta0))
    odom_drive = _Drive((0.0, 0.0, theta0))
    w_cmd = speed / radius
    # wheel velocities for the commanded twist (DiffDrive::IKin, diff_drive.cpp:30-38)
    vl = -TRACK_WIDTH / 2.0 / WHEEL_RADIUS * w_cmd + speed / WHEEL_RADIUS
    vr = TRACK_WIDTH / 2.0 / WHEEL_RADIUS * w_cmd + speed / WHEEL_RADIUS
    steps_per_scan = int(round(odom_hz / scan_hz))
    dt = 1.0 / odom_hz
    enc_l = enc_r = 0.0      # what the encoders report (drives the odometry)
    true_l = true_r = 0.0    # what the wheels really did (slip)
    scans = []
    for _ in range(n_scans):
        for _ in range(steps_per_scan):
            enc_l += vl * dt
            enc_r += vr * dt
            true_l += vl * dt * (1.0 + rng.uniform(-slip, slip))
            true_r += vr * dt * (1.0 + rng.uniform(-slip, slip))
            true_drive.fkin(true_l, true_r)
            odom_drive.fkin(enc_l, enc_r)
        pose = true_drive.cfg
        inv = _inverse(pose)
        xy, ids, act = [], [], []
        for i, (lx, ly) in enumerate(landmarks):
            rel = _compose(inv, (lx, ly, 0.0))
            dist = math.sqrt(rel[0] ** 2 + rel[1] ** 2)
            in_range = dist <= max_range
            nx, ny = rng.normal(0.0, sensor_std), rng.normal(0.0, sensor_std)
            if drop_out_of_range and not in_range:
                continue
            xy.append((rel[0] + nx, rel[1] + ny))
            ids.append(i)
            act.append(MARKER_ADD if in_range else MARKER_DELETE)
        if shuffle and len(xy) > 1:
            perm = rng.permutation(len(xy))
            xy = [xy[p] for p in perm]; ids = [ids[p] for p in perm]; act = [act[p] for p in perm]
        scans.append(Scan(odom=odom_drive.cfg, xy=np.array(xy, dtype=np.float64).reshape(-1, 2),
                          ids=np.array(ids, dtype=np.int32), action=np.array(act, dtype=np.uint8),
                          truth=pose, wheels=(enc_l, enc_r)))
    return scans


def scattered_field(n_landmarks, seed=0, extent=4.0, min_sep=0.5):
    """Config 3 field: landmarks uniform in an extent x extent box, >= min_sep apart."""
    rng = np.random.default_rng(seed)
    pts = []
    while len(pts) < n_landmarks:
        p = rng.uniform(-extent / 2, extent / 2, size=2)
        if all(np.hypot(*(p - q)) >= min_sep for q in pts) and np.hypot(*p) > 0.3:
            pts.append(p)
    return np.array(pts)


class MapHost:
    """Host-side bookkeeping of the node shell around the filter: composes the pose that
    predict() receives and maintains t_map_odom (slam.cpp:184, 273-277 / 321, 490-494)."""

    def __init__(self):
        self.t_map_odom = (0.0, 0.0, 0.0)

    def composed(self, odom):
        return _compose(self.t_map_odom, odom)      # slam.cpp:184: (x, y, theta)

    def after_update(self, filter_pose_theta_xy, odom):
        th, x, y = filter_pose_theta_xy
        self.t_map_odom = _compose((x, y, th), _inverse(odom))   # slam.cpp:273-277


# ---- batched Monte-Carlo workload (BASELINE config 3), vectorised over the filters ---------------
def batch_fields(n_filters, n_landmarks=20, seed=0, pitch=0.9, jitter=0.15):
    """[B, N, 2] landmark fields: a jittered grid around the origin (separation >= pitch - 2*jitter),
    a different field per filter."""
    rng = np.random.default_rng(seed)
    w = int(math.ceil(math.sqrt(n_landmarks)))
    k = np.arange(n_landmarks)
    gx = (k % w - (w - 1) / 2.0) * pitch + 0.37
    gy = (k // w - (math.ceil(n_landmarks / w) - 1) / 2.0) * pitch + 0.29
    base = np.stack([gx, gy], axis=1)[None]
    return base + rng.uniform(-jitter, jitter, size=(n_filters, n_landmarks, 2))


def batch_pose(n_filters, step, radius=0.25, dtheta=0.05):
    """[B, 3] (theta, x, y): every robot drives a small circle, phase-shifted per filter."""
    ph = dtheta * step + 0.001 * np.arange(n_filters)
    return np.stack([ph + math.pi / 2, radius * np.cos(ph), radius * np.sin(ph)], axis=1)


def batch_scan(fields, poses, rng, m=None, sensor_std=0.001):
    """Body-frame sightings [B, m, 2] of m landmarks per filter (a random subset and order per
    filter, nusim.cpp:325-340), and the true ids [B, m]."""
    B, N, _ = fields.shape
    m = N if m is None else m
    order = np.argsort(rng.random((B, N)), axis=1)[:, :m].astype(np.int32)
    lm = np.take_along_axis(fields, order[:, :, None].astype(np.int64), axis=1)
    th, x, y = poses[:, 0:1], poses[:, 1:2], poses[:, 2:3]
    dx, dy = lm[:, :, 0] - x, lm[:, :, 1] - y
    c, s = np.cos(th), np.sin(th)
    xy = np.stack([c * dx + s * dy, -s * dx + c * dy], axis=2) + rng.normal(0.0, sensor_std, (B, m, 2))
    return xy, order
