"""Shared test drivers: push one synthetic trace through an implementation of the boundary.

Every runner follows the node shell of the reference around the filter: compose
t_map_odom * t_odom_robot (slam.cpp:184/321), scan update, t_map_odom bookkeeping
(slam.cpp:273-277 / 490-494).  They return (assoc per scan, state, sigma, count).
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import ekf_slam_b200  # noqa: E402  (loader shim at the repo root)
from ekf_slam_b200 import synth  # noqa: E402


def run_literal(scans, known, n_landmarks=50, **kw):
    import ekf_literal as lit
    f = lit.LiteralEkf(n_landmarks=n_landmarks, **kw)
    assoc = []
    for s in scans:
        f.set_odometry(*s.odom)
        if known:
            f.fake_sensor_cb([(int(i), x, y, int(a)) for i, (x, y), a in zip(s.ids, s.xy, s.action)])
            assoc.append(list(s.ids))
        else:
            assoc.append(f.sensor_cb([(x, y) for (x, y) in s.xy]))
    return assoc, f.state.copy(), f.sigma.copy(), f.counter_obstacles


def _shell_run(f, scans, known):
    """f has predict / update_known / update_unknown / pose()-like state access."""
    host = synth.MapHost()
    assoc = []
    for s in scans:
        x, y, th = host.composed(s.odom)
        f.predict(th, x, y)
        if known:
            f.update_known(s.ids, s.xy, s.action)
            assoc.append(list(s.ids))
        else:
            assoc.append(list(f.update_unknown(s.xy)))
        pose = f.state[:3] if hasattr(f, "state") and not callable(f.state) else f.pose()
        host.after_update(tuple(pose), s.odom)
    return assoc


def run_struct(scans, known, n_landmarks=50, **kw):
    import struct_oracle as so
    f = so.StructEkf(n_landmarks=n_landmarks, **kw)
    assoc = _shell_run(f, scans, known)
    out = assoc, f.state.copy(), f.sigma.copy(), f.count
    f.close()
    return out


def run_model(scans, known, n_landmarks=50, **kw):
    from _deferred_model import DeferredModel
    f = DeferredModel(n_landmarks=n_landmarks, **kw)
    host = synth.MapHost()
    assoc = []
    for s in scans:
        x, y, th = host.composed(s.odom)
        a = f.scan((th, x, y), s.xy, s.ids if known else None, s.action if known else None)
        assoc.append(list(s.ids) if known else a)
        host.after_update(tuple(f.mu[:3]), s.odom)
    return assoc, f.mu.copy(), f.sigma(), f.cnt


def run_gpu(scans, known, n_landmarks=50, fused=False, overlap=False, no_action=False, probe=None, **kw):
    """overlap: ekf_set_overlap(1) (covariance pass of a scan behind the next scan's sequential side; fused scans
    only).  no_action: known association without the per-sighting action array (the overlapped path takes ids but
    not actions).  probe(f, scan_index): called between scans (accessor traffic in the middle of the pipeline)."""
    f = ekf_slam_b200.EkfSlam(n_landmarks=n_landmarks, **kw)
    if overlap:
        assert f.set_overlap(True)
    host = synth.MapHost()
    assoc = []
    for s in scans:
        x, y, th = host.composed(s.odom)
        if probe is not None:
            probe(f, len(assoc))
        if fused:
            a, _, pose = f.scan((th, x, y), s.xy, s.ids if known else None,
                                s.action if (known and not no_action) else None)
            assoc.append(list(s.ids) if known else list(a))
        else:
            f.predict(th, x, y)
            if known:
                f.update_known(s.ids, s.xy, s.action)
                assoc.append(list(s.ids))
            else:
                assoc.append(list(f.update_unknown(s.xy)))
            pose = f.pose()
        host.after_update(tuple(pose), s.odom)
    out = assoc, f.state(), f.sigma(), f.count
    f.close()
    return out


def rel_fro(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def active_block(sig, count):
    k = 3 + 2 * count
    return sig[:k, :k]
