"""Extended-precision "truth" for small filters (oracle construction plan): the reference's
sensor callbacks (nuslam/src/slam.cpp:180-316 known, :318-530 unknown association) evaluated in x87
80-bit long double (64-bit mantissa) on the SAME double inputs.  Its rounding error is ~2,000 x smaller
than any fp64 evaluation's, so the distance of oracle A, oracle B and the GPU path from it measures each
one's own floating-point error - in particular the cancellation floor the reference's 1e7 landmark prior
puts on the observed block (1e7 - ~1e7 at the first sighting of a landmark).

NOT an oracle for the product and not on the product path: tests only (tests/test_truth_floor.py).  The
covariance algebra is written as the rank-2 form Sigma - K (H Sigma), which equals the reference's
(I - K H) Sigma (slam.cpp:264-265) in exact arithmetic; at 64 mantissa bits the difference is invisible
at the scale being measured.
"""
from __future__ import annotations

import numpy as np

LD = np.longdouble
PI = LD(np.pi)            # turtlelib::PI is the double constant (geometry2d.hpp); keep ITS value
DELETE = 2


def normalize_angle(rad):                                     # geometry2d.cpp:5-14
    d = np.fmod(LD(rad) + PI, LD(2.0) * PI)
    return d + PI if d <= 0 else d - PI


class T2:
    def __init__(self, x=0.0, y=0.0, rot=0.0):
        self.x, self.y, self.rot = LD(x), LD(y), LD(rot)

    def inv(self):                                            # se2d.cpp:57-63
        c, s = np.cos(self.rot), np.sin(self.rot)
        return T2(-self.x * c - self.y * s, -self.y * c + self.x * s, -self.rot)

    def __mul__(self, r):                                     # se2d.cpp:66-74
        c, s = np.cos(self.rot), np.sin(self.rot)
        return T2(c * r.x - s * r.y + self.x, s * r.x + c * r.y + self.y, self.rot + r.rot)


class TruthEkf:
    def __init__(self, n_landmarks=50, q_noise=1.0e-2, r_noise=1.0e-2, mah_threshold=2.0, prior=10e6):
        self.N = n_landmarks
        self.n = n = 3 + 2 * n_landmarks
        self.thr = LD(mah_threshold)
        self.sigma = np.zeros((n, n), dtype=LD)
        self.sigma[np.arange(3, n), np.arange(3, n)] = LD(prior)        # slam.cpp:128-132
        self.q, self.r = LD(q_noise), LD(r_noise)
        self.state = np.zeros(n, dtype=LD)
        self.count = 0
        self.t_map_odom, self.prev, self.t_odom_robot = T2(), T2(), T2()

    def set_odometry(self, x, y, theta):
        self.t_odom_robot = T2(x, y, theta)

    def _predict(self):                                       # slam.cpp:184-198
        cur = self.t_map_odom * self.t_odom_robot
        s = self.state
        s[0], s[1], s[2] = normalize_angle(cur.rot), cur.x, cur.y
        dx, dy = cur.x - self.prev.x, cur.y - self.prev.y
        S = self.sigma
        S[1, :] += -dy * S[0, :]
        S[2, :] += dx * S[0, :]
        S[:, 1] += -dy * S[:, 0]
        S[:, 2] += dx * S[:, 0]
        S[0, 0] += self.q; S[1, 1] += self.q; S[2, 2] += self.q

    def _lin(self, j):                                        # slam.cpp:219-249
        s = self.state
        dx, dy = s[3 + 2 * j] - s[1], s[4 + 2 * j] - s[2]
        d = dx * dx + dy * dy
        sd = np.sqrt(d)
        H = np.zeros((2, 5), dtype=LD)                        # columns: theta, x, y, m_x, m_y
        H[0] = (0, -dx / sd, -dy / sd, dx / sd, dy / sd)
        H[1] = (-1, dy / d, -dx / d, -dy / d, dx / d)
        zhat = np.array([sd, normalize_angle(np.arctan2(dy, dx) - s[0])], dtype=LD)
        return H, zhat

    def _innovation(self, j, z):
        H, zhat = self._lin(j)
        idx = [0, 1, 2, 3 + 2 * j, 4 + 2 * j]
        HS = H @ self.sigma[idx, :]                           # 2 x n = H Sigma
        Sm = HS[:, idx] @ H.T
        Sm[0, 0] += self.r; Sm[1, 1] += self.r
        zd = z - zhat
        zd[1] = normalize_angle(zd[1])
        det = Sm[0, 0] * Sm[1, 1] - Sm[0, 1] * Sm[1, 0]
        Si = np.array([[Sm[1, 1], -Sm[0, 1]], [-Sm[1, 0], Sm[0, 0]]], dtype=LD) / det
        return H, HS, Si, zd, idx

    def _update(self, j, z):                                  # slam.cpp:252-267
        H, HS, Si, zd, idx = self._innovation(j, z)
        K = (self.sigma[:, idx] @ H.T) @ Si                   # Sigma H^T S^-1
        self.state = self.state + K @ zd
        self.sigma = self.sigma - K @ HS
        self.state[0] = normalize_angle(self.state[0])

    def _epilogue(self):                                      # slam.cpp:273-277,291
        filt = T2(self.state[1], self.state[2], self.state[0])
        self.t_map_odom = filt * self.t_odom_robot.inv()
        self.prev = filt

    def fake_sensor_cb(self, markers):
        self._predict()
        for (mid, mx, my, action) in markers:
            if action == DELETE:
                continue
            mx, my = LD(mx), LD(my)
            z = np.array([np.sqrt(mx * mx + my * my), np.arctan2(my, mx)], dtype=LD)
            s = self.state
            if s[3 + 2 * mid] == 0 and s[4 + 2 * mid] == 0:   # slam.cpp:213
                s[3 + 2 * mid] = s[1] + z[0] * np.cos(z[1] + s[0])
                s[4 + 2 * mid] = s[2] + z[0] * np.sin(z[1] + s[0])
            self._update(mid, z)
        self._epilogue()

    def sensor_cb(self, markers):
        """returns (assoc, margins): margins[i] = |d_best - threshold| / threshold of the decision, so a test
        can tell a genuine association difference from a knife-edge one"""
        self._predict()
        assoc, margins = [], []
        for (mx, my) in markers:
            mx, my = LD(mx), LD(my)
            z = np.array([np.sqrt(mx * mx + my * my), np.arctan2(my, mx)], dtype=LD)
            s = self.state
            best, kb = LD(np.inf), -1
            for k in range(self.count):                       # slam.cpp:361-404
                _, _, Si, zd, _ = self._innovation(k, z)
                d = zd @ Si @ zd
                if d < best:
                    best, kb = d, k
            margins.append(float(abs(best - self.thr) / self.thr) if kb >= 0 else 1.0)
            if kb >= 0 and best <= self.thr:                  # slam.cpp:406-440 (tentative slot holds thr)
                j = kb
            else:
                j = self.count
                s[3 + 2 * j] = s[1] + z[0] * np.cos(z[1] + s[0])      # slam.cpp:351-354
                s[4 + 2 * j] = s[2] + z[0] * np.sin(z[1] + s[0])
                self.count += 1
            assoc.append(j)
            self._update(j, z)
        self._epilogue()
        return assoc, margins


def run_truth(scans, known, n_landmarks=50, **kw):
    f = TruthEkf(n_landmarks=n_landmarks, **kw)
    assoc, margins = [], []
    for s in scans:
        f.set_odometry(*s.odom)
        if known:
            f.fake_sensor_cb([(int(i), x, y, int(a)) for i, (x, y), a in zip(s.ids, s.xy, s.action)])
            assoc.append(list(s.ids))
        else:
            a, mg = f.sensor_cb([(x, y) for (x, y) in s.xy])
            assoc.append(a); margins += mg
    return assoc, f.state, f.sigma, (4 if known else f.count), margins
