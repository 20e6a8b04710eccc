"""Oracle A - literal, dense, fp64 restatement of the reference EKF-SLAM callbacks.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path: only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import it, and only as the checker or the reported CPU baseline.

PARITY UNPINNED: the reference has no test, golden vector or fixture for this
path (nuslam/CMakeLists.txt:37-47 is lint only) and its own build (Armadillo +
ROS 2) is not available here, so this restatement cannot be checked against
reference outputs.  What IS pinned: `normalize_angle` and the SE(2) helpers,
against the reference's Catch2 vectors and against the reference's turtlelib
sources compiled into oracle/_ref (tests/test_oracle_pins.py).

Every method cites the lines of /root/reference/nuslam/src/slam.cpp it follows.
Dense matrices, the same expression order as the reference's Armadillo
expressions (three-operand products are evaluated left to right, which is what
Armadillo's glue_times picks for these shapes), an explicit 2x2 adjugate
inverse (Armadillo's tiny-matrix path), numpy/OpenBLAS dgemm for the products
(the BLAS family Armadillo links on Ubuntu).  O(n^3) per measurement: use for
n <= ~2,000 only; oracle B (ekf_struct.c) is the O(n^2) form for large maps.
"""
from __future__ import annotations

import math

import numpy as np

PI = 3.14159265358979323846  # turtlelib/include/turtlelib/geometry2d.hpp:14

DELETE = 2  # visualization_msgs/Marker::DELETE
ADD = 0


def normalize_angle(rad: float) -> float:
    """turtlelib/src/geometry2d.cpp:5-14 - wrap to (-PI, PI], -PI maps to +PI."""
    diff = math.fmod(rad + PI, 2 * PI)
    if diff <= 0.0:
        return diff + PI
    return diff - PI


class Transform2D:
    """turtlelib/src/se2d.cpp:20-84,123-125 (compose, inv, accessors)."""

    __slots__ = ("x", "y", "rot")

    def __init__(self, x: float = 0.0, y: float = 0.0, rot: float = 0.0):
        self.x, self.y, self.rot = float(x), float(y), float(rot)

    def inv(self) -> "Transform2D":  # se2d.cpp:60-66
        c, s = math.cos(self.rot), math.sin(self.rot)
        return Transform2D(-self.x * c - self.y * s, -self.y * c + self.x * s, -self.rot)

    def __mul__(self, rhs: "Transform2D") -> "Transform2D":  # se2d.cpp:69-77
        c, s = math.cos(self.rot), math.sin(self.rot)
        return Transform2D(c * rhs.x - s * rhs.y + self.x, s * rhs.x + c * rhs.y + self.y,
                           self.rot + rhs.rot)

    def apply_point(self, px: float, py: float):  # se2d.cpp:35-40
        c, s = math.cos(self.rot), math.sin(self.rot)
        return c * px - s * py + self.x, s * px + c * py + self.y


def integrate_twist(omega: float, vx: float, vy: float) -> Transform2D:
    """turtlelib/src/se2d.cpp:127-138."""
    if omega == 0.0:
        return Transform2D(vx, vy, 0.0)
    tsb = Transform2D(vy / omega, -vx / omega, 0.0)
    return tsb.inv() * Transform2D(0.0, 0.0, omega) * tsb


class DiffDrive:
    """turtlelib/src/diff_drive.cpp:5-28 (FKin only; that is what feeds the filter)."""

    def __init__(self, track: float, radius: float):
        self.track, self.radius = track, radius
        self.phi_l = self.phi_r = 0.0
        self.config = Transform2D()

    def fkin(self, rad_left: float, rad_right: float) -> Transform2D:
        dl, dr = rad_left - self.phi_l, rad_right - self.phi_r
        tw = (self.radius / self.track * (-dl + dr), self.radius / 2.0 * (dl + dr), 0.0)
        self.config = self.config * integrate_twist(*tw)
        self.phi_l, self.phi_r = rad_left, rad_right
        return self.config


def _inv2(a: np.ndarray) -> np.ndarray:
    """Explicit adjugate/determinant inverse of a 2x2 (arma::inv tiny-matrix path)."""
    det = a[0, 0] * a[1, 1] - a[0, 1] * a[1, 0]
    return np.array([[a[1, 1] / det, -a[0, 1] / det], [-a[1, 0] / det, a[0, 0] / det]])


def _index_min(d: np.ndarray) -> int:
    """arma::index_min: strict '<' scan starting from +inf, so the FIRST minimum wins
    and NaN entries are never selected (slam.cpp:418)."""
    best, idx = math.inf, 0
    for i, v in enumerate(d):
        if v < best:
            best, idx = v, i
    return idx


class LiteralEkf:
    """The filter members and the two sensor callbacks of class Slam.

    Members: slam.cpp:659-676; initialisation: slam.cpp:127-139.
    """

    def __init__(self, n_landmarks: int = 50, q_noise: float = 1.0e-2, r_noise: float = 1.0e-2,
                 mah_threshold: float = 2.0, prior: float = 10e6):
        self.n_landmarks = n_landmarks
        n = 2 * n_landmarks + 3
        self.n = n
        self.mah_threshold = mah_threshold
        # slam.cpp:128-132
        self.sigma = np.zeros((n, n))
        self.sigma[3:, 3:] = np.eye(n - 3) * prior
        # slam.cpp:135-136
        self.q_bar = np.zeros((n, n))
        self.q_bar[0:3, 0:3] = np.eye(3) * q_noise
        # slam.cpp:139
        self.R = np.eye(2) * r_noise
        self.state = np.zeros(n)  # slam.cpp:674
        self.counter_obstacles = 0  # slam.cpp:670
        self.t_map_odom = Transform2D()  # slam.cpp:659
        self.filter_previous_configuration = Transform2D()  # slam.cpp:660
        self.t_odom_robot = Transform2D()  # slam.cpp:650 (set by jointStateCallback :633)

    # -- shared pieces -------------------------------------------------------------------
    def _inject_and_predict(self):
        """slam.cpp:184-198 == slam.cpp:321-335."""
        cur = self.t_map_odom * self.t_odom_robot
        self.state[0] = normalize_angle(cur.rot)
        self.state[1] = cur.x
        self.state[2] = cur.y
        dx = cur.x - self.filter_previous_configuration.x
        dy = cur.y - self.filter_previous_configuration.y
        at = np.eye(self.n)
        at[1, 0] = -dy
        at[2, 0] = dx
        self.sigma = at @ self.sigma @ at.T + self.q_bar

    def _h_dense(self, j: int):
        """slam.cpp:234-249 / 377-392 / 458-473: dense 2 x n Jacobian with 9 non-zeros."""
        s = self.state
        delta_x = s[3 + 2 * j] - s[1]
        delta_y = s[3 + 2 * j + 1] - s[2]
        d = delta_x * delta_x + delta_y * delta_y
        h = np.zeros((2, self.n))
        h[0, 1] = -delta_x / math.sqrt(d)
        h[0, 2] = -delta_y / math.sqrt(d)
        h[1, 0] = -1.0
        h[1, 1] = delta_y / d
        h[1, 2] = -delta_x / d
        h[0, 3 + 2 * j] = delta_x / math.sqrt(d)
        h[0, 3 + 2 * j + 1] = delta_y / math.sqrt(d)
        h[1, 3 + 2 * j] = -delta_y / d
        h[1, 3 + 2 * j + 1] = delta_x / d
        return h

    def _z_hat(self, j: int):
        """slam.cpp:219-227 / 364-370 / 444-451."""
        s = self.state
        est_range = math.sqrt(math.pow(s[3 + 2 * j] - s[1], 2) + math.pow(s[3 + 2 * j + 1] - s[2], 2))
        est_bearing = math.atan2(s[3 + 2 * j + 1] - s[2], s[3 + 2 * j] - s[1]) - s[0]
        return np.array([est_range, normalize_angle(est_bearing)])

    def _measurement_update(self, j: int, z: np.ndarray):
        """slam.cpp:219-267 / 443-488: gain, state and covariance update for landmark j."""
        z_hat = self._z_hat(j)
        hj = self._h_dense(j)
        k = self.sigma @ hj.T @ _inv2(hj @ self.sigma @ hj.T + self.R)  # :252 / :476
        z_diff = z - z_hat
        z_diff[1] = normalize_angle(z_diff[1])  # :256 / :480
        self.state = self.state + k @ z_diff  # :261 / :482
        self.sigma = (np.eye(self.n) - k @ hj) @ self.sigma  # :264-265 / :485-486
        self.state[0] = normalize_angle(self.state[0])  # :267 / :488

    def _epilogue(self):
        """slam.cpp:273-277,291 / 490-494,508."""
        filt = Transform2D(self.state[1], self.state[2], self.state[0])
        self.t_map_odom = filt * self.t_odom_robot.inv()
        self.filter_previous_configuration = filt

    # -- known association ---------------------------------------------------------------
    def fake_sensor_cb(self, markers):
        """slam.cpp:180-316.  markers: iterable of (id, x, y, action)."""
        self._inject_and_predict()
        for (mid, mx, my, action) in markers:
            if action == DELETE:  # :205
                continue
            meas_range = math.sqrt(math.pow(mx, 2) + math.pow(my, 2))  # :208-209
            meas_bearing = math.atan2(my, mx)  # :210
            s = self.state
            if s[3 + mid * 2] == 0.0 and s[3 + mid * 2 + 1] == 0.0:  # :213
                s[3 + mid * 2] = s[1] + meas_range * math.cos(meas_bearing + s[0])
                s[3 + mid * 2 + 1] = s[2] + meas_range * math.sin(meas_bearing + s[0])
            self._measurement_update(mid, np.array([meas_range, meas_bearing]))
        self._epilogue()

    # -- unknown association -------------------------------------------------------------
    def sensor_cb(self, markers):
        """slam.cpp:318-530.  markers: iterable of (x, y).  Returns the landmark index each
        measurement was fused with (the value `min_index` takes at :443)."""
        self._inject_and_predict()
        assoc = []
        for (mx, my) in markers:
            meas_range = math.sqrt(math.pow(mx, 2) + math.pow(my, 2))  # :345-346
            meas_bearing = math.atan2(my, mx)  # :347
            s = self.state
            c = self.counter_obstacles
            if c >= self.n_landmarks:
                # the reference indexes past the end here (slam.cpp:351 with :665)
                raise IndexError("landmark capacity exceeded")
            s[3 + c * 2] = s[1] + meas_range * math.cos(meas_bearing + s[0])  # :351
            s[3 + c * 2 + 1] = s[2] + meas_range * math.sin(meas_bearing + s[0])  # :352-354
            self.counter_obstacles += 1  # :356
            dists = np.zeros(self.counter_obstacles)
            z = np.array([meas_range, meas_bearing])
            for k in range(self.counter_obstacles):  # :361
                z_hat = self._z_hat(k)
                hk = self._h_dense(k)
                psi = hk @ self.sigma @ hk.T + self.R  # :395
                z_diff = z - z_hat
                z_diff[1] = normalize_angle(z_diff[1])  # :399
                mah = z_diff @ _inv2(psi) @ z_diff  # :401
                dist = mah
                if k == self.counter_obstacles - 1:  # :406-408
                    dist = self.mah_threshold
                dists[k] = dist
            min_index = _index_min(dists)  # :418
            mn = dists[min_index]  # :419
            if min_index == self.counter_obstacles - 1 and mn >= self.mah_threshold:  # :421
                pass
            elif min_index == self.counter_obstacles - 1 and mn < self.mah_threshold:  # :423 (dead)
                dists[min_index] = 1000.0
                min_index = _index_min(dists)
                self.counter_obstacles -= 1
                s[3 + self.counter_obstacles * 2] = 0.0
                s[3 + self.counter_obstacles * 2 + 1] = 0.0
            else:  # :433-440
                self.counter_obstacles -= 1
                s[3 + self.counter_obstacles * 2] = 0.0
                s[3 + self.counter_obstacles * 2 + 1] = 0.0
            assoc.append(min_index)
            self._measurement_update(min_index, z)  # :443-488
        self._epilogue()
        return assoc

    # -- helpers for the boundary-level tests -----------------------------------------------
    def set_odometry(self, x: float, y: float, theta: float):
        """what jointStateCallback leaves in t_odom_robot (slam.cpp:604,633)."""
        self.t_odom_robot = Transform2D(x, y, theta)

    def composed_pose(self):
        """t_map_odom * t_odom_robot, the pose the C ABI's predict() receives (slam.cpp:184)."""
        cur = self.t_map_odom * self.t_odom_robot
        return cur.rot, cur.x, cur.y
