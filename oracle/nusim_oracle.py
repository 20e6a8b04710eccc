"""CPU restatement of nusim's two sensor models - TEST INFRASTRUCTURE ONLY.

Only tests/ may import this module; the product path (ekf-slam_b200/csrc/ekf_nusim.cuh) never does.

Follows, expression by expression:
  nusim/src/nusim.cpp:310-349   sensor_timer_callback: the fake landmark sensor
  nusim/src/nusim.cpp:559-710   publish_lidar_data: 360-beam ray cast against cylinders, then walls
  nusim/src/nusim.cpp:516-555   publish_arena_walls: the four wall markers the ray cast reads
  turtlelib/src/se2d.cpp:57-74  Transform2D::inv / operator*=
  turtlelib/src/geometry2d.cpp:52-59,102-110  normalize / dot / magnitude
  tf2 Quaternion::setRPY / getAngle (third-party, absent from /root/reference; the published formulas:
  setRPY(0,0,yaw) = (0, 0, sin(yaw/2), cos(yaw/2)), getAngle() = 2 acos(clamp(w, -1, 1)))

PINNED against the reference's own object code: oracle/_ref/libnusim_ref.so is nusim/src/nusim.cpp
compiled unmodified (oracle/Makefile, oracle/nusim_ref_shim.cpp, stub ROS headers under oracle/stubs/);
tests/test_nusim.py holds this restatement to it BIT FOR BIT (lidar ranges, fake-sensor coordinates, ids
and ADD/DELETE flags, wall markers), with and without noise.  The SE(2) algebra underneath is also
pinned against the reference's turtlelib (oracle/_ref/libturtlelib_ref.so).
The random draws (std::normal_distribution over mt19937, nusim.cpp:202,341-342,706) are an input here:
`noise` arrays are added where the reference adds `sensor_normal_dist(gen)`.  Draw order in the
reference binary (g++ evaluates the two noisy create_obstacle arguments right to left): per sensor tick
360 lidar draws in beam order, then per obstacle the Y draw before the X draw.
"""
from __future__ import annotations

import math

import numpy as np

N_BEAMS = 360
RANGE_MIN, RANGE_MAX = 0.11, 10.0          # nusim.cpp:570-571
LASER_OFFSET_X = -0.032                     # nusim.cpp:578
WALL_THICKNESS = 0.1                        # nusim.cpp:518


class T2D:
    """turtlelib::Transform2D: translation (x, y) then rotation `rot` (se2d.cpp)"""
    __slots__ = ("x", "y", "rot")

    def __init__(self, x=0.0, y=0.0, rot=0.0):
        self.x, self.y, self.rot = float(x), float(y), float(rot)

    def inv(self):                                        # se2d.cpp:57-63
        c, s = math.cos(self.rot), math.sin(self.rot)
        return T2D(-self.x * c - self.y * s, -self.y * c + self.x * s, -self.rot)

    def __mul__(self, r):                                 # se2d.cpp:66-74 (operator*= on a copy)
        c, s = math.cos(self.rot), math.sin(self.rot)
        return T2D(c * r.x - s * r.y + self.x, s * r.x + c * r.y + self.y, self.rot + r.rot)


def wall_markers(arena_x, arena_y):
    """(pos_x, pos_y, scale_x, yaw used by the ray cast) of the four walls, nusim.cpp:516-555.
    The ray cast recovers the yaw from the marker quaternion with tf2's getAngle(): the sign of a
    negative yaw is lost there (bottom wall: -3.14159 -> +3.14159)."""
    t = WALL_THICKNESS
    raw = [(0.0, -(arena_y + t) / 2.0, 3.0 / 4.0 * 2.0 * 3.1416),
           (0.0, (arena_y + t) / 2.0, 1.57),
           ((arena_x + t) / 2.0, 0.0, 0.0),
           (-(arena_x + t) / 2.0, 0.0, -3.14159)]
    out = []
    for px, py, yaw in raw:
        w = math.cos(yaw * 0.5)                            # setRPY(0, 0, yaw).w
        ang = 2.0 * math.acos(max(-1.0, min(1.0, w)))      # getAngle()
        out.append((px, py, t, ang))
    return out


def wall_lines(arena_x, arena_y):
    """two points (x3, y3, x4, y4) on the line the ray cast intersects, per wall: nusim.cpp:649-653"""
    lines = []
    for px, py, sx, ang in wall_markers(arena_x, arena_y):
        T = T2D(px, py) * T2D(0.0, 0.0, ang)
        T = T * T2D(-sx, 0.0)
        T2 = T * T2D(0.0, 0.1)
        lines.append((T.x, T.y, T2.x, T2.y))
    return lines


def fake_sensor(pose, obstacles_xy, max_range, noise=None):
    """nusim.cpp:317-346 for one robot pose (theta, x, y): body-frame (x, y) per obstacle and the id /
    DELETE rule: ids[j] = j when the TRUE distance <= max_range, else -(j+1) (Marker::DELETE)."""
    th, x, y = pose
    Tinv = T2D(x, y, th).inv()
    n = len(obstacles_xy)
    xy = np.zeros((n, 2))
    ids = np.zeros(n, dtype=np.int32)
    for j, (ox, oy) in enumerate(obstacles_xy):
        rel = Tinv * T2D(ox, oy, 0.0)
        dist = math.sqrt(rel.x ** 2 + rel.y ** 2)
        ids[j] = j if dist <= max_range else -(j + 1)
        nx, ny = (0.0, 0.0) if noise is None else noise[j]
        xy[j] = (rel.x + nx, rel.y + ny)
    return xy, ids


def lidar(pose, obstacles_xyr, arena_x, arena_y, max_range, noise=None):
    """nusim.cpp:559-710 for one robot pose (theta, x, y) -> float32 ranges[360]"""
    th, x, y = pose
    lines = wall_lines(arena_x, arena_y)
    out = np.zeros(N_BEAMS, dtype=np.float32)
    Twt = T2D(x, y, th)
    for i in range(N_BEAMS):
        Ttl = T2D(LASER_OFFSET_X, 0.0) * T2D(0.0, 0.0, float(i) * 2.0 * math.pi / 360.0)
        Twl = Twt * Ttl
        p1 = Twl
        p2 = Twl * T2D(1.0, 0.0)
        Tlw = Twl.inv()
        collision = False
        closest = max_range
        for (mx, my, r) in obstacles_xyr:
            Tlm = Tlw * T2D(mx, my, 0.0)
            if Tlm.x >= 0:
                ux, uy = mx - Twl.x, my - Twl.y
                dx, dy = p2.x - p1.x, p2.y - p1.y
                mod = math.sqrt(dx ** 2 + dy ** 2)
                rx, ry = dx / mod, dy / mod
                dot = ux * rx + uy * ry
                u1x, u1y = dot * rx, dot * ry
                u2x, u2y = ux - u1x, uy - u1y
                d = math.sqrt(u2x ** 2 + u2y ** 2)
                arg = (r * 2.0 / 2.0) ** 2 - d ** 2
                m = math.sqrt(arg) if arg >= 0.0 else math.nan
                q1x, q1y = Twl.x + u1x + m * rx, Twl.y + u1y + m * ry
                q2x, q2y = Twl.x + (-1.0 * u1x) + m * rx, Twl.y + (-1.0 * u1y) + m * ry
                d1 = math.sqrt((Twl.x - q1x) ** 2 + (Twl.y - q1y) ** 2)
                d2 = math.sqrt((Twl.x - q2x) ** 2 + (Twl.y - q2y) ** 2)
                dist = d1 if d1 < d2 else d2
                if dist < closest:
                    closest = dist
                    collision = True
        if not collision:
            x1, y1, x2, y2 = p1.x, p1.y, p2.x, p2.y
            for (x3, y3, x4, y4) in lines:
                denom = (x1 - x2) * (y3 - y4) - (y1 - y2) * (x3 - x4)
                num_x = (x1 * y2 - y1 * x2) * (x3 - x4) - (x1 - x2) * (x3 * y4 - y3 * x4)
                num_y = (x1 * y2 - y1 * x2) * (y3 - y4) - (y1 - y2) * (x3 * y4 - y3 * x4)
                if denom != 0.0:
                    px, py = num_x / denom, num_y / denom
                    Tl = Tlw * T2D(px, py)
                    if Tl.x > 0:
                        dist = math.sqrt((Twl.x - px) ** 2 + (Twl.y - py) ** 2)
                        if dist < closest:
                            closest = dist
        if closest > RANGE_MAX:
            closest = RANGE_MAX
        if closest < RANGE_MIN:
            closest = RANGE_MIN
        out[i] = np.float32(closest + (0.0 if noise is None else noise[i]))
    return out
