"""ctypes mirror of the rosbag2 / MCAP reader (include/ekf_slam_b200.h, section N4): the node's two input topics,
`/blue/joint_states` (slam.cpp:599-634) and `/green/detected_obstacles` (slam.cpp:318-530), as numpy arrays in the
order the reference's single-threaded executor would have dispatched them."""
import ctypes as C

import numpy as np

from . import load, status_string

JOINT_STATE, DETECTIONS = 0, 1


class BagError(RuntimeError):
    pass


class BagReader:
    def __init__(self, path, joint_topic=None, detections_topic=None):
        self._L = load()
        self._h = C.c_void_p()
        enc = lambda s: None if s is None else s.encode()
        st = self._L.ekf_bag_open(str(path).encode(), enc(joint_topic), enc(detections_topic), C.byref(self._h))
        if st != 0:
            self._h = C.c_void_p()
            raise BagError(f"{status_string(st)}: {self._L.ekf_bag_last_error().decode()}")
        c = [C.c_int64(0) for _ in range(5)]
        self._L.ekf_bag_counts(self._h, *[C.byref(x) for x in c])
        self.n_events, self.n_joint_states, self.n_detection_arrays, self.messages_total, self.chunks = \
            [int(x.value) for x in c]

    def close(self):
        if self._h:
            self._L.ekf_bag_close(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def event(self, e):
        """(kind, index, log_time_ns) of event e."""
        k, i, t = C.c_int32(0), C.c_int32(0), C.c_uint64(0)
        st = self._L.ekf_bag_event(self._h, e, C.byref(k), C.byref(i), C.byref(t))
        if st != 0:
            raise IndexError(e)
        return int(k.value), int(i.value), int(t.value)

    def joint_state(self, i):
        """(stamp_ns, positions[n_joints]) - positions[0] / [1] are the left / right wheel angles."""
        stamp, n = C.c_int64(0), C.c_int32(0)
        if self._L.ekf_bag_joint_state(self._h, i, C.byref(stamp), C.byref(n), 0, None) != 0:
            raise IndexError(i)
        pos = np.zeros(max(int(n.value), 1), dtype=np.float64)
        self._L.ekf_bag_joint_state(self._h, i, None, None, int(n.value), pos.ctypes.data_as(C.POINTER(C.c_double)))
        return int(stamp.value), pos[: int(n.value)]

    def detections(self, i):
        """(xy[m, 2], ids[m], actions[m]) of detection array i."""
        n = C.c_int32(0)
        if self._L.ekf_bag_detections(self._h, i, C.byref(n), 0, None, None, None) != 0:
            raise IndexError(i)
        m = int(n.value)
        xy = np.zeros((max(m, 1), 2), dtype=np.float64)
        ids = np.zeros(max(m, 1), dtype=np.int32)
        act = np.zeros(max(m, 1), dtype=np.int32)
        ip = C.POINTER(C.c_int32)
        self._L.ekf_bag_detections(self._h, i, None, m, xy.ctypes.data_as(C.POINTER(C.c_double)),
                                   ids.ctypes.data_as(ip), act.ctypes.data_as(ip))
        return xy[:m], ids[:m], act[:m]

    def events(self):
        for e in range(self.n_events):
            yield self.event(e)
