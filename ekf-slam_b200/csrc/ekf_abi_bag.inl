// C ABI of the rosbag2 / MCAP reader (include/ekf_slam_b200.h, "N4"): host-only glue over
// host/rosbag/mcap_reader.hpp.  Exceptions stop here: every failure becomes EKF_ERR_IO / EKF_ERR_INVALID plus a
// thread-local message.
#include "../host/rosbag/mcap_reader.hpp"

struct ekf_bag {
  rosbag::McapReader reader;
  ekf_bag(const char *path, const char *jt, const char *dt)
      : reader(path, jt ? jt : "/blue/joint_states", dt ? dt : "/green/detected_obstacles") {}
};

namespace {
thread_local std::string g_bag_error;
}

int ekf_bag_open(const char *path, const char *joint_topic, const char *detections_topic, ekf_bag **out) {
  if (!path || !out) return EKF_ERR_INVALID;
  *out = nullptr;
  g_bag_error.clear();
  try {
    *out = new ekf_bag(path, joint_topic, detections_topic);
  } catch (const std::bad_alloc &) {
    g_bag_error = "out of memory";
    return EKF_ERR_NOMEM;
  } catch (const std::exception &e) {
    g_bag_error = e.what();
    return EKF_ERR_IO;
  }
  return EKF_OK;
}

int ekf_bag_close(ekf_bag *b) {
  delete b;
  return EKF_OK;
}

const char *ekf_bag_last_error(void) { return g_bag_error.c_str(); }

int ekf_bag_counts(const ekf_bag *b, int64_t *events, int64_t *joint_states, int64_t *detection_arrays,
                   int64_t *messages_total, int64_t *chunks) {
  if (!b) return EKF_ERR_INVALID;
  if (events) *events = (int64_t)b->reader.events.size();
  if (joint_states) *joint_states = (int64_t)b->reader.joint_states.size();
  if (detection_arrays) *detection_arrays = (int64_t)b->reader.detections.size();
  if (messages_total) *messages_total = (int64_t)b->reader.messages_total;
  if (chunks) *chunks = (int64_t)b->reader.chunks;
  return EKF_OK;
}

int ekf_bag_event(const ekf_bag *b, int64_t e, int32_t *kind, int32_t *index, uint64_t *log_time_ns) {
  if (!b || e < 0 || e >= (int64_t)b->reader.events.size()) return EKF_ERR_INVALID;
  const auto &ev = b->reader.events[(size_t)e];
  if (kind) *kind = ev.kind;
  if (index) *index = ev.index;
  if (log_time_ns) *log_time_ns = ev.log_time;
  return EKF_OK;
}

int ekf_bag_joint_state(const ekf_bag *b, int32_t i, int64_t *stamp_ns, int32_t *n_joints, int32_t cap,
                        double *positions) {
  if (!b || i < 0 || i >= (int32_t)b->reader.joint_states.size() || cap < 0 || (cap > 0 && !positions))
    return EKF_ERR_INVALID;
  const auto &m = b->reader.joint_states[(size_t)i];
  if (stamp_ns) *stamp_ns = m.stamp_ns;
  if (n_joints) *n_joints = (int32_t)m.position.size();
  for (int32_t q = 0; q < cap && q < (int32_t)m.position.size(); ++q) positions[q] = m.position[(size_t)q];
  return EKF_OK;
}

int ekf_bag_detections(const ekf_bag *b, int32_t i, int32_t *n_markers, int32_t cap, double *xy, int32_t *ids,
                       int32_t *actions) {
  if (!b || i < 0 || i >= (int32_t)b->reader.detections.size() || cap < 0) return EKF_ERR_INVALID;
  const auto &m = b->reader.detections[(size_t)i];
  if (n_markers) *n_markers = (int32_t)m.markers.size();
  for (int32_t q = 0; q < cap && q < (int32_t)m.markers.size(); ++q) {
    if (xy) { xy[2 * q] = m.markers[(size_t)q].x; xy[2 * q + 1] = m.markers[(size_t)q].y; }
    if (ids) ids[q] = m.markers[(size_t)q].id;
    if (actions) actions[q] = m.markers[(size_t)q].action;
  }
  return EKF_OK;
}
