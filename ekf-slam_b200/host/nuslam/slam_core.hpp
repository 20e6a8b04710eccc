#ifndef EKF_SLAM_B200_HOST_NUSLAM_SLAM_CORE_HPP
#define EKF_SLAM_B200_HOST_NUSLAM_SLAM_CORE_HPP
// nuslam::SlamCore - the filter-side state and bookkeeping of the reference's `class Slam`
// (nuslam/src/slam.cpp) as a plain C++17 class over the C ABI (include/ekf_slam_b200.h), without ROS.
// with this class the node shell keeps its shape -
//
//   jointStateCallback  (slam.cpp:599-634)   ->  core.joint_state(left, right)
//   fake_sensor_cb      (slam.cpp:180-316)   ->  core.fake_sensor(markers)      known association
//   sensor_cb           (slam.cpp:318-530)   ->  core.sensor(markers)           Mahalanobis association
//   t_map_odom, t_odom_robot, the landmark markers (slam.cpp:297-314, 512-528) -> accessors
//
// - and only the ROS message plumbing (tf broadcast, marker publishing, parameters) stays outside.
// All EKF arithmetic runs on the GPU behind ekf_scan(); this header only does the SE(2) bookkeeping the
// reference does on the host (slam.cpp:184, 273-277, 291 / :321, 490-494, 508).
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "ekf_slam_b200.h"
#include "turtlelib/diff_drive.hpp"
#include "turtlelib/se2d.hpp"

namespace nuslam
{

/// one sighting as the callbacks receive it: marker.id, marker.pose.position.{x,y}, marker.action
struct Sighting
{
  int id = 0;
  double x = 0.0;
  double y = 0.0;
  int action = 0;   // visualization_msgs::msg::Marker::ADD = 0, DELETE = 2
};

class SlamCore
{
public:
  /// wheel_track / wheel_radius: the `track_width` / `wheel_radius` parameters (slam.cpp:57-58,86);
  /// the filter constants default to the reference's hard-coded members (slam.cpp:665-671,130)
  SlamCore(double wheel_track, double wheel_radius, int n_landmarks = 50, int device = -1)
  : ddrive_{wheel_track, wheel_radius}
  {
    ekf_config cfg;
    ekf_default_config(&cfg);
    cfg.n_landmarks = n_landmarks;
    cfg.device = device;
    const int rc = ekf_create(&cfg, &ekf_);
    if (rc != EKF_OK) {
      throw std::runtime_error(std::string("ekf_create: ") + ekf_status_string(rc));
    }
  }
  ~SlamCore() {if (ekf_) {ekf_destroy(ekf_);}}
  SlamCore(const SlamCore &) = delete;
  SlamCore & operator=(const SlamCore &) = delete;

  /// slam.cpp:604,633 - wheel angles in, odometry pose out
  turtlelib::Transform2D joint_state(double left, double right)
  {
    t_odom_robot_ = ddrive_.FKin(left, right);
    return t_odom_robot_;
  }
  /// for callers that track odometry themselves
  void set_odometry(const turtlelib::Transform2D & t_odom_robot) {t_odom_robot_ = t_odom_robot;}

  /// slam.cpp:180-291: known association (nusim's fake sensor).  Returns the filter configuration.
  turtlelib::Transform2D fake_sensor(const std::vector<Sighting> & markers)
  {
    pack(markers, true);
    run(markers.size(), true);
    return filter_configuration_;
  }

  /// slam.cpp:318-508: unknown association.  Returns, per sighting, the landmark it was fused with.
  std::vector<int32_t> sensor(const std::vector<Sighting> & markers)
  {
    pack(markers, false);
    run(markers.size(), false);
    return assoc_;
  }

  const turtlelib::Transform2D & t_map_odom() const {return t_map_odom_;}
  const turtlelib::Transform2D & t_odom_robot() const {return t_odom_robot_;}
  const turtlelib::Transform2D & filter_configuration() const {return filter_configuration_;}
  unsigned int counter_obstacles() const {return static_cast<unsigned int>(count_);}

  /// estimated landmark positions (slam.cpp:304-306 / :518-520): the first `count` landmarks, (x, y) pairs
  std::vector<double> landmarks(int count) const
  {
    std::vector<double> out(static_cast<size_t>(2 * (count > 0 ? count : 0)));
    int32_t got = 0;
    if (count > 0) {check(ekf_get_landmarks(ekf_, out.data(), count, &got), "ekf_get_landmarks");}
    out.resize(static_cast<size_t>(2 * got));
    return out;
  }
  ekf_filter * handle() {return ekf_;}

private:
  static void check(int rc, const char * where)
  {
    if (rc != EKF_OK) {
      throw std::runtime_error(std::string(where) + ": " + ekf_status_string(rc));
    }
  }
  void pack(const std::vector<Sighting> & markers, bool known)
  {
    xy_.clear(); ids_.clear(); act_.clear();
    for (const auto & mk : markers) {
      xy_.push_back(mk.x); xy_.push_back(mk.y);
      if (known) {ids_.push_back(mk.id); act_.push_back(static_cast<uint8_t>(mk.action));}
    }
    assoc_.assign(markers.size(), -1);
  }
  void run(size_t m, bool known)
  {
    // slam.cpp:184 / :321 - the pose the prediction is fed with
    const auto current = t_map_odom_ * t_odom_robot_;
    const double pose[3] = {current.rotation(), current.translation().x, current.translation().y};
    double out[3] = {0.0, 0.0, 0.0};
    if (m == 0) {
      // the reference throws at msg.markers.at(0) (slam.cpp:281,498) after running the prediction
      check(ekf_predict(ekf_, pose[0], pose[1], pose[2]), "ekf_predict");
      check(ekf_get_pose(ekf_, out), "ekf_get_pose");
    } else {
      check(
        ekf_scan(ekf_, pose, static_cast<int32_t>(m), xy_.data(), known ? ids_.data() : nullptr,
        known ? act_.data() : nullptr, known ? nullptr : assoc_.data(), &count_, out), "ekf_scan");
    }
    // slam.cpp:273-277 / :490-494
    filter_configuration_ = turtlelib::Transform2D{turtlelib::Vector2D{out[1], out[2]}, out[0]};
    t_map_odom_ = filter_configuration_ * t_odom_robot_.inv();
  }

  ekf_filter * ekf_ = nullptr;
  turtlelib::DiffDrive ddrive_;
  turtlelib::Transform2D t_odom_robot_{};
  turtlelib::Transform2D t_map_odom_{};
  turtlelib::Transform2D filter_configuration_{};
  int32_t count_ = 0;
  std::vector<double> xy_;
  std::vector<int32_t> ids_, assoc_;
  std::vector<uint8_t> act_;
};

}  // namespace nuslam
#endif
