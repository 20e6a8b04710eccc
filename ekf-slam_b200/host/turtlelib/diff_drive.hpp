// API-compatible differential-drive kinematics (odometry producer for the EKF).
//
// Mirrors turtlelib/include/turtlelib/diff_drive.hpp:12-55 and
// turtlelib/src/diff_drive.cpp:5-46 of the reference.  FKin integrates the
// wheel-angle deltas exactly (arc model) and is what feeds the prediction
// step: nuslam/src/slam.cpp:604,633.  Header-only.
#ifndef EKF_SLAM_B200_TURTLELIB_DIFF_DRIVE_HPP
#define EKF_SLAM_B200_TURTLELIB_DIFF_DRIVE_HPP

#include <exception>
#include <stdexcept>

#include "turtlelib/se2d.hpp"

namespace turtlelib
{
struct wheelSpeeds
{
  double left{};
  double right{};
};

class DiffDrive
{
public:
  DiffDrive(double track, double radius)
  : wheel_track{track}, wheel_radius{radius} {}

  /// absolute wheel angles in, world->body configuration out
  Transform2D FKin(double radLeft, double radRight)
  {
    const double d_left = radLeft - phi_left;
    const double d_right = radRight - phi_right;
    const Twist2D body{wheel_radius / wheel_track * (-d_left + d_right),
      wheel_radius / 2.0 * (d_left + d_right), 0.0};
    config *= integrate_twist(body);
    phi_left = radLeft;
    phi_right = radRight;
    return config;
  }

  /// body twist in, wheel velocities out; a lateral component cannot be realised
  wheelSpeeds IKin(Twist2D twist)
  {
    if (!almost_equal(twist.y, 0.0)) {
      throw std::logic_error("wheels would slip for input twist!");
    }
    const double spin = wheel_track / 2.0 / wheel_radius * twist.omega;
    const double roll = 1.0 / wheel_radius * twist.x;
    return {-spin + roll, spin + roll};
  }

  Transform2D getConfig() {return config;}
  void setConfig(Transform2D t) {config = t;}

private:
  double wheel_track;
  double wheel_radius;
  double phi_left{};
  double phi_right{};
  Transform2D config{};
};
}  // namespace turtlelib

#endif
