// API-compatible SE(2) types: Twist2D, Transform2D, integrate_twist.
//
// Mirrors turtlelib/include/turtlelib/se2d.hpp:15-127 and
// turtlelib/src/se2d.cpp:20-138 of the reference (same class/method names,
// same composition formula, rotation NOT re-normalised on compose - the EKF's
// pose injection relies on that: nuslam/src/slam.cpp:184-188).  Header-only.
#ifndef EKF_SLAM_B200_TURTLELIB_SE2D_HPP
#define EKF_SLAM_B200_TURTLELIB_SE2D_HPP

#include <cctype>
#include <cmath>
#include <iosfwd>
#include <iostream>

#include "turtlelib/geometry2d.hpp"

namespace turtlelib
{
struct Twist2D
{
  double omega = 0.0;
  double x = 0.0;
  double y = 0.0;
};

inline std::ostream & operator<<(std::ostream & os, const Twist2D & tw)
{
  return os << "[" << tw.omega << " " << tw.x << " " << tw.y << "]";
}
inline std::istream & operator>>(std::istream & is, Twist2D & tw)
{
  if (is.peek() == '[') {is.get();}
  is >> tw.omega >> tw.x >> tw.y;
  while (is.good() && is.get() != '\n') {}
  return is;
}

class Transform2D
{
public:
  Transform2D() = default;
  explicit Transform2D(Vector2D trans)
  : trans_{trans} {}
  explicit Transform2D(double radians)
  : rot_{radians} {}
  Transform2D(Vector2D trans, double radians)
  : trans_{trans}, rot_{radians} {}

  /// apply to a point (rotate, then translate) - se2d.cpp:35-40
  Point2D operator()(Point2D p) const
  {
    const double c = std::cos(rot_), s = std::sin(rot_);
    return Point2D{c * p.x - s * p.y + trans_.x, s * p.x + c * p.y + trans_.y};
  }
  /// apply to a free vector (rotation only) - se2d.cpp:43-48
  Vector2D operator()(Vector2D v) const
  {
    const double c = std::cos(rot_), s = std::sin(rot_);
    return Vector2D{c * v.x - s * v.y, s * v.x + c * v.y};
  }
  /// adjoint map of a twist - se2d.cpp:51-57
  Twist2D operator()(Twist2D v) const
  {
    const double c = std::cos(rot_), s = std::sin(rot_);
    return Twist2D{v.omega, trans_.y * v.omega + c * v.x - s * v.y,
      -trans_.x * v.omega + s * v.x + c * v.y};
  }
  /// inverse transform - se2d.cpp:60-66
  Transform2D inv() const
  {
    const double c = std::cos(rot_), s = std::sin(rot_);
    return Transform2D{Vector2D{-trans_.x * c - trans_.y * s, -trans_.y * c + trans_.x * s}, -rot_};
  }
  /// right-compose: this = this * rhs - se2d.cpp:69-77 (angle is a plain sum)
  Transform2D & operator*=(const Transform2D & rhs)
  {
    const double c = std::cos(rot_), s = std::sin(rot_);
    const double nx = c * rhs.trans_.x - s * rhs.trans_.y + trans_.x;
    const double ny = s * rhs.trans_.x + c * rhs.trans_.y + trans_.y;
    rot_ = rot_ + rhs.rot_;
    trans_.x = nx;
    trans_.y = ny;
    return *this;
  }
  Vector2D translation() const {return trans_;}
  double rotation() const {return rot_;}

  friend std::ostream & operator<<(std::ostream & os, const Transform2D & tf)
  {
    return os << "deg: " << rad2deg(tf.rot_) << " x: " << tf.trans_.x << " y: " << tf.trans_.y;
  }

private:
  Vector2D trans_{};
  double rot_{};
};

/// Accepts "deg: 90 x: 3 y: 5", "90 3 5" and newline-separated numbers (se2d.cpp:96-120).
inline std::istream & operator>>(std::istream & is, Transform2D & tf)
{
  double vals[3] = {0.0, 0.0, 0.0};
  const bool labelled = is.peek() == 'd';
  for (double & v : vals) {
    if (labelled) {
      while (is.good() && !std::isdigit(is.peek()) && is.peek() != '-') {is.get();}
    }
    is >> v;
  }
  tf = Transform2D{Vector2D{vals[1], vals[2]}, deg2rad(vals[0])};
  return is;
}

inline Transform2D operator*(Transform2D lhs, const Transform2D & rhs) {return lhs *= rhs;}

/// Exact integration of a constant body twist over unit time (se2d.cpp:127-138):
/// pure translation when omega == 0, otherwise rotate about the twist's centre.
inline Transform2D integrate_twist(Twist2D twist)
{
  if (twist.omega == 0.0) {
    return Transform2D{Vector2D{twist.x, twist.y}};
  }
  const Transform2D t_sb{Vector2D{twist.y / twist.omega, -twist.x / twist.omega}};
  return t_sb.inv() * Transform2D{twist.omega} * t_sb;
}
}  // namespace turtlelib

#endif
