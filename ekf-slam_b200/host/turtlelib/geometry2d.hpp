// API-compatible 2-D geometry primitives for the EKF-SLAM host side.
//
// Same names, signatures and numerical behaviour as the reference's
// turtlelib/include/turtlelib/geometry2d.hpp:14-198 and
// turtlelib/src/geometry2d.cpp:5-136, written header-only so the C++ host
// wrapper (ekf_slam.hpp), the harness and the tests need no separate library.
// Checked against the reference's own vectors (turtlelib/tests/test_geometry2d.cpp)
// in tests/test_host_types.py, and bit-for-bit against the reference sources
// compiled into oracle/_ref when /root/reference is present.
#ifndef EKF_SLAM_B200_TURTLELIB_GEOMETRY2D_HPP
#define EKF_SLAM_B200_TURTLELIB_GEOMETRY2D_HPP

#include <cmath>
#include <iosfwd>
#include <iostream>

namespace turtlelib
{
constexpr double PI = 3.14159265358979323846;

/// absolute-tolerance comparison (reference: geometry2d.hpp:22-30)
constexpr bool almost_equal(double d1, double d2, double epsilon = 1.0e-12)
{
  const double delta = d1 - d2;
  return (delta < 0 ? -delta : delta) < epsilon;
}

constexpr double deg2rad(double deg) {return deg * PI / 180.0;}
constexpr double rad2deg(double rad) {return rad / PI * 180.0;}

/// Wrap an angle to (-PI, PI].  The reference's quirk is kept on purpose:
/// every input congruent to PI (including -PI) maps to +PI, because the
/// shifted remainder is tested with `<= 0` (geometry2d.cpp:5-14).  The EKF
/// calls this 2*cnt+4 times per measurement, and the CUDA kernels use the
/// very same formula (csrc/ekf_math.cuh), so host and device agree.
inline double normalize_angle(double rad)
{
  const double shifted = std::fmod(rad + PI, 2 * PI);
  return shifted <= 0.0 ? shifted + PI : shifted - PI;
}

static_assert(almost_equal(0, 0), "almost_equal failed");
static_assert(almost_equal(-1e-13, 0.0), "almost_equal tolerance failed");
static_assert(almost_equal(deg2rad(rad2deg(2.1)), 2.1), "deg/rad round trip failed");

struct Point2D
{
  double x = 0.0;
  double y = 0.0;
};

struct Vector2D
{
  double x = 0.0;
  double y = 0.0;

  Vector2D & operator+=(const Vector2D & rhs) {x += rhs.x; y += rhs.y; return *this;}
  Vector2D & operator-=(const Vector2D & rhs) {x -= rhs.x; y -= rhs.y; return *this;}
  Vector2D & operator*=(const double & scalar) {x *= scalar; y *= scalar; return *this;}
};

inline Vector2D operator-(const Point2D & head, const Point2D & tail)
{
  return Vector2D{head.x - tail.x, head.y - tail.y};
}
inline Point2D operator+(const Point2D & tail, const Vector2D & disp)
{
  return Point2D{tail.x + disp.x, tail.y + disp.y};
}
inline Vector2D operator+(Vector2D lhs, const Vector2D & rhs) {return lhs += rhs;}
inline Vector2D operator-(Vector2D lhs, const Vector2D & rhs) {return lhs -= rhs;}
inline Vector2D operator*(Vector2D lhs, const double & scalar) {return lhs *= scalar;}
inline Vector2D operator*(const double & scalar, Vector2D rhs) {return rhs *= scalar;}

inline double dot(const Vector2D & lhs, const Vector2D & rhs) {return lhs.x * rhs.x + lhs.y * rhs.y;}
inline double magnitude(Vector2D vec) {return std::sqrt(std::pow(vec.x, 2) + std::pow(vec.y, 2));}
inline Vector2D normalize(const Vector2D & vec)
{
  const double len = magnitude(vec);
  return Vector2D{vec.x / len, vec.y / len};
}
inline double angle(const Vector2D & vec1, const Vector2D & vec2)
{
  return std::acos(dot(vec1, vec2) / magnitude(vec1) / magnitude(vec2));
}
/// after - before with wrap-around at +-PI (reference: geometry2d.cpp:119-132)
inline double angle_diff(const double before, const double after)
{
  if (after < -PI / 2.0 && before > PI / 2.0) {return 2.0 * PI - before + after;}
  if (before < -PI / 2.0 && after > PI / 2.0) {return -2.0 * PI - before + after;}
  return after - before;
}
inline double distance(const Point2D p1, const Point2D p2)
{
  return std::sqrt(std::pow(p1.x - p2.x, 2) + std::pow(p1.y - p2.y, 2));
}

// Stream I/O in the reference's "[x y]" format (geometry2d.cpp:16-50).
inline std::ostream & operator<<(std::ostream & os, const Point2D & p)
{
  return os << "[" << p.x << " " << p.y << "]";
}
inline std::ostream & operator<<(std::ostream & os, const Vector2D & v)
{
  return os << "[" << v.x << " " << v.y << "]";
}
namespace detail
{
inline void skip_open_bracket_read2(std::istream & is, double & a, double & b)
{
  if (is.peek() == '[') {is.get();}
  is >> a >> b;
  while (is.good() && is.get() != '\n') {}
}
}  // namespace detail
inline std::istream & operator>>(std::istream & is, Point2D & p)
{
  detail::skip_open_bracket_read2(is, p.x, p.y);
  return is;
}
inline std::istream & operator>>(std::istream & is, Vector2D & v)
{
  detail::skip_open_bracket_read2(is, v.x, v.y);
  return is;
}
}  // namespace turtlelib

#endif
