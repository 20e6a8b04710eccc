// extern "C" view of the repo's OWN API-compatible turtlelib headers, with the same entry
// points as oracle/ref_shim.cpp exposes for the reference's object code, so that
// tests/test_host_types.py can compare the two bit for bit.
#include "turtlelib/diff_drive.hpp"
extern "C" {
double ref_normalize_angle(double rad) {return turtlelib::normalize_angle(rad);}
void ref_compose(const double * a, const double * b, double * out) {
  const auto t = turtlelib::Transform2D{{a[0], a[1]}, a[2]} * turtlelib::Transform2D{{b[0], b[1]}, b[2]};
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}
void ref_inverse(const double * a, double * out) {
  const auto t = turtlelib::Transform2D{{a[0], a[1]}, a[2]}.inv();
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}
void ref_apply_point(const double * a, const double * p, double * out) {
  const auto q = turtlelib::Transform2D{{a[0], a[1]}, a[2]}(turtlelib::Point2D{p[0], p[1]});
  out[0] = q.x; out[1] = q.y;
}
void ref_integrate_twist(const double * tw, double * out) {
  const auto t = turtlelib::integrate_twist(turtlelib::Twist2D{tw[0], tw[1], tw[2]});
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}
void ref_fkin_sequence(double track, double radius, int count, const double * wheels, double * out) {
  turtlelib::DiffDrive dd{track, radius};
  for (int i = 0; i < count; ++i) {
    const auto t = dd.FKin(wheels[2 * i], wheels[2 * i + 1]);
    out[3 * i] = t.translation().x; out[3 * i + 1] = t.translation().y; out[3 * i + 2] = t.rotation();
  }
}
}
