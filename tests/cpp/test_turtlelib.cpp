// Known-answer tests for the API-compatible host types, using the reference's own vectors:
// turtlelib/tests/test_geometry2d.cpp:8-17,140-198, test_se2d.cpp:30-251, test_diff_drive.cpp:7-150.
// Plain asserts (Catch2 is not in this image).  Exit code 0 = all passed.
#include <cmath>
#include <cstdio>
#include <sstream>
#include <stdexcept>

#include "turtlelib/diff_drive.hpp"
#include "turtlelib/geometry2d.hpp"
#include "turtlelib/se2d.hpp"

using namespace turtlelib;
static int failures = 0;
#define NEAR(a, b, tol)                                                                     \
  do {                                                                                      \
    if (!(std::fabs((a) - (b)) <= (tol))) {                                                 \
      std::printf("FAIL %s:%d  %s = %.17g, expected %.17g\n", __FILE__, __LINE__, #a, (double)(a), (double)(b)); \
      ++failures;                                                                           \
    }                                                                                       \
  } while (0)
#define TRUE(c) do { if (!(c)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); ++failures; } } while (0)

int main() {
  // normalize_angle: test_geometry2d.cpp:8-17
  NEAR(normalize_angle(0), 0, 1e-8);
  NEAR(normalize_angle(PI), PI, 1e-8);
  NEAR(normalize_angle(-PI), PI, 1e-8);
  NEAR(normalize_angle(-PI / 4), -PI / 4, 1e-8);
  NEAR(normalize_angle(3 * PI / 2), -PI / 2, 1e-8);
  NEAR(normalize_angle(-3 * PI / 2), PI / 2, 1e-8);
  NEAR(normalize_angle(-5 * PI / 2), -PI / 2, 1e-8);
  NEAR(normalize_angle(5 * PI / 2), PI / 2, 1e-8);
  // vectors: test_geometry2d.cpp:140-198
  NEAR(dot(Vector2D{1, 0}, Vector2D{-1, -1}), -1.0, 1e-3);
  NEAR(magnitude(Vector2D{-1, -1}), 1.414, 1e-3);
  NEAR(angle(Vector2D{1, 0}, Vector2D{-1, -1}), 3 * PI / 4, 1e-3);
  NEAR(angle_diff(3 * PI / 4, -3 * PI / 4), PI / 2, 1e-3);
  NEAR(angle_diff(-3 * PI / 4, 3 * PI / 4), -PI / 2, 1e-3);
  NEAR(angle_diff(PI / 4, PI / 2), PI / 4, 1e-3);
  {
    std::stringstream ss; ss << Point2D{0, -203}; TRUE(ss.str() == "[0 -203]");
    std::stringstream in("[2.1 -3]\n"); Vector2D v; in >> v; NEAR(v.x, 2.1, 1e-12); NEAR(v.y, -3, 1e-12);
  }
  // Transform2D: test_se2d.cpp:30-194
  {
    Transform2D id; NEAR(id.rotation(), 0, 1e-8); NEAR(id.translation().x, 0, 1e-8);
    Transform2D t{Vector2D{0.2, -0.2}, 0.1};
    NEAR(t.rotation(), 0.1, 1e-8); NEAR(t.translation().x, 0.2, 1e-8); NEAR(t.translation().y, -0.2, 1e-8);
    Transform2D r{Vector2D{0.1, -0.1}, 3.8};
    Point2D p = r(Point2D{0.1, -0.1}); NEAR(p.x, -0.0403, 1e-3); NEAR(p.y, -0.0821, 1e-4);
    Vector2D v = r(Vector2D{0.1, -0.1}); NEAR(v.x, -0.1403, 1e-3); NEAR(v.y, 0.0179, 1e-4);
    Transform2D a{Vector2D{0.2, 1.1}, 0.34};
    Twist2D tw = a(Twist2D{1.0, 0.707, 0.707}); NEAR(tw.omega, 1.0, 1e-3); NEAR(tw.x, 1.531, 1e-3); NEAR(tw.y, 0.702, 1e-3);
    Transform2D b{Vector2D{1.2, -2.2}, 0.6};
    Transform2D bi = b.inv(); NEAR(bi.rotation(), -0.6, 1e-3); NEAR(bi.translation().x, 0.252, 1e-3); NEAR(bi.translation().y, 2.493, 1e-3);
    Transform2D c = b; c *= bi; NEAR(c.rotation(), 0, 1e-3); NEAR(c.translation().x, 0, 1e-3); NEAR(c.translation().y, 0, 1e-3);
    Transform2D d = b * Transform2D{Vector2D{0.3, 4.1}, -0.1};
    NEAR(d.rotation(), 0.5, 1e-3); NEAR(d.translation().x, -0.867, 1e-3); NEAR(d.translation().y, 1.353, 1e-3);
    std::stringstream ss; ss << Transform2D{Vector2D{1.2, -2.2}, deg2rad(90)}; TRUE(ss.str() == "deg: 90 x: 1.2 y: -2.2");
    for (const char *txt : {"deg: 320 x: -0.2 y: -4.1\n", "320 -0.2 -4.1\n", "320\n-0.2\n-4.1\n"}) {
      std::stringstream in(txt); Transform2D q; in >> q;
      NEAR(q.translation().x, -0.2, 1e-8); NEAR(q.translation().y, -4.1, 1e-8); NEAR(q.rotation(), deg2rad(320), 1e-8);
    }
  }
  // integrate_twist: test_se2d.cpp:226-251
  {
    auto t = integrate_twist(Twist2D{0.0, 1.0, -1.1}); NEAR(t.rotation(), 0, 1e-3); NEAR(t.translation().x, 1.0, 1e-3); NEAR(t.translation().y, -1.1, 1e-3);
    t = integrate_twist(Twist2D{1.0, 0, 0}); NEAR(t.rotation(), 1.0, 1e-3); NEAR(t.translation().x, 0, 1e-3);
    t = integrate_twist(Twist2D{1.0, 1.0, -1.0});
    NEAR(t.translation().y, -std::cos(1.0) - std::sin(1.0) + 1, 1e-3); NEAR(t.translation().x, -std::cos(1.0) + std::sin(1.0) + 1, 1e-3);
  }
  // DiffDrive: test_diff_drive.cpp:7-150
  {
    DiffDrive dd(0.1, 0.05); auto t = dd.FKin(PI, PI); NEAR(t.translation().x, PI * 0.05, 1e-8); NEAR(t.translation().y, 0, 1e-8); NEAR(t.rotation(), 0, 1e-8);
    DiffDrive rot(0.1, 0.2); const double a = (0.1 * PI / 4.0) / (2.0 * PI * 0.2) * 2.0 * PI;
    t = rot.FKin(a, -a); NEAR(t.rotation(), -PI / 2, 1e-8); NEAR(t.translation().x, 0, 1e-8);
    DiffDrive arc(0.1, 0.2);
    t = arc.FKin((0.4 * PI / 4.0) / (2.0 * PI * 0.2) * 2.0 * PI, (0.2 * PI / 4.0) / (2.0 * PI * 0.2) * 2.0 * PI);
    NEAR(t.translation().x, 0.15, 1e-8); NEAR(t.translation().y, -0.15, 1e-8); NEAR(t.rotation(), -PI / 2, 1e-8);
    t = arc.FKin(0.0, 0.0); NEAR(t.translation().x, 0, 1e-8); NEAR(t.translation().y, 0, 1e-8); NEAR(t.rotation(), 0, 1e-8);
    bool raised = false;
    try { arc.IKin(Twist2D{0.1, 1.0, 1.0}); } catch (std::logic_error const &) { raised = true; }
    TRUE(raised);
    auto sp = arc.IKin(Twist2D{0.0, 0.2 * 2.0 * PI, 0.0}); NEAR(sp.left, 2 * PI, 1e-8); NEAR(sp.right, 2 * PI, 1e-8);
    sp = arc.IKin(Twist2D{-PI / 2.0, 0, 0}); NEAR(sp.left, 0.1 * PI / 4.0 / (0.2 * PI) * PI, 1e-8); NEAR(sp.right, -0.1 * PI / 4.0 / (0.2 * PI) * PI, 1e-8);
  }
  std::printf(failures ? "%d FAILED\n" : "all host-type tests passed\n", failures);
  return failures ? 1 : 0;
}
