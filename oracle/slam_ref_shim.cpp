// TEST INFRASTRUCTURE ONLY.  extern "C" driver around the REFERENCE's own EKF source.
//
// oracle/Makefile compiles this file into oracle/_ref/libslam_ref.so.  It #includes
// /root/reference/nuslam/src/slam.cpp UNMODIFIED, from where it lies (nothing is copied
// into the repo), against the no-op ROS headers and the dense <armadillo> stand-in under
// oracle/stubs/.  `private` is re-defined to `public` and `main` renamed for that one
// inclusion only, so this driver can call Slam::fake_sensor_cb (slam.cpp:180),
// Slam::sensor_cb (:318) and Slam::jointStateCallback (:599) and read the filter members
// (:659-676).  Control flow, indices, angle wraps, the sentinel rule, the dead branch and
// the index_min decision are therefore the reference's object code; only the arithmetic
// inside the matrix operators is the stub's (oracle/stubs/armadillo states its order).
//
// tests/test_oracle_pins.py pins oracle A (oracle/ekf_literal.py) against this library and
// tests/golden/make_golden.py generates the golden fixtures FROM it.
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <string>

#include <armadillo>
#include "rclcpp/rclcpp.hpp"
#include "turtlelib/diff_drive.hpp"
#include "turtlelib/se2d.hpp"

namespace oracle_stub
{
std::map<std::string, ParamValue> & params()
{
  static std::map<std::string, ParamValue> table;
  return table;
}
void (* get_parameter_hook)(void *, const std::string &) = nullptr;
int64_t now_ns = 1000000000LL;
}  // namespace oracle_stub

#define private public
#define main slam_ref_main
#include "nuslam/src/slam.cpp"
#undef main
#undef private

namespace
{
int g_capacity = 50;

// runs inside Slam::Slam() at its first get_parameter (slam.cpp:63), i.e. after the default
// member initialisers (:665-674) and before sigma / q_bar are sized (:128-136)
void capacity_hook(void * node, const std::string & name)
{
  if (name != "wheel_radius") {return;}
  auto * s = static_cast<Slam *>(static_cast<rclcpp::Node *>(node));
  if (g_capacity != s->n_landmarks) {
    s->n_landmarks = g_capacity;
    s->state = arma::vec(2 * g_capacity + 3, arma::fill::zeros);    // what :674 does with the new value
  }
}

void set_param_d(const char * n, double v) {auto & p = oracle_stub::params()[n]; p.d = v; p.set = true;}
void set_param_s(const char * n, const char * v) {auto & p = oracle_stub::params()[n]; p.s = v; p.set = true;}

visualization_msgs::msg::MarkerArray make_markers(int m, const int * ids, const double * xy, const unsigned char * action)
{
  visualization_msgs::msg::MarkerArray arr;
  for (int i = 0; i < m; ++i) {
    visualization_msgs::msg::Marker mk;
    mk.id = ids ? ids[i] : i;
    mk.action = action ? action[i] : visualization_msgs::msg::Marker::ADD;
    mk.pose.position.x = xy[2 * i];
    mk.pose.position.y = xy[2 * i + 1];
    mk.header.stamp = rclcpp::Time(oracle_stub::now_ns);
    arr.markers.push_back(mk);
  }
  return arr;
}
}  // namespace

extern "C" {

// status: 0 ok, 1 the reference threw std::out_of_range (its bounds check / vector::at),
// 2 any other exception
struct SlamRef {std::shared_ptr<Slam> node; std::string error;};

SlamRef * slam_ref_create(int n_landmarks, const char * sensor_source, double wheel_radius, double track_width)
{
  oracle_stub::params().clear();
  set_param_d("wheel_radius", wheel_radius);
  set_param_d("track_width", track_width);
  set_param_s("body_id", "green/base_footprint");
  set_param_s("wheel_left", "wheel_left_joint");
  set_param_s("wheel_right", "wheel_right_joint");
  set_param_s("sensor_source", sensor_source);
  g_capacity = n_landmarks;
  oracle_stub::get_parameter_hook = capacity_hook;
  auto * h = new SlamRef;
  h->node = std::make_shared<Slam>();
  oracle_stub::get_parameter_hook = nullptr;
  return h;
}
void slam_ref_destroy(SlamRef * h) {delete h;}
const char * slam_ref_error(SlamRef * h) {return h->error.c_str();}

int slam_ref_n(SlamRef * h) {return 2 * h->node->n_landmarks + 3;}
int slam_ref_count(SlamRef * h) {return static_cast<int>(h->node->counter_obstacles);}
double slam_ref_threshold(SlamRef * h) {return h->node->mah_threshold;}
void slam_ref_set_noise(SlamRef * h, double q_noise, double r_noise, double mah_threshold)
{
  // re-does slam.cpp:135-139 with other constants (test facility for non-default filters)
  auto & s = *h->node;
  s.q_noise = q_noise; s.r_noise = r_noise; s.mah_threshold = mah_threshold;
  s.q_bar = arma::mat(2 * s.n_landmarks + 3, 2 * s.n_landmarks + 3, arma::fill::zeros);
  s.q_bar.submat(0, 0, 2, 2) = arma::mat(3, 3, arma::fill::eye) * q_noise;
  s.R = arma::mat(2, 2, arma::fill::eye) * r_noise;
}

// what jointStateCallback leaves in t_odom_robot (slam.cpp:633), set directly
void slam_ref_set_odom(SlamRef * h, double x, double y, double theta)
{
  h->node->t_odom_robot = turtlelib::Transform2D{{x, y}, theta};
}
// or through the reference's own wheel odometry: Slam::jointStateCallback (slam.cpp:599-634)
int slam_ref_joint_state(SlamRef * h, double left, double right)
{
  try {
    sensor_msgs::msg::JointState js;
    js.position = {left, right};
    oracle_stub::now_ns += 10000000LL;
    h->node->jointStateCallback(js);
    return 0;
  } catch (const std::exception & e) {h->error = e.what(); return 2;}
}
void slam_ref_get_odom(SlamRef * h, double * out)
{
  const auto & t = h->node->t_odom_robot;
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}
void slam_ref_get_map_odom(SlamRef * h, double * out)
{
  const auto & t = h->node->t_map_odom;
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}
// t_map_odom * t_odom_robot: the pose the C ABI's predict() receives (slam.cpp:184 / :321)
void slam_ref_composed_pose(SlamRef * h, double * out)
{
  const auto t = h->node->t_map_odom * h->node->t_odom_robot;
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}

// Slam::fake_sensor_cb (slam.cpp:180-316), known association
int slam_ref_fake_sensor(SlamRef * h, int m, const int * ids, const double * xy, const unsigned char * action)
{
  try {
    oracle_stub::now_ns += 200000000LL;
    h->node->fake_sensor_cb(make_markers(m, ids, xy, action));
    return 0;
  } catch (const std::out_of_range & e) {h->error = e.what(); return 1;
  } catch (const std::exception & e) {h->error = e.what(); return 2;}
}

// Slam::sensor_cb (slam.cpp:318-530), Mahalanobis association.  The reference does not
// return the association; it is read from the stub's index_min / min tap: per measurement
// the callback calls index_min then min (:418-419) and, in the branch at :423-432 only,
// index_min once more.  out_assoc[i] = the min_index used at :443; out_dist[i] = the value
// `min` (:419); *out_second_branch counts how often the :423 branch ran (it is unreachable).
int slam_ref_sensor_ex(
  SlamRef * h, int m, const double * xy, int * out_assoc, double * out_dist, int * out_second_branch,
  double * out_best, double * out_second);
int slam_ref_sensor(SlamRef * h, int m, const double * xy, int * out_assoc, double * out_dist, int * out_second_branch)
{
  return slam_ref_sensor_ex(h, m, xy, out_assoc, out_dist, out_second_branch, nullptr, nullptr);
}
// also: out_best[i] / out_second[i] = smallest / second-smallest distance of measurement i over the EXISTING
// landmarks (the decision margins: a decision can only flip when out_best is within arithmetic noise of the
// threshold or of out_second)
int slam_ref_sensor_ex(
  SlamRef * h, int m, const double * xy, int * out_assoc, double * out_dist, int * out_second_branch,
  double * out_best, double * out_second)
{
  int rc = 0;
  arma::reduce_log().clear();
  try {
    oracle_stub::now_ns += 200000000LL;
    h->node->sensor_cb(make_markers(m, nullptr, xy, nullptr));
  } catch (const std::out_of_range & e) {h->error = e.what(); rc = 1;
  } catch (const std::exception & e) {h->error = e.what(); rc = 2;}
  const auto & log = arma::reduce_log();
  int meas = 0, second = 0;
  for (size_t p = 0; p + 1 < log.size() && meas < m; ) {
    if (log[p].kind != 'I' || log[p + 1].kind != 'M') {rc = rc ? rc : 2; h->error = "unexpected reduce log"; break;}
    int idx = static_cast<int>(log[p].index);
    const double mn = log[p + 1].value;
    if (out_best) {out_best[meas] = log[p].head_best;}
    if (out_second) {out_second[meas] = log[p].head_second;}
    p += 2;
    const bool again = p < log.size() && log[p].kind == 'I' && (p + 1 >= log.size() || log[p + 1].kind == 'I');
    if (again) {idx = static_cast<int>(log[p].index); ++p; ++second;}
    if (out_assoc) {out_assoc[meas] = idx;}
    if (out_dist) {out_dist[meas] = mn;}
    ++meas;
  }
  for (int i = meas; i < m; ++i) {
    if (out_assoc) {out_assoc[i] = -1;}
  }
  if (out_second_branch) {*out_second_branch = second;}
  return rc;
}

void slam_ref_get_state(SlamRef * h, double * out)
{
  const auto & s = h->node->state;
  for (arma::uword i = 0; i < s.n_elem; ++i) {out[i] = s.at(i);}
}
// row-major n x n copy of the reference's (column-major, not exactly symmetric) sigma
void slam_ref_get_sigma(SlamRef * h, double * out)
{
  const auto & S = h->node->sigma;
  const arma::uword n = S.n_rows;
  for (arma::uword i = 0; i < n; ++i) {
    for (arma::uword j = 0; j < n; ++j) {out[i * n + j] = S.at(i, j);}
  }
}
void slam_ref_set_state(SlamRef * h, const double * in, int count)
{
  auto & s = h->node->state;
  for (arma::uword i = 0; i < s.n_elem; ++i) {s.at(i) = in[i];}
  h->node->counter_obstacles = static_cast<unsigned int>(count);
}
void slam_ref_set_sigma(SlamRef * h, const double * in)
{
  auto & S = h->node->sigma;
  const arma::uword n = S.n_rows;
  for (arma::uword i = 0; i < n; ++i) {
    for (arma::uword j = 0; j < n; ++j) {S.at(i, j) = in[i * n + j];}
  }
}
void slam_ref_set_previous(SlamRef * h, double x, double y, double theta)
{
  h->node->filter_previous_configuration = turtlelib::Transform2D{{x, y}, theta};
}
void slam_ref_set_map_odom(SlamRef * h, double x, double y, double theta)
{
  h->node->t_map_odom = turtlelib::Transform2D{{x, y}, theta};
}

// what the callbacks published: the map->odom tf (slam.cpp:281-289 / :498-506) as
// (x, y, qz, qw) and the number of markers on green/obstacles (:297-314 / :512-528)
void slam_ref_last_tf(SlamRef * h, double * out)
{
  const auto & t = h->node->tf_broadcaster_->last.transform;
  out[0] = t.translation.x; out[1] = t.translation.y; out[2] = t.rotation.z; out[3] = t.rotation.w;
}
int slam_ref_published_markers(SlamRef * h, double * xy, int * ids, int cap)
{
  const auto & arr = h->node->sensed_obstacles_publisher_->last.markers;
  const int k = static_cast<int>(arr.size());
  for (int i = 0; i < k && i < cap; ++i) {
    xy[2 * i] = arr[i].pose.position.x; xy[2 * i + 1] = arr[i].pose.position.y; ids[i] = arr[i].id;
  }
  return k;
}
}
