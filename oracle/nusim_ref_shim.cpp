// TEST INFRASTRUCTURE ONLY.  extern "C" driver around the REFERENCE's own simulator source.
//
// oracle/Makefile compiles this file into oracle/_ref/libnusim_ref.so.  It #includes
// /root/reference/nusim/src/nusim.cpp UNMODIFIED, from where it lies, against the no-op ROS
// headers under oracle/stubs/ (private -> public and main renamed for that one inclusion),
// so tests can run NuSim::sensor_timer_callback (nusim.cpp:310-349: fake landmark sensor),
// NuSim::publish_lidar_data (:559-710: 360-beam ray cast), NuSim::timer_callback (:211-289:
// simulation step with slip and collisions) and NuSim::wheel_cmd_cb (:353-370) as the
// reference wrote them, and read what they publish.  tests/test_nusim.py pins
// oracle/nusim_oracle.py (and through it the device generators) against this library.
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <random>
#include <string>

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
#define main nusim_ref_main
#include "nusim/src/nusim.cpp"
#undef main
#undef private

namespace
{
void set_d(const char * n, double v) {auto & p = oracle_stub::params()[n]; p.d = v; p.set = true;}
void set_v(const char * n, const double * v, int k)
{
  auto & p = oracle_stub::params()[n]; p.v.assign(v, v + k); p.set = true;
}
}  // namespace

extern "C" {
struct NusimRef {std::shared_ptr<NuSim> node;};

// obstacles as (x, y, r) triples.  sensor_std is what the reference calls
// basic_sensor_variance and passes as the STDDEV of its normal distribution (nusim.cpp:202).
NusimRef * nusim_ref_create(
  int n_obst, const double * obst_xyr, double arena_x, double arena_y, double max_range,
  double sensor_std, double input_noise, double slip_fraction, double collision_radius,
  double rate, unsigned seed)
{
  oracle_stub::params().clear();
  std::vector<double> ox(n_obst), oy(n_obst), orad(n_obst);
  for (int i = 0; i < n_obst; ++i) {ox[i] = obst_xyr[3 * i]; oy[i] = obst_xyr[3 * i + 1]; orad[i] = obst_xyr[3 * i + 2];}
  set_v("obstacles/x", ox.data(), n_obst);
  set_v("obstacles/y", oy.data(), n_obst);
  set_v("obstacles/r", orad.data(), n_obst);
  set_d("arena_x_length", arena_x);
  set_d("arena_y_length", arena_y);
  set_d("max_range", max_range);
  set_d("basic_sensor_variance", sensor_std);
  set_d("input_noise", input_noise);
  set_d("slip_fraction", slip_fraction);
  set_d("collision_radius", collision_radius);
  set_d("rate", rate);
  // nuturtle_description/config/diff_params.yaml:3-9
  set_d("wheel_radius", 0.033);
  set_d("track_width", 0.16);
  set_d("motor_cmd_per_rad_sec", 1.0 / 0.024);
  set_d("encoder_ticks_per_rad", 651.898646904);
  auto * h = new NusimRef;
  h->node = std::make_shared<NuSim>();
  // the constructor seeds from std::random_device (nusim.cpp:198); tests need a repeatable stream
  h->node->gen = std::mt19937{seed};
  return h;
}
void nusim_ref_destroy(NusimRef * h) {delete h;}

void nusim_ref_reseed(NusimRef * h, unsigned seed)
{
  auto & n = *h->node;
  n.gen = std::mt19937{seed};
  n.sensor_normal_dist.reset(); n.velocity_normal_dist.reset(); n.slip_uniform_dist.reset();
}

// the `count` values sensor_normal_dist would produce from a fresh mt19937{seed}
void nusim_ref_normal_draws(unsigned seed, double stddev, int count, double * out)
{
  std::mt19937 g{seed};
  std::normal_distribution<> d{0.0, stddev};
  for (int i = 0; i < count; ++i) {out[i] = d(g);}
}

// places the robot: what ddrive / turtle_transform hold after a simulation step
void nusim_ref_set_pose(NusimRef * h, double x, double y, double theta)
{
  h->node->ddrive.setConfig({{x, y}, theta});
  h->node->turtle_transform = turtlelib::Transform2D{{x, y}, theta};
}
void nusim_ref_get_pose(NusimRef * h, double * out)
{
  const auto & t = h->node->turtle_transform;
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
}

// NuSim::sensor_timer_callback (nusim.cpp:310-349): lidar first, then the fake sensor.
// ranges[360] float; xy[2*n_obst]; ids / action [n_obst]
int nusim_ref_sensor_tick(NusimRef * h, float * ranges, double * xy, int * ids, int * action)
{
  auto & n = *h->node;
  oracle_stub::now_ns += 200000000LL;
  n.sensor_timer_callback();
  const auto & scan = n.laser_publisher_->last;
  for (size_t i = 0; i < scan.ranges.size() && i < 360; ++i) {ranges[i] = scan.ranges[i];}
  const auto & arr = n.sensed_obstacles_publisher_->last.markers;
  for (size_t i = 0; i < arr.size(); ++i) {
    xy[2 * i] = arr[i].pose.position.x; xy[2 * i + 1] = arr[i].pose.position.y;
    ids[i] = arr[i].id; action[i] = arr[i].action;
  }
  return static_cast<int>(arr.size());
}

// NuSim::publish_lidar_data alone (nusim.cpp:559-710)
void nusim_ref_lidar(NusimRef * h, float * ranges)
{
  auto & n = *h->node;
  n.publish_lidar_data();
  const auto & scan = n.laser_publisher_->last;
  for (size_t i = 0; i < scan.ranges.size() && i < 360; ++i) {ranges[i] = scan.ranges[i];}
}

// the four wall markers publish_arena_walls left (nusim.cpp:516-555): (px, py, scale_x, qz, qw)
void nusim_ref_walls(NusimRef * h, double * out)
{
  const auto & w = h->node->walls_arr.markers;
  for (size_t i = 0; i < w.size() && i < 4; ++i) {
    out[5 * i] = w[i].pose.position.x; out[5 * i + 1] = w[i].pose.position.y; out[5 * i + 2] = w[i].scale.x;
    out[5 * i + 3] = w[i].pose.orientation.z; out[5 * i + 4] = w[i].pose.orientation.w;
  }
}

// NuSim::wheel_cmd_cb (nusim.cpp:353-370) then `steps` x NuSim::timer_callback (:211-289);
// out: pose (x, y, theta), wheel positions (l, r), encoder ticks (l, r)
void nusim_ref_drive(NusimRef * h, int left_cmd, int right_cmd, int steps, double * out)
{
  auto & n = *h->node;
  nuturtlebot_msgs::msg::WheelCommands c;
  c.left_velocity = left_cmd; c.right_velocity = right_cmd;
  n.wheel_cmd_cb(c);
  for (int i = 0; i < steps; ++i) {
    oracle_stub::now_ns += 5000000LL;
    n.timer_callback();
  }
  const auto & t = n.turtle_transform;
  out[0] = t.translation().x; out[1] = t.translation().y; out[2] = t.rotation();
  out[3] = n.wheel_pos_l; out[4] = n.wheel_pos_r;
  out[5] = n.sensor_publisher_->last.left_encoder; out[6] = n.sensor_publisher_->last.right_encoder;
}
}
