// TEST INFRASTRUCTURE ONLY (oracle/stubs): no-op stand-ins for the ROS 2 headers the
// reference's node sources include (rclcpp, tf2_ros, the message and service structs they
// touch).  ROS 2 is absent from this image; with these headers nuslam/src/slam.cpp and
// nusim/src/nusim.cpp compile UNMODIFIED from /root/reference (oracle/Makefile), so the
// oracle can be pinned to the reference's own object code.  Nothing here does arithmetic:
// publishers keep the last message, parameters come from a table the test shim fills,
// subscriptions / services / timers drop their callbacks (the shim calls them directly).
#ifndef ORACLE_STUB_ROS_CORE_HPP
#define ORACLE_STUB_ROS_CORE_HPP
#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

// ---------------------------------------------------------------- messages
namespace builtin_interfaces::msg
{
struct Time {int32_t sec{0}; uint32_t nanosec{0};};
}
namespace std_msgs::msg
{
struct Header {builtin_interfaces::msg::Time stamp; std::string frame_id;};
struct UInt64 {uint64_t data{0};};
struct ColorRGBA {float r{0}, g{0}, b{0}, a{0};};
}
namespace std_srvs::srv
{
struct Empty {struct Request {}; struct Response {};};
}
namespace geometry_msgs::msg
{
struct Vector3 {double x{0}, y{0}, z{0};};
struct Point {double x{0}, y{0}, z{0};};
struct Quaternion {double x{0}, y{0}, z{0}, w{1};};
struct Pose {Point position; Quaternion orientation;};
struct PoseStamped {std_msgs::msg::Header header; Pose pose;};
struct PoseWithCovariance {Pose pose; std::array<double, 36> covariance{};};
struct Twist {Vector3 linear; Vector3 angular;};
struct TwistWithCovariance {Twist twist; std::array<double, 36> covariance{};};
struct Transform {Vector3 translation; Quaternion rotation;};
struct TransformStamped {std_msgs::msg::Header header; std::string child_frame_id; Transform transform;};
}
namespace nav_msgs::msg
{
struct Odometry
{
  std_msgs::msg::Header header; std::string child_frame_id;
  geometry_msgs::msg::PoseWithCovariance pose; geometry_msgs::msg::TwistWithCovariance twist;
};
struct Path {std_msgs::msg::Header header; std::vector<geometry_msgs::msg::PoseStamped> poses;};
}
namespace sensor_msgs::msg
{
struct JointState
{
  std_msgs::msg::Header header; std::vector<std::string> name;
  std::vector<double> position, velocity, effort;
};
struct LaserScan
{
  std_msgs::msg::Header header;
  float angle_min{0}, angle_max{0}, angle_increment{0}, time_increment{0}, scan_time{0}, range_min{0}, range_max{0};
  std::vector<float> ranges, intensities;
};
}
namespace nuturtlebot_msgs::msg
{
struct WheelCommands {int32_t left_velocity{0}, right_velocity{0};};
struct SensorData {builtin_interfaces::msg::Time stamp; int32_t left_encoder{0}, right_encoder{0};};
}
namespace visualization_msgs::msg
{
struct Marker
{
  static constexpr int32_t ARROW = 0, CUBE = 1, SPHERE = 2, CYLINDER = 3;
  static constexpr int32_t ADD = 0, MODIFY = 0, DELETE = 2, DELETEALL = 3;
  std_msgs::msg::Header header; std::string ns; int32_t id{0}; int32_t type{0}; int32_t action{0};
  geometry_msgs::msg::Pose pose; geometry_msgs::msg::Vector3 scale; std_msgs::msg::ColorRGBA color;
};
struct MarkerArray {std::vector<Marker> markers;};
}
namespace nuturtle_control_interfaces::srv
{
struct InitialPose {struct Request {double x{0}, y{0}, theta{0};}; struct Response {};};
struct Control {struct Request {double velocity{0}, radius{0};}; struct Response {};};
}
namespace nusim_interfaces::srv
{
struct Teleport {struct Request {double x{0}, y{0}, theta{0};}; struct Response {};};
}

// ---------------------------------------------------------------- tf2 (quaternion only)
namespace tf2
{
class Quaternion
{
public:
  Quaternion() = default;
  Quaternion(double x, double y, double z, double w)
  : x_(x), y_(y), z_(z), w_(w) {}
  // tf2/LinearMath/Quaternion.h (third-party, absent here): the published formulas
  void setRPY(double roll, double pitch, double yaw)
  {
    const double hy = yaw * 0.5, hp = pitch * 0.5, hr = roll * 0.5;
    const double cy = std::cos(hy), sy = std::sin(hy), cp = std::cos(hp), sp = std::sin(hp);
    const double cr = std::cos(hr), sr = std::sin(hr);
    x_ = sr * cp * cy - cr * sp * sy;
    y_ = cr * sp * cy + sr * cp * sy;
    z_ = cr * cp * sy - sr * sp * cy;
    w_ = cr * cp * cy + sr * sp * sy;
  }
  double getAngle() const
  {
    double w = w_;
    if (w < -1.0) {w = -1.0;}
    if (w > 1.0) {w = 1.0;}
    return 2.0 * std::acos(w);
  }
  double x() const {return x_;}
  double y() const {return y_;}
  double z() const {return z_;}
  double w() const {return w_;}

private:
  double x_{0}, y_{0}, z_{0}, w_{1};
};
inline geometry_msgs::msg::Quaternion toMsg(const Quaternion & q)
{
  geometry_msgs::msg::Quaternion m; m.x = q.x(); m.y = q.y(); m.z = q.z(); m.w = q.w(); return m;
}
}  // namespace tf2

// ---------------------------------------------------------------- rclcpp
namespace oracle_stub
{
// parameter table and the construction hook, both owned by the test shim
struct ParamValue {double d{0}; std::string s; std::vector<double> v; bool b{false}; int64_t i{0}; bool set{false};};
std::map<std::string, ParamValue> & params();
// called from Node::get_parameter so a shim can adjust freshly default-initialised members
// (e.g. the hard-coded map capacity, slam.cpp:665) BEFORE the constructor body sizes its
// matrices from them; a test facility, not a change to the reference source
extern void (* get_parameter_hook)(void * node, const std::string & name);
// deterministic clock: nanoseconds, advanced by the shim
extern int64_t now_ns;
}  // namespace oracle_stub

namespace rclcpp
{
enum ParameterType {PARAMETER_NOT_SET, PARAMETER_BOOL, PARAMETER_INTEGER, PARAMETER_DOUBLE, PARAMETER_STRING,
  PARAMETER_DOUBLE_ARRAY};

class Duration
{
public:
  explicit Duration(int64_t ns = 0)
  : ns_(ns) {}
  double seconds() const {return static_cast<double>(ns_) * 1e-9;}
  int64_t nanoseconds() const {return ns_;}

private:
  int64_t ns_;
};
class Time
{
public:
  Time() = default;
  explicit Time(int64_t ns)
  : ns_(ns) {}
  Time(const builtin_interfaces::msg::Time & t)
  : ns_(static_cast<int64_t>(t.sec) * 1000000000LL + t.nanosec) {}
  double seconds() const {return static_cast<double>(ns_) * 1e-9;}
  int64_t nanoseconds() const {return ns_;}
  Duration operator-(const Time & o) const {return Duration(ns_ - o.ns_);}
  operator builtin_interfaces::msg::Time() const
  {
    builtin_interfaces::msg::Time t;
    t.sec = static_cast<int32_t>(ns_ / 1000000000LL);
    t.nanosec = static_cast<uint32_t>(ns_ % 1000000000LL);
    return t;
  }

private:
  int64_t ns_{0};
};
class Clock
{
public:
  Time now() const {return Time(oracle_stub::now_ns);}
};

class QoS
{
public:
  QoS(int depth = 10) {(void)depth;}
  QoS & transient_local() {return *this;}
  QoS & reliable() {return *this;}
  QoS & best_effort() {return *this;}
  QoS & durability_volatile() {return *this;}
  QoS & keep_last(int) {return *this;}
};
class SystemDefaultsQoS : public QoS {};
class SensorDataQoS : public QoS {};

class Parameter
{
public:
  Parameter(std::string name, const oracle_stub::ParamValue * v)
  : name_(std::move(name)), v_(v) {}
  double as_double() const {return need().d;}
  std::string as_string() const {return need().s;}
  std::vector<double> as_double_array() const {return need().v;}
  bool as_bool() const {return need().b;}
  int64_t as_int() const {return need().i;}

private:
  const oracle_stub::ParamValue & need() const
  {
    if (!v_ || !v_->set) {throw std::runtime_error("parameter '" + name_ + "' is not set");}
    return *v_;
  }
  std::string name_;
  const oracle_stub::ParamValue * v_;
};

template<class T>
class Publisher
{
public:
  using SharedPtr = std::shared_ptr<Publisher<T>>;
  void publish(const T & m) {last = m; ++count;}
  T last{};
  std::size_t count{0};
};
template<class T>
class Subscription
{
public:
  using SharedPtr = std::shared_ptr<Subscription<T>>;
};
template<class T>
class Service
{
public:
  using SharedPtr = std::shared_ptr<Service<T>>;
};
class TimerBase
{
public:
  using SharedPtr = std::shared_ptr<TimerBase>;
};
class Logger {};

class Node
{
public:
  explicit Node(const std::string & name)
  : name_(name) {}
  virtual ~Node() = default;

  // declare with a default: the table wins if the shim set the name, else the default is stored
  void declare_parameter(const std::string & n, ParameterType) {(void)n;}
  void declare_parameter(const std::string & n, const char * dflt) {dflt_s(n, dflt);}
  void declare_parameter(const std::string & n, const std::string & dflt) {dflt_s(n, dflt);}
  void declare_parameter(const std::string & n, double dflt)
  {
    auto & p = oracle_stub::params()[n];
    if (!p.set) {p.d = dflt; p.set = true;}
  }
  void declare_parameter(const std::string & n, int dflt)
  {
    auto & p = oracle_stub::params()[n];
    if (!p.set) {p.i = dflt; p.d = dflt; p.set = true;}
  }
  void declare_parameter(const std::string & n, bool dflt)
  {
    auto & p = oracle_stub::params()[n];
    if (!p.set) {p.b = dflt; p.set = true;}
  }
  void declare_parameter(const std::string & n, const std::vector<double> & dflt)
  {
    auto & p = oracle_stub::params()[n];
    if (!p.set) {p.v = dflt; p.set = true;}
  }
  Parameter get_parameter(const std::string & n)
  {
    if (oracle_stub::get_parameter_hook) {oracle_stub::get_parameter_hook(this, n);}
    auto it = oracle_stub::params().find(n);
    return Parameter(n, it == oracle_stub::params().end() ? nullptr : &it->second);
  }

  template<class T, class Q, class Cb>
  typename Subscription<T>::SharedPtr create_subscription(const std::string &, const Q &, Cb &&)
  {
    return std::make_shared<Subscription<T>>();
  }
  template<class T, class Q>
  typename Publisher<T>::SharedPtr create_publisher(const std::string &, const Q &)
  {
    return std::make_shared<Publisher<T>>();
  }
  template<class T, class Cb>
  typename Service<T>::SharedPtr create_service(const std::string &, Cb &&)
  {
    return std::make_shared<Service<T>>();
  }
  template<class D, class Cb>
  TimerBase::SharedPtr create_wall_timer(D, Cb &&) {return std::make_shared<TimerBase>();}
  std::shared_ptr<Clock> get_clock() {return clock_;}
  Logger get_logger() const {return Logger{};}
  const std::string & get_name() const {return name_;}

private:
  static void dflt_s(const std::string & n, const std::string & dflt)
  {
    auto & p = oracle_stub::params()[n];
    if (!p.set) {p.s = dflt; p.set = true;}
  }
  std::string name_;
  std::shared_ptr<Clock> clock_{std::make_shared<Clock>()};
};

inline void init(int, char **) {}
inline void shutdown() {}
template<class N>
inline void spin(std::shared_ptr<N>) {}
}  // namespace rclcpp

// logging macros swallow their stream expression
#define RCLCPP_INFO_STREAM(logger, expr) do {(void)(logger);} while (0)
#define RCLCPP_ERROR_STREAM(logger, expr) do {(void)(logger);} while (0)
#define RCLCPP_WARN_STREAM(logger, expr) do {(void)(logger);} while (0)
#define RCLCPP_DEBUG_STREAM(logger, expr) do {(void)(logger);} while (0)
#define RCLCPP_INFO(logger, ...) do {(void)(logger);} while (0)
#define RCLCPP_ERROR(logger, ...) do {(void)(logger);} while (0)

namespace tf2_ros
{
class TransformBroadcaster
{
public:
  template<class N>
  explicit TransformBroadcaster(N &) {}
  void sendTransform(const geometry_msgs::msg::TransformStamped & t) {last = t; ++count;}
  geometry_msgs::msg::TransformStamped last{};
  std::size_t count{0};
};
}  // namespace tf2_ros
#endif
