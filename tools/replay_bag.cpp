// replay_bag: drives nuslam::SlamCore (host/nuslam/slam_core.hpp) from a rosbag2 MCAP recording, the offline
// equivalent of `ros2 bag play` into the reference's slam node:
//   /blue/joint_states          -> jointStateCallback (slam.cpp:599-634)  -> core.joint_state(position[0], position[1])
//   /green/detected_obstacles   -> sensor_cb          (slam.cpp:318-530)  -> core.sensor(markers)
// (--known: the detections topic is a fake-sensor recording -> core.fake_sensor, slam.cpp:180-316.)
// Build: g++ -std=c++17 -O2 -Iinclude -Iekf-slam_b200/host tools/replay_bag.cpp -Lekf-slam_b200 -lekf_slam_b200
//        -Wl,-rpath,$PWD/ekf-slam_b200 -ldl -o replay_bag
// Usage: replay_bag FILE.mcap [--known] [--landmarks N] [--joint-topic T] [--detections-topic T]
// Output: one line per detection array "log_time theta x y count assoc...", finally "MAPODOM x y theta".
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "nuslam/slam_core.hpp"
#include "rosbag/mcap_reader.hpp"

int main(int argc, char ** argv)
{
  if (argc < 2) {
    std::fprintf(stderr, "usage: %s FILE.mcap [--known] [--landmarks N] [--joint-topic T] [--detections-topic T]\n", argv[0]);
    return 2;
  }
  bool known = false;
  int landmarks = 50;                                    // slam.cpp:665
  std::string joint = "/blue/joint_states", det = "/green/detected_obstacles";
  for (int i = 2; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--known")) {known = true;} else if (!std::strcmp(argv[i], "--landmarks") && i + 1 < argc) {
      landmarks = std::atoi(argv[++i]);
    } else if (!std::strcmp(argv[i], "--joint-topic") && i + 1 < argc) {joint = argv[++i];} else if (
      !std::strcmp(argv[i], "--detections-topic") && i + 1 < argc)
    {
      det = argv[++i];
    } else {std::fprintf(stderr, "unknown option %s\n", argv[i]); return 2;}
  }
  try {
    const rosbag::McapReader bag(argv[1], joint, det);
    std::fprintf(stderr, "%s: %zu joint states, %zu detection arrays, %llu messages in %llu chunks\n", argv[1],
      bag.joint_states.size(), bag.detections.size(), static_cast<unsigned long long>(bag.messages_total),
      static_cast<unsigned long long>(bag.chunks));
    nuslam::SlamCore core{0.160, 0.033, landmarks};      // diff_params.yaml:3-4
    for (const auto & ev : bag.events) {
      if (ev.kind == 0) {
        const auto & js = bag.joint_states[static_cast<size_t>(ev.index)];
        if (js.position.size() < 2) {continue;}
        core.joint_state(js.position[0], js.position[1]);
        continue;
      }
      const auto & arr = bag.detections[static_cast<size_t>(ev.index)];
      if (arr.markers.empty()) {continue;}               // the reference throws at msg.markers.at(0): slam.cpp:498
      std::vector<nuslam::Sighting> mk;
      for (const auto & m : arr.markers) {mk.push_back({m.id, m.x, m.y, m.action});}
      std::vector<int32_t> assoc;
      if (known) {core.fake_sensor(mk);} else {assoc = core.sensor(mk);}
      const auto & fc = core.filter_configuration();
      std::printf("%llu %.17g %.17g %.17g %u", static_cast<unsigned long long>(ev.log_time), fc.rotation(),
        fc.translation().x, fc.translation().y, core.counter_obstacles());
      for (auto a : assoc) {std::printf(" %d", a);}
      std::printf("\n");
    }
    const auto & t = core.t_map_odom();
    std::printf("MAPODOM %.17g %.17g %.17g\n", t.translation().x, t.translation().y, t.rotation());
  } catch (const std::exception & e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
  return 0;
}
