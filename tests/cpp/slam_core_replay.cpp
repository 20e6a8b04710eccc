// Replays a trace through nuslam::SlamCore (host/nuslam/slam_core.hpp) - the C++ host path over the C ABI.
// Input (stdin, text): known(0/1) N n_scans, then per scan: left right m, then m lines "id x y action".
// Output (stdout): per scan "theta x y count assoc...", finally "MAPODOM x y theta".
#include <cstdio>
#include <iostream>
#include <vector>

#include "nuslam/slam_core.hpp"

int main()
{
  int known = 0, N = 0, n_scans = 0;
  if (!(std::cin >> known >> N >> n_scans)) {return 2;}
  try {
    nuslam::SlamCore core{0.160, 0.033, N};
    for (int s = 0; s < n_scans; ++s) {
      double left = 0.0, right = 0.0;
      int m = 0;
      std::cin >> left >> right >> m;
      std::vector<nuslam::Sighting> mk(static_cast<size_t>(m));
      for (auto & k : mk) {std::cin >> k.id >> k.x >> k.y >> k.action;}
      core.joint_state(left, right);
      std::vector<int32_t> assoc;
      if (known) {core.fake_sensor(mk);} else {assoc = core.sensor(mk);}
      const auto & fc = core.filter_configuration();
      std::printf("%.17g %.17g %.17g %u", fc.rotation(), fc.translation().x, fc.translation().y,
        core.counter_obstacles());
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
