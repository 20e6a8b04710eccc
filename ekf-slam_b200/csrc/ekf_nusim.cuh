// nusim's two sensor models as device-side generators: inputs for the batched
// Monte-Carlo loop without a host round trip.
//
//   nusim/src/nusim.cpp:310-349   fake landmark sensor: every obstacle in the robot's body frame, plus
//                                 per-axis Gaussian noise, ADD / DELETE by true range (:331-336)
//   nusim/src/nusim.cpp:559-710   360-beam lidar: ray against every cylinder (:590-643, the closest hit
//                                 wins), against the four wall LINES only when no cylinder was hit
//                                 (:648-697, the reference intersects infinite lines), clamp to
//                                 [0.11, 10] (:699-704), plus noise (:706); float ranges as in LaserScan
//   nusim/src/nusim.cpp:516-555   the wall markers (positions, thickness, yaw through tf2 getAngle)
//   turtlelib/src/se2d.cpp:57-74  Transform2D inv / compose, restated below as T2
//
// Many robots at once: one thread per (robot, obstacle) / per (robot, beam).  Outputs use the layouts the
// consumers take: xy[robot][obstacle][2] + ids[robot][obstacle] feed ekf_batch_step_dev (a negative id is
// the DELETE rule of slam.cpp:205), ranges[robot][360] feed ekf_landmarks_detect_dev.
// Noise: the reference draws std::normal_distribution over one mt19937 stream; here either the caller
// supplies the draws (parity tests), or a counter-based generator keyed by (seed, robot, index) makes them.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace ekf {

constexpr int kLidarBeams = 360;
constexpr int kNusimMaxObstacles = 64;

struct T2 { double x, y, rot; };
__host__ __device__ inline T2 t2_inv(const T2 a) {                 // se2d.cpp:57-63
  const double c = cos(a.rot), s = sin(a.rot);
  return T2{-a.x * c - a.y * s, -a.y * c + a.x * s, -a.rot};
}
__host__ __device__ inline T2 t2_mul(const T2 a, const T2 b) {     // se2d.cpp:66-74
  const double c = cos(a.rot), s = sin(a.rot);
  return T2{c * b.x - s * b.y + a.x, s * b.x + c * b.y + a.y, a.rot + b.rot};
}

// the same with cos / sin of a.rot supplied by the caller (the lidar evaluates each angle's pair once)
__host__ __device__ inline T2 t2_inv_cs(const T2 a, double c, double s) {
  return T2{-a.x * c - a.y * s, -a.y * c + a.x * s, -a.rot};
}
__host__ __device__ inline T2 t2_mul_cs(const T2 a, double c, double s, const T2 b) {
  return T2{c * b.x - s * b.y + a.x, s * b.x + c * b.y + a.y, a.rot + b.rot};
}

// splitmix64 -> two uniforms -> Box-Muller: one standard normal per (seed, stream, index)
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ double normal_draw(uint64_t seed, uint64_t stream, uint64_t index) {
  const uint64_t k = mix64(seed ^ mix64(stream * 0xD1342543DE82EF95ull + index));
  const uint64_t a = mix64(k), b = mix64(k ^ 0xA0761D6478BD642Full);
  const double u1 = ((double)(a >> 11) + 1.0) * (1.0 / 9007199254740992.0);   // (0, 1]
  const double u2 = (double)(b >> 11) * (1.0 / 9007199254740992.0);           // [0, 1)
  return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

struct NusimWorld {
  int n_obstacles;
  double wall[4][4];     // (x3, y3, x4, y4) per wall: two points of the line the ray cast intersects
};

// ---- fake sensor: nusim.cpp:317-346 ---------------------------------------------------------------
__global__ void nusim_fake_sensor_kernel(const double *poses, int n_robots, const double *obstacles, int n_obs,
                                         double max_range, double noise_std, uint64_t seed, const double *noise,
                                         double *out_xy, int *out_ids) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)n_robots * n_obs) return;
  const int r = (int)(e / n_obs), j = (int)(e % n_obs);
  const T2 Twr{poses[3 * r + 1], poses[3 * r + 2], poses[3 * r]};
  const T2 rel = t2_mul(t2_inv(Twr), T2{obstacles[3 * j], obstacles[3 * j + 1], 0.0});   // :324-327
  const double dist = sqrt(rel.x * rel.x + rel.y * rel.y);                              // :329-331
  double nx = 0.0, ny = 0.0;
  if (noise != nullptr) { nx = noise[2 * e]; ny = noise[2 * e + 1]; }
  else if (noise_std > 0.0) {
    nx = noise_std * normal_draw(seed, (uint64_t)r, 2 * (uint64_t)j);
    ny = noise_std * normal_draw(seed, (uint64_t)r, 2 * (uint64_t)j + 1);
  }
  out_xy[2 * e] = rel.x + nx;                                                           // :341-342
  out_xy[2 * e + 1] = rel.y + ny;
  out_ids[e] = dist <= max_range ? j : -(j + 1);                                        // :332-336
}

// ---- lidar: nusim.cpp:559-710 ---------------------------------------------------------------------
__global__ void __launch_bounds__(128) nusim_lidar_kernel(const double *poses, int n_robots, const double *obstacles,
                                                          NusimWorld world, double max_range, double noise_std,
                                                          uint64_t seed, const double *noise, float *out_ranges) {
  __shared__ double s_obs[kNusimMaxObstacles * 3];
  for (int i = threadIdx.x; i < world.n_obstacles * 3; i += blockDim.x) s_obs[i] = obstacles[i];
  __syncthreads();
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)n_robots * kLidarBeams) return;
  const int r = (int)(e / kLidarBeams), i = (int)(e % kLidarBeams);
  const T2 Twt{poses[3 * r + 1], poses[3 * r + 2], poses[3 * r]};
  // Four transforms of the reference, two angles: cos / sin of the robot's heading and of the beam's world heading are
  // evaluated once each; the inverse transform's are (c, -s) - cos is even and sin odd bit for bit, so this is what
  // se2d.cpp:57-74 computes with its own calls
  const T2 Ttl = t2_mul_cs(T2{-0.032, 0.0, 0.0}, 1.0, 0.0, T2{0.0, 0.0, (double)i * 2.0 * 3.14159265358979323846 / 360.0});  // :577-580
  double c_r, s_r, c_b, s_b;
  sincos(Twt.rot, &s_r, &c_r);
  const T2 Twl = t2_mul_cs(Twt, c_r, s_r, Ttl);                                         // :581
  sincos(Twl.rot, &s_b, &c_b);
  const T2 p2 = t2_mul_cs(Twl, c_b, s_b, T2{1.0, 0.0, 0.0});                            // :585
  const T2 Tlw = t2_inv_cs(Twl, c_b, s_b);
  const double x1 = Twl.x, y1 = Twl.y, x2 = p2.x, y2 = p2.y;
  // beam direction: normalize(p2 - p1), :604-606
  const double ddx = x2 - x1, ddy = y2 - y1;
  const double mod = sqrt(ddx * ddx + ddy * ddy);
  const double rx = ddx / mod, ry = ddy / mod;
  bool collision = false;
  double closest = max_range;                                                           // :588
  for (int j = 0; j < world.n_obstacles; ++j) {
    const double mx = s_obs[3 * j], my = s_obs[3 * j + 1], rad = s_obs[3 * j + 2];
    const T2 Tlm = t2_mul_cs(Tlw, c_b, -s_b, T2{mx, my, 0.0});                          // :596
    if (Tlm.x >= 0) {                                                                   // :598
      const double ux = mx - x1, uy = my - y1;                                          // :600-601
      const double dt = ux * rx + uy * ry;
      const double u1x = dt * rx, u1y = dt * ry;                                        // :608
      const double u2x = ux - u1x, u2y = uy - u1y;                                      // :610
      const double d = sqrt(u2x * u2x + u2y * u2y);                                     // :612
      const double half = rad * 2.0 / 2.0;                                              // marker.scale.x / 2
      const double m = sqrt(half * half - d * d);                                       // :614, NaN on a miss
      const double q1x = x1 + u1x + m * rx, q1y = y1 + u1y + m * ry;                    // :616-618
      const double q2x = x1 + (-1.0 * u1x) + m * rx, q2y = y1 + (-1.0 * u1y) + m * ry;  // :619-621
      const double d1 = sqrt((x1 - q1x) * (x1 - q1x) + (y1 - q1y) * (y1 - q1y));        // :624-626
      const double d2 = sqrt((x1 - q2x) * (x1 - q2x) + (y1 - q2y) * (y1 - q2y));        // :628-630
      const double dist = d1 < d2 ? d1 : d2;                                            // :632-637
      if (dist < closest) { closest = dist; collision = true; }                         // :639-643 (NaN never hits)
    }
  }
  if (!collision) {                                                                     // :649
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const double x3 = world.wall[w][0], y3 = world.wall[w][1], x4 = world.wall[w][2], y4 = world.wall[w][3];
      const double denom = (x1 - x2) * (y3 - y4) - (y1 - y2) * (x3 - x4);               // :678-680
      const double num_x = (x1 * y2 - y1 * x2) * (x3 - x4) - (x1 - x2) * (x3 * y4 - y3 * x4);
      const double num_y = (x1 * y2 - y1 * x2) * (y3 - y4) - (y1 - y2) * (x3 * y4 - y3 * x4);
      if (denom != 0.0) {                                                               // :682
        const double px = num_x / denom, py = num_y / denom;
        const T2 Tl = t2_mul_cs(Tlw, c_b, -s_b, T2{px, py, 0.0});                       // :685-686
        if (Tl.x > 0) {                                                                 // :687
          const double dist = sqrt((x1 - px) * (x1 - px) + (y1 - py) * (y1 - py));      // :688-690
          if (dist < closest) closest = dist;                                           // :691-693
        }
      }
    }
  }
  if (closest > 10.0) closest = 10.0;                                                   // :699-701 (range_max)
  if (closest < 0.11) closest = 0.11;                                                   // :702-704 (range_min)
  double nz = 0.0;
  if (noise != nullptr) nz = noise[e];
  else if (noise_std > 0.0) nz = noise_std * normal_draw(seed, (uint64_t)r + 0x100000000ull, (uint64_t)i);
  out_ranges[e] = (float)(closest + nz);                                                // :706
}

}  // namespace ekf
