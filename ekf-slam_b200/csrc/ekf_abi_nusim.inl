// Part of ekf_abi.cu: nusim's sensor models as device-side generators (ekf_nusim.cuh).

struct ekf_nusim {
  int device = 0;
  int n_obs = 0;
  NusimWorld world{};
  double *obstacles_dev = nullptr;   // [n_obs][3] = (x, y, r)
  // staging of the host-buffer calls, grown on demand
  int cap_robots = 0;
  double *poses_dev = nullptr, *noise_dev = nullptr, *xy_dev = nullptr;
  int *ids_dev = nullptr;
  float *ranges_dev = nullptr;
  cudaStream_t stream = nullptr;
  int64_t launches = 0;
};

namespace {

// the four wall lines the ray cast intersects: nusim.cpp:516-555 (markers) and :649-653 (two points)
void nusim_wall_lines(double arena_x, double arena_y, NusimWorld *w) {
  const double t = 0.1;                                       // wall_thickness = marker.scale.x
  const double pos[4][2] = {{0.0, -(arena_y + t) / 2.0}, {0.0, (arena_y + t) / 2.0},
                            {(arena_x + t) / 2.0, 0.0}, {-(arena_x + t) / 2.0, 0.0}};
  const double yaw[4] = {3.0 / 4.0 * 2.0 * 3.1416, 1.57, 0.0, -3.14159};
  for (int i = 0; i < 4; ++i) {
    double qw = std::cos(yaw[i] * 0.5);                       // tf2 setRPY(0, 0, yaw).w
    if (qw > 1.0) qw = 1.0;
    if (qw < -1.0) qw = -1.0;
    const double ang = 2.0 * std::acos(qw);                   // tf2 getAngle(): the yaw's sign is lost
    T2 T = t2_mul(T2{pos[i][0], pos[i][1], 0.0}, T2{0.0, 0.0, ang});
    T = t2_mul(T, T2{-t, 0.0, 0.0});
    const T2 Tb = t2_mul(T, T2{0.0, 0.1, 0.0});
    w->wall[i][0] = T.x; w->wall[i][1] = T.y; w->wall[i][2] = Tb.x; w->wall[i][3] = Tb.y;
  }
}

int nusim_reserve(ekf_nusim *h, int n_robots) {
  if (n_robots <= h->cap_robots) return EKF_OK;
  cudaFree(h->poses_dev); cudaFree(h->noise_dev); cudaFree(h->xy_dev); cudaFree(h->ids_dev); cudaFree(h->ranges_dev);
  h->poses_dev = h->noise_dev = h->xy_dev = nullptr; h->ids_dev = nullptr; h->ranges_dev = nullptr;
  h->cap_robots = 0;
  const size_t per_noise = (size_t)(2 * h->n_obs > kLidarBeams ? 2 * h->n_obs : kLidarBeams);
  bool ok = cudaMalloc(&h->poses_dev, (size_t)n_robots * 3 * sizeof(double)) == cudaSuccess &&
            cudaMalloc(&h->noise_dev, (size_t)n_robots * per_noise * sizeof(double)) == cudaSuccess &&
            cudaMalloc(&h->xy_dev, (size_t)n_robots * h->n_obs * 2 * sizeof(double)) == cudaSuccess &&
            cudaMalloc(&h->ids_dev, (size_t)n_robots * h->n_obs * sizeof(int)) == cudaSuccess &&
            cudaMalloc(&h->ranges_dev, (size_t)n_robots * kLidarBeams * sizeof(float)) == cudaSuccess;
  if (!ok) return EKF_ERR_NOMEM;
  h->cap_robots = n_robots;
  return EKF_OK;
}

}  // namespace

extern "C" {

int ekf_nusim_destroy(ekf_nusim *h) {
  if (!h) return EKF_ERR_INVALID;
  {
    DeviceGuard guard(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->obstacles_dev); cudaFree(h->poses_dev); cudaFree(h->noise_dev); cudaFree(h->xy_dev);
    cudaFree(h->ids_dev); cudaFree(h->ranges_dev);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
  return EKF_OK;
}

int ekf_nusim_create(int32_t n_obstacles, const double *obstacles_xyr, double arena_x_length, double arena_y_length,
                     int32_t device, ekf_nusim **out) {
  if (!out || n_obstacles < 1 || n_obstacles > kNusimMaxObstacles || !obstacles_xyr) return EKF_ERR_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return EKF_ERR_NO_DEVICE;
  int dev = device;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return EKF_ERR_NO_DEVICE;
  if (dev >= ndev) return EKF_ERR_INVALID;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major < 10) return EKF_ERR_NO_DEVICE;
  ekf_nusim *h = new (std::nothrow) ekf_nusim();
  if (!h) return EKF_ERR_NOMEM;
  h->device = dev; h->n_obs = n_obstacles;
  h->world.n_obstacles = n_obstacles;
  nusim_wall_lines(arena_x_length, arena_y_length, &h->world);
  DeviceGuard guard(dev);
  bool ok = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaMalloc(&h->obstacles_dev, (size_t)n_obstacles * 3 * sizeof(double)) == cudaSuccess &&
            cudaMemcpy(h->obstacles_dev, obstacles_xyr, (size_t)n_obstacles * 3 * sizeof(double),
                       cudaMemcpyHostToDevice) == cudaSuccess;
  if (!ok) { ekf_nusim_destroy(h); return EKF_ERR_CUDA; }
  *out = h;
  return EKF_OK;
}

int ekf_nusim_wall_lines(ekf_nusim *h, double *out16) {
  if (!h || !out16) return EKF_ERR_INVALID;
  for (int i = 0; i < 4; ++i)
    for (int q = 0; q < 4; ++q) out16[4 * i + q] = h->world.wall[i][q];
  return EKF_OK;
}

int ekf_nusim_fake_sensor_dev(ekf_nusim *h, int32_t n_robots, const double *poses_dev, double max_range,
                              double noise_std, uint64_t seed, const double *noise_dev, double *out_xy_dev,
                              int32_t *out_ids_dev, void *cuda_stream) {
  if (!h || n_robots < 1 || !poses_dev || !out_xy_dev || !out_ids_dev) return EKF_ERR_INVALID;
  DeviceGuard guard(h->device);
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
  const int64_t total = (int64_t)n_robots * h->n_obs;
  nusim_fake_sensor_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
      poses_dev, n_robots, h->obstacles_dev, h->n_obs, max_range, noise_std, seed, noise_dev, out_xy_dev, out_ids_dev);
  h->launches++;
  return cudaGetLastError() == cudaSuccess ? EKF_OK : EKF_ERR_CUDA;
}

int ekf_nusim_lidar_dev(ekf_nusim *h, int32_t n_robots, const double *poses_dev, double max_range, double noise_std,
                        uint64_t seed, const double *noise_dev, float *out_ranges_dev, void *cuda_stream) {
  if (!h || n_robots < 1 || !poses_dev || !out_ranges_dev) return EKF_ERR_INVALID;
  DeviceGuard guard(h->device);
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
  const int64_t total = (int64_t)n_robots * kLidarBeams;
  nusim_lidar_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(poses_dev, n_robots, h->obstacles_dev, h->world,
                                                                      max_range, noise_std, seed, noise_dev,
                                                                      out_ranges_dev);
  h->launches++;
  return cudaGetLastError() == cudaSuccess ? EKF_OK : EKF_ERR_CUDA;
}

int ekf_nusim_fake_sensor(ekf_nusim *h, int32_t n_robots, const double *poses, double max_range, double noise_std,
                          uint64_t seed, const double *noise, double *out_xy, int32_t *out_ids) {
  if (!h || n_robots < 1 || !poses || !out_xy || !out_ids) return EKF_ERR_INVALID;
  DeviceGuard guard(h->device);
  int rc = nusim_reserve(h, n_robots);
  if (rc != EKF_OK) return rc;
  const size_t ne = (size_t)n_robots * h->n_obs;
  bool ok = cudaMemcpyAsync(h->poses_dev, poses, (size_t)n_robots * 3 * sizeof(double), cudaMemcpyHostToDevice,
                            h->stream) == cudaSuccess;
  if (ok && noise)
    ok = cudaMemcpyAsync(h->noise_dev, noise, ne * 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream) == cudaSuccess;
  if (!ok) return EKF_ERR_CUDA;
  rc = ekf_nusim_fake_sensor_dev(h, n_robots, h->poses_dev, max_range, noise_std, seed, noise ? h->noise_dev : nullptr,
                                 h->xy_dev, h->ids_dev, nullptr);
  if (rc != EKF_OK) return rc;
  ok = cudaMemcpyAsync(out_xy, h->xy_dev, ne * 2 * sizeof(double), cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
       cudaMemcpyAsync(out_ids, h->ids_dev, ne * sizeof(int), cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
       cudaStreamSynchronize(h->stream) == cudaSuccess;
  return ok ? EKF_OK : EKF_ERR_CUDA;
}

int ekf_nusim_lidar(ekf_nusim *h, int32_t n_robots, const double *poses, double max_range, double noise_std,
                    uint64_t seed, const double *noise, float *out_ranges) {
  if (!h || n_robots < 1 || !poses || !out_ranges) return EKF_ERR_INVALID;
  DeviceGuard guard(h->device);
  int rc = nusim_reserve(h, n_robots);
  if (rc != EKF_OK) return rc;
  const size_t ne = (size_t)n_robots * kLidarBeams;
  bool ok = cudaMemcpyAsync(h->poses_dev, poses, (size_t)n_robots * 3 * sizeof(double), cudaMemcpyHostToDevice,
                            h->stream) == cudaSuccess;
  if (ok && noise)
    ok = cudaMemcpyAsync(h->noise_dev, noise, ne * sizeof(double), cudaMemcpyHostToDevice, h->stream) == cudaSuccess;
  if (!ok) return EKF_ERR_CUDA;
  rc = ekf_nusim_lidar_dev(h, n_robots, h->poses_dev, max_range, noise_std, seed, noise ? h->noise_dev : nullptr,
                           h->ranges_dev, nullptr);
  if (rc != EKF_OK) return rc;
  ok = cudaMemcpyAsync(out_ranges, h->ranges_dev, ne * sizeof(float), cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
       cudaStreamSynchronize(h->stream) == cudaSuccess;
  return ok ? EKF_OK : EKF_ERR_CUDA;
}

}  // extern "C"
