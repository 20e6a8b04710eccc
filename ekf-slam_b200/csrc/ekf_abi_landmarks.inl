// Part of ekf_abi.cu: landmark extraction (ekf_landmarks.cuh), the producer of the EKF's sightings.

struct ekf_landmarks {
  int device = 0;
  int max_scans = 0, n_ranges = 0, max_out = 0;
  float *ranges_dev = nullptr;
  double2 *beam_dev = nullptr;   // (cos, sin) of the beam angles of the current call
  double *out_dev = nullptr;
  int *count_dev = nullptr, *status_dev = nullptr;
  cudaStream_t stream = nullptr;
  int64_t launches = 0;
};

extern "C" {

int ekf_landmarks_destroy(ekf_landmarks *h) {
  if (!h) return EKF_ERR_INVALID;
  {
    DeviceGuard guard(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->ranges_dev); cudaFree(h->beam_dev); cudaFree(h->out_dev); cudaFree(h->count_dev); cudaFree(h->status_dev);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
  return EKF_OK;
}

int ekf_landmarks_create(int32_t max_scans, int32_t n_ranges, int32_t max_out, int32_t device, ekf_landmarks **out) {
  if (!out || max_scans < 1 || n_ranges < 2 || n_ranges > 4096 || max_out < 1) return EKF_ERR_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return EKF_ERR_NO_DEVICE;
  int dev = device;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return EKF_ERR_NO_DEVICE;
  if (dev >= ndev) return EKF_ERR_INVALID;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major < 10) return EKF_ERR_NO_DEVICE;
  ekf_landmarks *h = new (std::nothrow) ekf_landmarks();
  if (!h) return EKF_ERR_NOMEM;
  h->device = dev; h->max_scans = max_scans; h->n_ranges = n_ranges; h->max_out = max_out;
  DeviceGuard guard(dev);
  bool ok = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaMalloc(&h->ranges_dev, (size_t)max_scans * n_ranges * sizeof(float)) == cudaSuccess &&
            cudaMalloc(&h->out_dev, (size_t)max_scans * max_out * 3 * sizeof(double)) == cudaSuccess &&
            cudaMalloc(&h->count_dev, (size_t)max_scans * sizeof(int)) == cudaSuccess &&
            cudaMalloc(&h->status_dev, (size_t)max_scans * sizeof(int)) == cudaSuccess &&
            cudaMalloc(&h->beam_dev, (size_t)n_ranges * sizeof(double2)) == cudaSuccess &&
            cudaFuncSetAttribute(ekf_landmarks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)lm_smem_bytes(4096)) == cudaSuccess &&
            cudaFuncSetAttribute(ekf_landmarks_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 cudaSharedmemCarveoutMaxShared) == cudaSuccess;
  if (!ok) { ekf_landmarks_destroy(h); return EKF_ERR_CUDA; }
  *out = h;
  return EKF_OK;
}

int ekf_landmarks_detect_dev(ekf_landmarks *h, int32_t n_scans, const float *ranges_dev, float angle_min,
                             float angle_increment, double cluster_threshold, double *out_xyr_dev,
                             int32_t *out_count_dev, int32_t *out_status_dev, void *cuda_stream) {
  if (!h || n_scans < 1 || !ranges_dev || !out_xyr_dev || !out_count_dev || !out_status_dev) return EKF_ERR_INVALID;
  DeviceGuard guard(h->device);
  LmParams p{};
  p.ranges = ranges_dev; p.beam = h->beam_dev; p.n_scans = n_scans; p.n_ranges = h->n_ranges; p.max_out = h->max_out;
  p.thr = cluster_threshold;
  p.out = out_xyr_dev; p.count = out_count_dev; p.status = out_status_dev;
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
  lm_beam_table_kernel<<<(h->n_ranges + 127) / 128, 128, 0, st>>>(h->beam_dev, h->n_ranges, angle_min, angle_increment);
  ekf_landmarks_kernel<<<n_scans, kLmThreads, lm_smem_bytes(h->n_ranges), st>>>(p);
  h->launches++;
  h->launches++;
  return cudaGetLastError() == cudaSuccess ? EKF_OK : EKF_ERR_CUDA;
}

int ekf_landmarks_detect(ekf_landmarks *h, int32_t n_scans, const float *ranges, float angle_min,
                         float angle_increment, double cluster_threshold, double *out_xyr, int32_t *out_count,
                         int32_t *out_status) {
  if (!h || n_scans < 1 || n_scans > h->max_scans || !ranges || !out_xyr || !out_count) return EKF_ERR_INVALID;
  DeviceGuard guard(h->device);
  const size_t nb = (size_t)n_scans * h->n_ranges * sizeof(float);
  if (cudaMemcpyAsync(h->ranges_dev, ranges, nb, cudaMemcpyHostToDevice, h->stream) != cudaSuccess) return EKF_ERR_CUDA;
  if (cudaMemsetAsync(h->out_dev, 0, (size_t)n_scans * h->max_out * 3 * sizeof(double), h->stream) != cudaSuccess)
    return EKF_ERR_CUDA;
  int rc = ekf_landmarks_detect_dev(h, n_scans, h->ranges_dev, angle_min, angle_increment, cluster_threshold,
                                    h->out_dev, h->count_dev, h->status_dev, nullptr);
  if (rc != EKF_OK) return rc;
  bool ok = cudaMemcpyAsync(out_xyr, h->out_dev, (size_t)n_scans * h->max_out * 3 * sizeof(double),
                            cudaMemcpyDeviceToHost, h->stream) == cudaSuccess &&
            cudaMemcpyAsync(out_count, h->count_dev, (size_t)n_scans * sizeof(int), cudaMemcpyDeviceToHost,
                            h->stream) == cudaSuccess;
  if (ok && out_status)
    ok = cudaMemcpyAsync(out_status, h->status_dev, (size_t)n_scans * sizeof(int), cudaMemcpyDeviceToHost,
                         h->stream) == cudaSuccess;
  ok = ok && cudaStreamSynchronize(h->stream) == cudaSuccess;
  return ok ? EKF_OK : EKF_ERR_CUDA;
}

}  // extern "C"
