// Part of ekf_abi.cu: landmark extraction (ekf_landmarks.cuh), the producer of the EKF's sightings.

struct ekf_landmarks {
  int device = 0;
  int max_scans = 0, n_ranges = 0, max_out = 0;
  float *ranges_dev = nullptr;
  double2 *beam_dev = nullptr;   // (cos, sin) of the beam angles of the current call
  // between the kernels of a call (ekf_landmarks.cuh): the batch's candidate clusters and their fits
  LmTask *tasks_dev = nullptr;
  double4 *res_dev = nullptr;
  int *n_tasks_dev = nullptr, *first_task_dev = nullptr, *n_task_of_dev = nullptr;
  int task_cap = 0, fit_blocks = 0;
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
    cudaFree(h->ranges_dev); cudaFree(h->beam_dev); cudaFree(h->out_dev);
    cudaFree(h->tasks_dev); cudaFree(h->res_dev); cudaFree(h->n_tasks_dev); cudaFree(h->first_task_dev); cudaFree(h->n_task_of_dev); cudaFree(h->count_dev); cudaFree(h->status_dev);
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
            cudaMalloc(&h->beam_dev, (size_t)n_ranges * sizeof(double2)) == cudaSuccess;
  const long long cap = (long long)max_scans * lm_max_tasks_per_scan(n_ranges);
  if (cap > 0x7fffffffLL) { ekf_landmarks_destroy(h); return EKF_ERR_INVALID; }
  h->task_cap = (int)cap;
  h->fit_blocks = prop.multiProcessorCount * 4;
  ok = ok && cudaMalloc(&h->tasks_dev, (size_t)cap * sizeof(LmTask)) == cudaSuccess &&
       cudaMalloc(&h->res_dev, (size_t)cap * sizeof(double4)) == cudaSuccess &&
       cudaMalloc(&h->n_tasks_dev, sizeof(int)) == cudaSuccess &&
       cudaMalloc(&h->first_task_dev, (size_t)max_scans * sizeof(int)) == cudaSuccess &&
       cudaMalloc(&h->n_task_of_dev, (size_t)max_scans * sizeof(int)) == cudaSuccess &&
       cudaFuncSetAttribute(ekf_landmarks_segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (int)lm_smem_bytes(4096)) == cudaSuccess &&
       cudaFuncSetAttribute(ekf_landmarks_segment_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
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
  if (n_scans > h->max_scans) return EKF_ERR_INVALID;            // the task arrays are sized for max_scans
  p.tasks = h->tasks_dev; p.res = h->res_dev; p.n_tasks = h->n_tasks_dev; p.first_task = h->first_task_dev;
  p.n_task_of = h->n_task_of_dev; p.task_cap = h->task_cap;
  // table of beam directions -> segmentation (a warp per scan) -> circle test + fit (a thread per candidate cluster of
  // the whole batch) -> ordered detections (a thread per scan)
  lm_beam_table_kernel<<<(h->n_ranges + 127) / 128, 128, 0, st>>>(h->beam_dev, h->n_ranges, angle_min, angle_increment);
  if (cudaMemsetAsync(h->n_tasks_dev, 0, sizeof(int), st) != cudaSuccess) return EKF_ERR_CUDA;
  ekf_landmarks_segment_kernel<<<n_scans, kLmThreads, lm_smem_bytes(h->n_ranges), st>>>(p);
  ekf_landmarks_fit_kernel<<<h->fit_blocks, kLmFitThreads, 0, st>>>(p);
  ekf_landmarks_collect_kernel<<<(n_scans + 127) / 128, 128, 0, st>>>(p);
  h->launches += 4;
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
