// Part of ekf_abi.cu: configuration, lifetime, stream and status entry points.

extern "C" {

void ekf_default_config(ekf_config *cfg) {
  if (!cfg) return;
  memset(cfg, 0, sizeof(*cfg));
  cfg->n_landmarks = 50;       // slam.cpp:665
  cfg->q_noise = 1.0e-2;       // slam.cpp:666
  cfg->r_noise = 1.0e-2;       // slam.cpp:667
  cfg->mah_threshold = 2.0;    // slam.cpp:671
  cfg->prior = 10e6;           // slam.cpp:130
  cfg->device = -1;
  cfg->max_scan = 0;
  cfg->shard_rank = 0;
  cfg->shard_count = 1;
}

const char *ekf_status_string(int status) {
  switch (status) {
    case EKF_OK: return "ok";
    case EKF_ERR_INVALID: return "invalid argument";
    case EKF_ERR_CAPACITY: return "landmark capacity exceeded";
    case EKF_ERR_EMPTY_SCAN: return "empty scan";
    case EKF_ERR_BAD_ID: return "landmark id out of range";
    case EKF_ERR_CUDA: return "CUDA error";
    case EKF_ERR_NO_DEVICE: return "no usable CUDA device";
    case EKF_ERR_NOMEM: return "out of memory";
    case EKF_ERR_IO: return "bag file unreadable or unsupported";
    default: return "unknown status";
  }
}

const char *ekf_version(void) { return "ekf_slam_b200 0.1 (sm_100a)"; }

namespace {
int selftest_run(int which, int32_t device, uint64_t seed, int32_t count, uint64_t *mismatches, uint64_t *outside_guard) {
  if (!mismatches || !outside_guard || count < 1) return EKF_ERR_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return EKF_ERR_NO_DEVICE;
  int dev = device;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return EKF_ERR_NO_DEVICE;
  if (dev >= ndev) return EKF_ERR_INVALID;
  DeviceGuard guard(dev);
  unsigned long long *d = nullptr, h[2] = {0, 0};
  if (cudaMalloc(&d, sizeof(h)) != cudaSuccess) return EKF_ERR_NOMEM;
  cudaError_t e = cudaMemset(d, 0, sizeof(h));
  if (e == cudaSuccess) {
    if (which == 0) ekf::selftest_division_kernel<<<(count + 255) / 256, 256>>>(seed, count, d, d + 1);
    else ekf::selftest_atan2_kernel<<<(count + 255) / 256, 256>>>(seed, count, d, d + 1);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) { fprintf(stderr, "ekf_selftest: %s\n", cudaGetErrorString(e)); return EKF_ERR_CUDA; }
  *mismatches = h[0]; *outside_guard = h[1];
  return EKF_OK;
}
}  // namespace

int ekf_selftest_division(int32_t device, uint64_t seed, int32_t count, uint64_t *mismatches, uint64_t *outside_guard) {
  return selftest_run(0, device, seed, count, mismatches, outside_guard);
}
int ekf_selftest_atan2(int32_t device, uint64_t seed, int32_t count, uint64_t *mismatches, uint64_t *outside_guard) {
  return selftest_run(1, device, seed, count, mismatches, outside_guard);
}

int ekf_destroy(ekf_filter *f) {
  if (!f) return EKF_ERR_INVALID;
  {
    DeviceGuard guard(f->device);
    if (f->stream) cudaStreamSynchronize(f->stream);
    if (f->shard_stream) {
      cudaStreamSynchronize(f->shard_stream);
      cudaStreamDestroy(f->shard_stream);
      cudaEventDestroy(f->shard_ev_seq); cudaEventDestroy(f->shard_ev_pass); cudaEventDestroy(f->shard_ev_push);
    }
    pipe_release(f);
    cudaFree(f->sigma); cudaFree(f->mu); cudaFree(f->W); cudaFree(f->V);
    cudaFree(f->Pp); cudaFree(f->cnt); cudaFree(f->status);   // D lives in the Pp allocation
    cudaFree(f->cols); cudaFree(f->guess); cudaFree(f->done_dev);
    cudaFree(f->guess_part_d); cudaFree(f->guess_part_k); cudaFree(f->guess_ticket);
    for (int q = 0; q < 16; ++q)
      if (f->peer_ipc[q]) cudaIpcCloseMemHandle(f->peer_ipc[q]);
    cudaFree(f->xchg); cudaFree(f->push_done);
    cudaFree(f->prev_xy); cudaFree(f->bar); cudaFree(f->slot_d); cudaFree(f->slot_k); cudaFree(f->slot_rec);
    cudaFree(f->hot_hdr); cudaFree(f->hot_rec); cudaFree(f->hot_w);
    cudaFree(f->io_dev); cudaFree(f->res_dev); cudaFree(f->scratch_dev);
    if (f->io_host) cudaFreeHost(f->io_host);
    if (f->res_host) cudaFreeHost(f->res_host);
    if (f->res_event) cudaEventDestroy(f->res_event);
    for (auto &e : f->ev) if (e) cudaEventDestroy(e);
    f->ev.clear();
    for (auto &e : f->ev_x) if (e) cudaEventDestroy(e);
    f->ev_x.clear();
    if (f->own_stream) cudaStreamDestroy(f->own_stream);
  }
  delete f;
  return EKF_OK;
}

int ekf_create(const ekf_config *cfg, ekf_filter **out) {
  if (!cfg || !out) return EKF_ERR_INVALID;
  *out = nullptr;
  if (cfg->n_landmarks < 1 || cfg->n_landmarks > 500000000) return EKF_ERR_INVALID;
  if (cfg->shard_count > 1 && (cfg->shard_rank < 0 || cfg->shard_rank >= cfg->shard_count)) return EKF_ERR_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return EKF_ERR_NO_DEVICE;
  int dev = cfg->device;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return EKF_ERR_NO_DEVICE;
  if (dev >= ndev) return EKF_ERR_INVALID;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return EKF_ERR_NO_DEVICE;
  if (prop.major < 10) return EKF_ERR_NO_DEVICE;  // sm_100a code only

  ekf_filter *f = new (std::nothrow) ekf_filter();
  if (!f) return EKF_ERR_NOMEM;
  f->cfg = *cfg;
  f->device = dev;
  f->num_sms = prop.multiProcessorCount;
  f->N = cfg->n_landmarks;
  f->n = 3 + 2 * f->N;
  f->n_int = ekf_n_int(f->N);
  f->n_pad = ekf_n_pad(f->N);
  f->sharded = cfg->shard_count > 1;
  f->lay = ekf_lay_full(f->n_pad);
  if (f->sharded) {
    if (cfg->shard_count > f->n_pad / EKF_STRIP) return (delete f, EKF_ERR_INVALID);
    f->lay.rb = ekf_shard_bound(f->n_pad, cfg->shard_rank, cfg->shard_count);
    f->lay.re = ekf_shard_bound(f->n_pad, cfg->shard_rank + 1, cfg->shard_count);
    if (f->lay.re <= f->lay.rb) return (delete f, EKF_ERR_INVALID);
  }
  f->sigma_elems = ekf_lay_elems(f->lay);
  DeviceGuard guard(dev);

  auto fail = [&](cudaError_t e, const char *what) {
    fprintf(stderr, "ekf_create: %s: %s\n", what, cudaGetErrorString(e));
    ekf_destroy(f);
    return e == cudaErrorMemoryAllocation ? EKF_ERR_NOMEM : EKF_ERR_CUDA;
  };
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(e_, #call); } while (0)

  CKC(cudaStreamCreateWithFlags(&f->own_stream, cudaStreamNonBlocking));
  CKC(cudaEventCreateWithFlags(&f->res_event, cudaEventDisableTiming));
  f->stream = f->own_stream;
  const size_t panel_bytes = (size_t)f->n_pad * f->kmax * sizeof(double);
  CKC(cudaMalloc(&f->sigma, (size_t)f->sigma_elems * sizeof(double)));
  CKC(cudaMalloc(&f->mu, (size_t)f->n_pad * sizeof(double)));
  CKC(cudaMalloc(&f->W, panel_bytes));
  CKC(cudaMalloc(&f->V, panel_bytes));
  // Pp (n_pad rows) and D (N blocks) share one allocation: a sharded handle all-reduces both at once
  CKC(cudaMalloc(&f->Pp, ((size_t)f->n_pad + f->N) * sizeof(double4)));
  f->D = f->Pp + f->n_pad;
  // the exchange buffers of the shard protocol are allocated on first use (ekf_abi_shard.inl)
  CKC(cudaMalloc(&f->cnt, sizeof(int)));
  CKC(cudaMalloc(&f->status, sizeof(int)));
  CKC(cudaMalloc(&f->prev_xy, 2 * sizeof(double)));
  CKC(cudaMalloc(&f->bar, sizeof(unsigned long long)));

  // grids: sequential kernel = cooperative (shape chosen from N, see ekf_seq.cuh); covariance pass = one CTA per SM
  CKC(cudaFuncSetAttribute(ekf_seq_kernel<kSeqThreadsSmall>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)seq_smem_bytes(kSeqThreadsSmall, true)));
  CKC(cudaFuncSetAttribute(ekf_seq_kernel<kSeqThreadsLarge>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)seq_smem_bytes(kSeqThreadsLarge)));
  // co-resident CTAs the cooperative launch may use (occupancy of the kernel with its dynamic smem), capped by
  // what one warp can scan in the arg-min slot read
  int seq_occ = 1;
  f->seq_hot = env_int("EKF_SEQ_HOT", 1) != 0;
  CKC(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&seq_occ, ekf_seq_kernel<kSeqThreadsSmall>, kSeqThreadsSmall,
                                                    seq_smem_bytes(kSeqThreadsSmall, f->seq_hot)));
  if (seq_occ < 1) seq_occ = 1;
  if (seq_occ > kSeqCtasPerSm) seq_occ = kSeqCtasPerSm;
  int seq_cap = f->num_sms * seq_occ;
  if (seq_cap > 32 * kMaxSlotsPerLane) seq_cap = 32 * kMaxSlotsPerLane;
  f->seq_threads = env_int("EKF_SEQ_THREADS", f->N <= seq_cap * kSeqThreadsSmall ? kSeqThreadsSmall : kSeqThreadsLarge);
  if (f->seq_threads != kSeqThreadsSmall) {
    f->seq_threads = kSeqThreadsLarge;
    if (seq_cap > f->num_sms) seq_cap = f->num_sms;      // one CTA per SM
  }
  int seq_blocks = (f->N + f->seq_threads - 1) / f->seq_threads;
  int max_seq = env_int("EKF_SEQ_BLOCKS", seq_cap);
  if (max_seq < 1) max_seq = 1;
  if (max_seq > seq_cap) max_seq = seq_cap;
  f->seq_grid = seq_blocks < max_seq ? seq_blocks : max_seq;
  f->dd_grid = env_int("EKF_DD_BLOCKS", f->num_sms);
  if (f->dd_grid < 1) f->dd_grid = 1;
  f->dd_prefetch = env_int("EKF_DD_PREFETCH", 0);
  f->dd_variant = env_int("EKF_DD_VARIANT", 2);
  f->dd_ctas = env_int("EKF_DD_CTAS", 2);
  if (f->dd_ctas < 1) f->dd_ctas = 1;
  if (f->dd_prefetch < 0) f->dd_prefetch = 0;
  f->slot_cap = seq_cap;
  CKC(cudaMalloc(&f->slot_d, (size_t)kSlotReps * 2 * seq_cap * sizeof(double)));
  CKC(cudaMalloc(&f->slot_k, (size_t)kSlotReps * 2 * seq_cap * sizeof(int)));
  CKC(cudaMalloc(&f->slot_rec, (size_t)kSlotReps * 2 * seq_cap * kRecStride * sizeof(double)));
  CKC(cudaMalloc(&f->hot_hdr, (size_t)kSlotReps * 2 * sizeof(unsigned long long)));
  CKC(cudaMemset(f->hot_hdr, 0, (size_t)kSlotReps * 2 * sizeof(unsigned long long)));
  CKC(cudaMalloc(&f->hot_rec, (size_t)kSlotReps * 2 * kRecStride * sizeof(double)));
  CKC(cudaMemset(f->hot_rec, 0, (size_t)kSlotReps * 2 * kRecStride * sizeof(double)));
  CKC(cudaMalloc(&f->hot_w, (size_t)kSlotReps * 2 * kMaxK2 * sizeof(double2)));
  CKC(cudaMemset(f->hot_w, 0, (size_t)kSlotReps * 2 * kMaxK2 * sizeof(double2)));
  CKC(cudaFuncSetAttribute(ekf_downdate_kernel<64, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)dd_smem_bytes(64, kMaxK2)));
  CKC(cudaFuncSetAttribute(ekf_downdate_kernel<64, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)dd_smem_bytes(64, kMaxK2)));
  CKC(cudaFuncSetAttribute(ekf_downdate_kernel<32, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)dd_smem_bytes(32, kMaxK2)));
  CKC(cudaFuncSetAttribute(ekf_downdate_kernel<32, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)dd_smem_bytes(32, kMaxK2)));
  CKC(cudaFuncSetAttribute(ekf_downdate_kernel<32, 2>, cudaFuncAttributePreferredSharedMemoryCarveout,
                           cudaSharedmemCarveoutMaxShared));
  // several CTAs of the <32> shape share an SM only if the SM is configured with (nearly) all of its
  // L1 / shared storage as shared memory
  CKC(cudaFuncSetAttribute(ekf_downdate_kernel<32, 0>, cudaFuncAttributePreferredSharedMemoryCarveout,
                           cudaSharedmemCarveoutMaxShared));
  if (env_int("EKF_DEBUG", 0)) {
    int occ = 0;
    for (int k2 = 8; k2 <= kMaxK2; k2 += 8) {
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ekf_downdate_kernel<32, 0>, dd_threads(32),
                                                    dd_smem_bytes(32, k2));
      fprintf(stderr, "ekf: downdate<32> k2=%d smem=%zu B -> %d CTAs / SM\n", k2, dd_smem_bytes(32, k2), occ);
    }
  }

  CKC(cudaMemsetAsync(f->sigma, 0, (size_t)f->sigma_elems * sizeof(double), f->stream));
  CKC(cudaMemsetAsync(f->mu, 0, (size_t)f->n_pad * sizeof(double), f->stream));
  CKC(cudaMemsetAsync(f->W, 0, panel_bytes, f->stream));
  CKC(cudaMemsetAsync(f->V, 0, panel_bytes, f->stream));
  CKC(cudaMemsetAsync(f->Pp, 0, ((size_t)f->n_pad + f->N) * sizeof(double4), f->stream));
  CKC(cudaMemsetAsync(f->cnt, 0, sizeof(int), f->stream));
  CKC(cudaMemsetAsync(f->status, 0, sizeof(int), f->stream));
  CKC(cudaMemsetAsync(f->prev_xy, 0, 2 * sizeof(double), f->stream));
  CKC(cudaMemsetAsync(f->bar, 0, sizeof(unsigned long long), f->stream));
  init_prior_kernel<<<(2 * f->N + 255) / 256, 256, 0, f->stream>>>(f->sigma, f->lay, f->n_int, cfg->prior);
  f->launches++;
  CKC(cudaGetLastError());
  if (ensure_io(f, cfg->max_scan > 0 ? cfg->max_scan : 64) != EKF_OK) { ekf_destroy(f); return EKF_ERR_CUDA; }
  CKC(cudaStreamSynchronize(f->stream));
#undef CKC
  *out = f;
  return EKF_OK;
}

int ekf_set_stream(ekf_filter *f, void *cuda_stream) {
  if (!f) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }
  CK(f, cudaStreamSynchronize(f->stream));
  f->stream = cuda_stream ? (cudaStream_t)cuda_stream : f->own_stream;
  return EKF_OK;
}

int ekf_synchronize(ekf_filter *f) {
  if (!f) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  if (f->back_stream) CK(f, cudaStreamSynchronize(f->back_stream));   // a pass of the overlapped mode (no drain)
  if (f->shard_stream) CK(f, cudaStreamSynchronize(f->shard_stream));  // a pass of the shard protocol
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_last_status(ekf_filter *f) {
  if (!f) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaStreamSynchronize(f->stream));
  int dev_status = 0;
  CK(f, cudaMemcpyAsync(&dev_status, f->status, sizeof(int), cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaMemsetAsync(f->status, 0, sizeof(int), f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  int s = f->latched != EKF_OK ? f->latched : dev_status;
  f->latched = EKF_OK;
  return s;
}

int ekf_read_timing_shard(ekf_filter *f, float *assoc_ms, float *push_ms, int32_t max_scans, int32_t *n_scans) {
  if (!f || max_scans < 0) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaStreamSynchronize(f->stream));
  if (f->shard_stream) CK(f, cudaStreamSynchronize(f->shard_stream));
  int64_t have = f->ev_count < f->ev_cap ? f->ev_count : f->ev_cap;
  if (have > max_scans) have = max_scans;
  for (int64_t i = 0; i < have; ++i) {
    const size_t slot = (size_t)((f->ev_count - have + i) % f->ev_cap) * 4;
    float a = 0.f, b = 0.f;
    if (cudaEventElapsedTime(&a, f->ev_x[slot], f->ev_x[slot + 1]) != cudaSuccess) { cudaGetLastError(); a = 0.f; }
    if (cudaEventElapsedTime(&b, f->ev_x[slot + 2], f->ev_x[slot + 3]) != cudaSuccess) { cudaGetLastError(); b = 0.f; }
    if (assoc_ms) assoc_ms[i] = a;
    if (push_ms) push_ms[i] = b;
  }
  if (n_scans) *n_scans = (int32_t)have;
  return EKF_OK;
}

const char *ekf_last_cuda_error(ekf_filter *f) { return f ? f->cuda_err.c_str() : ""; }

int ekf_enable_timing(ekf_filter *f, int32_t ring_scans) {
  if (!f || ring_scans < 0) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaStreamSynchronize(f->stream));
  if (f->back_stream) CK(f, cudaStreamSynchronize(f->back_stream));
  if (f->shard_stream) CK(f, cudaStreamSynchronize(f->shard_stream));
  for (auto &e : f->ev) if (e) cudaEventDestroy(e);
  f->ev.assign((size_t)ring_scans * 4, nullptr);   // sequential start / end, pass start / end
  for (auto &e : f->ev) CK(f, cudaEventCreate(&e));
  for (auto &e : f->ev_x) if (e) cudaEventDestroy(e);
  f->ev_x.assign((size_t)ring_scans * 4, nullptr);
  for (auto &e : f->ev_x) CK(f, cudaEventCreate(&e));
  f->ev_cap = ring_scans;
  f->ev_count = 0;
  return EKF_OK;
}

int ekf_read_timing(ekf_filter *f, float *seq_ms, float *downdate_ms, int32_t max_scans, int32_t *n_scans) {
  if (!f || max_scans < 0) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaStreamSynchronize(f->stream));
  if (f->back_stream) CK(f, cudaStreamSynchronize(f->back_stream));
  if (f->shard_stream) CK(f, cudaStreamSynchronize(f->shard_stream));
  int64_t have = f->ev_count < f->ev_cap ? f->ev_count : f->ev_cap;
  if (have > max_scans) have = max_scans;
  // the most recent `have` scans, oldest first
  for (int64_t i = 0; i < have; ++i) {
    const int64_t scan = f->ev_count - have + i;
    const size_t slot = (size_t)(scan % f->ev_cap) * 4;
    float a = 0.f, b = 0.f;
    CK(f, cudaEventElapsedTime(&a, f->ev[slot], f->ev[slot + 1]));
    CK(f, cudaEventElapsedTime(&b, f->ev[slot + 2], f->ev[slot + 3]));
    if (seq_ms) seq_ms[i] = a;
    if (downdate_ms) downdate_ms[i] = b;
  }
  if (n_scans) *n_scans = (int32_t)have;
  return EKF_OK;
}

int ekf_set_overlap(ekf_filter *f, int32_t on) {
  if (!f) return EKF_ERR_INVALID;
  if (f->sharded) return EKF_ERR_INVALID;               // a row shard already overlaps across ranks
  DeviceGuard guard(f->device);
  if (!on) {
    const int rcd = pipe_drain(f);
    if (rcd != EKF_OK) return rcd;
    f->overlap = false;
    return EKF_OK;
  }
  const int rc = pipe_resources(f);                     // second covariance buffer: EKF_ERR_NOMEM if it does not fit
  if (rc != EKF_OK) return rc;
  f->overlap = true;
  return EKF_OK;
}

int ekf_launch_count(ekf_filter *f, int64_t *launches) {
  if (!f || !launches) return EKF_ERR_INVALID;
  *launches = f->launches;
  return EKF_OK;
}

int ekf_sigma_storage(ekf_filter *f, int64_t *stored_doubles, int64_t *n_pad) {
  if (!f) return EKF_ERR_INVALID;
  if (stored_doubles) *stored_doubles = f->sigma_elems;
  if (n_pad) *n_pad = f->n_pad;
  return EKF_OK;
}

}  // extern "C"
