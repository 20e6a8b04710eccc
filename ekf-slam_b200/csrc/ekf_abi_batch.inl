// Part of ekf_abi.cu: batched mode (B independent small filters, ekf_batch.cuh).

struct ekf_batch {
  ekf_config cfg{};
  int device = 0;
  int B = 0, N = 0, n = 0, tri = 0;
  double *sigma = nullptr, *mu = nullptr, *prev = nullptr;
  int *cnt = nullptr, *status = nullptr;
  uchar2 *tri_map = nullptr;     // (row, column) of every packed lower-triangle element
  ushort4 *blk_map = nullptr;    // 2x2 blocks of the stored triangle (BatchParams::blk_map)
  int nblk = 0;
  unsigned short *soff = nullptr;   // shared-memory offset of every packed element (layout of this handle's variant)
  int variant = 0;               // bits 0-1: downdate shape DD (0, 1, 2), bit 2: one-pass association, bit 3: 20 CTAs/SM
  // (7: covariance in registers for the whole scan, ekf_batch_reg.cuh - measured slower, kept selectable)
  // staging for the host-pointer step
  unsigned char *io_dev = nullptr, *io_host = nullptr, *res_dev = nullptr, *res_host = nullptr;
  size_t io_cap = 0, res_cap = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  // host-buffer steps on page-locked caller memory are split into chunks of filters that alternate between two
  // side streams, so that the H2D copy of chunk c+1, the kernel of chunk c and the D2H copy of chunk c-1 overlap
  cudaStream_t side[2] = {nullptr, nullptr};
  cudaEvent_t ev_begin = nullptr, ev_side[2] = {nullptr, nullptr};
  int chunks = 1;
  int64_t launches = 0;
  int latched = EKF_OK;
};

namespace {

constexpr int kBatchVariantReg = 7;        // covariance in registers for the whole scan (ekf_batch_reg.cuh): n = 41 .. 48, m <= 32
constexpr int kBatchRegMinBlocks = 12;     // resident one-warp CTAs per SM (3 warps per scheduler: 168 registers each)
constexpr int BATCH_DEFAULT_VARIANT = 6;   // one-pass association + DMMA 8x8-tile downdate, 16 CTAs/SM (measured best)

inline int batch_dd(const ekf_batch *b) { return b->variant == kBatchVariantReg ? 2 : (b->variant & 3); }

#define CKB(b, call)                                                               \
  do {                                                                             \
    cudaError_t e_ = (call);                                                       \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "ekf_batch: %s: %s\n", #call, cudaGetErrorString(e_));       \
      (b)->latched = EKF_ERR_CUDA;                                                 \
      return EKF_ERR_CUDA;                                                         \
    }                                                                              \
  } while (0)

int batch_launch(ekf_batch *b, const double *poses, int m, const double *xy, const int32_t *ids,
                 const int32_t *m_count, int32_t *assoc, int32_t *count, double *pose_out, int f0 = 0, int nf = -1,
                 cudaStream_t stream = nullptr) {
  if (nf < 0) nf = b->B - f0;
  if (!stream) stream = b->stream;
  BatchParams p{};
  p.B = b->B; p.N = b->N; p.n = b->n; p.tri = b->tri; p.f0 = f0;
  p.q_noise = b->cfg.q_noise; p.r_noise = b->cfg.r_noise; p.thr = b->cfg.mah_threshold;
  p.sigma = b->sigma; p.mu = b->mu; p.prev = b->prev; p.cnt = b->cnt; p.status = b->status;
  p.tri_map = b->tri_map; p.blk_map = b->blk_map; p.nblk = b->nblk; p.soff = b->soff;
  p.pose = poses; p.xy = xy; p.ids = ids; p.m_count = m_count; p.m = m;
  p.assoc_out = assoc; p.cnt_out = count; p.pose_out = pose_out;
  const size_t smem = batch_smem_bytes(b->n, m, batch_dd(b));
  if (smem > 200 * 1024) return EKF_ERR_INVALID;   // m beyond the per-CTA shared memory
  if (b->variant == kBatchVariantReg && m <= 32) {  // covariance in registers (ekf_batch_reg.cuh)
    ekf_batch_reg_kernel<kBatchRegMinBlocks><<<nf, kBatchThreads, batch_reg_smem_bytes(), stream>>>(p);
    b->launches++;
    CKB(b, cudaGetLastError());
    return EKF_OK;
  }
#define EKF_BATCH_CASE(V, MINB, ONE, DD) \
  case V: ekf_batch_kernel<MINB, ONE, DD><<<nf, kBatchThreads, smem, stream>>>(p); break;
  switch (b->variant == kBatchVariantReg ? 6 : b->variant) {   // (longer scans of a register-shape handle: tiles)
    EKF_BATCH_CASE(0, 16, false, 0)  EKF_BATCH_CASE(1, 16, false, 1)  EKF_BATCH_CASE(2, 16, false, 2)
    EKF_BATCH_CASE(4, 16, true, 0)   EKF_BATCH_CASE(5, 16, true, 1)   EKF_BATCH_CASE(6, 16, true, 2)
    EKF_BATCH_CASE(8, 20, false, 0)  EKF_BATCH_CASE(9, 20, false, 1)  EKF_BATCH_CASE(10, 20, false, 2)
    EKF_BATCH_CASE(12, 20, true, 0)  EKF_BATCH_CASE(13, 20, true, 1)  EKF_BATCH_CASE(14, 20, true, 2)
    default: return EKF_ERR_INVALID;
  }
#undef EKF_BATCH_CASE
  b->launches++;
  CKB(b, cudaGetLastError());
  return EKF_OK;
}

struct BatchIo {
  size_t pose, xy, ids, mc, total;
  size_t r_assoc, r_cnt, r_pose, r_status, r_total;
  BatchIo(int B, int m) {
    pose = 0; xy = align_up((size_t)24 * B, 16); ids = xy + (size_t)16 * B * m;
    mc = ids + (size_t)4 * B * m; total = align_up(mc + (size_t)4 * B, 16);
    r_pose = 0; r_assoc = align_up((size_t)24 * B, 16); r_cnt = r_assoc + (size_t)4 * B * m;
    r_status = r_cnt + (size_t)4 * B; r_total = align_up(r_status + 4, 16);
  }
};

int batch_ensure_io(ekf_batch *b, int m) {
  const BatchIo io(b->B, m);
  if (io.total > b->io_cap) {
    if (b->io_dev) cudaFree(b->io_dev);
    if (b->io_host) cudaFreeHost(b->io_host);
    b->io_dev = b->io_host = nullptr; b->io_cap = 0;
    CKB(b, cudaMalloc(&b->io_dev, io.total));
    CKB(b, cudaMallocHost(&b->io_host, io.total));
    b->io_cap = io.total;
  }
  if (io.r_total > b->res_cap) {
    if (b->res_dev) cudaFree(b->res_dev);
    if (b->res_host) cudaFreeHost(b->res_host);
    b->res_dev = b->res_host = nullptr; b->res_cap = 0;
    CKB(b, cudaMalloc(&b->res_dev, io.r_total));
    CKB(b, cudaMallocHost(&b->res_host, io.r_total));
    b->res_cap = io.r_total;
  }
  return EKF_OK;
}

}  // namespace

extern "C" {

int ekf_batch_destroy(ekf_batch *b) {
  if (!b) return EKF_ERR_INVALID;
  {
    DeviceGuard guard(b->device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    cudaFree(b->sigma); cudaFree(b->mu); cudaFree(b->prev); cudaFree(b->cnt); cudaFree(b->status); cudaFree(b->tri_map); cudaFree(b->blk_map); cudaFree(b->soff);
    cudaFree(b->io_dev); cudaFree(b->res_dev);
    if (b->io_host) cudaFreeHost(b->io_host);
    if (b->res_host) cudaFreeHost(b->res_host);
    if (b->own_stream) cudaStreamDestroy(b->own_stream);
    for (int q = 0; q < 2; ++q) {
      if (b->side[q]) cudaStreamDestroy(b->side[q]);
      if (b->ev_side[q]) cudaEventDestroy(b->ev_side[q]);
    }
    if (b->ev_begin) cudaEventDestroy(b->ev_begin);
  }
  delete b;
  return EKF_OK;
}

int ekf_batch_create(const ekf_config *cfg, int32_t n_filters, ekf_batch **out) {
  if (!cfg || !out || n_filters < 1) return EKF_ERR_INVALID;
  *out = nullptr;
  if (cfg->n_landmarks < 1 || 3 + 2 * cfg->n_landmarks > 131) return EKF_ERR_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return EKF_ERR_NO_DEVICE;
  int dev = cfg->device;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return EKF_ERR_NO_DEVICE;
  if (dev >= ndev) return EKF_ERR_INVALID;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major < 10) return EKF_ERR_NO_DEVICE;
  ekf_batch *b = new (std::nothrow) ekf_batch();
  if (!b) return EKF_ERR_NOMEM;
  b->cfg = *cfg; b->device = dev; b->B = n_filters; b->N = cfg->n_landmarks;
  b->n = 3 + 2 * b->N; b->tri = b->n * (b->n + 1) / 2;
  DeviceGuard guard(dev);
  auto fail = [&](cudaError_t e, const char *what) {
    fprintf(stderr, "ekf_batch_create: %s: %s\n", what, cudaGetErrorString(e));
    ekf_batch_destroy(b);
    return e == cudaErrorMemoryAllocation ? EKF_ERR_NOMEM : EKF_ERR_CUDA;
  };
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(e_, #call); } while (0)
  CKC(cudaStreamCreateWithFlags(&b->own_stream, cudaStreamNonBlocking));
  b->stream = b->own_stream;
  for (int q = 0; q < 2; ++q) {
    CKC(cudaStreamCreateWithFlags(&b->side[q], cudaStreamNonBlocking));
    CKC(cudaEventCreateWithFlags(&b->ev_side[q], cudaEventDisableTiming));
  }
  CKC(cudaEventCreateWithFlags(&b->ev_begin, cudaEventDisableTiming));
  b->chunks = 1;
  if (b->B >= 4096) b->chunks = 8;
  if (const char *v = getenv("EKF_BATCH_CHUNKS")) b->chunks = atoi(v);
  if (b->chunks < 1) b->chunks = 1;
  if (b->chunks > 64) b->chunks = 64;
  if (b->chunks > b->B) b->chunks = b->B;
  const size_t sig_bytes = (size_t)b->B * b->tri * sizeof(double);
  CKC(cudaMalloc(&b->sigma, sig_bytes));
  CKC(cudaMalloc(&b->mu, (size_t)b->B * b->n * sizeof(double)));
  CKC(cudaMalloc(&b->prev, (size_t)b->B * 2 * sizeof(double)));
  CKC(cudaMalloc(&b->cnt, (size_t)b->B * sizeof(int)));
  CKC(cudaMalloc(&b->status, sizeof(int)));
  {
    std::vector<uchar2> map((size_t)b->tri);
    size_t e = 0;
    for (int i = 0; i < b->n; ++i)
      for (int c = 0; c <= i; ++c) map[e++] = make_uchar2((unsigned char)i, (unsigned char)c);
    CKC(cudaMalloc(&b->tri_map, (size_t)b->tri * sizeof(uchar2)));
    CKC(cudaMemcpy(b->tri_map, map.data(), (size_t)b->tri * sizeof(uchar2), cudaMemcpyHostToDevice));
    std::vector<ushort4> blk;
    for (int i0 = 0; i0 < b->n; i0 += 2)
      for (int j0 = 0; j0 <= i0; j0 += 2) {
        const bool row1 = i0 + 1 < b->n;
        const int mask = 1 | (j0 + 1 <= i0 ? 2 : 0) | (row1 ? 4 | 8 : 0);   // (i0+1, j0+1) is stored whenever row i0+1 is
        blk.push_back(make_ushort4((unsigned short)i0, (unsigned short)j0,
                                   (unsigned short)(batch_row_off(i0) + j0), (unsigned short)mask));
      }
    b->nblk = (int)blk.size();
    b->variant = BATCH_DEFAULT_VARIANT;
    if (const char *v = getenv("EKF_BATCH_VARIANT")) b->variant = atoi(v) & 15;   // development override
    if (b->variant == kBatchVariantReg && !batch_reg_supported(b->N, b->n)) b->variant = 6;
    if (b->variant != kBatchVariantReg && (b->variant & 3) == 3) b->variant &= ~1;
    if (b->N > 32) b->variant &= ~4;                                               // one-pass needs a lane per landmark
    std::vector<unsigned short> soff((size_t)b->tri);
    for (size_t q = 0; q < soff.size(); ++q)
      soff[q] = (unsigned short)(batch_dd(b) == 2 ? batch_tile_idx(map[q].x, map[q].y)
                                                       : batch_row_off(map[q].x) + map[q].y);
    CKC(cudaMalloc(&b->soff, soff.size() * sizeof(unsigned short)));
    CKC(cudaMemcpy(b->soff, soff.data(), soff.size() * sizeof(unsigned short), cudaMemcpyHostToDevice));
    CKC(cudaMalloc(&b->blk_map, blk.size() * sizeof(ushort4)));
    CKC(cudaMemcpy(b->blk_map, blk.data(), blk.size() * sizeof(ushort4), cudaMemcpyHostToDevice));
  }
  CKC(cudaMemsetAsync(b->sigma, 0, sig_bytes, b->stream));
  CKC(cudaMemsetAsync(b->mu, 0, (size_t)b->B * b->n * sizeof(double), b->stream));
  CKC(cudaMemsetAsync(b->prev, 0, (size_t)b->B * 2 * sizeof(double), b->stream));
  CKC(cudaMemsetAsync(b->cnt, 0, (size_t)b->B * sizeof(int), b->stream));
  CKC(cudaMemsetAsync(b->status, 0, sizeof(int), b->stream));
  batch_init_prior_kernel<<<(unsigned)(((size_t)b->B * b->n + 255) / 256), 256, 0, b->stream>>>(
      b->sigma, b->B, b->n, b->tri, cfg->prior);   // slam.cpp:128-132
  b->launches++;
  CKC(cudaGetLastError());
#define EKF_BATCH_ATTR(MINB, ONE, DD)                                                                          \
  CKC(cudaFuncSetAttribute(ekf_batch_kernel<MINB, ONE, DD>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                           200 * 1024));                                                                       \
  CKC(cudaFuncSetAttribute(ekf_batch_kernel<MINB, ONE, DD>, cudaFuncAttributePreferredSharedMemoryCarveout,    \
                           cudaSharedmemCarveoutMaxShared));
  EKF_BATCH_ATTR(16, false, 0) EKF_BATCH_ATTR(16, false, 1) EKF_BATCH_ATTR(16, false, 2)
  EKF_BATCH_ATTR(16, true, 0)  EKF_BATCH_ATTR(16, true, 1)  EKF_BATCH_ATTR(16, true, 2)
  EKF_BATCH_ATTR(20, false, 0) EKF_BATCH_ATTR(20, false, 1) EKF_BATCH_ATTR(20, false, 2)
  EKF_BATCH_ATTR(20, true, 0)  EKF_BATCH_ATTR(20, true, 1)  EKF_BATCH_ATTR(20, true, 2)
#undef EKF_BATCH_ATTR
  CKC(cudaFuncSetAttribute(ekf_batch_reg_kernel<kBatchRegMinBlocks>, cudaFuncAttributePreferredSharedMemoryCarveout,
                           cudaSharedmemCarveoutMaxShared));
  CKC(cudaStreamSynchronize(b->stream));
#undef CKC
  *out = b;
  return EKF_OK;
}

int ekf_batch_set_stream(ekf_batch *b, void *cuda_stream) {
  if (!b) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  b->stream = cuda_stream ? (cudaStream_t)cuda_stream : b->own_stream;
  return EKF_OK;
}

int ekf_batch_synchronize(ekf_batch *b) {
  if (!b) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  return EKF_OK;
}

int ekf_batch_last_status(ekf_batch *b) {
  if (!b) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  int dev_status = 0;
  CKB(b, cudaMemcpyAsync(&dev_status, b->status, sizeof(int), cudaMemcpyDeviceToHost, b->stream));
  CKB(b, cudaMemsetAsync(b->status, 0, sizeof(int), b->stream));
  CKB(b, cudaStreamSynchronize(b->stream));
  const int s = b->latched != EKF_OK ? b->latched : dev_status;
  b->latched = EKF_OK;
  return s;
}

int ekf_batch_launch_count(ekf_batch *b, int64_t *launches) {
  if (!b || !launches) return EKF_ERR_INVALID;
  *launches = b->launches;
  return EKF_OK;
}

int ekf_batch_step_bytes(ekf_batch *b, int32_t m, int64_t *bytes) {
  if (!b || !bytes || m < 0) return EKF_ERR_INVALID;
  *bytes = 8LL * b->n * (b->n + 1) + 16LL * b->n + 16LL * m + 24;
  return EKF_OK;
}

int ekf_batch_step_dev(ekf_batch *b, const double *poses_dev, int32_t m, const double *xy_dev,
                       const int32_t *ids_dev, const int32_t *m_count_dev, int32_t *assoc_dev,
                       int32_t *count_dev, double *pose_dev) {
  if (!b || m < 0 || !poses_dev) return EKF_ERR_INVALID;
  if (m == 0) return EKF_ERR_EMPTY_SCAN;
  if (!xy_dev) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  return batch_launch(b, poses_dev, m, xy_dev, ids_dev, m_count_dev, assoc_dev, count_dev, pose_dev);
}

int ekf_batch_step(ekf_batch *b, const double *poses, int32_t m, const double *xy, const int32_t *ids,
                   const int32_t *m_count, int32_t *out_assoc, int32_t *out_count, double *out_pose) {
  if (!b || m < 0 || !poses) return EKF_ERR_INVALID;
  if (m == 0) return EKF_ERR_EMPTY_SCAN;
  if (!xy) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  int rc = batch_ensure_io(b, m);
  if (rc != EKF_OK) return rc;
  const BatchIo io(b->B, m);
  const size_t B = (size_t)b->B;
  // Caller buffers that are page-locked (cudaHostAlloc / cudaHostRegister; torch pin_memory) are copied from / to
  // directly; pageable ones go through the handle's pinned staging (one extra host memcpy each way).
  auto pinned = [](const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
  };
  const bool in_direct = pinned(poses) && pinned(xy) && (!ids || pinned(ids)) && (!m_count || pinned(m_count));
  const bool out_direct = (!out_assoc || pinned(out_assoc)) && (!out_count || pinned(out_count)) &&
                          (!out_pose || pinned(out_pose));
  double *const d_pose = reinterpret_cast<double *>(b->io_dev + io.pose);
  double *const d_xy = reinterpret_cast<double *>(b->io_dev + io.xy);
  int32_t *const d_ids = ids ? reinterpret_cast<int32_t *>(b->io_dev + io.ids) : nullptr;
  int32_t *const d_mc = m_count ? reinterpret_cast<int32_t *>(b->io_dev + io.mc) : nullptr;
  int32_t *const r_assoc = reinterpret_cast<int32_t *>(b->res_dev + io.r_assoc);
  int32_t *const r_cnt = reinterpret_cast<int32_t *>(b->res_dev + io.r_cnt);
  double *const r_pose = reinterpret_cast<double *>(b->res_dev + io.r_pose);
  if (in_direct && out_direct && b->chunks > 1) {
    // chunked pipeline over two side streams (copy engines for the two directions and the SMs all busy at once);
    // every per-filter array is contiguous in filters, so a chunk is one slice of each
    CKB(b, cudaEventRecord(b->ev_begin, b->stream));
    for (int q = 0; q < 2; ++q) CKB(b, cudaStreamWaitEvent(b->side[q], b->ev_begin, 0));
    for (int c = 0; c < b->chunks; ++c) {
      const size_t f0 = B * c / b->chunks, f1 = B * (c + 1) / b->chunks, nf = f1 - f0;
      cudaStream_t st = b->side[c & 1];
      CKB(b, cudaMemcpyAsync(d_pose + 3 * f0, poses + 3 * f0, 24 * nf, cudaMemcpyHostToDevice, st));
      CKB(b, cudaMemcpyAsync(d_xy + 2 * f0 * m, xy + 2 * f0 * m, 16 * nf * m, cudaMemcpyHostToDevice, st));
      if (ids) CKB(b, cudaMemcpyAsync(d_ids + f0 * m, ids + f0 * m, 4 * nf * m, cudaMemcpyHostToDevice, st));
      if (m_count) CKB(b, cudaMemcpyAsync(d_mc + f0, m_count + f0, 4 * nf, cudaMemcpyHostToDevice, st));
      rc = batch_launch(b, d_pose, m, d_xy, d_ids, d_mc, r_assoc, r_cnt, r_pose, (int)f0, (int)nf, st);
      if (rc != EKF_OK) return rc;
      if (out_assoc) CKB(b, cudaMemcpyAsync(out_assoc + f0 * m, r_assoc + f0 * m, 4 * nf * m, cudaMemcpyDeviceToHost, st));
      if (out_count) CKB(b, cudaMemcpyAsync(out_count + f0, r_cnt + f0, 4 * nf, cudaMemcpyDeviceToHost, st));
      if (out_pose) CKB(b, cudaMemcpyAsync(out_pose + 3 * f0, r_pose + 3 * f0, 24 * nf, cudaMemcpyDeviceToHost, st));
    }
    for (int q = 0; q < 2; ++q) {
      CKB(b, cudaEventRecord(b->ev_side[q], b->side[q]));
      CKB(b, cudaStreamWaitEvent(b->stream, b->ev_side[q], 0));
    }
    CKB(b, cudaMemcpyAsync(b->res_host + io.r_status, b->status, sizeof(int), cudaMemcpyDeviceToHost, b->stream));
    CKB(b, cudaMemsetAsync(b->status, 0, sizeof(int), b->stream));
    CKB(b, cudaStreamSynchronize(b->stream));
    int st_chunked = 0;
    memcpy(&st_chunked, b->res_host + io.r_status, 4);
    return st_chunked;
  }
  if (in_direct) {
    CKB(b, cudaMemcpyAsync(b->io_dev + io.pose, poses, 24 * B, cudaMemcpyHostToDevice, b->stream));
    CKB(b, cudaMemcpyAsync(b->io_dev + io.xy, xy, 16 * B * m, cudaMemcpyHostToDevice, b->stream));
    if (ids) CKB(b, cudaMemcpyAsync(b->io_dev + io.ids, ids, 4 * B * m, cudaMemcpyHostToDevice, b->stream));
    if (m_count) CKB(b, cudaMemcpyAsync(b->io_dev + io.mc, m_count, 4 * B, cudaMemcpyHostToDevice, b->stream));
  } else {
    memcpy(b->io_host + io.pose, poses, 24 * B);
    memcpy(b->io_host + io.xy, xy, 16 * B * m);
    if (ids) memcpy(b->io_host + io.ids, ids, 4 * B * m);
    if (m_count) memcpy(b->io_host + io.mc, m_count, 4 * B);
    CKB(b, cudaMemcpyAsync(b->io_dev, b->io_host, io.total, cudaMemcpyHostToDevice, b->stream));
  }
  rc = batch_launch(b, d_pose, m, d_xy, d_ids, d_mc, r_assoc, r_cnt, r_pose);
  if (rc != EKF_OK) return rc;
  CKB(b, cudaMemcpyAsync(b->res_dev + io.r_status, b->status, sizeof(int), cudaMemcpyDeviceToDevice, b->stream));
  CKB(b, cudaMemsetAsync(b->status, 0, sizeof(int), b->stream));
  if (out_direct) {
    if (out_assoc) CKB(b, cudaMemcpyAsync(out_assoc, b->res_dev + io.r_assoc, 4 * B * m, cudaMemcpyDeviceToHost, b->stream));
    if (out_count) CKB(b, cudaMemcpyAsync(out_count, b->res_dev + io.r_cnt, 4 * B, cudaMemcpyDeviceToHost, b->stream));
    if (out_pose) CKB(b, cudaMemcpyAsync(out_pose, b->res_dev + io.r_pose, 24 * B, cudaMemcpyDeviceToHost, b->stream));
    CKB(b, cudaMemcpyAsync(b->res_host + io.r_status, b->res_dev + io.r_status, sizeof(int), cudaMemcpyDeviceToHost,
                           b->stream));
    CKB(b, cudaStreamSynchronize(b->stream));
  } else {
    CKB(b, cudaMemcpyAsync(b->res_host, b->res_dev, io.r_total, cudaMemcpyDeviceToHost, b->stream));
    CKB(b, cudaStreamSynchronize(b->stream));
    if (out_assoc) memcpy(out_assoc, b->res_host + io.r_assoc, 4 * B * m);
    if (out_count) memcpy(out_count, b->res_host + io.r_cnt, 4 * B);
    if (out_pose) memcpy(out_pose, b->res_host + io.r_pose, 24 * B);
  }
  int dev_status = 0;
  memcpy(&dev_status, b->res_host + io.r_status, 4);
  return dev_status;
}

// ---- accessors / fixtures (host-side pack / unpack of one filter; never on the timed path) ----
int ekf_batch_get_state(ekf_batch *b, int32_t index, double *out, int32_t n) {
  if (!b || !out || index < 0 || index >= b->B || n < 0 || n > b->n) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  CKB(b, cudaMemcpy(out, b->mu + (size_t)index * b->n, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
  return EKF_OK;
}

int ekf_batch_set_state(ekf_batch *b, int32_t index, const double *in, int32_t n) {
  if (!b || !in || index < 0 || index >= b->B || n < 0 || n > b->n) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  CKB(b, cudaMemcpy(b->mu + (size_t)index * b->n, in, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
  return EKF_OK;
}

int ekf_batch_get_sigma(ekf_batch *b, int32_t index, double *out) {
  if (!b || !out || index < 0 || index >= b->B) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  std::vector<double> tri((size_t)b->tri);
  CKB(b, cudaMemcpy(tri.data(), b->sigma + (size_t)index * b->tri, tri.size() * sizeof(double),
                    cudaMemcpyDeviceToHost));
  const int n = b->n;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j) out[(size_t)i * n + j] = out[(size_t)j * n + i] = tri[(size_t)i * (i + 1) / 2 + j];
  return EKF_OK;
}

int ekf_batch_set_sigma(ekf_batch *b, int32_t index, const double *in) {
  if (!b || !in || index < 0 || index >= b->B) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  std::vector<double> tri((size_t)b->tri);
  const int n = b->n;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j) tri[(size_t)i * (i + 1) / 2 + j] = in[(size_t)i * n + j];
  CKB(b, cudaMemcpy(b->sigma + (size_t)index * b->tri, tri.data(), tri.size() * sizeof(double),
                    cudaMemcpyHostToDevice));
  return EKF_OK;
}

int ekf_batch_get_count(ekf_batch *b, int32_t index, int32_t *count) {
  if (!b || !count || index < 0 || index >= b->B) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  CKB(b, cudaMemcpy(count, b->cnt + index, sizeof(int), cudaMemcpyDeviceToHost));
  return EKF_OK;
}

int ekf_batch_set_count(ekf_batch *b, int32_t index, int32_t count) {
  if (!b || index < 0 || index >= b->B || count < 0 || count > b->N) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  CKB(b, cudaMemcpy(b->cnt + index, &count, sizeof(int), cudaMemcpyHostToDevice));
  return EKF_OK;
}

int ekf_batch_set_previous_xy(ekf_batch *b, int32_t index, double x, double y) {
  if (!b || index < 0 || index >= b->B) return EKF_ERR_INVALID;
  DeviceGuard guard(b->device);
  CKB(b, cudaStreamSynchronize(b->stream));
  const double v[2] = {x, y};
  CKB(b, cudaMemcpy(b->prev + 2 * (size_t)index, v, sizeof(v), cudaMemcpyHostToDevice));
  return EKF_OK;
}

}  // extern "C"
