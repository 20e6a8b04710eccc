// Part of ekf_abi.cu: the scan update (prediction + measurement loop + covariance pass).

namespace {

// status + results -> pinned host staging, then the event ekf_scan's caller waits on.  Enqueued BEFORE the
// covariance pass of the scan's last chunk: association, pose and count do not depend on it.
int results_out(ekf_filter *f) {
  const size_t total = f->res_pending;
  f->res_pending = 0;
  if (!f->res_direct) {
    CK(f, cudaMemcpyAsync(f->res_dev + ResLayout(0).status, f->status, sizeof(int), cudaMemcpyDeviceToDevice, f->stream));
    CK(f, cudaMemsetAsync(f->status, 0, sizeof(int), f->stream));
    CK(f, cudaMemcpyAsync(f->res_host, f->res_dev, total, cudaMemcpyDeviceToHost, f->stream));
  }   // else: the finishing sequential launch stored pose / count / association / status in res_host itself
  CK(f, cudaEventRecord(f->res_event, f->stream));
  return EKF_OK;
}

// Launches the sequential kernel for measurements [t0, t0+mc) of a scan and, when mc > 0, the
// covariance pass for the 2*mc panel columns it produced.  All pointers are device pointers.
int launch_chunk(ekf_filter *f, const double *pose_dev, int do_predict, int mc, const double *xy_dev,
                 const int32_t *ids_dev, const uint8_t *action_dev, int32_t *assoc_dev, int finish,
                 double *pose_out_dev, int *cnt_out_dev) {
  SeqParams sp{};
  sp.N = f->N; sp.n_pad = f->n_pad; sp.kmax = f->kmax; sp.lay = f->lay;
  sp.side_mode = 0; sp.cols = nullptr; sp.guess = nullptr; sp.done_out = nullptr;
  sp.q_noise = f->cfg.q_noise; sp.r_noise = f->cfg.r_noise; sp.thr = f->cfg.mah_threshold;
  sp.sigma = f->sigma; sp.mu = f->mu; sp.Pp = f->Pp; sp.D = f->D; sp.W = f->W; sp.V = f->V;
  sp.cnt = f->cnt; sp.prev_xy = f->prev_xy; sp.status = f->status;
  sp.pose = pose_dev; sp.xy = xy_dev; sp.ids = ids_dev; sp.action = action_dev;
  sp.m = mc; sp.do_predict = do_predict; sp.finish = finish;
  sp.assoc_out = assoc_dev; sp.pose_out = pose_out_dev; sp.cnt_out = cnt_out_dev;
  sp.status_out = f->res_direct ? reinterpret_cast<int *>(f->res_host_dev + ResLayout(0).status) : nullptr;
  sp.bar = f->bar; sp.bar_base = f->bar_base;
  sp.slot_d = f->slot_d; sp.slot_k = f->slot_k; sp.slot_rec = f->slot_rec;
  f->bar_base += (unsigned long long)(mc > 0 ? mc : 1) * (unsigned long long)f->seq_grid;
  const bool hot = f->seq_hot && f->seq_threads == kSeqThreadsSmall;
  if (hot) { sp.hot_hdr = f->hot_hdr; sp.hot_rec = f->hot_rec; sp.hot_w = f->hot_w; }

  const bool timed = f->ev_cap > 0 && mc > 0;
  const size_t slot = timed ? (size_t)(f->ev_count % f->ev_cap) * 4 : 0;
  if (timed) CK(f, cudaEventRecord(f->ev[slot], f->stream));
  void *args[] = {(void *)&sp};
  const void *seq_fn = f->seq_threads == kSeqThreadsSmall ? (const void *)ekf_seq_kernel<kSeqThreadsSmall>
                                                          : (const void *)ekf_seq_kernel<kSeqThreadsLarge>;
  CK(f, cudaLaunchCooperativeKernel(seq_fn, dim3(f->seq_grid), dim3(f->seq_threads), args,
                                    seq_smem_bytes(f->seq_threads, hot), f->stream));
  f->launches++;
  if (timed) { CK(f, cudaEventRecord(f->ev[slot + 1], f->stream)); CK(f, cudaEventRecord(f->ev[slot + 2], f->stream)); }
  if (finish && f->res_pending) {                        // everything the caller reads back is final here
    const int rco = results_out(f);
    if (rco != EKF_OK) return rco;
  }
  if (mc > 0) {
    DdParams dp{};
    dp.sigma = f->sigma; dp.V = f->V; dp.W = f->W;
    dp.n_pad = f->n_pad; dp.kmax = f->kmax; dp.k2 = (2 * mc + 3) & ~3;  // whole k-steps; pad columns are zero
    dp.lay = f->lay; dp.row_begin = f->lay.rb; dp.row_end = f->lay.re;
    dp.prefetch_ahead = f->dd_prefetch;
    if (f->dd_variant == 0 || f->dd_variant == 2) {
      dp.total_chunks = dd_total_chunks(64, dp.row_begin, dp.row_end);
      long long grid = f->dd_grid;
      if (grid > dp.total_chunks) grid = dp.total_chunks;
      if (f->dd_variant == 0)
        ekf_downdate_kernel<64, 1><<<(unsigned)grid, dd_threads(64), dd_smem_bytes(64, dp.k2), f->stream>>>(dp);
      else
        ekf_downdate_kernel<64, 2><<<(unsigned)grid, dd_threads(64), dd_smem_bytes(64, dp.k2), f->stream>>>(dp);
    } else {
      dp.total_chunks = dd_total_chunks(32, dp.row_begin, dp.row_end);
      long long grid = (long long)f->dd_grid * f->dd_ctas;
      if (grid > dp.total_chunks) grid = dp.total_chunks;
      if (f->dd_variant == 3)
        ekf_downdate_kernel<32, 2><<<(unsigned)grid, dd_threads(32), dd_smem_bytes(32, dp.k2), f->stream>>>(dp);
      else
        ekf_downdate_kernel<32, 0><<<(unsigned)grid, dd_threads(32), dd_smem_bytes(32, dp.k2), f->stream>>>(dp);
    }
    f->launches++;
    CK(f, cudaGetLastError());
  }
  if (timed) { CK(f, cudaEventRecord(f->ev[slot + 3], f->stream)); f->ev_count++; }
  return EKF_OK;
}

// One round (at most kMaxChunkM sightings) of the overlapped path (ekf_abi_overlap.inl): sequential side on
// `stream`, its covariance pass queued on `back_stream`.  Returns the number of sightings the round processed
// in *done (a failed speculation stops the round early; the caller runs another one for the rest).
int overlap_round(ekf_filter *f, const double *pose_dev, int mc, const double *xy_dev, const int32_t *ids_dev,
                  int32_t *assoc_dev, int finish, double *pose_out_dev, int *cnt_out_dev, size_t res_bytes, int *done) {
  const int64_t k = f->pipe_step;
  double *const buf[2] = {f->sigma, f->sigma2};
  double *const Wp[2] = {f->W, f->W2}, *const Vp[2] = {f->V, f->V2};
  const int cur = (int)(k & 1), prv = cur ^ 1;
  if (k >= 2) CK(f, cudaStreamWaitEvent(f->stream, f->ev_pass[cur], 0));   // pass k-2: its output is read below
  const bool timed = f->ev_cap > 0;
  const size_t slot = timed ? (size_t)(f->ev_count % f->ev_cap) * 4 : 0;
  if (timed) CK(f, cudaEventRecord(f->ev[slot], f->stream));
  if (pose_dev) {
    shard_predict_kernel<<<f->num_sms, 256, 0, f->stream>>>(f->Pp, f->mu, pose_dev, f->prev_xy, f->N, f->cfg.q_noise);
    f->launches++;
  }
  // speculative association on 128-thread CTAs: 128 x 96 registers fit beside the pass's CTA, 256 would wait for it
  constexpr int kPipeGuessThreads = 128;
  int slices = (f->N + kPipeGuessThreads - 1) / kPipeGuessThreads;
  if (slices > kGuessMaxSlices) slices = kGuessMaxSlices;
  if (slices < 1) slices = 1;
  shard_guess_kernel<<<dim3(mc, slices), kPipeGuessThreads, 0, f->stream>>>(
      f->Pp, f->D, f->mu, f->cnt, xy_dev, ids_dev, f->guess, f->cfg.r_noise, f->cfg.mah_threshold, f->guess_part_d,
      f->guess_part_k, f->guess_ticket);
  const int drop = env_int("EKF_SHARD_DROP_GUESS", -1);                      // test-only fault injection
  if (drop >= 0 && drop < mc) CK(f, cudaMemsetAsync(f->guess + drop, 0xFF, sizeof(int), f->stream));
  // raw columns of the covariance as pass k-1 will leave it: its input minus its update (k2 = 0: nothing pending)
  pipe_extract_cols_kernel<<<f->num_sms * 4, 256, 0, f->stream>>>(
      k >= 1 ? buf[prv] : buf[0], f->lay, f->n_pad, f->N, f->guess, mc, Vp[prv], Wp[prv], f->kmax,
      k >= 1 ? f->pipe_k2_prev : 0, f->cols);
  f->launches += 2;
  CK(f, cudaGetLastError());

  SeqParams sp{};
  sp.N = f->N; sp.n_pad = f->n_pad; sp.kmax = f->kmax; sp.lay = f->lay;
  sp.q_noise = f->cfg.q_noise; sp.r_noise = f->cfg.r_noise; sp.thr = f->cfg.mah_threshold;
  sp.sigma = buf[cur]; sp.mu = f->mu; sp.Pp = f->Pp; sp.D = f->D; sp.W = Wp[cur]; sp.V = Vp[cur];
  sp.cnt = f->cnt; sp.prev_xy = f->prev_xy; sp.status = f->status;
  sp.pose = nullptr; sp.xy = xy_dev; sp.ids = ids_dev; sp.action = nullptr;
  sp.m = mc; sp.do_predict = 0; sp.finish = finish;
  sp.assoc_out = assoc_dev; sp.pose_out = pose_out_dev; sp.cnt_out = cnt_out_dev;
  sp.side_mode = 1; sp.cols = f->cols; sp.guess = f->guess; sp.done_out = f->done_dev;
  sp.bar = f->bar; sp.bar_base = f->bar_base;
  sp.slot_d = f->slot_d; sp.slot_k = f->slot_k; sp.slot_rec = f->slot_rec;
  f->bar_base += (unsigned long long)mc * (unsigned long long)f->pipe_grid;
  // plain launch below: the hand-written grid barrier then has no residency guarantee from the driver (another
  // handle's kernels, another process or a long-running kernel may hold SMs), so its wait is bounded - a grid that
  // is only partly resident latches EKF_ERR_CUDA after this long instead of hanging the GPU
  sp.bar_timeout_ns = (unsigned long long)env_int("EKF_PIPE_BARRIER_TIMEOUT_MS", 2000) * 1000000ULL;
  // NOT a cooperative launch: the driver holds a cooperative grid back until the device is otherwise idle, which
  // would put it behind the very pass it is meant to overlap.  Co-residency - all the hand-written grid barrier
  // needs - holds by construction: at most one CTA per SM, and a CTA fits beside the pass's CTA (registers, shared
  // memory); the only other kernels that can be resident (the pass on the back stream, other handles' kernels) never
  // wait for this one.
  if (env_int("EKF_PIPE_COOP", 0)) {
    sp.bar_timeout_ns = 0;
    void *args[] = {(void *)&sp};
    CK(f, cudaLaunchCooperativeKernel((const void *)ekf_seq_kernel<kSeqThreadsSmall>, dim3(f->pipe_grid),
                                      dim3(kSeqThreadsSmall), args, seq_smem_bytes(kSeqThreadsSmall), f->stream));
  } else {
    ekf_seq_kernel<kSeqThreadsSmall><<<f->pipe_grid, kSeqThreadsSmall, seq_smem_bytes(kSeqThreadsSmall), f->stream>>>(sp);
    CK(f, cudaGetLastError());
  }
  f->launches++;
  if (timed) CK(f, cudaEventRecord(f->ev[slot + 1], f->stream));
  CK(f, cudaEventRecord(f->ev_seq[cur], f->stream));
  // what the host needs of this round: how far it got, and (host-buffer scans) the results
  CK(f, cudaMemcpyAsync(f->done_host, f->done_dev, sizeof(int), cudaMemcpyDeviceToHost, f->stream));
  if (res_bytes) {
    CK(f, cudaMemcpyAsync(f->res_dev + ResLayout(0).status, f->status, sizeof(int), cudaMemcpyDeviceToDevice, f->stream));
    CK(f, cudaMemcpyAsync(f->res_host, f->res_dev, res_bytes, cudaMemcpyDeviceToHost, f->stream));
  }
  CK(f, cudaEventRecord(f->done_event, f->stream));

  CK(f, cudaStreamWaitEvent(f->back_stream, f->ev_seq[cur], 0));
  DdParams dp{};
  dp.sigma = buf[cur]; dp.sigma_out = buf[prv]; dp.V = Vp[cur]; dp.W = Wp[cur];
  dp.n_pad = f->n_pad; dp.kmax = f->kmax; dp.k2 = (2 * mc + 3) & ~3;
  dp.lay = f->lay; dp.row_begin = f->lay.rb; dp.row_end = f->lay.re;
  dp.prefetch_ahead = f->dd_prefetch;
  dp.total_chunks = dd_total_chunks(64, dp.row_begin, dp.row_end);
  long long grid = f->dd_grid;
  if (grid > dp.total_chunks) grid = dp.total_chunks;
  if (timed) CK(f, cudaEventRecord(f->ev[slot + 2], f->back_stream));
  ekf_downdate_kernel<64, 2><<<(unsigned)grid, dd_threads(64), dd_smem_bytes(64, dp.k2), f->back_stream>>>(dp);
  f->launches++;
  CK(f, cudaGetLastError());
  if (timed) { CK(f, cudaEventRecord(f->ev[slot + 3], f->back_stream)); f->ev_count++; }
  CK(f, cudaEventRecord(f->ev_pass[cur], f->back_stream));
  f->pipe_k2_prev = dp.k2;
  f->pipe_step = k + 1;
  f->pipe_active = true;

  CK(f, cudaEventSynchronize(f->done_event));          // the pass keeps running behind this
  *done = *f->done_host;
  return EKF_OK;
}

// A whole scan on the overlapped path.  res_bytes > 0: host-buffer scan, results land in res_host every round.
int scan_overlap(ekf_filter *f, const double *pose_dev, int m, const double *xy_dev, const int32_t *ids_dev,
                 int32_t *assoc_dev, double *pose_out_dev, int *cnt_out_dev, size_t res_bytes) {
  if (!f->side_valid) {                                  // Pp / D from the (drained, current) covariance
    shard_side_extract_kernel<<<f->num_sms, 256, 0, f->stream>>>(f->sigma, f->lay, f->n_pad, f->N, f->Pp, f->D);
    shard_side_fix_pose_kernel<<<1, 32, 0, f->stream>>>(f->Pp);
    f->launches += 2;
    CK(f, cudaGetLastError());
    f->side_valid = true;
  }
  int t0 = 0;
  bool first = true;
  while (t0 < m) {
    const int mc = (m - t0) < kMaxChunkM ? (m - t0) : kMaxChunkM;
    int done = 0;
    const int rc = overlap_round(f, first ? pose_dev : nullptr, mc, xy_dev + 2 * t0, ids_dev ? ids_dev + t0 : nullptr,
                                 assoc_dev ? assoc_dev + t0 : nullptr, t0 + mc >= m, pose_out_dev, cnt_out_dev, res_bytes,
                                 &done);
    if (rc != EKF_OK) return rc;
    if (done < 0 || done > mc) {
      // the bounded grid barrier gave up (done == -1) or the device wrote nonsense: wait for every CTA of the launch
      // to leave, then start the arrival counter afresh so that the handle stays usable on the classic path
      cudaStreamSynchronize(f->stream);
      if (f->back_stream) cudaStreamSynchronize(f->back_stream);
      cudaMemsetAsync(f->bar, 0, sizeof(unsigned long long), f->stream);
      f->bar_base = 0;
      f->cuda_err = "overlapped sequential kernel: grid barrier timed out (grid not co-resident)";
      return EKF_ERR_CUDA;
    }
    first = false;
    t0 += done;
    if (done == 0) break;       // a round that cannot get past its first sighting: finish the scan classically
  }
  if (res_bytes) {
    CK(f, cudaMemsetAsync(f->status, 0, sizeof(int), f->stream));   // the latched status went out with the results
    f->res_pending = 0;
  }
  if (t0 < m) {
    // the prediction lives in Pp / mu only; the drain writes it into the covariance, then the classic path
    // (which re-extracts its side state) takes the remaining sightings without predicting again
    int rc = pipe_drain(f);
    if (rc != EKF_OK) return rc;
    f->side_valid = false;
    f->res_pending = res_bytes;
    while (t0 < m) {
      const int mc = (m - t0) < kMaxChunkM ? (m - t0) : kMaxChunkM;
      rc = launch_chunk(f, nullptr, 0, mc, xy_dev + 2 * t0, ids_dev ? ids_dev + t0 : nullptr, nullptr,
                        assoc_dev ? assoc_dev + t0 : nullptr, t0 + mc >= m, pose_out_dev, cnt_out_dev);
      if (rc != EKF_OK) return rc;
      t0 += mc;
    }
  }
  return EKF_OK;
}

// One whole scan from device inputs; splits into chunks of kMaxChunkM measurements.
int scan_device(ekf_filter *f, const double *pose_dev, int m, const double *xy_dev, const int32_t *ids_dev,
                const uint8_t *action_dev, int32_t *assoc_dev, double *pose_out_dev, int *cnt_out_dev) {
  if (f->sharded) return EKF_ERR_INVALID;   // a row shard cannot run a scan alone: ekf_shard_scan_*
  // overlapped mode: whole scans (pose + sightings) without per-sighting actions; everything else runs classically
  if (f->overlap && pose_dev != nullptr && m > 0 && action_dev == nullptr) {
    const int rcp = pipe_resources(f);
    if (rcp == EKF_OK)
      return scan_overlap(f, pose_dev, m, xy_dev, ids_dev, assoc_dev, pose_out_dev, cnt_out_dev, f->res_pending);
    if (rcp != EKF_ERR_NOMEM) return rcp;
    f->overlap = false;                     // no room for the second covariance buffer: classic path from here on
  }
  {
    const int rcd = pipe_drain(f);
    if (rcd != EKF_OK) return rcd;
  }
  f->side_valid = false;                    // this path re-extracts Pp / D from Sigma every scan
  int t0 = 0;
  bool first = true;
  do {
    const int mc = (m - t0) < kMaxChunkM ? (m - t0) : kMaxChunkM;
    const int last = (t0 + mc >= m);
    int rc = launch_chunk(f, pose_dev, first && pose_dev != nullptr, mc, xy_dev ? xy_dev + 2 * t0 : nullptr,
                          ids_dev ? ids_dev + t0 : nullptr, action_dev ? action_dev + t0 : nullptr,
                          assoc_dev ? assoc_dev + t0 : nullptr, last && m > 0, pose_out_dev, cnt_out_dev);
    if (rc != EKF_OK) return rc;
    first = false;
    t0 += mc;
  } while (t0 < m);
  return EKF_OK;
}

// Host-pointer scan: pack inputs into pinned staging, one H2D, kernels, one D2H, sync.
int scan_host(ekf_filter *f, const double *pose, int m, const double *xy, const int32_t *ids, const uint8_t *action,
              int32_t *out_assoc, int32_t *out_count, double *out_pose) {
  int rc = ensure_io(f, m);
  if (rc != EKF_OK) return rc;
  const IoLayout in(m);
  const ResLayout res(m);
  if (pose) memcpy(f->io_host + in.pose, pose, 24);
  if (m > 0) {
    memcpy(f->io_host + in.xy, xy, (size_t)16 * m);
    if (ids) memcpy(f->io_host + in.ids, ids, (size_t)4 * m);
    if (action) memcpy(f->io_host + in.action, action, (size_t)m);
  }
  CK(f, cudaMemcpyAsync(f->io_dev, f->io_host, in.total, cudaMemcpyHostToDevice, f->stream));
  f->res_pending = res.total;
  // classic path with sightings: the sequential kernel stores its few result words straight into the mapped pinned
  // buffer (posted PCIe writes) - no copy operations between it and the covariance pass (they cost ~3 us each of
  // GPU time line); the caller still waits for the event recorded right after that launch
  f->res_direct = m > 0 && !f->overlap && f->res_host_dev != nullptr;
  unsigned char *rb = f->res_direct ? f->res_host_dev : f->res_dev;
  rc = scan_device(f, pose ? reinterpret_cast<const double *>(f->io_dev + in.pose) : nullptr, m,
                   reinterpret_cast<const double *>(f->io_dev + in.xy),
                   ids ? reinterpret_cast<const int32_t *>(f->io_dev + in.ids) : nullptr,
                   action ? reinterpret_cast<const uint8_t *>(f->io_dev + in.action) : nullptr,
                   reinterpret_cast<int32_t *>(rb + res.assoc), reinterpret_cast<double *>(rb + res.pose),
                   reinterpret_cast<int *>(rb + res.cnt));
  if (rc != EKF_OK) { f->res_pending = 0; f->res_direct = false; return rc; }
  if (f->res_pending) {                           // a scan without measurements has no finishing launch
    rc = results_out(f);
    if (rc != EKF_OK) return rc;
  }
  // the covariance pass of this scan may still be running: it is stream-ordered before anything that
  // reads Sigma, and the next call's kernels queue up behind it
  CK(f, cudaEventSynchronize(f->res_event));
  f->res_direct = false;
  if (out_assoc && m > 0) memcpy(out_assoc, f->res_host + res.assoc, (size_t)4 * m);
  if (out_count) memcpy(out_count, f->res_host + res.cnt, 4);
  if (out_pose) memcpy(out_pose, f->res_host + res.pose, 24);
  int dev_status = 0;
  memcpy(&dev_status, f->res_host + res.status, 4);
  return dev_status;
}

}  // namespace

extern "C" {

int ekf_predict(ekf_filter *f, double theta, double x, double y) {
  if (!f) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  const double pose[3] = {theta, x, y};
  return scan_host(f, pose, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
}

int ekf_update_known(ekf_filter *f, int32_t m, const int32_t *ids, const double *xy, const uint8_t *action) {
  if (!f || m < 0) return EKF_ERR_INVALID;
  if (m == 0) return EKF_ERR_EMPTY_SCAN;
  if (!ids || !xy) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  return scan_host(f, nullptr, m, xy, ids, action, nullptr, nullptr, nullptr);
}

int ekf_update_unknown(ekf_filter *f, int32_t m, const double *xy, int32_t *out_assoc, int32_t *out_count) {
  if (!f || m < 0) return EKF_ERR_INVALID;
  if (m == 0) return EKF_ERR_EMPTY_SCAN;
  if (!xy) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  return scan_host(f, nullptr, m, xy, nullptr, nullptr, out_assoc, out_count, nullptr);
}

int ekf_scan(ekf_filter *f, const double pose[3], int32_t m, const double *xy, const int32_t *ids,
             const uint8_t *action, int32_t *out_assoc, int32_t *out_count, double out_pose[3]) {
  if (!f || m < 0 || !pose) return EKF_ERR_INVALID;
  if (m == 0) return EKF_ERR_EMPTY_SCAN;
  if (!xy) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  return scan_host(f, pose, m, xy, ids, action, out_assoc, out_count, out_pose);
}

int ekf_scan_dev(ekf_filter *f, const double *pose_dev, int32_t m, const double *xy_dev, const int32_t *ids_dev,
                 const uint8_t *action_dev, int32_t *assoc_dev) {
  if (!f || m < 0 || !pose_dev) return EKF_ERR_INVALID;
  if (m == 0) return EKF_ERR_EMPTY_SCAN;
  if (!xy_dev) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  int rc = ensure_io(f, 1);
  if (rc != EKF_OK) return rc;
  const ResLayout res(1);
  return scan_device(f, pose_dev, m, xy_dev, ids_dev, action_dev, assoc_dev,
                     reinterpret_cast<double *>(f->res_dev + res.pose), reinterpret_cast<int *>(f->res_dev + res.cnt));
}

}  // extern "C"
