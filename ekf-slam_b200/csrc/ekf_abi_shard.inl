// Part of ekf_abi.cu: row-sharded single large map (kernels and protocol: ekf_shard.cuh).

namespace {

// the shard protocol also runs on an unsharded handle (one shard that owns every row)
int shard_ok(ekf_filter *f) { return f != nullptr; }

// the peer-visible exchange region: two column buffers (round parity, poisoned with kColsSentinel) + per-rank
// completion flags [kMaxPeers] + per-rank credits [kMaxPeers]
constexpr size_t kXchgFlagBytes = 2 * kMaxPeers * sizeof(unsigned long long);

__global__ void shard_poison_kernel(double *buf, int64_t doubles) {
  const double v = __longlong_as_double((long long)kColsSentinel);
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < doubles; e += (int64_t)gridDim.x * blockDim.x)
    buf[e] = v;
}
int shard_xchg_region(ekf_filter *f) {
  if (f->xchg) return EKF_OK;
  f->xchg_cols_bytes = (size_t)kMaxChunkM * f->n_pad * 2 * sizeof(double);
  const size_t total = 2 * f->xchg_cols_bytes + kXchgFlagBytes;
  CK(f, cudaMalloc(&f->xchg, total));
  CK(f, cudaMemset(f->xchg, 0, total));
  shard_poison_kernel<<<f->num_sms * 4, 256>>>(reinterpret_cast<double *>(f->xchg),
                                               (int64_t)(2 * f->xchg_cols_bytes / sizeof(double)));
  CK(f, cudaGetLastError());
  CK(f, cudaDeviceSynchronize());           // poisoned before any peer can be told about the region
  CK(f, cudaMalloc(&f->push_done, 2 * sizeof(unsigned int)));   // tickets: push kernel, recycle kernel
  CK(f, cudaMemset(f->push_done, 0, 2 * sizeof(unsigned int)));
  f->peer_base[f->cfg.shard_rank] = f->xchg;
  f->peers_attached = 1;
  return EKF_OK;
}
bool shard_peer_mode(const ekf_filter *f) { return f->xchg != nullptr && f->peers_attached == f->cfg.shard_count; }

}  // namespace

extern "C" {

int ekf_shard_peer_export(ekf_filter *f, void *ipc_handle_64, void **local_base, int64_t *region_bytes) {
  if (!shard_ok(f) || f->cfg.shard_count > kMaxPeers) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  int rc = shard_xchg_region(f);
  if (rc != EKF_OK) return rc;
  if (ipc_handle_64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
    cudaIpcMemHandle_t h;
    CK(f, cudaIpcGetMemHandle(&h, f->xchg));
    memcpy(ipc_handle_64, &h, 64);
  }
  if (local_base) *local_base = f->xchg;
  if (region_bytes) *region_bytes = (int64_t)(2 * f->xchg_cols_bytes + kXchgFlagBytes);
  return EKF_OK;
}

int ekf_shard_peer_attach(ekf_filter *f, int32_t peer_rank, const void *ipc_handle_64, void *same_process_base) {
  if (!shard_ok(f) || peer_rank < 0 || peer_rank >= f->cfg.shard_count || peer_rank == f->cfg.shard_rank)
    return EKF_ERR_INVALID;
  if ((ipc_handle_64 == nullptr) == (same_process_base == nullptr)) return EKF_ERR_INVALID;   // exactly one
  DeviceGuard guard(f->device);
  int rc = shard_xchg_region(f);
  if (rc != EKF_OK) return rc;
  if (f->peer_base[peer_rank]) return EKF_ERR_INVALID;        // already attached
  if (ipc_handle_64) {
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle_64, 64);
    void *p = nullptr;
    CK(f, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    f->peer_ipc[peer_rank] = p;
    f->peer_base[peer_rank] = static_cast<unsigned char *>(p);
  } else {
    f->peer_base[peer_rank] = static_cast<unsigned char *>(same_process_base);
  }
  f->peers_attached++;
  return EKF_OK;
}

int ekf_shard_info(ekf_filter *f, int32_t *row_begin, int32_t *row_end, int64_t *stored_doubles) {
  if (!f) return EKF_ERR_INVALID;
  if (row_begin) *row_begin = f->lay.rb;
  if (row_end) *row_end = f->lay.re;
  if (stored_doubles) *stored_doubles = f->sigma_elems;
  return EKF_OK;
}

int ekf_shard_sync_begin(ekf_filter *f, double **xchg_dev, int64_t *xchg_doubles) {
  if (!shard_ok(f) || !xchg_dev || !xchg_doubles) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }
  int rcb = shard_buffers(f);
  if (rcb != EKF_OK) return rcb;
  shard_side_extract_kernel<<<f->num_sms, 256, 0, f->stream>>>(f->sigma, f->lay, f->n_pad, f->N, f->Pp, f->D);
  f->launches++;
  CK(f, cudaGetLastError());
  *xchg_dev = reinterpret_cast<double *>(f->Pp);
  *xchg_doubles = 4 * ((int64_t)f->n_pad + f->N);
  return EKF_OK;
}

int ekf_shard_sync_finish(ekf_filter *f) {
  if (!shard_ok(f)) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  shard_side_fix_pose_kernel<<<1, 32, 0, f->stream>>>(f->Pp);
  f->launches++;
  CK(f, cudaGetLastError());
  f->side_valid = true;
  return EKF_OK;
}

int ekf_shard_scan_begin(ekf_filter *f, const double *pose_dev, int32_t m, const double *xy_dev,
                         const int32_t *ids_dev, int32_t last_round, double **xchg_dev, int64_t *xchg_doubles) {
  if (!shard_ok(f) || m < 1 || !xy_dev || !xchg_dev || !xchg_doubles) return EKF_ERR_INVALID;
  if (!f->side_valid) return EKF_ERR_INVALID;        // ekf_shard_sync_begin / _finish first
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f, /*keep_side=*/true); if (rcd != EKF_OK) return rcd; }
  const int mc = m < kMaxChunkM ? m : kMaxChunkM;    // a round handles at most kMaxChunkM sightings
  int slices = (2 * f->num_sms + mc - 1) / mc;        // about two CTAs per SM in total
  const int useful = (f->N + kGuessThreads - 1) / kGuessThreads;
  if (slices > useful) slices = useful;
  if (slices > kGuessMaxSlices) slices = kGuessMaxSlices;
  if (slices < 1) slices = 1;
  f->pend_peer = shard_peer_mode(f) ? 1 : 0;
  { const int rcb = shard_buffers(f); if (rcb != EKF_OK) return rcb; }
  // The previous round's covariance pass runs on the shard stream; what the caller enqueued on the handle's stream
  // since (the input copies of this scan) ran behind it.  The kernels of this round do not: measured (2 GPUs,
  // N = 10,000) a speculative association co-resident with the pass slows the pass by a third - its CTAs take the
  // registers the pass's CTA needs on the same SM - for 16 us of hidden work.
  if (f->shard_pass_pending) {
    CK(f, cudaStreamWaitEvent(f->stream, f->shard_ev_pass, 0));
    f->shard_pass_pending = false;
  }
  // (1) speculative association, prediction applied on the fly (the sequential kernel applies it to the replicated Pp
  //     when it starts)
  BeginParams bp{};
  bp.Pp = f->Pp; bp.D = f->D; bp.mu = f->mu; bp.cnt = f->cnt;
  bp.pose = pose_dev; bp.prev_xy = f->prev_xy;
  bp.q_noise = f->cfg.q_noise; bp.r_noise = f->cfg.r_noise; bp.thr = f->cfg.mah_threshold;
  bp.xy = xy_dev; bp.ids = ids_dev; bp.guess = f->guess;
  bp.part_d = f->guess_part_d; bp.part_k = f->guess_part_k; bp.ticket = f->guess_ticket;
  bp.drop = env_int("EKF_SHARD_DROP_GUESS", -1);       // test-only fault injection (stop / resume path)
  if (bp.drop >= mc) bp.drop = -1;
  const bool timed_x = f->ev_cap > 0;
  const size_t xs = timed_x ? (size_t)(f->ev_count % f->ev_cap) * 4 : 0;
  if (timed_x) CK(f, cudaEventRecord(f->ev_x[xs], f->stream));
  shard_begin_kernel<<<dim3(mc, slices), kGuessThreads, 0, f->stream>>>(bp);
  f->launches++;
  if (timed_x) CK(f, cudaEventRecord(f->ev_x[xs + 1], f->stream));
  // (2) the columns, from Sigma as the previous round's pass left it
  if (f->pend_peer) {
    // extraction + delivery + arrival flags in one kernel over peer memory: nothing left for the caller
    f->xchg_round++;
    PeerTable pt{};
    pt.count = f->cfg.shard_count; pt.self = f->cfg.shard_rank;
    const size_t par_off = (size_t)(f->xchg_round & 1) * f->xchg_cols_bytes;
    for (int q = 0; q < pt.count; ++q) {
      pt.cols[q] = reinterpret_cast<double *>(f->peer_base[q] + par_off);
      pt.flags[q] = reinterpret_cast<unsigned long long *>(f->peer_base[q] + 2 * f->xchg_cols_bytes);
    }
    // ... on the SHARD stream, ordered after the speculation: the sequential kernel (handle's stream) starts at once and
    // polls the elements it needs as they land (flag-in-data), so the delivery hides behind its first measurements
    CK(f, cudaEventRecord(f->shard_ev_seq, f->stream));
    CK(f, cudaStreamWaitEvent(f->shard_stream, f->shard_ev_seq, 0));
    if (timed_x) CK(f, cudaEventRecord(f->ev_x[xs + 2], f->shard_stream));
    shard_push_cols_kernel<<<(f->n_pad + kPushThreads - 1) / kPushThreads, kPushThreads, 0, f->shard_stream>>>(
        f->sigma, f->lay, f->n_pad, f->N, f->guess, mc, pt,
        reinterpret_cast<const unsigned long long *>(f->xchg + 2 * f->xchg_cols_bytes) + kMaxPeers, f->xchg_round,
        f->push_done, f->status, env_int("EKF_SHARD_TWO_HOP", pt.count > 2 ? 1 : 0));
    if (timed_x) CK(f, cudaEventRecord(f->ev_x[xs + 3], f->shard_stream));
    CK(f, cudaEventRecord(f->shard_ev_push, f->shard_stream));
  } else {
    shard_extract_cols_kernel<<<dim3((f->n_pad + 255) / 256, mc), 256, 0, f->stream>>>(f->sigma, f->lay, f->n_pad,
                                                                                    f->N, f->guess, mc, f->cols);
  }
  f->launches++;
  CK(f, cudaGetLastError());
  f->pend_pose = pose_dev;
  f->pend_m = mc; f->pend_xy = xy_dev; f->pend_ids = ids_dev;
  f->pend_finish = (last_round && mc == m) ? 1 : 0;
  *xchg_dev = f->pend_peer ? nullptr : f->cols;
  *xchg_doubles = f->pend_peer ? 0 : (int64_t)mc * f->n_pad * 2;
  return EKF_OK;
}

int ekf_shard_scan_finish(ekf_filter *f, int32_t *assoc_dev, int32_t *done_dev, double *pose_out_dev,
                          int32_t *count_dev) {
  return ekf_shard_scan_finish_ex(f, assoc_dev, done_dev, pose_out_dev, count_dev, nullptr, nullptr, 0, nullptr);
}

int ekf_shard_scan_finish_ex(ekf_filter *f, int32_t *assoc_dev, int32_t *done_dev, double *pose_out_dev,
                             int32_t *count_dev, const void *results_dev, void *results_host, int64_t results_bytes,
                             void *results_event) {
  if (!shard_ok(f) || f->pend_m < 1) return EKF_ERR_INVALID;
  if (results_bytes < 0 || (results_bytes > 0 && (!results_dev || !results_host))) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  int rc = ensure_io(f, 1);
  if (rc != EKF_OK) return rc;
  const ResLayout res(1);
  const int mc = f->pend_m;
  SeqParams sp{};
  sp.N = f->N; sp.n_pad = f->n_pad; sp.kmax = f->kmax; sp.lay = f->lay;
  sp.q_noise = f->cfg.q_noise; sp.r_noise = f->cfg.r_noise; sp.thr = f->cfg.mah_threshold;
  sp.sigma = f->sigma; sp.mu = f->mu; sp.Pp = f->Pp; sp.D = f->D; sp.W = f->W; sp.V = f->V;
  sp.cnt = f->cnt; sp.prev_xy = f->prev_xy; sp.status = f->status;
  sp.pose = f->pend_pose; sp.xy = f->pend_xy; sp.ids = f->pend_ids; sp.action = nullptr;
  sp.m = mc; sp.do_predict = f->pend_pose != nullptr ? 1 : 0; sp.finish = f->pend_finish;
  sp.assoc_out = assoc_dev;
  sp.pose_out = pose_out_dev ? pose_out_dev : reinterpret_cast<double *>(f->res_dev + res.pose);
  sp.cnt_out = count_dev ? count_dev : reinterpret_cast<int *>(f->res_dev + res.cnt);
  sp.side_mode = 1; sp.cols = f->cols; sp.guess = f->guess;
  if (f->pend_peer) {
    // every rank's part of this round must have landed in this rank's buffer before it is read
    sp.cols = reinterpret_cast<const double *>(f->xchg + (size_t)(f->xchg_round & 1) * f->xchg_cols_bytes);
    // ... element by element, as the sequential kernel reads them (no wait kernel, no flags on the critical path)
    sp.cols_poll = 1;
  }
  sp.done_out = done_dev ? done_dev : f->done_dev;
  sp.bar = f->bar; sp.bar_base = f->bar_base;
  sp.slot_d = f->slot_d; sp.slot_k = f->slot_k; sp.slot_rec = f->slot_rec;
  const bool hot = f->seq_hot && f->seq_threads == kSeqThreadsSmall;
  if (hot) { sp.hot_hdr = f->hot_hdr; sp.hot_rec = f->hot_rec; sp.hot_w = f->hot_w; }
  f->bar_base += (unsigned long long)mc * (unsigned long long)f->seq_grid;
  f->pend_m = -1;
  f->pend_pose = nullptr;

  const bool timed = f->ev_cap > 0;
  const size_t slot = timed ? (size_t)(f->ev_count % f->ev_cap) * 4 : 0;
  if (timed) CK(f, cudaEventRecord(f->ev[slot], f->stream));
  void *args[] = {(void *)&sp};
  const void *seq_fn = f->seq_threads == kSeqThreadsSmall ? (const void *)ekf_seq_kernel<kSeqThreadsSmall>
                                                          : (const void *)ekf_seq_kernel<kSeqThreadsLarge>;
  // The overlap needs a push CTA (128 threads x 104 registers) to fit on an SM BESIDE the resident sequential CTAs:
  // true for the 64-thread shape (2 x 64 x 244 registers), not for the 256-thread shape of large maps, whose CTAs
  // fill the register file - there the push kernel would wait for the sequential kernel that polls for its data.
  // Large maps therefore order the sequential kernel after the push (their passes take milliseconds; the exchange
  // is noise), and keep the cooperative launch.
  const bool overlap_push = f->pend_peer && f->seq_threads == kSeqThreadsSmall && env_int("EKF_SHARD_COOP_SEQ", 0) == 0;
  if (f->pend_peer && !overlap_push) CK(f, cudaStreamWaitEvent(f->stream, f->shard_ev_push, 0));
  if (overlap_push) {
    // PLAIN launch: a cooperative grid is held back until the device is otherwise idle, i.e. until the push kernel on
    // the shard stream has finished - the very thing it is meant to overlap.  Co-residency, all the hand-written grid
    // barrier needs, holds once the (short, never-waiting) push CTAs have left; the barrier's wait is bounded anyway:
    // a grid that never becomes resident latches EKF_ERR_CUDA (and reports done = -1) instead of hanging the GPU.
    sp.bar_timeout_ns = 12000000000ull;                      // above the 10 s bound of the peer wait inside
    if (f->seq_threads == kSeqThreadsSmall)
      ekf_seq_kernel<kSeqThreadsSmall><<<f->seq_grid, kSeqThreadsSmall, seq_smem_bytes(kSeqThreadsSmall, hot), f->stream>>>(sp);
    else
      ekf_seq_kernel<kSeqThreadsLarge><<<f->seq_grid, kSeqThreadsLarge, seq_smem_bytes(kSeqThreadsLarge, hot), f->stream>>>(sp);
    CK(f, cudaGetLastError());
  } else {
    CK(f, cudaLaunchCooperativeKernel(seq_fn, dim3(f->seq_grid), dim3(f->seq_threads), args,
                                      seq_smem_bytes(f->seq_threads, hot), f->stream));
  }
  f->launches++;
  if (timed) CK(f, cudaEventRecord(f->ev[slot + 1], f->stream));
  // the covariance pass rewrites Sigma in place: the push kernel of this round (shard stream) must have read its
  // columns first - the sequential kernel only waited for the elements it used
  const bool pushed = f->pend_peer != 0;
  // association / done / pose / count are final here: hand them to the host before the covariance pass
  if (results_bytes > 0)
    CK(f, cudaMemcpyAsync(results_host, results_dev, (size_t)results_bytes, cudaMemcpyDeviceToHost, f->stream));
  if (results_event) CK(f, cudaEventRecord((cudaEvent_t)results_event, f->stream));
  // A caller that takes the results early (results_event / results_host: the host-buffer loop) gets the covariance
  // pass on the shard stream, ordered after the sequential kernel: what it enqueues next on the handle's stream (the
  // next scan's input copies) then runs behind the pass.  A device-resident pipeline keeps everything on one stream:
  // the two cross-stream hops cost it ~10 us per scan and hide nothing.
  const bool side = (results_event != nullptr || results_bytes > 0) && env_int("EKF_SHARD_SIDE_PASS", 1) != 0;
  cudaStream_t pass_stream = f->stream;
  if (side) {
    CK(f, cudaEventRecord(f->shard_ev_seq, f->stream));
    CK(f, cudaStreamWaitEvent(f->shard_stream, f->shard_ev_seq, 0));
    pass_stream = f->shard_stream;
  } else if (pushed) {
    CK(f, cudaStreamWaitEvent(f->stream, f->shard_ev_push, 0));
  }
  DdParams dp{};
  dp.sigma = f->sigma; dp.V = f->V; dp.W = f->W;
  dp.n_pad = f->n_pad; dp.kmax = f->kmax; dp.k2 = (2 * mc + 3) & ~3;
  dp.lay = f->lay; dp.row_begin = f->lay.rb; dp.row_end = f->lay.re;
  dp.prefetch_ahead = f->dd_prefetch;
  dp.total_chunks = dd_total_chunks(64, dp.row_begin, dp.row_end);   // shard bounds are multiples of 128 rows
  long long grid = f->dd_grid;
  if (grid > dp.total_chunks) grid = dp.total_chunks;
  if (timed) CK(f, cudaEventRecord(f->ev[slot + 2], pass_stream));
  ekf_downdate_kernel<64, 2><<<(unsigned)grid, dd_threads(64), dd_smem_bytes(64, dp.k2), pass_stream>>>(dp);
  // Pp / D now hold the current pose panel / diagonal blocks; the Sigma entries they mirror (which the pass has
  // just recomputed with its own rounding) are written back lazily, by whatever reads or rewrites Sigma next
  f->side_dirty = true;
  f->launches += 1;
  CK(f, cudaGetLastError());
  if (timed) { CK(f, cudaEventRecord(f->ev[slot + 3], pass_stream)); f->ev_count++; }
  if (side) {
    CK(f, cudaEventRecord(f->shard_ev_pass, f->shard_stream));
    f->shard_pass_pending = true;
  }
  if (pushed) {
    // recycling of this round's column buffer, on the shard stream behind the sequential kernel (and behind the pass
    // when that runs there too): off the critical path
    if (!side) {
      CK(f, cudaEventRecord(f->shard_ev_seq, f->stream));
      CK(f, cudaStreamWaitEvent(f->shard_stream, f->shard_ev_seq, 0));
    }
    PeerTable pt{};
    pt.count = f->cfg.shard_count; pt.self = f->cfg.shard_rank;
    for (int q = 0; q < pt.count; ++q)
      pt.flags[q] = reinterpret_cast<unsigned long long *>(f->peer_base[q] + 2 * f->xchg_cols_bytes);
    shard_recycle_kernel<<<64, 256, 0, f->shard_stream>>>(
        reinterpret_cast<const unsigned long long *>(f->xchg + 2 * f->xchg_cols_bytes), pt, f->xchg_round,
        reinterpret_cast<double *>(f->xchg + (size_t)(f->xchg_round & 1) * f->xchg_cols_bytes),
        (int64_t)(f->xchg_cols_bytes / sizeof(double)), f->push_done + 1, f->status);
    f->launches++;
    CK(f, cudaGetLastError());
  }
  return EKF_OK;
}

}  // extern "C"
