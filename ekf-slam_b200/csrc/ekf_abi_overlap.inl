// Part of ekf_abi.cu: overlapped scans on one GPU (ekf_set_overlap).
//
// The classic path runs, per scan, the sequential kernel (latency-bound: one dependent chain per measurement) and
// then the covariance pass (streams the whole triangle) back to back on one stream.  The two use disjoint parts of
// an SM - the pass 288 threads x 168 registers + 143 KB, a 64-thread sequential CTA 244 registers + 51 KB - so here
// they run CONCURRENTLY: round k's pass on `back_stream`, round k+1's sequential side on `stream`.
//
//   * the pass is out of place: pass k reads covariance buffer k & 1 and writes the other one;
//   * the sequential side never reads the covariance itself (persistent Pp / D side state, like a row shard);
//     the raw columns of the guessed landmarks come from pipe_extract_cols_kernel = input of the pending pass minus
//     that pass's rank-2m update, which is what the pass is about to store;
//   * panels are double-buffered the same way (round k writes / pass k reads pair k & 1).
//
// Ordering (F = stream, B = back_stream):
//   F: wait pass k-2 | predict, guess, extract (reads buffer (k-1) & 1 and panels (k-1) & 1), seq k | record ev_seq
//   B: wait ev_seq k | pass k (buffer k & 1 -> (k+1) & 1, panels k & 1)                             | record ev_pass
// Buffer (k-1) & 1 is next written by pass k (after seq k), panels (k-1) & 1 by seq k+1 (after the wait for pass
// k-1), so nothing read on F can be overwritten underneath it.  Everything else the library does with the filter
// (accessors, fixtures, the classic path, the shard protocol) first DRAINS: waits for the last pass, writes Pp / D
// back into the covariance entries they mirror and makes `sigma` the current buffer again.

namespace {

int shard_buffers(ekf_filter *f) {
  if (!f->cols) CK(f, cudaMalloc(&f->cols, (size_t)kMaxChunkM * f->n_pad * 2 * sizeof(double)));
  if (!f->guess) CK(f, cudaMalloc(&f->guess, kMaxChunkM * sizeof(int)));
  if (!f->done_dev) CK(f, cudaMalloc(&f->done_dev, sizeof(int)));
  if (!f->shard_stream) {
    CK(f, cudaStreamCreateWithFlags(&f->shard_stream, cudaStreamNonBlocking));
    CK(f, cudaEventCreateWithFlags(&f->shard_ev_seq, cudaEventDisableTiming));
    CK(f, cudaEventCreateWithFlags(&f->shard_ev_pass, cudaEventDisableTiming));
    CK(f, cudaEventCreateWithFlags(&f->shard_ev_push, cudaEventDisableTiming));
    // the speculative association runs beside the covariance pass of the previous round: an SM changes its L1 /
    // shared-memory split only when idle, so it asks for the pass's carve-out
    CK(f, cudaFuncSetAttribute(shard_begin_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                               cudaSharedmemCarveoutMaxShared));
  }
  if (!f->guess_part_d) {
    CK(f, cudaMalloc(&f->guess_part_d, (size_t)kMaxChunkM * kGuessMaxSlices * sizeof(double)));
    CK(f, cudaMalloc(&f->guess_part_k, (size_t)kMaxChunkM * kGuessMaxSlices * sizeof(int)));
    CK(f, cudaMalloc(&f->guess_ticket, kMaxChunkM * sizeof(unsigned int)));
    CK(f, cudaMemsetAsync(f->guess_ticket, 0, kMaxChunkM * sizeof(unsigned int), f->stream));
  }
  return EKF_OK;
}

// Back to the single-buffer state every other entry point expects.  Cheap when nothing is in flight.
int pipe_drain(ekf_filter *f, bool keep_side = false) {
  // a covariance pass of the shard protocol still running on its own stream: order the handle's stream after it
  // (keep_side: the next shard round, which waits for it itself - after its speculative association)
  if (f->shard_pass_pending && !keep_side) {
    CK(f, cudaStreamWaitEvent(f->stream, f->shard_ev_pass, 0));
    f->shard_pass_pending = false;
  }
  // the shard protocol keeps the pose panel / diagonal blocks current in Pp / D only; everything that reads or
  // rewrites Sigma gets them written back first (keep_side: the next shard round, which does not read them there)
  if (f->side_dirty && !keep_side && !f->pipe_active) {
    shard_side_writeback_kernel<<<f->num_sms, 256, 0, f->stream>>>(f->sigma, f->lay, f->n_pad, f->N, f->Pp, f->D);
    f->launches++;
    CK(f, cudaGetLastError());
    f->side_dirty = false;
  }
  if (!f->pipe_active) return EKF_OK;
  const int64_t k = f->pipe_step;                       // rounds issued: passes 0 .. k-1 are queued on B
  CK(f, cudaStreamWaitEvent(f->stream, f->ev_pass[(k - 1) & 1], 0));
  if (k & 1) {                                          // the last pass wrote the second buffer
    double *t = f->sigma; f->sigma = f->sigma2; f->sigma2 = t;
  }
  // the covariance entries mirrored by the side state are stale in both buffers; the current one gets them now
  shard_side_writeback_kernel<<<f->num_sms, 256, 0, f->stream>>>(f->sigma, f->lay, f->n_pad, f->N, f->Pp, f->D);
  f->launches++;
  CK(f, cudaGetLastError());
  f->side_dirty = false;
  f->pipe_active = false;
  f->pipe_step = 0;
  f->pipe_k2_prev = 0;
  return EKF_OK;
}

// Allocations of the overlapped mode (second covariance buffer, second panel pair, stream, events).
void pipe_release(ekf_filter *f);

int pipe_resources_alloc(ekf_filter *f);

// idempotent; a partial failure releases what it had got, so that a later call starts from scratch
int pipe_resources(ekf_filter *f) {
  if (f->pipe_ready) return EKF_OK;
  const int rc = pipe_resources_alloc(f);
  if (rc != EKF_OK) { pipe_release(f); return rc; }
  f->pipe_ready = true;
  return EKF_OK;
}

int pipe_resources_alloc(ekf_filter *f) {
  const size_t panel_bytes = (size_t)f->n_pad * f->kmax * sizeof(double);
  cudaError_t e = cudaMalloc(&f->sigma2, (size_t)f->sigma_elems * sizeof(double));
  if (e != cudaSuccess) { cudaGetLastError(); f->sigma2 = nullptr; return EKF_ERR_NOMEM; }
  // the pass copies every stored element; what it never touches (nothing, today) would be zero
  CK(f, cudaMemsetAsync(f->sigma2, 0, (size_t)f->sigma_elems * sizeof(double), f->stream));
  CK(f, cudaMalloc(&f->W2, panel_bytes));
  CK(f, cudaMalloc(&f->V2, panel_bytes));
  CK(f, cudaMemsetAsync(f->W2, 0, panel_bytes, f->stream));
  CK(f, cudaMemsetAsync(f->V2, 0, panel_bytes, f->stream));
  CK(f, cudaStreamCreateWithFlags(&f->back_stream, cudaStreamNonBlocking));
  for (int q = 0; q < 2; ++q) {
    CK(f, cudaEventCreateWithFlags(&f->ev_seq[q], cudaEventDisableTiming));
    CK(f, cudaEventCreateWithFlags(&f->ev_pass[q], cudaEventDisableTiming));
  }
  CK(f, cudaEventCreateWithFlags(&f->done_event, cudaEventDisableTiming));
  CK(f, cudaMallocHost(&f->done_host, sizeof(int)));
  // a sequential CTA and the pass's CTA must fit on one SM together: both ask for the large shared-memory carve-out
  // ... and so must every other kernel of the sequential side: an SM only changes its L1 / shared-memory split when it
  // is idle, so a kernel that prefers another split than the resident pass waits for the pass to leave
  CK(f, cudaFuncSetAttribute(ekf_seq_kernel<kSeqThreadsSmall>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared));
  CK(f, cudaFuncSetAttribute(shard_predict_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared));
  CK(f, cudaFuncSetAttribute(shard_guess_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared));
  CK(f, cudaFuncSetAttribute(pipe_extract_cols_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared));
  CK(f, cudaFuncSetAttribute(ekf_downdate_kernel<64, 2>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared));
  int blocks = (f->N + kSeqThreadsSmall - 1) / kSeqThreadsSmall;
  f->pipe_grid = blocks < f->num_sms ? blocks : f->num_sms;   // one 64-thread CTA per SM beside the pass
  // the arg-min slot arrays (and the one-warp slot scan) are sized for the classic grid: never launch more CTAs
  // than they hold, whatever the SM count of the device
  if (f->pipe_grid > f->slot_cap) f->pipe_grid = f->slot_cap;
  if (f->pipe_grid < 1) f->pipe_grid = 1;
  // residency check for the plain launch: the kernel alone must fit pipe_grid CTAs (one per SM at least); the bounded
  // barrier covers what this cannot see (the pass and other clients of the device)
  int occ = 0;
  CK(f, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ekf_seq_kernel<kSeqThreadsSmall>, kSeqThreadsSmall,
                                                      seq_smem_bytes(kSeqThreadsSmall)));
  if (occ < 1) return EKF_ERR_INVALID;
  return shard_buffers(f);
}

void pipe_release(ekf_filter *f) {
  if (f->back_stream) cudaStreamSynchronize(f->back_stream);
  cudaFree(f->sigma2); cudaFree(f->W2); cudaFree(f->V2);
  for (int q = 0; q < 2; ++q) {
    if (f->ev_seq[q]) cudaEventDestroy(f->ev_seq[q]);
    if (f->ev_pass[q]) cudaEventDestroy(f->ev_pass[q]);
  }
  if (f->done_event) cudaEventDestroy(f->done_event);
  if (f->done_host) cudaFreeHost(f->done_host);
  if (f->back_stream) cudaStreamDestroy(f->back_stream);
  f->sigma2 = nullptr; f->W2 = nullptr; f->V2 = nullptr; f->back_stream = nullptr; f->done_host = nullptr;
  f->done_event = nullptr;
  for (int q = 0; q < 2; ++q) { f->ev_seq[q] = nullptr; f->ev_pass[q] = nullptr; }
  f->pipe_ready = false;
}

}  // namespace
