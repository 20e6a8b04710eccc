// C ABI of the B200-native EKF-SLAM core (include/ekf_slam_b200.h): host-side orchestration of
// the sm_100a kernels.  No CPU fallback exists anywhere in this file: every entry point drives
// CUDA kernels and reports EKF_ERR_CUDA / EKF_ERR_NO_DEVICE when it cannot.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/ekf_slam_b200.h"
#include "ekf_batch.cuh"
#include "ekf_batch_reg.cuh"
#include "ekf_downdate.cuh"
#include "ekf_landmarks.cuh"
#include "ekf_layout.h"
#include "ekf_nusim.cuh"
#include "ekf_seq.cuh"
#include "ekf_shard.cuh"
#include "ekf_util.cuh"

using namespace ekf;

struct ekf_filter {
  ekf_config cfg{};
  int device = 0;
  int num_sms = 0;
  int N = 0, n = 0, n_int = 0, n_pad = 0, kmax = kMaxK2;
  int64_t sigma_elems = 0;
  EkfLay lay{0, 0};        // rows of Sigma this handle stores (whole matrix unless row-sharded)
  bool sharded = false;
  bool side_valid = false; // sharded: Pp / D replicated and current
  bool side_dirty = false; // sharded: Pp / D are NEWER than the Sigma entries they mirror (written back lazily)
  const double *pend_pose = nullptr;   // the round opened by ekf_shard_scan_begin predicts from this pose (or nullptr)
  // shard protocol: the covariance pass of a round runs on its own stream, so that what the caller enqueues next on
  // the handle's stream (input copies, the next round's speculative association) is ordered after the SEQUENTIAL
  // kernel only and runs behind the pass; everything that touches Sigma or the panels waits for shard_ev_pass
  cudaStream_t shard_stream = nullptr;
  cudaEvent_t shard_ev_seq = nullptr, shard_ev_pass = nullptr, shard_ev_push = nullptr;
  bool shard_pass_pending = false;
  std::vector<cudaEvent_t> ev_x;   // timing ring of the shard begin phase: association start / end, push start / end
  double *cols = nullptr;  // sharded: exchange buffer, kMaxChunkM x n_pad x 2
  int *guess = nullptr;    // sharded: speculative association of the round
  int *done_dev = nullptr; // sharded: measurements processed by the last finish
  double *guess_part_d = nullptr;        // sharded: per-slice partial candidates of the speculative association
  int *guess_part_k = nullptr;
  unsigned int *guess_ticket = nullptr;
  // sharded, peer-memory exchange (ekf_shard_peer_*): one allocation = [cols round-parity 0][cols parity 1][flags]
  unsigned char *xchg = nullptr;
  size_t xchg_cols_bytes = 0;             // bytes of ONE column buffer
  unsigned char *peer_base[16] = {};      // the same region of every rank (own entry = xchg), peer-mapped
  void *peer_ipc[16] = {};                // regions opened with cudaIpcOpenMemHandle (closed at destroy)
  int peers_attached = 0;
  unsigned long long xchg_round = 0;      // rounds pushed so far
  unsigned int *push_done = nullptr;      // last-CTA counter of the push kernel
  int pend_peer = 0;                      // the pending round was exchanged through peer memory
  // the round opened by ekf_shard_scan_begin
  int pend_m = -1, pend_finish = 0;
  const double *pend_xy = nullptr;
  const int32_t *pend_ids = nullptr;
  // device storage
  double *sigma = nullptr, *mu = nullptr, *W = nullptr, *V = nullptr;
  double4 *Pp = nullptr, *D = nullptr;
  int *cnt = nullptr, *status = nullptr;
  double *prev_xy = nullptr;
  unsigned long long *bar = nullptr;
  unsigned long long bar_base = 0;
  double *slot_d = nullptr, *slot_rec = nullptr;
  int *slot_k = nullptr;
  // hot-winner area of the sequential kernel (SeqParams::hot_hdr); classic path, 64-thread shape only
  unsigned long long *hot_hdr = nullptr;
  double *hot_rec = nullptr;
  double2 *hot_w = nullptr;
  bool seq_hot = false;
  // staging for the host-pointer entry points
  unsigned char *io_dev = nullptr, *io_host = nullptr;   // inputs: pose | xy | ids | action
  unsigned char *res_dev = nullptr, *res_host = nullptr; // results: pose | cnt | status | assoc
  unsigned char *res_host_dev = nullptr;   // device-side address of res_host (mapped pinned memory)
  bool res_direct = false;                 // this scan's kernels write their results straight into res_host
  // host-buffer scans hand their results back as soon as the sequential phase is done: the copy is
  // enqueued between the last sequential launch and its covariance pass, the caller waits on this event
  cudaEvent_t res_event = nullptr;
  size_t res_pending = 0;    // bytes of results the next finishing sequential launch should copy out (0 = none)
  size_t io_cap = 0, res_cap = 0;
  double *scratch_dev = nullptr;  // block get/set staging
  size_t scratch_cap = 0;
  // streams / events / bookkeeping
  cudaStream_t own_stream = nullptr, stream = nullptr;
  // optional per-scan timing: a ring of (start, after sequential phase, after covariance pass) events
  std::vector<cudaEvent_t> ev;
  int ev_cap = 0;          // scans the ring holds; 0 = timing off
  int64_t ev_count = 0;    // scans recorded since timing was enabled
  int64_t launches = 0;
  int seq_grid = 1;
  int seq_threads = 64;    // shape of the sequential kernel (ekf_seq.cuh)
  int dd_grid = 1;
  int dd_prefetch = 0;
  int dd_ctas = 2;
  int dd_variant = 2;      // 0: <64 rows, register prefetch>, 1 CTA / SM;  1: <32 rows>, 3 CTAs / SM
  int latched = EKF_OK;
  std::string cuda_err;
  // overlapped scans (ekf_set_overlap, ekf_abi_overlap.inl): the covariance pass of scan s runs out of place on a
  // second stream while the sequential side of scan s+1 runs on `stream`
  bool overlap = false;          // requested
  bool pipe_active = false;      // rounds issued since the last drain
  bool pipe_ready = false;       // every allocation of the mode is in place
  double *sigma2 = nullptr;      // the other covariance buffer (pass k reads buffer k & 1, writes the other)
  double *W2 = nullptr, *V2 = nullptr;   // the other panel pair (round k writes / pass k reads pair k & 1)
  cudaStream_t back_stream = nullptr;
  cudaEvent_t ev_seq[2] = {nullptr, nullptr}, ev_pass[2] = {nullptr, nullptr};
  int64_t pipe_step = 0;         // rounds issued since the pipeline was entered
  int pipe_k2_prev = 0;          // panel columns of the previous round's pass
  int slot_cap = 1;              // CTAs the arg-min slot arrays hold (<= 32 * kMaxSlotsPerLane)
  int pipe_grid = 1;             // CTAs of the sequential kernel that fit beside the pass (<= one per SM)
  int *done_host = nullptr;      // pinned: measurements the last round processed
  cudaEvent_t done_event = nullptr;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int cuda_fail(ekf_filter *f, cudaError_t e, const char *what) {
  if (f) {
    f->cuda_err = std::string(what) + ": " + cudaGetErrorString(e);
    f->latched = EKF_ERR_CUDA;
  }
  return EKF_ERR_CUDA;
}

#define CK(f, call)                                             \
  do {                                                          \
    cudaError_t e_ = (call);                                    \
    if (e_ != cudaSuccess) return cuda_fail((f), e_, #call);    \
  } while (0)

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// layout of the packed input / result staging buffers for a scan of m measurements
struct IoLayout {
  size_t pose, xy, ids, action, total;
  explicit IoLayout(int m) {
    pose = 0; xy = 24; ids = xy + (size_t)16 * m; action = ids + (size_t)4 * m;
    total = align_up(action + (size_t)m, 16);
  }
};
struct ResLayout {
  size_t pose, cnt, status, assoc, total;
  explicit ResLayout(int m) {
    pose = 0; cnt = 24; status = 28; assoc = 32; total = align_up(assoc + (size_t)4 * m, 16);
  }
};

int ensure_io(ekf_filter *f, int m) {
  const IoLayout in(m);
  const ResLayout res(m);
  if (in.total > f->io_cap) {
    if (f->io_dev) cudaFree(f->io_dev);
    if (f->io_host) cudaFreeHost(f->io_host);
    f->io_dev = nullptr; f->io_host = nullptr; f->io_cap = 0;
    const size_t cap = align_up(in.total * 2, 256);
    CK(f, cudaMalloc(&f->io_dev, cap));
    CK(f, cudaMallocHost(&f->io_host, cap));
    f->io_cap = cap;
  }
  if (res.total > f->res_cap) {
    if (f->res_dev) cudaFree(f->res_dev);
    if (f->res_host) cudaFreeHost(f->res_host);
    f->res_dev = nullptr; f->res_host = nullptr; f->res_host_dev = nullptr; f->res_cap = 0;
    const size_t cap = align_up(res.total * 2, 256);
    CK(f, cudaMalloc(&f->res_dev, cap));
    CK(f, cudaMallocHost(&f->res_host, cap));
    void *mapped = nullptr;
    f->res_host_dev = cudaHostGetDevicePointer(&mapped, f->res_host, 0) == cudaSuccess ? (unsigned char *)mapped : nullptr;
    if (!f->res_host_dev) cudaGetLastError();
    f->res_cap = cap;
  }
  return EKF_OK;
}

int ensure_scratch(ekf_filter *f, size_t bytes) {
  if (bytes > f->scratch_cap) {
    if (f->scratch_dev) cudaFree(f->scratch_dev);
    f->scratch_dev = nullptr; f->scratch_cap = 0;
    CK(f, cudaMalloc(&f->scratch_dev, bytes));
    f->scratch_cap = bytes;
  }
  return EKF_OK;
}

int env_int(const char *name, int dflt) {
  const char *v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

}  // namespace

#include "ekf_abi_overlap.inl"
#include "ekf_abi_lifetime.inl"
#include "ekf_abi_scan.inl"
#include "ekf_abi_access.inl"
#include "ekf_abi_noalloc.inl"
#include "ekf_abi_batch.inl"
#include "ekf_abi_shard.inl"
#include "ekf_abi_landmarks.inl"
#include "ekf_abi_nusim.inl"
#include "ekf_abi_bag.inl"
