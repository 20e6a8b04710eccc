// Sequential phase of one scan: prediction + m x (associate, gain, state update) with every
// covariance rank-2 downdate DEFERRED into the W / V panels.
//
// Replaces, for one sensor callback of the reference (nuslam/src/slam.cpp):
//   :186-198 / :323-335   pose injection and Sigma <- A Sigma A^T + Qbar  (only the n x 3 pose
//                         panel of Sigma changes; two dense n^3 products in the reference)
//   :345-356              polar conversion, tentative landmark
//   :358-419              Mahalanobis distance to every mapped landmark + index_min
//   :421-440              new-landmark decision
//   :443-482,488          expected measurement, Jacobian, gain K, state update
// The covariance update (:264-265 / :485-486, `sigma = (I - K*Hj) * sigma`) is NOT applied
// here: for measurement t the kernel stores G_t = Sigma_cur H_t^T (n x 2) in panel W and
// K_t = G_t S_t^-1 in panel V, so that after the scan
//       Sigma_new = Sigma_0 - sum_t K_t G_t^T            (== V W^T, rank 2m)
// which ekf_downdate.cuh applies in ONE pass over the stored triangle.  While the scan runs,
// the entries of the *current* covariance that later measurements need are kept up to date
// without touching Sigma_0:
//   * the n x 3 pose panel  Pp  and the N 2x2 diagonal blocks  D  (all association needs),
//     updated by O(n) work per measurement;
//   * the two landmark columns of the associated landmark, gathered from Sigma_0 and corrected
//     with the panels written so far:  Sigma_cur[:,c] = Sigma_0[:,c] - sum_{s<t} K_s G_s[c,:]^T.
//
// Parallel decomposition: a cooperative grid; thread g owns landmarks g, g+T, ... (their two
// state entries, pose-panel rows, diagonal block, panel rows).  The 3 pose rows are tracked
// redundantly by every CTA in shared memory.  Exactly ONE grid barrier per measurement.  The
// phase is latency-bound (one dependent chain per measurement), so the chain is kept short:
//   * association already forms, for every candidate, the linearisation, S and S^-1 the update
//     needs; each CTA publishes its best candidate TOGETHER with that record (22 doubles) in its
//     slot before the barrier, so after the barrier nothing is recomputed serially: every warp
//     reduces the slots (one L2 round trip), then warp 0 fetches the winner's record while warp 1
//     fetches the winner's panel-row history (a second, shared L2 round trip);
//   * the gather of the two landmark columns from Sigma_0 (HBM latency) and the panel-history
//     loads of the correction are issued together, 256 bits at a time;
//   * a new landmark / known association takes the (rare) serial path through thread 0.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_layout.h"
#include "ekf_math.cuh"

namespace ekf {

// Two shapes of the sequential kernel (template parameter NT = threads per CTA), chosen per handle:
//   NT =  64, up to 2 CTAs / SM: every thread owns ONE landmark while N <= 160 * 64; the phase is a chain
//            of latencies, and small CTAs keep the CTA-level reductions and barriers short
//   NT = 256, 1 CTA / SM: larger maps, where a thread owns several landmarks and throughput matters
constexpr int kSeqThreadsSmall = 64, kSeqThreadsLarge = 256;
constexpr int kSeqCtasPerSm = 2;          // cooperative grid: at most this many CTAs per SM (NT = 64)
#ifndef EKF_SEQ_MINB
#define EKF_SEQ_MINB 1
#endif
constexpr int kMaxChunkM = 24;            // measurements per sequential launch / covariance pass
constexpr int kMaxK2 = 2 * kMaxChunkM;    // panel columns in use per pass
// dynamic shared memory of ekf_seq_kernel: the own-landmark V rows, [kMaxChunkM][2][NT] double2
// (+ the same for the W rows when the launch uses the hot-winner area, see SeqParams::hot_hdr)
constexpr size_t seq_smem_bytes(int nt, bool hot = false) {
  return (size_t)kMaxChunkM * 2 * nt * sizeof(double2) * (hot ? 2 : 1);
}

// status codes latched on the device (values of ekf_status in include/ekf_slam_b200.h)
constexpr int kErrCapacity = 2;
constexpr int kErrBadId = 4;
constexpr int kErrCuda = 5;

struct SeqParams {
  // geometry / constants
  int N, n_pad, kmax;
  double q_noise, r_noise, thr;
  // persistent filter storage
  double *sigma;   // column strips (ekf_layout.h)
  double *mu;      // n_pad, internal indexing
  double4 *Pp;     // n_pad rows: (Sigma_cur[i,theta], [i,x], [i,y], 0)
  double4 *D;      // N: (Sigma_cur[a,a], [b,a], [b,b], 0)
  double *W, *V;   // panels, blocked (ekf_panel_elem)
  int *cnt;        // counter_obstacles
  double *prev_xy; // filter_previous_configuration translation
  int *status;     // latched error
  // this call
  const double *pose;     // device (theta, x, y) or nullptr when do_predict == 0
  const double *xy;       // device 2m
  const int *ids;         // device m (known association) or nullptr (unknown)
  const uint8_t *action;  // device m or nullptr
  int m;
  int do_predict;
  int finish;             // 1: last chunk of the scan -> latch filter_previous_configuration
  int *assoc_out;         // device m or nullptr
  double *pose_out;       // device 3 (always written)
  int *cnt_out;           // device 1 (always written)
  int *status_out;        // nullptr, or where the finishing launch hands over (and clears) the latched status
  // storage geometry of this handle's rows (whole matrix, or one row shard)
  EkfLay lay;
  // row-sharded handles (ekf_shard.cuh): side_mode = 1 means Pp / D are persistent and current (Sigma
  // is not read at all here), `cols` holds the all-reduced raw Sigma_0 columns of the guessed
  // landmarks, [slot][row][2]; guess[s] = landmark whose columns slot s holds (-1 = none)
  int side_mode;
  const double *cols;
  const int *guess;
  int *done_out;          // measurements actually processed (a failed speculation stops the scan early)
  // peer-memory exchange: `cols` is being filled by every rank's shard_push_cols_kernel WHILE this kernel runs; an
  // element has arrived when neither of its halves is kColsSentinel any more (the buffer was poisoned before it was
  // handed out).  0: `cols` is complete before the launch (caller's all-reduce, overlapped single-GPU mode).
  int cols_poll;
  // grid barrier + arg-min slots
  unsigned long long *bar;
  unsigned long long bar_base;
  unsigned long long bar_timeout_ns;   // 0: cooperative launch, wait for ever; > 0: plain launch, bounded wait
  // every CTA publishes its best candidate kSlotReps times; CTA b reads replica b % kSlotReps.  A single
  // copy made ~300 warps read the same few L2 lines right after the barrier, and the L2 slice serving
  // a line handles those requests one after the other (3.6 us per measurement, tools/seq_phase_probe.py)
  double *slot_d;  // [kSlotReps][2][gridDim]
  int *slot_k;     // [kSlotReps][2][gridDim]
  double *slot_rec;  // [kSlotReps][2][gridDim][kRecStride]: the slot candidate's update record
  // Hot-winner area (nullptr = off).  After the barrier the decision needs the slots (one L2 round trip) and THEN the
  // winner's record and panel-row history (a second, dependent one).  A CTA whose best candidate is under the
  // threshold - normally exactly one CTA per matched sighting - therefore also pushes that record and its W-row
  // history here before the barrier, and the readers load this area TOGETHER with the slots.  The header counts the
  // claimants of this measurement (tag = the barrier target, unique and increasing): the preloaded copy is used
  // only when there is exactly one and it is the arg-min winner; anything else takes the two-round-trip path.
  unsigned long long *hot_hdr;   // [kSlotReps][2]: tag * 256 + claimants
  double *hot_rec;               // [kSlotReps][2][kRecStride]: record, [kRecDoubles] = landmark index
  double2 *hot_w;                // [kSlotReps][2][kMaxK2]: (W[a][q], W[a+1][q]) of the claimant's landmark
};
constexpr int kSlotReps = 8;
constexpr int kMaxSlotsPerLane = 5;       // gridDim <= 32 * kMaxSlotsPerLane (one CTA per SM, 148 SMs)

// update record of one (measurement, landmark) pair: everything the row update needs
//   [0..3] h0[1..4]  [4..7] h1[1..4]  [8..13] T[0..2][0..1] (pose rows of G)  [14..16] S^-1
//   [17..18] innovation  [19..21] the landmark's current diagonal block (aa, ba, bb)
constexpr int kRecDoubles = 22;
constexpr int kRecStride = 24;

// One arrival counter for the whole grid.  Arrival = release add (orders the CTA's earlier stores - its slot, its
// hot-winner push - before the count; the CTA's own threads are ordered before it by the bar.sync above), wait =
// acquire loads.  SASS: MEMBAR.ALL.GPU + REDG (no return value) / LDG.STRONG.GPU + CCTL.IVALL per poll.  The first
// version bracketed an ATOMG with two __threadfence() (MEMBAR.SC.GPU each, the second one on the critical path
// right after the last arrival): -DEKF_SEQ_FENCE_BARRIER keeps it for A/B runs (tools/seq_soak.py).
__device__ __forceinline__ void grid_barrier(unsigned long long *bar, unsigned long long target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long v;
#ifdef EKF_SEQ_FENCE_BARRIER
    __threadfence();
    atomicAdd(bar, 1ULL);
    do {
      asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(bar) : "memory");
    } while (v < target);
    __threadfence();
#else
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(bar), "l"(1ULL) : "memory");
    do {
      asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(bar) : "memory");
    } while (v < target);
#endif
  }
  __syncthreads();
}

// The same barrier for launches that are NOT cooperative (the overlapped path starts the kernel plainly so that it
// may run beside the covariance pass): nothing guarantees that every CTA is resident, so the wait is bounded.
// Returns true (to every thread of the CTA) when the count did not arrive within timeout_ns; the caller latches
// EKF_ERR_CUDA and leaves the kernel instead of hanging the device.
__device__ __forceinline__ bool grid_barrier_bounded(unsigned long long *bar, unsigned long long target,
                                                     unsigned long long timeout_ns) {
  __syncthreads();
  int late = 0;
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(bar), "l"(1ULL) : "memory");
    unsigned long long v, t0 = 0;
    unsigned polls = 0;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(bar) : "memory");
      if (v >= target) break;
      if ((++polls & 1023u) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > timeout_ns) { late = 1; break; }
      }
    }
  }
  return __syncthreads_or(late) != 0;
}

__device__ __forceinline__ double ldcg_d(const double *p) { return __ldcg(p); }
__device__ __forceinline__ double4 ldcg_d4(const double4 *p) {
  double4 v;
  asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(v.z), "=d"(v.w) : "l"((const double *)p + 2));
  return v;
}

// Flag-in-data arrival of the exchanged columns (row-sharded path, ekf_shard.cuh): a column buffer is filled with this
// pattern before it is handed out - a NaN no arithmetic produces (the hardware's NaNs are canonical) - and the owner's
// stores replace it.  Both 32-bit halves are equal so that any fill granularity works.
constexpr unsigned long long kColsSentinel = 0xFFF1FFF1FFF1FFF1ull;
// the four doubles (rows a, a + 1 of one slot) at src, once they have arrived; bounded (a peer that never delivers
// latches EKF_ERR_CUDA after ~10 s and lets the kernel run on with the NaNs instead of hanging the GPU)
__device__ __forceinline__ void poll_cols(const double *src, double2 &r0, double2 &r1, int *status) {
  unsigned long long b0, b1, b2, b3, t0 = 0;
  unsigned polls = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(b0), "=l"(b1) : "l"(src) : "memory");
    asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(b2), "=l"(b3) : "l"(src + 2) : "memory");
    if (b0 != kColsSentinel && b1 != kColsSentinel && b2 != kColsSentinel && b3 != kColsSentinel) break;
    if ((++polls & 255u) == 0) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 >= 10000000000ull) { atomicMax(status, 5 /* EKF_ERR_CUDA */); break; }
    }
  }
  r0 = make_double2(__longlong_as_double((long long)b0), __longlong_as_double((long long)b1));
  r1 = make_double2(__longlong_as_double((long long)b2), __longlong_as_double((long long)b3));
}

// Shared per-measurement broadcast: the chosen landmark and its update record.
struct Bcast {
  int j;        // associated landmark, or -1 when the measurement is skipped
  int is_new;   // the landmark's state is (re)initialised from the sighting
  int slot;     // sharded: column slot of landmark j (-1: fresh landmark, columns are zero; -2: missing)
  double lx, ly;             // landmark j position used when is_new
  double rec[kRecStride];    // update record (layout above)
};

__device__ __forceinline__ double ld_plain256(const double *p, double &a, double &b, double &c) {
  double d;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p) : "memory");
  return d;
}

// -DEKF_SEQ_PROF (tools/seq_phase_probe.py only): thread 0 of the first and the last CTA leave clock64()
// stamps of the phase boundaries behind the association output, which the probe over-allocates
#ifdef EKF_SEQ_PROF
#define SEQ_STAMP(ph)                                                                                   \
  do {                                                                                                  \
    if (tid == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) && p.assoc_out != nullptr)         \
      reinterpret_cast<long long *>(p.assoc_out + 32)[((blockIdx.x != 0) * kMaxChunkM + t) * 8 + (ph)] = clock64(); \
  } while (0)
#else
#define SEQ_STAMP(ph) do { } while (0)
#endif

template <int NT>
__global__ void __launch_bounds__(NT, EKF_SEQ_MINB) ekf_seq_kernel(const SeqParams p) {
  constexpr int kSeqThreads = NT;
  __shared__ double s_ps[3];      // pose state (theta, x, y)
  __shared__ double s_pb[3][3];   // pose block of Sigma_cur, full symmetric
  __shared__ Cand s_red[kSeqThreads / 32];
  __shared__ double s_wrec[kSeqThreads / 32][kRecStride];   // each warp's best candidate record
  __shared__ Bcast s_b;
  __shared__ double s_wa[kMaxK2], s_wb[kMaxK2];
  __shared__ int s_cnt;
  __shared__ int s_cnt0;          // counter_obstacles when the launch began
  __shared__ int s_fast;          // this measurement matched an existing landmark (decision of warp 0)
  __shared__ double s_pred[2];    // (-dy, dx)
  __shared__ double s_pb0[3][3];  // pose block right after the prediction (what Sigma_0 holds)
  __shared__ double s_zr[kMaxChunkM], s_zphi[kMaxChunkM];   // polar form of the scan's sightings
  // V rows (a, a+1) of the thread's FIRST landmark, [measurement][row][thread] as (column 2t, 2t+1)
  // pairs: the deferred-downdate correction of phase D reads them back every measurement
  extern __shared__ double2 s_vown[];
  // the same for its W rows (hot-winner launches only: the claimant pushes this history, see SeqParams::hot_hdr)
  const bool hot = p.hot_hdr != nullptr;
  double2 *const s_wown = s_vown + kMaxChunkM * 2 * kSeqThreads;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int nthreads_total = gridDim.x * blockDim.x;
  const int gtid = blockIdx.x * blockDim.x + tid;
  const int n_pad = p.n_pad;
  const int nks = p.kmax >> 2;     // k-steps per panel block
  double *const strip0 = p.sigma;  // strip 0 starts at offset 0: row i at i*128

  // ---------------- phase 0: pose injection, prediction, extraction of Pp / D -------------
  if (tid == 0) {
    s_cnt = *p.cnt;
    s_cnt0 = s_cnt;
    double P[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j)
        P[i][j] = P[j][i] = p.side_mode ? (j == 0 ? p.Pp[i].x : j == 1 ? p.Pp[i].y : p.Pp[i].z)
                                        : strip0[i * EKF_STRIP + j];
    if (p.do_predict) {
      const double th = normalize_angle(p.pose[0]);          // slam.cpp:186
      const double x = p.pose[1], y = p.pose[2];             // slam.cpp:187-188
      const double dx = x - p.prev_xy[0], dy = y - p.prev_xy[1];  // slam.cpp:190-191
      s_ps[0] = th; s_ps[1] = x; s_ps[2] = y;
      s_pred[0] = -dy; s_pred[1] = dx;
      // B = A P (rows 1,2 += (-dy, dx) * row 0), C = B A^T (cols likewise): slam.cpp:194-198
      double Bm[3][3];
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        Bm[0][j] = P[0][j];
        Bm[1][j] = (-dy) * P[0][j] + P[1][j];
        Bm[2][j] = dx * P[0][j] + P[2][j];
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        P[i][0] = Bm[i][0];
        P[i][1] = Bm[i][0] * (-dy) + Bm[i][1];
        P[i][2] = Bm[i][0] * dx + Bm[i][2];
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) P[i][i] += p.q_noise;      // slam.cpp:135-136
    } else {
      s_ps[0] = p.mu[0]; s_ps[1] = p.mu[1]; s_ps[2] = p.mu[2];
      s_pred[0] = 0.0; s_pred[1] = 0.0;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) s_pb[i][j] = s_pb[j][i] = P[i][j];  // lower triangle rules
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) s_pb0[i][j] = s_pb[i][j];
    // the pose block / pose state in global memory are rewritten in the epilogue, AFTER a grid
    // barrier: every CTA reads the old values above, in no particular order
  }
  __syncthreads();
  // The thread's first landmark (k = gtid; the only one when N <= grid threads) lives in registers for
  // the whole launch: pose-panel rows, diagonal block, position.  Global copies are still written after
  // every update (thread 0's serial path and the next launch read them) but never re-read by the owner.
  double4 o_pa = make_double4(0.0, 0.0, 0.0, 0.0), o_pb = o_pa, o_dk = o_pa;
  double2 o_lm = make_double2(0.0, 0.0);
  if (p.side_mode && p.do_predict) {
    // the prediction on the replicated pose panel (no separate kernel on the sharded path): slam.cpp:194-198
    const double ndy = s_pred[0], pdx = s_pred[1];
    for (int k = gtid; k < p.N; k += nthreads_total) {
      const int a = 4 + 2 * k;
      double4 va = p.Pp[a], vb = p.Pp[a + 1];
      va.y = va.x * ndy + va.y; va.z = va.x * pdx + va.z;
      vb.y = vb.x * ndy + vb.y; vb.z = vb.x * pdx + vb.z;
      p.Pp[a] = va; p.Pp[a + 1] = vb;
      if (k == gtid) { o_pa = va; o_pb = vb; }
    }
    if (gtid < p.N) o_dk = p.D[gtid];
  } else if (p.side_mode && gtid < p.N) {
    o_pa = p.Pp[4 + 2 * gtid]; o_pb = p.Pp[5 + 2 * gtid]; o_dk = p.D[gtid];
  }
  if (gtid < p.N) o_lm = *reinterpret_cast<const double2 *>(p.mu + 4 + 2 * gtid);
  if (!p.side_mode) {
    const double ndy = s_pred[0], pdx = s_pred[1];
    for (int k = gtid; k < p.N; k += nthreads_total) {
      const int a = 4 + 2 * k;
      double4 *ra = reinterpret_cast<double4 *>(strip0 + (int64_t)a * EKF_STRIP);
      double4 *rb = reinterpret_cast<double4 *>(strip0 + (int64_t)(a + 1) * EKF_STRIP);
      double4 va = *ra, vb = *rb;
      if (p.do_predict) {
        va.y = va.x * ndy + va.y; va.z = va.x * pdx + va.z;
        vb.y = vb.x * ndy + vb.y; vb.z = vb.x * pdx + vb.z;
        va.w = 0.0; vb.w = 0.0;
        *ra = va; *rb = vb;
      }
      p.Pp[a] = va; p.Pp[a + 1] = vb;
      const double *dg = p.sigma + ekf_lay_elem(p.lay, a, a);
      const double2 d0 = *reinterpret_cast<const double2 *>(dg);                // (a,a), (a,b) scratch
      const double2 d1 = *reinterpret_cast<const double2 *>(dg + EKF_STRIP);    // (b,a), (b,b)
      p.D[k] = make_double4(d0.x, d1.x, d1.y, 0.0);
      if (k == gtid) { o_pa = va; o_pb = vb; o_dk = make_double4(d0.x, d1.x, d1.y, 0.0); }
    }
  }

  // ---------------- measurement loop --------------------------------------------------------
  const bool unknown = (p.ids == nullptr);
  const int G = (int)gridDim.x;
  int done = p.m;
  // slam.cpp:345-347 for every sighting at once (it does not depend on the state): keeps sqrt / atan2
  // out of the per-measurement dependent chain
  if (tid < p.m) {
    const Polar zt = to_polar(p.xy[2 * tid], p.xy[2 * tid + 1]);
    s_zr[tid] = zt.r; s_zphi[tid] = zt.phi;
  }
  __syncthreads();
  for (int t = 0; t < p.m; ++t) {
    const Polar z{s_zr[t], s_zphi[t]};
    const double th = s_ps[0], px = s_ps[1], py = s_ps[2];
    const int cnt = s_cnt;
    const int par = t & 1;
    SEQ_STAMP(0);

    // ---- A: Mahalanobis distance to every mapped landmark (own ones); the best candidate keeps
    //         its update record in registers --------------------------------------------------------
    if (unknown) {
      Cand best{INFINITY, 0x7fffffff};
      double rec[kRecDoubles];
#pragma unroll
      for (int i = 0; i < kRecDoubles; ++i) rec[i] = 0.0;
      for (int k = gtid; k < cnt; k += nthreads_total) {
        const int a = 4 + 2 * k;
        double2 lm = o_lm;
        double4 pa = o_pa, pbr = o_pb, dk = o_dk;
        if (k != gtid) {
          lm = *reinterpret_cast<const double2 *>(p.mu + a);
          pa = p.Pp[a]; pbr = p.Pp[a + 1];
          dk = p.D[k];
        }
        const Lin L = linearize(th, px, py, lm.x, lm.y);     // slam.cpp:364-392
        const double par3[3] = {pa.x, pa.y, pa.z}, pbr3[3] = {pbr.x, pbr.y, pbr.z};
        const Block5 B = make_block5(s_pb, par3, pbr3, dk.x, dk.y, dk.z);
        const Innov I = innovation(B, L, p.r_noise);         // slam.cpp:395
        const double zd0 = z.r - L.zhat_r;
        const double zd1 = normalize_angle(z.phi - L.zhat_b);  // slam.cpp:398-399
        double i00, i01, i11;
        const double d = mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, i00, i01, i11);  // slam.cpp:401
        const Cand c{d, k};
        const Cand nb = cand_min(best, c);
        if (nb.k == k) {                                     // this candidate is the thread's best so far
#pragma unroll
          for (int q = 0; q < 4; ++q) { rec[q] = L.h0[q + 1]; rec[4 + q] = L.h1[q + 1]; }
#pragma unroll
          for (int q = 0; q < 3; ++q) { rec[8 + 2 * q] = I.T[q][0]; rec[9 + 2 * q] = I.T[q][1]; }
          rec[14] = i00; rec[15] = i01; rec[16] = i11;
          rec[17] = zd0; rec[18] = zd1;
          rec[19] = dk.x; rec[20] = dk.y; rec[21] = dk.z;
        }
        best = nb;
      }
      int wsrc;
      const Cand wbest = warp_argmin_redux(best, wsrc);
      if (lane == wsrc && best.k != 0x7fffffff) {
#pragma unroll
        for (int i = 0; i < kRecDoubles; ++i) s_wrec[warp][i] = rec[i];
      }
      if (lane == 0) s_red[warp] = wbest;
      __syncthreads();
      if (warp == 0) {
        Cand c = s_red[0];
        int bw = 0;
#pragma unroll
        for (int w = 1; w < kSeqThreads / 32; ++w) {
          const Cand o = s_red[w];
          const Cand n2 = cand_min(c, o);
          if (n2.k != c.k) bw = w;
          c = n2;
        }
        if (lane < kRecDoubles) {
          const double v = s_wrec[bw][lane];
#pragma unroll
          for (int r = 0; r < kSlotReps; ++r)
            p.slot_rec[(int64_t)((r * 2 + par) * G + blockIdx.x) * kRecStride + lane] = v;
        }
        if (lane < kSlotReps) {
          p.slot_d[(lane * 2 + par) * G + blockIdx.x] = c.d;
          p.slot_k[(lane * 2 + par) * G + blockIdx.x] = c.k;
        }
        // hot-winner claim: the candidate is a match on its own evidence and it is its thread's first landmark
        // (the one whose panel rows live in shared memory)
        if (hot && c.d <= p.thr && c.k < cnt && c.k / kSeqThreads == (int)blockIdx.x) {
          const int otid = c.k - (int)blockIdx.x * kSeqThreads;
          if (lane <= kRecDoubles) {
            const double v = lane < kRecDoubles ? s_wrec[bw][lane] : (double)c.k;
#pragma unroll
            for (int r = 0; r < kSlotReps; ++r) p.hot_rec[(r * 2 + par) * kRecStride + lane] = v;
          }
          for (int q = lane; q < 2 * t; q += 32) {
            // panel column q = 2 s2 + h of rows a, a+1: component h of the (column 2 s2, 2 s2 + 1) pairs
            const double2 ra = s_wown[(q & ~1) * kSeqThreads + otid], rb = s_wown[((q & ~1) + 1) * kSeqThreads + otid];
            const double2 v = (q & 1) ? make_double2(ra.y, rb.y) : make_double2(ra.x, rb.x);
#pragma unroll
            for (int r = 0; r < kSlotReps; ++r) p.hot_w[(r * 2 + par) * kMaxK2 + q] = v;
          }
          if (lane < kSlotReps) {
            const unsigned long long tag = p.bar_base + (unsigned long long)(t + 1) * gridDim.x;
            atomicMax(p.hot_hdr + lane * 2 + par, tag << 8);
            atomicAdd(p.hot_hdr + lane * 2 + par, 1ULL);
          }
        }
      }
    }

    // ---- B: the one grid barrier of this measurement ----------------------------------------
    SEQ_STAMP(2);
    if (p.bar_timeout_ns == 0) {
      grid_barrier(p.bar, p.bar_base + (unsigned long long)(t + 1) * gridDim.x);
    } else if (grid_barrier_bounded(p.bar, p.bar_base + (unsigned long long)(t + 1) * gridDim.x, p.bar_timeout_ns)) {
      // part of the grid never became resident: report instead of spinning (the host resets the counter)
      if (tid == 0) {
        atomicMax(p.status, kErrCuda);
        if (p.done_out) *(volatile int *)p.done_out = -1;   // (plain store: done_out may be mapped pinned host memory)
      }
      return;
    }
    SEQ_STAMP(3);

    // ---- C: decision + the winner's record: warps 0 and 1 only (two requesters per CTA on the slot
    //         lines), broadcast to the rest of the CTA through shared memory ----------------------------
    bool fast = false;   // matched an existing landmark: record comes from the winner's slot
    int jfast = -1;
    if (unknown) {
      if (warp < 2) {
        const int rep = (int)(blockIdx.x % kSlotReps);
        const double *sd = p.slot_d + (rep * 2 + par) * G;
        const int *sk = p.slot_k + (rep * 2 + par) * G;
        double od[kMaxSlotsPerLane];
        int ok[kMaxSlotsPerLane];
#pragma unroll
        for (int i = 0; i < kMaxSlotsPerLane; ++i) {     // all loads in flight together
          const int b = lane + 32 * i;
          od[i] = INFINITY; ok[i] = 0x7fffffff;
          if (b < G) { od[i] = ldcg_d(sd + b); ok[i] = __ldcg(sk + b); }
        }
        // hot-winner area, in the same round trip: header + record (warp 0) / panel-row history (warp 1)
        unsigned long long hhdr = 0;
        double hval = 0.0;
        double2 hw0 = make_double2(0.0, 0.0), hw1 = hw0;
        if (hot) {
          const int hs = rep * 2 + par;
          asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(hhdr) : "l"(p.hot_hdr + hs));
          if (lane <= kRecDoubles) hval = ldcg_d(p.hot_rec + hs * kRecStride + (warp == 0 ? lane : kRecDoubles));
          if (warp == 1) {
            if (lane < 2 * t) hw0 = __ldcg(p.hot_w + hs * kMaxK2 + lane);
            if (lane + 32 < 2 * t) hw1 = __ldcg(p.hot_w + hs * kMaxK2 + lane + 32);
          }
        }
        Cand c{INFINITY, 0x7fffffff};
        int bstar = 0;
#ifdef EKF_SEQ_PROF
        if (od[0] + od[1] + od[2] + od[3] + od[4] == -1.0) bstar = 1;   // wait for the loads
        SEQ_STAMP(1);
#endif
#pragma unroll
        for (int i = 0; i < kMaxSlotsPerLane; ++i) {
          const int b = lane + 32 * i;
          if (b < G) {
            const Cand o{od[i], ok[i]};
            const Cand n2 = cand_min(c, o);
            if (n2.k != c.k) bstar = b;
            c = n2;
          }
        }
        {
          int csrc;
          c = warp_argmin_redux(c, csrc);
          bstar = __shfl_sync(0xffffffffu, bstar, csrc);
        }
        SEQ_STAMP(7);
        // the tentative landmark sits last with distance == threshold and wins only on a
        // strictly smaller value: slam.cpp:406-408,418-440
        fast = (c.d <= p.thr && c.k < cnt);
        if (tid == 0) s_fast = fast ? 1 : 0;
        // the preloaded copy is the winner's iff this measurement had exactly one claimant and it is the winner
        bool hot_ok = false;
        if (hot && fast) {
          const unsigned long long tag = p.bar_base + (unsigned long long)(t + 1) * gridDim.x;
          const double hk = __shfl_sync(0xffffffffu, hval, warp == 0 ? kRecDoubles : 0);
          hot_ok = (hhdr == ((tag << 8) | 1ULL)) && (hk == (double)c.k);
        }
        if (fast) {
          jfast = c.k;
          if (warp == 0) {
            if (lane < kRecDoubles)
              s_b.rec[lane] = hot_ok ? hval
                                     : ldcg_d(p.slot_rec + (int64_t)((rep * 2 + par) * G + bstar) * kRecStride + lane);
            if (lane == 0) {
              s_b.j = jfast; s_b.is_new = 0;
              if (blockIdx.x == 0 && p.assoc_out) p.assoc_out[t] = jfast;
            }
          } else if (hot_ok) {
            if (lane < 2 * t) { s_wa[lane] = hw0.x; s_wb[lane] = hw0.y; }
            if (lane + 32 < 2 * t) { s_wa[lane + 32] = hw1.x; s_wb[lane + 32] = hw1.y; }
          } else {
            // panel history of the chosen landmark's two rows (column operand of the correction):
            // 8 doubles per k-step, columns q < 2t; at most 3 per lane, all in flight together
            const int a = 4 + 2 * jfast;
            const double *wrow = p.W + (((int64_t)(a >> 5) * nks) << 7) + (ekf_panel_slot(a & 31) << 2);
            double wv[3];
#pragma unroll
            for (int u = 0; u < 3; ++u) {
              const int i = lane + 32 * u;
              const int ks = i >> 3, w = i & 7, q = 4 * ks + (w & 3);
              wv[u] = 0.0;
              if (q < 2 * t) wv[u] = ldcg_d(wrow + ((int64_t)ks << 7) + w);
            }
#pragma unroll
            for (int u = 0; u < 3; ++u) {
              const int i = lane + 32 * u;
              const int ks = i >> 3, w = i & 7, q = 4 * ks + (w & 3);
              if (q < 2 * t) { if (w < 4) s_wa[q] = wv[u]; else s_wb[q] = wv[u]; }
            }
          }
        }
      }
      __syncthreads();
      fast = s_fast != 0;
    }
    if (!fast) {
      // serial path (thread 0): new landmark, known association, skipped sighting, errors
      if (tid == 0) {
        int j = -1, is_new = 0, err = 0;
        if (unknown) {
          if (cnt < p.N) { j = cnt; is_new = 1; }
          else err = kErrCapacity;
        } else {
          const int id = p.ids[t];
          const bool del = p.action != nullptr && p.action[t] == 2;   // slam.cpp:205
          if (!del) {
            if (id < 0 || id >= p.N) err = kErrBadId;
            else {
              j = id;
              // first sighting: both state slots exactly zero (slam.cpp:213)
              is_new = (ldcg_d(p.mu + 4 + 2 * id) == 0.0 && ldcg_d(p.mu + 5 + 2 * id) == 0.0);
            }
          }
        }
        s_b.j = j; s_b.is_new = is_new;
        if (err && blockIdx.x == 0) atomicMax(p.status, err);
        if (blockIdx.x == 0 && p.assoc_out) p.assoc_out[t] = j;
        if (j >= 0) {
          const int a = 4 + 2 * j;
          double lx, ly;
          if (is_new) {                                      // slam.cpp:214-215 / :351-354
            lx = px + z.r * cos(z.phi + th);
            ly = py + z.r * sin(z.phi + th);
            if (unknown) s_cnt = cnt + 1;                    // slam.cpp:356,421
          } else {
            lx = ldcg_d(p.mu + a); ly = ldcg_d(p.mu + a + 1);
          }
          const double4 pa = ldcg_d4(p.Pp + a), pbr = ldcg_d4(p.Pp + a + 1), dj = ldcg_d4(p.D + j);
          const Lin L = linearize(th, px, py, lx, ly);       // slam.cpp:444-473
          const double par3[3] = {pa.x, pa.y, pa.z}, pbr3[3] = {pbr.x, pbr.y, pbr.z};
          const Block5 B = make_block5(s_pb, par3, pbr3, dj.x, dj.y, dj.z);
          const Innov I = innovation(B, L, p.r_noise);       // S = H Sigma H^T + R, slam.cpp:476
          const double zd0 = z.r - L.zhat_r;
          const double zd1 = normalize_angle(z.phi - L.zhat_b);   // slam.cpp:479-480
          double i00, i01, i11;
          (void)mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, i00, i01, i11);
          s_b.lx = lx; s_b.ly = ly;
#pragma unroll
          for (int q = 0; q < 4; ++q) { s_b.rec[q] = L.h0[q + 1]; s_b.rec[4 + q] = L.h1[q + 1]; }
#pragma unroll
          for (int q = 0; q < 3; ++q) { s_b.rec[8 + 2 * q] = I.T[q][0]; s_b.rec[9 + 2 * q] = I.T[q][1]; }
          s_b.rec[14] = i00; s_b.rec[15] = i01; s_b.rec[16] = i11;
          s_b.rec[17] = zd0; s_b.rec[18] = zd1;
          s_b.rec[19] = dj.x; s_b.rec[20] = dj.y; s_b.rec[21] = dj.z;
        }
      }
      __syncthreads();
      const int js = s_b.j;
      if (js >= 0) {
        const int a = 4 + 2 * js;
        const double *wrow = p.W + (((int64_t)(a >> 5) * nks) << 7) + (ekf_panel_slot(a & 31) << 2);
        for (int i = tid; i < 8 * ((t + 1) >> 1); i += blockDim.x) {
          const int ks = i >> 3, w = i & 7, q = 4 * ks + (w & 3);
          if (q < 2 * t) {
            const double v = ldcg_d(wrow + ((int64_t)ks << 7) + w);
            if (w < 4) s_wa[q] = v; else s_wb[q] = v;
          }
        }
      }
      __syncthreads();
    }
    SEQ_STAMP(4);
    const int j = s_b.j;
    // sharded handle: which slot of the exchanged column buffer holds landmark j's raw Sigma_0 columns
    int slot = -1;
    // (a landmark created during this launch has an all-zero Sigma_0 column: nothing to look up)
    if (p.cols != nullptr && j >= 0 && !(unknown && j >= s_cnt0)) {
      slot = -2;
      if (__ldg(p.guess + t) == j) slot = t;
      else
        for (int s2 = 0; s2 < p.m; ++s2)
          if (__ldg(p.guess + s2) == j) { slot = s2; break; }
    }
    if (slot == -2) {
      // failed speculation (identical decision in every thread of the grid): stop before measurement
      // t is applied; the host runs another round for the rest.  Pay the barrier arrivals the host
      // has already accounted for.
      if (tid == 0 && p.m - t - 1 > 0) atomicAdd(p.bar, (unsigned long long)(p.m - t - 1));
      done = t;
      break;
    }

    // ---- pose rows: gain, state and pose-block update (slam.cpp:476,482,485-486 restricted to
    //      rows/cols 0..2), redundantly in every CTA, one lane per pose row ------------------------
    if (warp == 0 && lane < 3) {
      const int c = lane;
      if (j >= 0) {
        const double m00 = s_b.rec[14], m01 = s_b.rec[15], m11 = s_b.rec[16];
        const double zd0 = s_b.rec[17], zd1 = s_b.rec[18];
        const double t0 = s_b.rec[8 + 2 * c], t1 = s_b.rec[9 + 2 * c];
        const double k0 = t0 * m00 + t1 * m01, k1 = t0 * m01 + t1 * m11;
        double psn = s_ps[c] + (k0 * zd0 + k1 * zd1);
        if (c == 0) psn = normalize_angle(psn);              // slam.cpp:488
        for (int c2 = 0; c2 <= c; ++c2) {
          const double v = s_pb[c][c2] - (k0 * s_b.rec[8 + 2 * c2] + k1 * s_b.rec[9 + 2 * c2]);
          s_pb[c][c2] = v; s_pb[c2][c] = v;
        }
        s_ps[c] = psn;
        if (blockIdx.x == 0) {
          p.W[ekf_panel_elem(p.kmax, c, 2 * t)] = t0;
          p.W[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = t1;
          p.V[ekf_panel_elem(p.kmax, c, 2 * t)] = k0;
          p.V[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = k1;
        }
      } else if (blockIdx.x == 0) {
        p.W[ekf_panel_elem(p.kmax, c, 2 * t)] = 0.0; p.W[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = 0.0;
        p.V[ekf_panel_elem(p.kmax, c, 2 * t)] = 0.0; p.V[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = 0.0;
      }
    }

    // ---- D: own rows: gather the two landmark columns, gain, state, Pp, D, panels ------------
    if (j < 0) {
      for (int k = gtid; k < p.N; k += nthreads_total) {
        const int a = 4 + 2 * k;
        const double2 zero = make_double2(0.0, 0.0);
        double *wp = p.W + ekf_panel_elem(p.kmax, a, 2 * t);
        double *vp = p.V + ekf_panel_elem(p.kmax, a, 2 * t);
        *reinterpret_cast<double2 *>(wp) = zero; *reinterpret_cast<double2 *>(wp + 4) = zero;
        *reinterpret_cast<double2 *>(vp) = zero; *reinterpret_cast<double2 *>(vp + 4) = zero;
      }
      s_vown[(2 * t) * kSeqThreads + tid] = make_double2(0.0, 0.0);
      s_vown[(2 * t + 1) * kSeqThreads + tid] = make_double2(0.0, 0.0);
      if (hot) {
        s_wown[(2 * t) * kSeqThreads + tid] = make_double2(0.0, 0.0);
        s_wown[(2 * t + 1) * kSeqThreads + tid] = make_double2(0.0, 0.0);
      }
    } else {
      Lin L;
      L.h0[0] = 0.0; L.h1[0] = -1.0;
#pragma unroll
      for (int q = 0; q < 4; ++q) { L.h0[q + 1] = s_b.rec[q]; L.h1[q + 1] = s_b.rec[4 + q]; }
      const double tp00 = s_b.rec[8], tp01 = s_b.rec[9], tp10 = s_b.rec[10], tp11 = s_b.rec[11];
      const double tp20 = s_b.rec[12], tp21 = s_b.rec[13];
      const double m00 = s_b.rec[14], m01 = s_b.rec[15], m11 = s_b.rec[16];
      const double zd0 = s_b.rec[17], zd1 = s_b.rec[18];
      const int aj = 4 + 2 * j;
      const int full_ks = (2 * t) >> 2;          // complete k-steps of panel history
      const bool tail = ((2 * t) & 2) != 0;      // + two more columns
      for (int k = gtid; k < p.N; k += nthreads_total) {
        const int a = 4 + 2 * k;
        double q00, q01, q10, q11;  // Sigma_cur[{a,b}_k, {a,b}_j]
        double *vrow = p.V + (((int64_t)(a >> 5) * nks) << 7) + (ekf_panel_slot(a & 31) << 2);
        double *wrow = p.W + (((int64_t)(a >> 5) * nks) << 7) + (ekf_panel_slot(a & 31) << 2);
        if (k == j) {
          q00 = s_b.rec[19]; q01 = s_b.rec[20]; q10 = s_b.rec[20]; q11 = s_b.rec[21];
        } else {
          if (p.cols != nullptr) {
            if (slot >= 0) {             // exchanged raw columns, [slot][row][2]
              const double *src = p.cols + ((int64_t)slot * n_pad + a) * 2;
              double2 r0, r1;
              if (p.cols_poll) {
                poll_cols(src, r0, r1, p.status);
              } else {
                r0 = __ldcg(reinterpret_cast<const double2 *>(src));
                r1 = __ldcg(reinterpret_cast<const double2 *>(src + 2));
              }
              q00 = r0.x; q01 = r0.y; q10 = r1.x; q11 = r1.y;
            } else {                     // landmark created in this scan: Sigma_0 column is zero
              q00 = 0.0; q01 = 0.0; q10 = 0.0; q11 = 0.0;
            }
          } else if (k > j) {
            const double *src = p.sigma + ekf_lay_elem(p.lay, a, aj);
            const double2 r0 = __ldg(reinterpret_cast<const double2 *>(src));
            const double2 r1 = __ldg(reinterpret_cast<const double2 *>(src + EKF_STRIP));
            q00 = r0.x; q01 = r0.y; q10 = r1.x; q11 = r1.y;
          } else {
            const double *src = p.sigma + ekf_lay_elem(p.lay, aj, a);
            const double2 c0 = __ldg(reinterpret_cast<const double2 *>(src));
            const double2 c1 = __ldg(reinterpret_cast<const double2 *>(src + EKF_STRIP));
            q00 = c0.x; q10 = c0.y; q01 = c1.x; q11 = c1.y;
          }
          // deferred downdates of earlier measurements of this scan.  First landmark: its V rows are in
          // shared memory; further ones: rows a, a+1 of V are 8 contiguous doubles per k-step (this
          // thread wrote them itself).  Same accumulation order (panel column ascending) either way.
          if (k == gtid) {
            for (int s2 = 0; s2 < t; ++s2) {
              const double2 va = s_vown[(2 * s2) * kSeqThreads + tid];
              const double2 vb = s_vown[(2 * s2 + 1) * kSeqThreads + tid];
              const double wa0 = s_wa[2 * s2], wa1 = s_wa[2 * s2 + 1], wb0 = s_wb[2 * s2], wb1 = s_wb[2 * s2 + 1];
              q00 -= va.x * wa0; q01 -= va.x * wb0; q10 -= vb.x * wa0; q11 -= vb.x * wb0;
              q00 -= va.y * wa1; q01 -= va.y * wb1; q10 -= vb.y * wa1; q11 -= vb.y * wb1;
            }
          } else {
          for (int ks = 0; ks < full_ks; ++ks) {
            double va0, va1, va2, va3, vb0, vb1, vb2, vb3;
            va3 = ld_plain256(vrow + ((int64_t)ks << 7), va0, va1, va2);
            vb3 = ld_plain256(vrow + ((int64_t)ks << 7) + 4, vb0, vb1, vb2);
            const double *wa = s_wa + 4 * ks, *wb = s_wb + 4 * ks;
            q00 -= va0 * wa[0]; q01 -= va0 * wb[0]; q10 -= vb0 * wa[0]; q11 -= vb0 * wb[0];
            q00 -= va1 * wa[1]; q01 -= va1 * wb[1]; q10 -= vb1 * wa[1]; q11 -= vb1 * wb[1];
            q00 -= va2 * wa[2]; q01 -= va2 * wb[2]; q10 -= vb2 * wa[2]; q11 -= vb2 * wb[2];
            q00 -= va3 * wa[3]; q01 -= va3 * wb[3]; q10 -= vb3 * wa[3]; q11 -= vb3 * wb[3];
          }
          if (tail) {
            const double2 va = *reinterpret_cast<const double2 *>(vrow + ((int64_t)full_ks << 7));
            const double2 vb = *reinterpret_cast<const double2 *>(vrow + ((int64_t)full_ks << 7) + 4);
            const double *wa = s_wa + 4 * full_ks, *wb = s_wb + 4 * full_ks;
            q00 -= va.x * wa[0]; q01 -= va.x * wb[0]; q10 -= vb.x * wa[0]; q11 -= vb.x * wb[0];
            q00 -= va.y * wa[1]; q01 -= va.y * wb[1]; q10 -= vb.y * wa[1]; q11 -= vb.y * wb[1];
          }
          }
        }
        const bool own = (k == gtid);
        double4 pa = o_pa, pbr = o_pb;
        if (!own) { pa = p.Pp[a]; pbr = p.Pp[a + 1]; }
        const double rowa[5] = {pa.x, pa.y, pa.z, q00, q01};
        const double rowb[5] = {pbr.x, pbr.y, pbr.z, q10, q11};
        double ga0, ga1, gb0, gb1;
        row_times_ht(rowa, L, ga0, ga1);                     // G = Sigma H^T
        row_times_ht(rowb, L, gb0, gb1);
        const double ka0 = ga0 * m00 + ga1 * m01, ka1 = ga0 * m01 + ga1 * m11;   // K = G S^-1
        const double kb0 = gb0 * m00 + gb1 * m01, kb1 = gb0 * m01 + gb1 * m11;
        double2 lm = o_lm;
        if (!own) lm = *reinterpret_cast<const double2 *>(p.mu + a);
        if (k == j && s_b.is_new) { lm.x = s_b.lx; lm.y = s_b.ly; }
        lm.x = lm.x + (ka0 * zd0 + ka1 * zd1);               // slam.cpp:482
        lm.y = lm.y + (kb0 * zd0 + kb1 * zd1);
        *reinterpret_cast<double2 *>(p.mu + a) = lm;
        // Sigma_cur[i, pose c] -= K[i,:] . G[c,:]
        pa.x -= ka0 * tp00 + ka1 * tp01;
        pa.y -= ka0 * tp10 + ka1 * tp11;
        pa.z -= ka0 * tp20 + ka1 * tp21;
        pbr.x -= kb0 * tp00 + kb1 * tp01;
        pbr.y -= kb0 * tp10 + kb1 * tp11;
        pbr.z -= kb0 * tp20 + kb1 * tp21;
        p.Pp[a] = pa; p.Pp[a + 1] = pbr;
        double4 dk = o_dk;
        if (!own) dk = p.D[k];
        dk.x -= ka0 * ga0 + ka1 * ga1;
        dk.y -= kb0 * ga0 + kb1 * ga1;
        dk.z -= kb0 * gb0 + kb1 * gb1;
        p.D[k] = dk;
        // columns 2t, 2t+1 of rows a (slot s) and a+1 (slot s+1) of both panels
        double *wp = wrow + ((int64_t)((2 * t) >> 2) << 7) + ((2 * t) & 3);
        double *vp = vrow + ((int64_t)((2 * t) >> 2) << 7) + ((2 * t) & 3);
        *reinterpret_cast<double2 *>(wp) = make_double2(ga0, ga1);
        *reinterpret_cast<double2 *>(wp + 4) = make_double2(gb0, gb1);
        *reinterpret_cast<double2 *>(vp) = make_double2(ka0, ka1);
        *reinterpret_cast<double2 *>(vp + 4) = make_double2(kb0, kb1);
        if (own) {
          o_pa = pa; o_pb = pbr; o_dk = dk; o_lm = lm;
          s_vown[(2 * t) * kSeqThreads + tid] = make_double2(ka0, ka1);
          s_vown[(2 * t + 1) * kSeqThreads + tid] = make_double2(kb0, kb1);
          if (hot) {
            s_wown[(2 * t) * kSeqThreads + tid] = make_double2(ga0, ga1);
            s_wown[(2 * t + 1) * kSeqThreads + tid] = make_double2(gb0, gb1);
          }
        }
      }
    }
    SEQ_STAMP(5);
    __syncthreads();  // s_b / s_wa / s_ps are rewritten by the next measurement
    SEQ_STAMP(6);
  }

  // columns [2 * done, round-up-to-4(2m)) of both panels are zeroed: the covariance pass works on whole
  // k-steps of 4 columns (and, after a stopped scan, on the columns the host sized it for)
  for (int qz = 2 * done; qz < ((2 * p.m + 3) & ~3); qz += 2) {
    const double2 zero = make_double2(0.0, 0.0);
    for (int k = gtid; k < p.N; k += nthreads_total) {
      const int a = 4 + 2 * k;
      double *wp = p.W + ekf_panel_elem(p.kmax, a, qz);
      double *vp = p.V + ekf_panel_elem(p.kmax, a, qz);
      *reinterpret_cast<double2 *>(wp) = zero; *reinterpret_cast<double2 *>(wp + 4) = zero;
      *reinterpret_cast<double2 *>(vp) = zero; *reinterpret_cast<double2 *>(vp + 4) = zero;
    }
    if (blockIdx.x == 0 && tid < 3) {
      p.W[ekf_panel_elem(p.kmax, tid, qz)] = 0.0; p.W[ekf_panel_elem(p.kmax, tid, qz + 1)] = 0.0;
      p.V[ekf_panel_elem(p.kmax, tid, qz)] = 0.0; p.V[ekf_panel_elem(p.kmax, tid, qz + 1)] = 0.0;
    }
  }

  // ---------------- epilogue: slam.cpp:273-274,291 / :490-491,508 ---------------------------------
  // a scan without measurements has passed no grid barrier yet: the global pose block / pose state
  // below may only be rewritten once every CTA has read the old ones
  if (p.m == 0) {
    if (p.bar_timeout_ns == 0) {
      grid_barrier(p.bar, p.bar_base + gridDim.x);
    } else if (grid_barrier_bounded(p.bar, p.bar_base + gridDim.x, p.bar_timeout_ns)) {
      if (tid == 0) {
        atomicMax(p.status, kErrCuda);
        if (p.done_out) *(volatile int *)p.done_out = -1;   // (plain store: done_out may be mapped pinned host memory)
      }
      return;
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      if (p.side_mode) {
        p.Pp[i] = make_double4(s_pb[i][0], s_pb[i][1], s_pb[i][2], 0.0);     // current block, persistent
      } else {
        p.Pp[i] = make_double4(s_pb0[i][0], s_pb0[i][1], s_pb0[i][2], 0.0);
#pragma unroll
        for (int j = 0; j <= i; ++j) strip0[i * EKF_STRIP + j] = s_pb0[i][j];  // Sigma_0 after prediction
      }
    }
    p.mu[0] = s_ps[0]; p.mu[1] = s_ps[1]; p.mu[2] = s_ps[2];
    *p.cnt = s_cnt;
    *p.cnt_out = s_cnt;
    p.pose_out[0] = s_ps[0]; p.pose_out[1] = s_ps[1]; p.pose_out[2] = s_ps[2];
    if (p.finish && done == p.m) { p.prev_xy[0] = s_ps[1]; p.prev_xy[1] = s_ps[2]; }
    if (p.done_out) *p.done_out = done;
    if (p.status_out && p.finish) *p.status_out = atomicExch(p.status, 0);
  }
}

}  // namespace ekf
