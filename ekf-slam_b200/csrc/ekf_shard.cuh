// Row-sharded single large map (BASELINE configs 4-5 on 2/4/8 GPUs): the small kernels around the
// exchange step.  One process per GPU; shard g owns the rows [rb, re) of Sigma's lower triangle
// (equal-area split, ekf_layout.h).  Everything O(n) is REPLICATED and computed redundantly and
// identically on every rank: state, pose panel Pp, diagonal blocks D, the W / V panels, the
// association decisions.  Per scan there is ONE exchange:
//
//   begin   (this file)  prediction on the replicated pose panel (slam.cpp:186-198 / :323-335);
//                        speculative association of all m sightings against the scan-start state
//                        -> guess[t]; every rank writes the part it owns of the raw Sigma_0 columns
//                        of the guessed landmarks into the exchange buffer (zeros elsewhere)
//   exchange             every used element of the buffer has exactly one owner, so nothing is reduced, only
//                        delivered: by shard_push_cols_kernel over peer memory (NVLink / NVSwitch stores,
//                        flag-in-data arrival, two-hop relay; concurrent with the sequential kernel), or by the
//                        caller's sum-all-reduce (NCCL; m * n_pad * 2 doubles = 6.4 MB at N = 10,000, m = 20; exact,
//                        because all but one contributor are zero)
//   finish  (ekf_seq.cuh + ekf_downdate.cuh)  the unchanged sequential phase, replicated, reading
//                        the columns from the buffer instead of Sigma; it re-derives every association
//                        with the reference's sequential semantics and STOPS if a decision ever needs
//                        a landmark that was not guessed (the caller then runs another round for the
//                        rest of the scan).  Then each rank applies the rank-2m downdate to its own rows
//                        and writes Pp / D back into the Sigma entries they mirror.
//
// Pp and D are persistent between scans on a sharded handle (no per-scan extraction); after a
// fixture call (fill / set) they are rebuilt once by side_extract + one more all-reduce.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_layout.h"
#include "ekf_math.cuh"

namespace ekf {

// ---- prediction on the replicated side state ------------------------------------------------------
__global__ void shard_predict_kernel(double4 *Pp, double *mu, const double *pose, const double *prev_xy, int N,
                                     double q_noise) {
  const double th = normalize_angle(pose[0]);              // slam.cpp:186
  const double x = pose[1], y = pose[2];
  const double dx = x - prev_xy[0], dy = y - prev_xy[1];   // slam.cpp:190-191
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  for (int k = gtid; k < N; k += gridDim.x * blockDim.x) {
    const int a = 4 + 2 * k;
    double4 va = Pp[a], vb = Pp[a + 1];
    va.y = va.x * (-dy) + va.y; va.z = va.x * dx + va.z;
    vb.y = vb.x * (-dy) + vb.y; vb.z = vb.x * dx + vb.z;
    Pp[a] = va; Pp[a + 1] = vb;
  }
  if (gtid == 0) {
    double P[3][3], Bm[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double4 r = Pp[i];
      P[i][0] = r.x; P[i][1] = r.y; P[i][2] = r.z;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < i; ++j) P[j][i] = P[i][j];        // lower triangle rules
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
    for (int i = 0; i < 3; ++i) P[i][i] += q_noise;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < i; ++j) P[j][i] = P[i][j];
#pragma unroll
    for (int i = 0; i < 3; ++i) Pp[i] = make_double4(P[i][0], P[i][1], P[i][2], 0.0);
    mu[0] = th; mu[1] = x; mu[2] = y;
  }
}

// ---- speculative association against the scan-start state: grid (m sightings, P slices of the map) ----
// every CTA reduces its slice to one candidate; the last CTA of a sighting to finish (ticket counter)
// reduces the P partial candidates and writes guess[t].  m * P is sized to about two CTAs per SM.
constexpr int kGuessThreads = 256;
constexpr int kGuessMaxSlices = 32;
__global__ void __launch_bounds__(kGuessThreads) shard_guess_kernel(const double4 *Pp, const double4 *D,
                                                                   const double *mu, const int *cnt_p,
                                                                   const double *xy, const int *ids, int *guess,
                                                                   double r_noise, double thr, double *part_d,
                                                                   int *part_k, unsigned int *ticket) {
  __shared__ Cand s_red[kGuessThreads / 32];
  __shared__ int s_last;
  const int t = blockIdx.x, slice = blockIdx.y, P = gridDim.y;
  if (ids != nullptr) {                                    // known association: nothing to speculate
    if (threadIdx.x == 0 && slice == 0) guess[t] = ids[t];
    return;
  }
  const int cnt = *cnt_p;
  const Polar z = to_polar(xy[2 * t], xy[2 * t + 1]);
  const double th = mu[0], px = mu[1], py = mu[2];
  double pb[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double4 r = Pp[i];
    pb[i][0] = r.x; pb[i][1] = r.y; pb[i][2] = r.z;
  }
  Cand best{INFINITY, 0x7fffffff};
  for (int k = slice * (int)blockDim.x + threadIdx.x; k < cnt; k += P * (int)blockDim.x) {   // blockDim <= kGuessThreads
    const int a = 4 + 2 * k;
    const double2 lm = *reinterpret_cast<const double2 *>(mu + a);
    const double4 pa = Pp[a], pbr = Pp[a + 1], dk = D[k];
    const Lin L = linearize(th, px, py, lm.x, lm.y);
    const double par3[3] = {pa.x, pa.y, pa.z}, pbr3[3] = {pbr.x, pbr.y, pbr.z};
    const Block5 B = make_block5(pb, par3, pbr3, dk.x, dk.y, dk.z);
    const Innov I = innovation(B, L, r_noise);
    const double zd0 = z.r - L.zhat_r, zd1 = normalize_angle(z.phi - L.zhat_b);
    double i00, i01, i11;
    const double d = mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, i00, i01, i11);
    best = cand_min(best, Cand{d, k});
  }
  best = warp_argmin(best);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    Cand c = s_red[0];
    for (int w = 1; w < (int)blockDim.x / 32; ++w) c = cand_min(c, s_red[w]);
    part_d[t * kGuessMaxSlices + slice] = c.d;
    part_k[t * kGuessMaxSlices + slice] = c.k;
    __threadfence();
    const unsigned int prev = atomicAdd(ticket + t, 1u);
    s_last = (prev == (unsigned int)(P - 1));
    if (s_last) {
      __threadfence();
      ticket[t] = 0;                                       // ready for the next launch
      Cand g{INFINITY, 0x7fffffff};
      for (int q = 0; q < P; ++q)
        g = cand_min(g, Cand{__ldcg(part_d + t * kGuessMaxSlices + q), __ldcg(part_k + t * kGuessMaxSlices + q)});
      // generous acceptance (4 x threshold): a sighting that is borderline now may match after the
      // earlier updates of the scan; an unnecessary column costs nothing but bandwidth
      guess[t] = (g.k < cnt && g.d <= 4.0 * thr) ? g.k : -1;
    }
  }
}

// ---- exchange buffer: this rank's part of the raw Sigma_0 columns of the guessed landmarks --------
// cols[slot][row][2] = (Sigma_0[row, a_j], Sigma_0[row, b_j]) for row outside the landmark's own 2x2
// block; every element is written by every rank (zero unless owned).
__global__ void shard_extract_cols_kernel(const double *sigma, EkfLay lay, int n_pad, int N, const int *guess,
                                          int m, double *cols) {
  // grid = (ceil(n_pad / blockDim), m), one thread per (slot, row)
  const int slot = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad || slot >= m) return;
  const int j = guess[slot];
  double2 v = make_double2(0.0, 0.0);
  if (j >= 0 && j < N) {
    const int a = 4 + 2 * j;
    if (i > a + 1) {
      if (ekf_lay_owns_row(lay, i)) v = *reinterpret_cast<const double2 *>(sigma + ekf_lay_elem(lay, i, a));
    } else if (i < a) {
      if (ekf_lay_owns_row(lay, a)) {
        v.x = sigma[ekf_lay_elem(lay, a, i)];
        v.y = sigma[ekf_lay_elem(lay, a + 1, i)];
      }
    }
  }
  *reinterpret_cast<double2 *>(cols + ((int64_t)slot * n_pad + i) * 2) = v;
}

// ---- overlapped scans on one GPU (ekf_set_overlap): the columns of the covariance the PENDING pass will produce ---
// While the covariance pass of scan s is still streaming Sigma_old -> Sigma_new (out of place), scan s+1 already
// needs the raw columns of Sigma_new for its guessed landmarks.  They are formed here from the pass's INPUT and its
// panels:  Sigma_new[r, c] = Sigma_old[r, c] - sum_q V[r][q] W[c][q]  for the stored (row r >= column c) element,
// exactly the element the pass writes (it sums the same products on the tensor path, so the two agree to rounding).
// k2 = 0: no pass pending, plain extraction.  Same output layout as shard_extract_cols_kernel (whole matrix owned).
__global__ void pipe_extract_cols_kernel(const double *sigma, EkfLay lay, int n_pad, int N, const int *guess, int m,
                                         const double *V, const double *W, int kmax, int k2, double *cols) {
  const int64_t total = (int64_t)m * n_pad;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int slot = (int)(e / n_pad), i = (int)(e % n_pad);
    const int j = guess[slot];
    double2 v = make_double2(0.0, 0.0);
    if (j >= 0 && j < N) {
      const int a = 4 + 2 * j;
      if (i > a + 1) {                 // stored elements (i, a), (i, a + 1): V row i, W rows a / a + 1
        v = *reinterpret_cast<const double2 *>(sigma + ekf_lay_elem(lay, i, a));
        for (int q = 0; q < k2; ++q) {
          const double vi = V[ekf_panel_elem(kmax, i, q)];
          v.x -= vi * __ldg(W + ekf_panel_elem(kmax, a, q));
          v.y -= vi * __ldg(W + ekf_panel_elem(kmax, a + 1, q));
        }
      } else if (i < a) {              // stored elements (a, i), (a + 1, i): V rows a / a + 1, W row i
        v.x = sigma[ekf_lay_elem(lay, a, i)];
        v.y = sigma[ekf_lay_elem(lay, a + 1, i)];
        for (int q = 0; q < k2; ++q) {
          const double wi = W[ekf_panel_elem(kmax, i, q)];
          v.x -= __ldg(V + ekf_panel_elem(kmax, a, q)) * wi;
          v.y -= __ldg(V + ekf_panel_elem(kmax, a + 1, q)) * wi;
        }
      }
    }
    *reinterpret_cast<double2 *>(cols + e * 2) = v;
  }
}

// ---- the same exchange as ONE kernel over peer memory (NVLink / NVSwitch): no collective call -------
// Every used element of the column buffer has exactly one owner (the rank that stores the Sigma row it
// comes from).  The owner reads it and STORES it straight into the column buffer of every rank - its own
// and, through peer-mapped pointers (CUDA IPC between processes), the others' - so the "all-reduce" is
// R - 1 remote stores per owned element and no reduction at all.  The last CTA to finish publishes this
// rank's arrival flag in every rank's flag array (system-scope release after the data); the consumer
// side is shard_wait_peers_kernel right before the sequential kernel.  Buffers alternate by round parity:
// a fast rank may already push round r + 1 while a slow one still reads round r.
constexpr int kMaxPeers = 16;
constexpr int kErrPeerTimeout = 5;         // EKF_ERR_CUDA: the exchange did not complete
struct PeerTable {
  int count, self;
  double *cols[kMaxPeers];                 // this round's column buffer of every rank
  unsigned long long *flags[kMaxPeers];    // flag array of every rank: flags[q][g] = rounds rank g has pushed to q
};
// One thread per ROW, all m slots, 8 at a time: a thread first gathers its owned elements of a batch (loads in flight
// together - the compiler cannot hoist them over the peer stores itself: the pointers may alias), then stores them to
// every rank.  The CONSUMER does not wait for this kernel: the sequential kernel runs concurrently (other stream) and
// polls the very elements it needs - every column buffer is filled with kColsSentinel before it is handed out, and an
// element is there when neither of its halves is the sentinel any more (flag-in-data, the low-latency protocol of
// collective libraries).  What the fences, tickets and flags at the END of this kernel are for is buffer RECYCLING,
// off the critical path: shard_recycle_kernel re-poisons a buffer only after every rank's push of that round has
// completely landed, and this kernel writes into a peer's buffer only after that peer has released it (credit).
//
// TWO HOPS for the row segment (two_hop != 0; more than two ranks).  The rank that stores row a of a sighted landmark
// owns the whole segment (i < a) of that slot: delivered directly it ships ~22 MB per scan at 8 GPUs, N = 10,000,
// m = 20, while the other ranks ship a few hundred KB - 56 us for that one push kernel, and every sequential kernel
// waits for it (bench.py run.per_rank, profiles/r2l_bench8_flag_in_data_direct.json).  So the segment travels the way an all-gather
// balances its links: 128-row blocks are assigned to RELAY ranks block-cyclically (block b -> rank b % R); the owner
// stores block b to itself and to the relay only; the relay's CTA of that block polls its own buffer (flag-in-data
// again) and forwards each element to the remaining ranks as it lands.  The column part (i > a + 1) is spread over
// the ranks already and goes out directly.
constexpr int kPushThreads = 128;
static_assert(kPushThreads == EKF_STRIP, "one CTA per 128-row block (relay assignment)");
__global__ void __launch_bounds__(kPushThreads) shard_push_cols_kernel(
    const double *sigma, EkfLay lay, int n_pad, int N, const int *guess, int m, PeerTable peers,
    const unsigned long long *credits, unsigned long long round_no, unsigned int *done_ctas, int *status, int two_hop) {
  __shared__ int s_guess[32], s_owner[32];
  __shared__ unsigned int s_prev;
  const int R = peers.count, self = peers.self;
  if (threadIdx.x < 32) {
    const int j = (int)threadIdx.x < m ? guess[threadIdx.x] : -1;
    s_guess[threadIdx.x] = j;
    int own = -1;
    if (j >= 0 && j < N) {                                   // the rank that stores rows a, a + 1 of landmark j
      const int a = 4 + 2 * j;
      for (int g = 0; g < R; ++g)
        if (a >= ekf_shard_bound(n_pad, g, R) && a < ekf_shard_bound(n_pad, g + 1, R)) own = g;
    }
    s_owner[threadIdx.x] = own;
  }
  // credit: rank q has re-poisoned the buffer this round writes into (the one of round_no - 2); two rounds old, so
  // this never waits in practice.  Bounded like every cross-GPU wait here.
  if ((int)threadIdx.x < R && round_no > 2) {
    unsigned long long t0, now, v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(credits + threadIdx.x) : "memory");
      if (v + 2 >= round_no) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (now - t0 >= 10000000000ull) { atomicMax(status, kErrPeerTimeout); break; }
    }
  }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int relay = two_hop ? (int)(blockIdx.x % (unsigned)R) : self;
  const bool own_i = i < n_pad && ekf_lay_owns_row(lay, i);
  bool any = false;
  for (int s0 = 0; s0 < m; s0 += 8) {
    double2 v[8];
    int kind[8];                                             // 0 nothing, 1 column part, 2 row segment
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      kind[u] = 0;
      const int slot = s0 + u;
      const int j = slot < m ? s_guess[slot] : -1;
      if (j < 0 || j >= N || i >= n_pad) continue;
      const int a = 4 + 2 * j;
      if (i > a + 1) {
        if (own_i) { v[u] = *reinterpret_cast<const double2 *>(sigma + ekf_lay_elem(lay, i, a)); kind[u] = 1; }
      } else if (i < a) {
        if (s_owner[slot] == self) {
          v[u].x = sigma[ekf_lay_elem(lay, a, i)];
          v[u].y = sigma[ekf_lay_elem(lay, a + 1, i)];
          kind[u] = 2;
        }
      }                                    // (the landmark's own 2 x 2 block comes from D, not from here)
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (kind[u] == 0) continue;
      any = true;
      const int64_t e = ((int64_t)(s0 + u) * n_pad + i) * 2;
      if (kind[u] == 1 || !two_hop || relay == self) {       // direct: every rank (an owner relays its own blocks itself)
        for (int q = 0; q < R; ++q) *reinterpret_cast<double2 *>(peers.cols[q] + e) = v[u];
      } else {                                               // hop 1: own copy + the relay's
        *reinterpret_cast<double2 *>(peers.cols[self] + e) = v[u];
        *reinterpret_cast<double2 *>(peers.cols[relay] + e) = v[u];
      }
    }
  }
  // hop 2: this rank relays the block: forward what the other owners leave in its buffer as it lands.  All slots of a
  // batch are polled together (independent loads in flight), not one after the other: a poll is an L2 round trip
  if (two_hop && relay == self && i < n_pad) {
    for (int s0 = 0; s0 < m; s0 += 8) {
      unsigned pending = 0;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int slot = s0 + u;
        if (slot >= m) continue;
        const int j = s_guess[slot], own = s_owner[slot];
        if (j < 0 || j >= N || own < 0 || own == self || i >= 4 + 2 * j) continue;
        pending |= 1u << u;
      }
      unsigned long long t0 = 0;
      unsigned polls = 0;
      while (pending) {
        unsigned long long b0[8], b1[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (!((pending >> u) & 1u)) continue;
          const double *src = peers.cols[self] + ((int64_t)(s0 + u) * n_pad + i) * 2;
          asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(b0[u]), "=l"(b1[u]) : "l"(src) : "memory");
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          if (!((pending >> u) & 1u)) continue;
          if (b0[u] == kColsSentinel || b1[u] == kColsSentinel) continue;
          const int64_t e = ((int64_t)(s0 + u) * n_pad + i) * 2;
          const int own = s_owner[s0 + u];
          const double2 w = make_double2(__longlong_as_double((long long)b0[u]), __longlong_as_double((long long)b1[u]));
          for (int q = 0; q < R; ++q)
            if (q != self && q != own) *reinterpret_cast<double2 *>(peers.cols[q] + e) = w;
          pending &= ~(1u << u);
          any = true;
        }
        if (pending && (++polls & 63u) == 0) {
          unsigned long long now;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
          if (t0 == 0) t0 = now;
          else if (now - t0 >= 10000000000ull) { atomicMax(status, kErrPeerTimeout); break; }
        }
      }
    }
  }
  // completion (for the recycling only): every CTA's stores are visible system-wide before the flags move
  if (__syncthreads_or(any ? 1 : 0)) __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_prev = atomicAdd(done_ctas, 1u);
  __syncthreads();
  if (s_prev == gridDim.x - 1) {
    if (threadIdx.x == 0) *done_ctas = 0;                    // ready for the next launch
    __threadfence_system();
    if ((int)threadIdx.x < R) {
      unsigned long long *dst = peers.flags[threadIdx.x] + self;
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(round_no) : "memory");
    }
  }
}

// Recycling of this rank's column buffer of round `round_no`, on the shard stream after the sequential kernel that
// read it: wait until EVERY rank's push of that round has completely landed here (their completion flags; a late
// store after the re-poisoning would look like fresh data two rounds later), fill the buffer with the sentinel again,
// then tell every rank that it may write round_no + 2 into it (credit).
__global__ void __launch_bounds__(256) shard_recycle_kernel(const unsigned long long *flags, PeerTable peers,
                                                            unsigned long long round_no, double *buf, int64_t doubles,
                                                            unsigned int *ticket, int *status) {
  __shared__ unsigned int s_prev;
  if ((int)threadIdx.x < peers.count) {
    unsigned long long t0, now, v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + threadIdx.x) : "memory");
      if (v >= round_no) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (now - t0 >= 10000000000ull) { atomicMax(status, kErrPeerTimeout); break; }
    }
  }
  __syncthreads();
  const double2 poison = make_double2(__longlong_as_double((long long)kColsSentinel),
                                      __longlong_as_double((long long)kColsSentinel));
  double2 *b2 = reinterpret_cast<double2 *>(buf);
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < doubles / 2; e += (int64_t)gridDim.x * blockDim.x)
    b2[e] = poison;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_prev = atomicAdd(ticket, 1u);
  __syncthreads();
  if (s_prev == gridDim.x - 1) {
    if (threadIdx.x == 0) *ticket = 0;
    __threadfence_system();
    if ((int)threadIdx.x < peers.count) {
      // credit array of rank q sits behind its completion flags: credits[self] = last round this rank has released
      unsigned long long *dst = peers.flags[threadIdx.x] + kMaxPeers + peers.self;
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(round_no) : "memory");
    }
  }
}
// consumer side: wait until every rank has pushed round `round_no` into this rank's buffer.  Bounded: a peer
// that never arrives (crashed process) latches kErrPeerTimeout after ~10 s instead of hanging the GPU.
__global__ void shard_wait_peers_kernel(const unsigned long long *flags, int count, unsigned long long round_no,
                                        int *status) {
  if ((int)threadIdx.x >= count) return;
  unsigned long long t0, now, v;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + threadIdx.x) : "memory");
    if (v >= round_no) return;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
  } while (now - t0 < 10000000000ull);
  atomicMax(status, kErrPeerTimeout);
}

// ---- begin phase of the per-scan path: prediction on the fly + speculative association -------------------------
// Replaces shard_predict_kernel + shard_guess_kernel on the sharded path:
//   * the prediction is not applied here at all: the candidates' pose-panel rows and the pose block are transformed
//     in registers (slam.cpp:194-198 restricted to the n x 3 panel) - the sequential kernel applies the same
//     transform to the replicated Pp when it starts (SeqParams::do_predict with side_mode);
//   * every slice is reduced to its landmark nearest to the sighting and only that one gets a Mahalanobis distance;
//     the last CTA of a sighting to finish (ticket) reduces the slices' candidates to guess[t].
// Delivery of the columns stays a separate, fully parallel kernel (shard_push_cols_kernel / shard_extract_cols_kernel):
// fused into the last CTA of each sighting it was one CTA walking 20,096 rows with dependent gathers - 23-36 us
// instead of 5 (profiles/r2g_shard_launches_*.csv).
struct BeginParams {
  const double4 *Pp, *D;
  const double *mu;
  const int *cnt;
  const double *pose;        // device (theta, x, y); nullptr = continuation round, no prediction
  const double *prev_xy;
  double q_noise, r_noise, thr;
  const double *xy;
  const int *ids;
  int *guess;
  double *part_d;
  int *part_k;
  unsigned int *ticket;
  int drop;                  // test-only fault injection: forget the speculation of this sighting (-1 = none)
};

// A P A^T + Qbar on the 3 x 3 pose block, lower triangle rules (the arithmetic of shard_predict_kernel / phase 0 of
// ekf_seq_kernel); ndy = -dy, pdx = dx
__device__ __forceinline__ void predict_block3(double P[3][3], double ndy, double pdx, double q_noise) {
  double Bm[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < i; ++j) P[j][i] = P[i][j];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    Bm[0][j] = P[0][j];
    Bm[1][j] = ndy * P[0][j] + P[1][j];
    Bm[2][j] = pdx * P[0][j] + P[2][j];
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    P[i][0] = Bm[i][0];
    P[i][1] = Bm[i][0] * ndy + Bm[i][1];
    P[i][2] = Bm[i][0] * pdx + Bm[i][2];
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) P[i][i] += q_noise;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < i; ++j) P[j][i] = P[i][j];
}

__global__ void __launch_bounds__(kGuessThreads) shard_begin_kernel(const BeginParams p) {
  __shared__ Cand s_red[kGuessThreads / 32];
  __shared__ int s_last;
  const int t = blockIdx.x, slice = blockIdx.y, P = gridDim.y;
  if (p.ids != nullptr) {                                  // known association: nothing to speculate
    if (slice == 0 && threadIdx.x == 0) p.guess[t] = p.ids[t];
  } else {
    const int cnt = *p.cnt;
    const Polar z = to_polar(p.xy[2 * t], p.xy[2 * t + 1]);
    double th = p.mu[0], px = p.mu[1], py = p.mu[2], ndy = 0.0, pdx = 0.0;
    double pb[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double4 r = p.Pp[i];
      pb[i][0] = r.x; pb[i][1] = r.y; pb[i][2] = r.z;
    }
    const bool pred = p.pose != nullptr;
    if (pred) {
      th = normalize_angle(p.pose[0]); px = p.pose[1]; py = p.pose[2];          // slam.cpp:186-188
      ndy = -(py - p.prev_xy[1]); pdx = px - p.prev_xy[0];                        // slam.cpp:190-196
      predict_block3(pb, ndy, pdx, p.q_noise);
    }
    // This is a SPECULATION (it only decides which columns travel; the sequential kernel re-derives every association
    // exactly and stops if it ever needs a landmark that was not fetched), so the slice is first reduced to its
    // candidate NEAREST to the sighting's world position (two loads and three flops per landmark instead of the ~400
    // instructions of a Mahalanobis distance), and only that one is evaluated in full.  The landmark with the smallest
    // Mahalanobis distance is the nearest one unless the uncertainties of neighbouring landmarks differ wildly.
    const double wx = px + z.r * cos(z.phi + th), wy = py + z.r * sin(z.phi + th);
    Cand best{INFINITY, 0x7fffffff};
    for (int k = slice * (int)blockDim.x + threadIdx.x; k < cnt; k += P * (int)blockDim.x) {
      const double2 lm = *reinterpret_cast<const double2 *>(p.mu + 4 + 2 * k);
      const double ex = lm.x - wx, ey = lm.y - wy;
      best = cand_min(best, Cand{ex * ex + ey * ey, k});
    }
    best = warp_argmin(best);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
      Cand c = s_red[0];
      for (int w = 1; w < (int)blockDim.x / 32; ++w) c = cand_min(c, s_red[w]);
      if (c.k < cnt) {                                     // the slice's nearest landmark, in full: slam.cpp:364-404
        const int k = c.k, a = 4 + 2 * k;
        const double2 lm = *reinterpret_cast<const double2 *>(p.mu + a);
        double4 pa = p.Pp[a], pbr = p.Pp[a + 1];
        const double4 dk = p.D[k];
        if (pred) {
          pa.y = pa.x * ndy + pa.y; pa.z = pa.x * pdx + pa.z;
          pbr.y = pbr.x * ndy + pbr.y; pbr.z = pbr.x * pdx + pbr.z;
        }
        const Lin L = linearize(th, px, py, lm.x, lm.y);
        const double par3[3] = {pa.x, pa.y, pa.z}, pbr3[3] = {pbr.x, pbr.y, pbr.z};
        const Block5 B = make_block5(pb, par3, pbr3, dk.x, dk.y, dk.z);
        const Innov I = innovation(B, L, p.r_noise);
        const double zd0 = z.r - L.zhat_r, zd1 = normalize_angle(z.phi - L.zhat_b);
        double i00, i01, i11;
        c.d = mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, i00, i01, i11);
        if (!(c.d < INFINITY)) c = Cand{INFINITY, 0x7fffffff};
      }
      p.part_d[t * kGuessMaxSlices + slice] = c.d;
      p.part_k[t * kGuessMaxSlices + slice] = c.k;
      __threadfence();
      const unsigned int prev = atomicAdd(p.ticket + t, 1u);
      s_last = (prev == (unsigned int)(P - 1));
      if (s_last) {
        __threadfence();
        p.ticket[t] = 0;                                     // ready for the next launch
        Cand g{INFINITY, 0x7fffffff};
        for (int q = 0; q < P; ++q)
          g = cand_min(g, Cand{__ldcg(p.part_d + t * kGuessMaxSlices + q), __ldcg(p.part_k + t * kGuessMaxSlices + q)});
        // generous acceptance (4 x threshold), as in shard_guess_kernel
        int gj = (g.k < cnt && g.d <= 4.0 * p.thr) ? g.k : -1;
        if (t == p.drop) gj = -1;
        p.guess[t] = gj;
      }
    }
  }
}

// ---- side state <-> Sigma -----------------------------------------------------------------------
// side[0 .. 4 n_pad) = Pp rows (x, y, z, 0), side[4 n_pad ..) = D blocks (aa, ba, bb, 0): this rank's
// owned rows, zero elsewhere (summed over the ranks afterwards)
__global__ void shard_side_extract_kernel(const double *sigma, EkfLay lay, int n_pad, int N, double4 *Pp, double4 *D) {
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  for (int i = gtid; i < n_pad; i += gridDim.x * blockDim.x) {
    double4 v = make_double4(0.0, 0.0, 0.0, 0.0);
    if (ekf_lay_owns_row(lay, i) && i != 3) {
      const double *r = sigma + ekf_lay_elem(lay, i, 0);
      v.x = r[0];
      if (i >= 1) v.y = r[1];
      if (i >= 2) v.z = r[2];
      if (i < 3) {                                         // pose block: mirror the stored lower part
        if (i == 0) { v.y = 0.0; v.z = 0.0; }
        if (i == 1) v.z = 0.0;
      }
    }
    Pp[i] = v;
  }
  for (int k = gtid; k < N; k += gridDim.x * blockDim.x) {
    const int a = 4 + 2 * k;
    double4 v = make_double4(0.0, 0.0, 0.0, 0.0);
    if (ekf_lay_owns_row(lay, a)) {
      const double *dg = sigma + ekf_lay_elem(lay, a, a);
      v.x = dg[0]; v.y = dg[EKF_STRIP]; v.z = dg[EKF_STRIP + 1];
    }
    D[k] = v;
  }
}
// after the all-reduce the pose block rows hold only their lower part: make it fully symmetric
__global__ void shard_side_fix_pose_kernel(double4 *Pp) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    double4 r0 = Pp[0], r1 = Pp[1], r2 = Pp[2];
    r0.y = r1.x; r0.z = r2.x; r1.z = r2.y;
    Pp[0] = r0; Pp[1] = r1;
  }
}
// Pp / D -> the Sigma entries they mirror (owned rows only); run after the covariance pass
__global__ void shard_side_writeback_kernel(double *sigma, EkfLay lay, int n_pad, int N, const double4 *Pp,
                                            const double4 *D) {
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  for (int i = gtid; i < n_pad; i += gridDim.x * blockDim.x) {
    if (!ekf_lay_owns_row(lay, i) || i == 3) continue;
    const double4 v = Pp[i];
    double *r = sigma + ekf_lay_elem(lay, i, 0);
    r[0] = v.x;
    if (i >= 1) r[1] = v.y;
    if (i >= 2) r[2] = v.z;
  }
  for (int k = gtid; k < N; k += gridDim.x * blockDim.x) {
    const int a = 4 + 2 * k;
    if (!ekf_lay_owns_row(lay, a)) continue;
    const double4 v = D[k];
    double *dg = sigma + ekf_lay_elem(lay, a, a);
    dg[0] = v.x; dg[EKF_STRIP] = v.y; dg[EKF_STRIP + 1] = v.z;
  }
}

}  // namespace ekf
