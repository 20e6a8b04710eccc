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
// redundantly by every CTA in shared memory.  Exactly ONE grid barrier per measurement: local
// arg-min -> slot -> barrier -> every CTA reduces the slots identically, reads the chosen
// landmark's data through L2 (ld.cg) and proceeds on its own rows.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_layout.h"
#include "ekf_math.cuh"

namespace ekf {

constexpr int kSeqThreads = 128;
constexpr int kMaxChunkM = 24;            // measurements per sequential launch / covariance pass
constexpr int kMaxK2 = 2 * kMaxChunkM;    // panel columns in use per pass

// status codes latched on the device (values of ekf_status in include/ekf_slam_b200.h)
constexpr int kErrCapacity = 2;
constexpr int kErrBadId = 4;

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
  // grid barrier + arg-min slots
  unsigned long long *bar;
  unsigned long long bar_base;
  double *slot_d;  // 2 * gridDim
  int *slot_k;     // 2 * gridDim
};

__device__ __forceinline__ void grid_barrier(unsigned long long *bar, unsigned long long target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1ULL);
    unsigned long long v;
    do {
      asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(bar) : "memory");
    } while (v < target);
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ double ldcg_d(const double *p) { return __ldcg(p); }
__device__ __forceinline__ double4 ldcg_d4(const double4 *p) {
  double4 v;
  asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(v.z), "=d"(v.w) : "l"((const double *)p + 2));
  return v;
}

// Everything a CTA's threads need to process their rows for the current measurement.
struct Bcast {
  int j;        // associated landmark, or -1 when the measurement is skipped
  int is_new;   // the landmark's state is (re)initialised from the sighting
  Lin L;        // linearisation at landmark j
  double lx, ly;        // landmark j position used (after initialisation)
  double m00, m01, m11; // S^-1
  double zd0, zd1;      // innovation, bearing wrapped
  double Tp[3][2];      // pose rows of G
  double daa, dba, dbb; // diagonal block of landmark j (current)
};

__global__ void __launch_bounds__(kSeqThreads) ekf_seq_kernel(const SeqParams p) {
  __shared__ double s_ps[3];      // pose state (theta, x, y)
  __shared__ double s_pb[3][3];   // pose block of Sigma_cur, full symmetric
  __shared__ Cand s_red[kSeqThreads / 32];
  __shared__ Bcast s_b;
  __shared__ double s_wa[kMaxK2], s_wb[kMaxK2];
  __shared__ int s_cnt;
  __shared__ double s_pred[2];    // (-dy, dx)

  const int tid = threadIdx.x;
  const int nthreads_total = gridDim.x * blockDim.x;
  const int gtid = blockIdx.x * blockDim.x + tid;
  const int n_pad = p.n_pad;
  double *const strip0 = p.sigma;  // strip 0 starts at offset 0: row i at i*128

  // ---------------- phase 0: pose injection, prediction, extraction of Pp / D -------------
  if (tid == 0) {
    s_cnt = *p.cnt;
    double P[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) P[i][j] = P[j][i] = strip0[i * EKF_STRIP + j];
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
    if (blockIdx.x == 0) {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) strip0[i * EKF_STRIP + j] = P[i][j];
        p.Pp[i] = make_double4(s_pb[i][0], s_pb[i][1], s_pb[i][2], 0.0);
        p.mu[i] = s_ps[i];
      }
    }
  }
  __syncthreads();
  {
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
      const double *dg = p.sigma + ekf_elem(n_pad, a, a);
      const double2 d0 = *reinterpret_cast<const double2 *>(dg);                // (a,a), (a,b) scratch
      const double2 d1 = *reinterpret_cast<const double2 *>(dg + EKF_STRIP);    // (b,a), (b,b)
      p.D[k] = make_double4(d0.x, d1.x, d1.y, 0.0);
    }
  }

  // ---------------- measurement loop --------------------------------------------------------
  for (int t = 0; t < p.m; ++t) {
    const double mx = p.xy[2 * t], my = p.xy[2 * t + 1];
    const Polar z = to_polar(mx, my);                        // slam.cpp:345-347
    const double th = s_ps[0], px = s_ps[1], py = s_ps[2];
    const int cnt = s_cnt;
    const bool unknown = (p.ids == nullptr);
    const int par = t & 1;

    // ---- A: Mahalanobis distance to every mapped landmark (own ones), local arg-min ------
    if (unknown) {
      Cand best{INFINITY, 0x7fffffff};
      for (int k = gtid; k < cnt; k += nthreads_total) {
        const int a = 4 + 2 * k;
        const double2 lm = *reinterpret_cast<const double2 *>(p.mu + a);
        const double4 pa = p.Pp[a], pbr = p.Pp[a + 1];
        const double4 dk = p.D[k];
        const Lin L = linearize(th, px, py, lm.x, lm.y);     // slam.cpp:364-392
        const double par3[3] = {pa.x, pa.y, pa.z}, pbr3[3] = {pbr.x, pbr.y, pbr.z};
        const Block5 B = make_block5(s_pb, par3, pbr3, dk.x, dk.y, dk.z);
        const Innov I = innovation(B, L, p.r_noise);         // slam.cpp:395
        const double zd0 = z.r - L.zhat_r;
        const double zd1 = normalize_angle(z.phi - L.zhat_b);  // slam.cpp:398-399
        const double d = mahalanobis(I.s00, I.s01, I.s11, zd0, zd1);  // slam.cpp:401
        best = cand_min(best, Cand{d, k});
      }
      best = warp_argmin(best);
      if ((tid & 31) == 0) s_red[tid >> 5] = best;
      __syncthreads();
      if (tid == 0) {
        Cand c = s_red[0];
#pragma unroll
        for (int w = 1; w < kSeqThreads / 32; ++w) c = cand_min(c, s_red[w]);
        p.slot_d[par * gridDim.x + blockIdx.x] = c.d;
        p.slot_k[par * gridDim.x + blockIdx.x] = c.k;
      }
    }

    // ---- B: the one grid barrier of this measurement ----------------------------------------
    grid_barrier(p.bar, p.bar_base + (unsigned long long)(t + 1) * gridDim.x);

    // ---- C: decision + linearisation at the chosen landmark (identical in every CTA) --------
    if (tid < 32) {
      int j = -1, is_new = 0, err = 0;
      if (unknown) {
        Cand c{INFINITY, 0x7fffffff};
        for (int b = tid; b < (int)gridDim.x; b += 32)
          c = cand_min(c, Cand{ldcg_d(p.slot_d + par * gridDim.x + b), __ldcg(p.slot_k + par * gridDim.x + b)});
        c = warp_argmin(c);
        // the tentative landmark sits last with distance == threshold and wins only on a
        // strictly smaller value: slam.cpp:406-408,418-440
        if (c.d <= p.thr && c.k < cnt) j = c.k;
        else if (cnt < p.N) { j = cnt; is_new = 1; }
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
      if (tid == 0) {
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
          const double det = I.s00 * I.s11 - I.s01 * I.s01;
          const double m00 = I.s11 / det, m01 = -I.s01 / det, m11 = I.s00 / det;
          const double zd0 = z.r - L.zhat_r;
          const double zd1 = normalize_angle(z.phi - L.zhat_b);   // slam.cpp:479-480
          s_b.L = L; s_b.lx = lx; s_b.ly = ly;
          s_b.m00 = m00; s_b.m01 = m01; s_b.m11 = m11;
          s_b.zd0 = zd0; s_b.zd1 = zd1;
          s_b.daa = dj.x; s_b.dba = dj.y; s_b.dbb = dj.z;
          // pose rows: gain, state and pose-block update (slam.cpp:476,482,485-486 restricted
          // to rows/cols 0..2), done redundantly by every CTA
          double Kp[3][2];
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            s_b.Tp[c][0] = I.T[c][0]; s_b.Tp[c][1] = I.T[c][1];
            Kp[c][0] = I.T[c][0] * m00 + I.T[c][1] * m01;
            Kp[c][1] = I.T[c][0] * m01 + I.T[c][1] * m11;
          }
          double ps_new[3];
#pragma unroll
          for (int c = 0; c < 3; ++c) ps_new[c] = s_ps[c] + (Kp[c][0] * zd0 + Kp[c][1] * zd1);
          ps_new[0] = normalize_angle(ps_new[0]);            // slam.cpp:488
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int c2 = 0; c2 <= c; ++c2) {
              const double v = s_pb[c][c2] - (Kp[c][0] * I.T[c2][0] + Kp[c][1] * I.T[c2][1]);
              s_pb[c][c2] = v; s_pb[c2][c] = v;
            }
#pragma unroll
          for (int c = 0; c < 3; ++c) s_ps[c] = ps_new[c];
          if (blockIdx.x == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              p.W[ekf_panel_elem(p.kmax, c, 2 * t)] = I.T[c][0];
              p.W[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = I.T[c][1];
              p.V[ekf_panel_elem(p.kmax, c, 2 * t)] = Kp[c][0];
              p.V[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = Kp[c][1];
              p.Pp[c] = make_double4(s_pb[c][0], s_pb[c][1], s_pb[c][2], 0.0);
            }
          }
        } else if (blockIdx.x == 0) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            p.W[ekf_panel_elem(p.kmax, c, 2 * t)] = 0.0; p.W[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = 0.0;
            p.V[ekf_panel_elem(p.kmax, c, 2 * t)] = 0.0; p.V[ekf_panel_elem(p.kmax, c, 2 * t + 1)] = 0.0;
          }
        }
      }
    }
    __syncthreads();
    const int j = s_b.j;
    // panel history of the chosen landmark's two rows (column operand of the correction)
    if (j >= 0) {
      const int a = 4 + 2 * j;
      for (int q = tid; q < 2 * t; q += blockDim.x) {
        s_wa[q] = ldcg_d(p.W + ekf_panel_elem(p.kmax, a, q));
        s_wb[q] = ldcg_d(p.W + ekf_panel_elem(p.kmax, a + 1, q));
      }
    }
    __syncthreads();

    // ---- D: own rows: gather the two landmark columns, gain, state, Pp, D, panels ------------
    if (j < 0) {
      for (int k = gtid; k < p.N; k += nthreads_total) {
        const int a = 4 + 2 * k;
        const double2 zero = make_double2(0.0, 0.0);
        *reinterpret_cast<double2 *>(p.W + ekf_panel_elem(p.kmax, a, 2 * t)) = zero;
        *reinterpret_cast<double2 *>(p.W + ekf_panel_elem(p.kmax, a, 2 * t + 1)) = zero;
        *reinterpret_cast<double2 *>(p.V + ekf_panel_elem(p.kmax, a, 2 * t)) = zero;
        *reinterpret_cast<double2 *>(p.V + ekf_panel_elem(p.kmax, a, 2 * t + 1)) = zero;
      }
    } else {
      const Lin L = s_b.L;
      const double m00 = s_b.m00, m01 = s_b.m01, m11 = s_b.m11;
      const double zd0 = s_b.zd0, zd1 = s_b.zd1;
      const int aj = 4 + 2 * j;
      for (int k = gtid; k < p.N; k += nthreads_total) {
        const int a = 4 + 2 * k;
        double q00, q01, q10, q11;  // Sigma_cur[{a,b}_k, {a,b}_j]
        if (k == j) {
          q00 = s_b.daa; q01 = s_b.dba; q10 = s_b.dba; q11 = s_b.dbb;
        } else {
          if (k > j) {
            const double *src = p.sigma + ekf_elem(n_pad, a, aj);
            const double2 r0 = __ldg(reinterpret_cast<const double2 *>(src));
            const double2 r1 = __ldg(reinterpret_cast<const double2 *>(src + EKF_STRIP));
            q00 = r0.x; q01 = r0.y; q10 = r1.x; q11 = r1.y;
          } else {
            const double *src = p.sigma + ekf_elem(n_pad, aj, a);
            const double2 c0 = __ldg(reinterpret_cast<const double2 *>(src));
            const double2 c1 = __ldg(reinterpret_cast<const double2 *>(src + EKF_STRIP));
            q00 = c0.x; q10 = c0.y; q01 = c1.x; q11 = c1.y;
          }
          // deferred downdates of earlier measurements of this scan
          const double *vrow = p.V + ekf_panel_elem(p.kmax, a, 0);
          for (int q = 0; q < 2 * t; ++q) {
            const double2 v = *reinterpret_cast<const double2 *>(vrow + (int64_t)q * EKF_PBLK);
            const double wa = s_wa[q], wb = s_wb[q];
            q00 -= v.x * wa; q01 -= v.x * wb;
            q10 -= v.y * wa; q11 -= v.y * wb;
          }
        }
        double4 pa = p.Pp[a], pbr = p.Pp[a + 1];
        const double rowa[5] = {pa.x, pa.y, pa.z, q00, q01};
        const double rowb[5] = {pbr.x, pbr.y, pbr.z, q10, q11};
        double ga0, ga1, gb0, gb1;
        row_times_ht(rowa, L, ga0, ga1);                     // G = Sigma H^T
        row_times_ht(rowb, L, gb0, gb1);
        const double ka0 = ga0 * m00 + ga1 * m01, ka1 = ga0 * m01 + ga1 * m11;   // K = G S^-1
        const double kb0 = gb0 * m00 + gb1 * m01, kb1 = gb0 * m01 + gb1 * m11;
        double2 lm = *reinterpret_cast<const double2 *>(p.mu + a);
        if (k == j && s_b.is_new) { lm.x = s_b.lx; lm.y = s_b.ly; }
        lm.x = lm.x + (ka0 * zd0 + ka1 * zd1);               // slam.cpp:482
        lm.y = lm.y + (kb0 * zd0 + kb1 * zd1);
        *reinterpret_cast<double2 *>(p.mu + a) = lm;
        // Sigma_cur[i, pose c] -= K[i,:] . G[c,:]
        pa.x -= ka0 * s_b.Tp[0][0] + ka1 * s_b.Tp[0][1];
        pa.y -= ka0 * s_b.Tp[1][0] + ka1 * s_b.Tp[1][1];
        pa.z -= ka0 * s_b.Tp[2][0] + ka1 * s_b.Tp[2][1];
        pbr.x -= kb0 * s_b.Tp[0][0] + kb1 * s_b.Tp[0][1];
        pbr.y -= kb0 * s_b.Tp[1][0] + kb1 * s_b.Tp[1][1];
        pbr.z -= kb0 * s_b.Tp[2][0] + kb1 * s_b.Tp[2][1];
        p.Pp[a] = pa; p.Pp[a + 1] = pbr;
        double4 dk = p.D[k];
        dk.x -= ka0 * ga0 + ka1 * ga1;
        dk.y -= kb0 * ga0 + kb1 * ga1;
        dk.z -= kb0 * gb0 + kb1 * gb1;
        p.D[k] = dk;
        *reinterpret_cast<double2 *>(p.W + ekf_panel_elem(p.kmax, a, 2 * t)) = make_double2(ga0, gb0);
        *reinterpret_cast<double2 *>(p.W + ekf_panel_elem(p.kmax, a, 2 * t + 1)) = make_double2(ga1, gb1);
        *reinterpret_cast<double2 *>(p.V + ekf_panel_elem(p.kmax, a, 2 * t)) = make_double2(ka0, kb0);
        *reinterpret_cast<double2 *>(p.V + ekf_panel_elem(p.kmax, a, 2 * t + 1)) = make_double2(ka1, kb1);
      }
    }
    __syncthreads();  // s_b / s_wa are rewritten by the next measurement
  }

  // ---------------- epilogue: slam.cpp:273-274,291 / :490-491,508 ---------------------------------
  if (blockIdx.x == 0 && tid == 0) {
    p.mu[0] = s_ps[0]; p.mu[1] = s_ps[1]; p.mu[2] = s_ps[2];
    *p.cnt = s_cnt;
    *p.cnt_out = s_cnt;
    p.pose_out[0] = s_ps[0]; p.pose_out[1] = s_ps[1]; p.pose_out[2] = s_ps[2];
    if (p.finish) { p.prev_xy[0] = s_ps[1]; p.prev_xy[1] = s_ps[2]; }
  }
}

}  // namespace ekf
