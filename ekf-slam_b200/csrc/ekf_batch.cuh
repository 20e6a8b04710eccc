// Batched mode: B independent small EKF-SLAM filters, ONE CTA (one warp) per filter, the filter's
// whole covariance resident in shared memory for the duration of a scan update.
//
// Per filter this is the complete sensor callback of the reference (nuslam/src/slam.cpp):
//   :186-198 / :323-335  pose injection + Sigma <- A Sigma A^T + Qbar
//   :201-271             known-association loop            (ids != nullptr)
//   :338-489             Mahalanobis association + update  (ids == nullptr)
// applied measurement by measurement exactly in the reference's order (no deferral: with Sigma in
// shared memory the rank-2 downdate `sigma = (I - K*Hj) * sigma`, :264-265 / :485-486, costs
// n(n+1)/2 shared-memory updates).  BASELINE config 3: 65,536 filters x 20 landmarks (n = 43).
//
// HBM layout (struct of arrays, everything per filter contiguous so a CTA's loads coalesce):
//   sigma[B][tri]   packed lower triangle, row-major: (i, j<=i) at i(i+1)/2 + j, EXTERNAL indices
//                   (theta, x, y, m1x, m1y, ...), tri = n(n+1)/2
//   mu[B][n]        state;   cnt[B] counter_obstacles;   prev[B][2] previous filter translation
// HBM traffic per filter-step: read + write of sigma and mu, the sightings, the pose:
//   8 n (n+1) + 16 n + 16 m + 24 bytes  (16,168 B at n = 43, m = 20).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_math.cuh"

namespace ekf {

struct BatchParams {
  int B, N, n, tri;
  double q_noise, r_noise, thr;
  double *sigma, *mu, *prev;
  int *cnt;
  int *status;             // latched error (max over filters)
  const uchar2 *tri_map;   // [tri] (row, column) of packed lower-triangle element e (n <= 131 < 256)
  // this step (device pointers)
  const double *pose;      // [B][3] (theta, x, y) composed odometry pose
  const double *xy;        // [B][m][2]
  const int *ids;          // [B][m] or nullptr (unknown association)
  const int *m_count;      // [B] or nullptr (every filter has m sightings)
  int m;
  int *assoc_out;          // [B][m] or nullptr
  int *cnt_out;            // [B] or nullptr
  double *pose_out;        // [B][3] or nullptr
};

constexpr int kBatchThreads = 32;
constexpr int kBatchErrCapacity = 2, kBatchErrBadId = 4;   // ekf_status values

__host__ __device__ inline size_t batch_smem_bytes(int n, int m) {
  const size_t tri = (size_t)n * (n + 1) / 2;
  return (tri + (size_t)n + 4 * (size_t)n + 2 * (size_t)m + 8) * sizeof(double);
}

// Sigma_0 = diag(0, 0, 0, prior * I_2N) for every filter: slam.cpp:128-132
__global__ void batch_init_prior_kernel(double *sigma, int B, int n, int tri, double prior) {
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)B * n) return;
  const int f = (int)(g / n), i = (int)(g % n);
  if (i >= 3) sigma[(size_t)f * tri + (size_t)i * (i + 1) / 2 + i] = prior;
}

__device__ __forceinline__ int tri_idx(int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; }

__global__ void __launch_bounds__(kBatchThreads) ekf_batch_kernel(const BatchParams p) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.n, tri = p.tri;
  double *S = sm;                 // packed lower triangle
  double *mu = S + tri;           // state
  double2 *G = reinterpret_cast<double2 *>(mu + n + (n & 1));   // G = Sigma H^T   (n x 2)
  double2 *K = G + n;                                           // K = G S^-1      (n x 2)
  double *ZP = reinterpret_cast<double *>(K + n);               // polar sightings (m x 2)
  const int f = blockIdx.x;
  const int lane = threadIdx.x;
  if (f >= p.B) return;

  // ---- load the filter (coalesced) ---------------------------------------------------------
  {
    const double *gs = p.sigma + (size_t)f * tri;
    for (int e = lane; e < tri; e += 32) S[e] = gs[e];
    const double *gm = p.mu + (size_t)f * n;
    for (int i = lane; i < n; i += 32) mu[i] = gm[i];
  }
  int cnt = p.cnt[f];
  __syncwarp();

  // ---- pose injection + prediction: slam.cpp:186-198 -------------------------------------------
  {
    const double th = normalize_angle(p.pose[3 * f]);
    const double x = p.pose[3 * f + 1], y = p.pose[3 * f + 2];
    const double dx = x - p.prev[2 * f], dy = y - p.prev[2 * f + 1];
    for (int i = 3 + lane; i < n; i += 32) {       // rows >= 3 of the pose panel: C = B A^T
      const int r = i * (i + 1) / 2;
      const double s0 = S[r];
      S[r + 1] = s0 * (-dy) + S[r + 1];
      S[r + 2] = s0 * dx + S[r + 2];
    }
    if (lane == 0) {
      double P[3][3], Bm[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) P[i][j] = P[j][i] = S[i * (i + 1) / 2 + j];
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
      for (int i = 0; i < 3; ++i) P[i][i] += p.q_noise;
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) S[i * (i + 1) / 2 + j] = P[i][j];
      mu[0] = th; mu[1] = x; mu[2] = y;
    }
  }
  __syncwarp();

  // ---- measurement loop ----------------------------------------------------------------------
  const int mf = p.m_count ? min(p.m_count[f], p.m) : p.m;
  const bool unknown = (p.ids == nullptr);
  int err = 0;
  // polar form of every sighting of the scan, one lane each (slam.cpp:345-347 does not depend on the state)
  for (int t = lane; t < mf; t += 32) {
    const Polar z = to_polar(p.xy[((size_t)f * p.m + t) * 2], p.xy[((size_t)f * p.m + t) * 2 + 1]);
    ZP[2 * t] = z.r; ZP[2 * t + 1] = z.phi;
  }
  __syncwarp();
  for (int t = 0; t < mf; ++t) {
    const Polar z{ZP[2 * t], ZP[2 * t + 1]};
    const double th = mu[0], px = mu[1], py = mu[2];
    double pb[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) pb[i][j] = pb[j][i] = S[i * (i + 1) / 2 + j];

    int j = -1, is_new = 0;
    // linearisation of the landmark being fused: in the matched case it is the winning lane's own
    // candidate evaluation, broadcast with shuffles instead of being recomputed by every lane
    Lin L;
    double m00 = 0.0, m01 = 0.0, m11 = 0.0, zd0 = 0.0, zd1 = 0.0;
    bool have = false;
    if (unknown) {
      Cand best{INFINITY, 0x7fffffff};
      // the lane's best candidate so far and what the update needs of it: the four distinct Jacobian
      // quotients (dx/sd, dy/sd, dx/d, dy/d), S^-1 and the innovation
      double r_xs = 0.0, r_ys = 0.0, r_xd = 0.0, r_yd = 0.0, r_m00 = 0.0, r_m01 = 0.0, r_m11 = 0.0, r_z0 = 0.0,
             r_z1 = 0.0;
      for (int k = lane; k < cnt; k += 32) {                   // slam.cpp:361-404
        const int a = 3 + 2 * k, ra = a * (a + 1) / 2, rb = (a + 1) * (a + 2) / 2;
        const double pa[3] = {S[ra], S[ra + 1], S[ra + 2]}, pbr[3] = {S[rb], S[rb + 1], S[rb + 2]};
        const Lin Lk = linearize(th, px, py, mu[a], mu[a + 1]);
        const Block5 Bk = make_block5(pb, pa, pbr, S[ra + a], S[rb + a], S[rb + a + 1]);
        const Innov I = innovation(Bk, Lk, p.r_noise);
        const double z0 = z.r - Lk.zhat_r, z1 = normalize_angle(z.phi - Lk.zhat_b);
        double i00, i01, i11;
        const double d = mahalanobis(I.s00, I.s01, I.s11, z0, z1, i00, i01, i11);
        const Cand nb = cand_min(best, Cand{d, k});
        if (nb.k == k) {
          r_xs = Lk.h0[3]; r_ys = Lk.h0[4]; r_xd = Lk.h1[4]; r_yd = -Lk.h1[3];
          r_m00 = i00; r_m01 = i01; r_m11 = i11; r_z0 = z0; r_z1 = z1;
        }
        best = nb;
      }
      best = warp_argmin(best);
      if (best.d <= p.thr && best.k < cnt) {                   // slam.cpp:406-408,418-440
        j = best.k;
        const int src = j & 31;                                // candidate k lives in lane k % 32
        const double xs = __shfl_sync(0xffffffffu, r_xs, src), ys = __shfl_sync(0xffffffffu, r_ys, src);
        const double xd = __shfl_sync(0xffffffffu, r_xd, src), yd = __shfl_sync(0xffffffffu, r_yd, src);
        m00 = __shfl_sync(0xffffffffu, r_m00, src); m01 = __shfl_sync(0xffffffffu, r_m01, src);
        m11 = __shfl_sync(0xffffffffu, r_m11, src);
        zd0 = __shfl_sync(0xffffffffu, r_z0, src); zd1 = __shfl_sync(0xffffffffu, r_z1, src);
        L.h0[0] = 0.0;  L.h0[1] = -xs; L.h0[2] = -ys; L.h0[3] = xs;  L.h0[4] = ys;
        L.h1[0] = -1.0; L.h1[1] = yd;  L.h1[2] = -xd; L.h1[3] = -yd; L.h1[4] = xd;
        have = true;
      }
      else if (cnt < p.N) { j = cnt; is_new = 1; }
      else err = kBatchErrCapacity;
    } else {
      const int id = p.ids[(size_t)f * p.m + t];
      if (id < 0) { /* padding / DELETE: skipped (slam.cpp:205) */ }
      else if (id >= p.N) err = kBatchErrBadId;
      else { j = id; is_new = (mu[3 + 2 * id] == 0.0 && mu[4 + 2 * id] == 0.0); }   // slam.cpp:213
    }
    if (lane == 0 && p.assoc_out) p.assoc_out[(size_t)f * p.m + t] = j;
    if (j < 0) continue;                                       // uniform across the warp

    // ---- update with landmark j: slam.cpp:443-488 ----------------------------------------------------
    const int a = 3 + 2 * j, ra = a * (a + 1) / 2, rb = (a + 1) * (a + 2) / 2;
    double lx, ly;
    if (is_new) {                                              // slam.cpp:214-215 / :351-354
      lx = px + z.r * cos(z.phi + th);
      ly = py + z.r * sin(z.phi + th);
      if (unknown) cnt += 1;
    } else { lx = mu[a]; ly = mu[a + 1]; }
    if (!have) {                                               // new landmark / known association
      L = linearize(th, px, py, lx, ly);
      const double pa[3] = {S[ra], S[ra + 1], S[ra + 2]}, pbr[3] = {S[rb], S[rb + 1], S[rb + 2]};
      const Block5 Bk = make_block5(pb, pa, pbr, S[ra + a], S[rb + a], S[rb + a + 1]);
      const Innov I = innovation(Bk, L, p.r_noise);
      zd0 = z.r - L.zhat_r; zd1 = normalize_angle(z.phi - L.zhat_b);
      (void)mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, m00, m01, m11);
    }
    __syncwarp();
    if (is_new && lane == 0) { mu[a] = lx; mu[a + 1] = ly; }
    __syncwarp();
    for (int i = lane; i < n; i += 32) {                       // G = Sigma H^T, K = G S^-1, state
      const double row[5] = {S[tri_idx(i, 0)], S[tri_idx(i, 1)], S[tri_idx(i, 2)], S[tri_idx(i, a)],
                             S[tri_idx(i, a + 1)]};
      double g0, g1;
      row_times_ht(row, L, g0, g1);
      const double k0 = g0 * m00 + g1 * m01, k1 = g0 * m01 + g1 * m11;
      G[i] = make_double2(g0, g1);
      K[i] = make_double2(k0, k1);
      double v = mu[i] + (k0 * zd0 + k1 * zd1);                // slam.cpp:482
      if (i == 0) v = normalize_angle(v);                      // slam.cpp:488
      mu[i] = v;
    }
    __syncwarp();
    // Sigma -= K G^T on the stored triangle, element by element in packed order: 32 consecutive
    // elements per step keep every lane busy (a column-per-lane walk idles half of them on the triangle);
    // (row, column) of an element comes from a small table that stays in L1
    for (int e = lane; e < tri; e += 32) {
      const uchar2 ic = __ldg(p.tri_map + e);
      const double2 ki = K[ic.x], gc = G[ic.y];
      S[e] = S[e] - (ki.x * gc.x + ki.y * gc.y);
    }
    __syncwarp();
  }

  // ---- store back ----------------------------------------------------------------------------------
  {
    double *gs = p.sigma + (size_t)f * tri;
    for (int e = lane; e < tri; e += 32) gs[e] = S[e];
    double *gm = p.mu + (size_t)f * n;
    for (int i = lane; i < n; i += 32) gm[i] = mu[i];
    if (lane == 0) {
      p.cnt[f] = cnt;
      p.prev[2 * f] = mu[1]; p.prev[2 * f + 1] = mu[2];       // slam.cpp:291 / :508
      if (p.cnt_out) p.cnt_out[f] = cnt;
      if (p.pose_out) { p.pose_out[3 * f] = mu[0]; p.pose_out[3 * f + 1] = mu[1]; p.pose_out[3 * f + 2] = mu[2]; }
      if (err) atomicMax(p.status, err);
    }
  }
}

}  // namespace ekf
