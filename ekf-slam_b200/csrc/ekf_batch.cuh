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
  int f0;                  // first filter of this launch (CTA b works on filter f0 + b; chunked host-buffer steps)
  double q_noise, r_noise, thr;
  double *sigma, *mu, *prev;
  int *cnt;
  int *status;             // latched error (max over filters)
  const uchar2 *tri_map;   // [tri] (row, column) of packed lower-triangle element e (n <= 131 < 256)
  const ushort4 *blk_map;  // [nblk] 2x2 blocks of the stored triangle: (row i0, column j0, packed index of (i0, j0),
                           // mask of the block's stored elements: 1 (i0,j0) 2 (i0,j0+1) 4 (i0+1,j0) 8 (i0+1,j0+1))
  int nblk;
  const unsigned short *soff;   // [tri] shared-memory offset (batch_row_off(row) + column) of packed element e
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
// Kernel variants (template parameters; ekf_abi_batch.inl picks one per handle, EKF_BATCH_VARIANT overrides):
//   MINB     resident CTAs per SM the register allocation is sized for (16: 128 registers, 20: 96)
//   ONEPASS  every lane evaluates at most one candidate landmark (N <= 32): no loop-carried best-candidate record
//   DD       rank-2 downdate Sigma -= K G^T of the stored triangle in shared memory:
//            0  element by element in packed order (row / column table)
//            1  2x2 blocks, K rows / G columns loaded once per block, 128-bit accesses
//            2  8x8 tiles on the fp64 tensor path (DMMA m8n8k4, k = 2 of 4 used), Sigma stored tile by tile
//               in accumulator-fragment order: one conflict-free 128-bit load + store per lane per tile
constexpr int kBatchErrCapacity = 2, kBatchErrBadId = 4;   // ekf_status values

// Shared-memory layout of one filter's covariance: the stored triangle row by row, every row starting on a
// 16-byte boundary (rows 2k and 2k+1 both occupy 2k+2 doubles), so that the two adjacent columns of a 2x2
// block are one aligned 128-bit access and consecutive lanes' blocks are consecutive.  Row i starts at
// batch_row_off(i); the HBM layout stays the plain packed triangle.
__host__ __device__ __forceinline__ int batch_row_off(int i) {
  const int k = i >> 1;
  return (i & 1) ? 2 * (k + 1) * (k + 1) : 2 * k * (k + 1);
}
// G (n x 2) is kept as two arrays, even columns then odd columns, (n+1)/2 entries each: the two G rows of a
// 2x2 block are then conflict-free 128-bit loads across the warp.
__host__ __device__ __forceinline__ int batch_g_idx(int j, int half) { return (j & 1) * half + (j >> 1); }

// Tiled layout (DD = 2): tile (I, J <= I) of 8x8 doubles at (I (I + 1) / 2 + J) * 64.  Inside a tile the unit is the
// 16-byte chunk a lane holds of the DMMA accumulator (lane l: row l / 4, columns 2 (l % 4), +1).  Row pair h = r / 2
// keeps its own 128 bytes (so a quarter-warp of the tile load / store still covers all 32 banks), but inside it the
// chunks are permuted: column chunk ^ f(h), and the two rows swap halves in odd block-rows.  A walk down one COLUMN (the
// G = Sigma H^T gather, the pose-panel rows of the candidates) then spreads over all eight chunk positions instead of
// hitting one or two banks.  Diagonal tiles are stored whole; their strictly upper part and rows / columns >= n are
// padding nobody reads.
__host__ __device__ __forceinline__ int batch_tile_chunk(int I, int r, int c2) {
  // f = multiplication by a primitive element of GF(4) (0, 1, 2, 3 -> 0, 2, 3, 1): both h -> f(h) and h -> h ^ f(h)
  // are bijections, so a column walk AND the diagonal (the candidates' own 2x2 blocks) spread over the chunks
  const int h = r >> 1, f = (0x78 >> (2 * h)) & 3;
  return (h << 3) | ((((r ^ I) & 1)) << 2) | (c2 ^ f);
}
__host__ __device__ __forceinline__ int batch_tile_idx(int i, int c) {
  const int I = i >> 3, J = c >> 3;
  return ((I * (I + 1) / 2 + J) << 6) + (batch_tile_chunk(I, i & 7, (c & 7) >> 1) << 1) + (c & 1);
}
// offset of stored element (i, c <= i) / doubles of Sigma storage / entries of the G and K arrays per layout
template <int DD> __host__ __device__ __forceinline__ int batch_s_idx(int i, int c) {
  return DD == 2 ? batch_tile_idx(i, c) : batch_row_off(i) + c;
}
__host__ __device__ inline int batch_s_doubles(int n, int dd) {
  const int nt = (n + 7) >> 3;
  return dd == 2 ? nt * (nt + 1) / 2 * 64 : batch_row_off(n);
}
__host__ __device__ inline int batch_gk_entries(int n, int dd) { return dd == 2 ? ((n + 7) & ~7) : 2 * ((n + 1) / 2) + 1; }

__host__ __device__ inline size_t batch_smem_bytes(int n, int m, int dd) {
  //      Sigma                          mu (even count)         G + K as double2                         polar sightings
  return ((size_t)batch_s_doubles(n, dd) + (size_t)(n + (n & 1)) + 4 * (size_t)batch_gk_entries(n, dd) + 2 * (size_t)m + 8) *
         sizeof(double);
}

// Sigma_0 = diag(0, 0, 0, prior * I_2N) for every filter: slam.cpp:128-132
__global__ void batch_init_prior_kernel(double *sigma, int B, int n, int tri, double prior) {
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)B * n) return;
  const int f = (int)(g / n), i = (int)(g % n);
  if (i >= 3) sigma[(size_t)f * tri + (size_t)i * (i + 1) / 2 + i] = prior;
}

template <int DD> __device__ __forceinline__ int tri_idx(int i, int j) {
  return i >= j ? batch_s_idx<DD>(i, j) : batch_s_idx<DD>(j, i);
}
__device__ __forceinline__ void dmma884_b(double &c0, double &c1, double a, double b) {   // D = A (8x4) B (4x8) + C
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Sigma -= K G^T over the NT (NT + 1) / 2 stored 8x8 tiles, fully unrolled for a compile-time tile count: the G
// fragments of all block-columns stay in registers, every tile is one 128-bit load, one DMMA and one 128-bit store at
// a constant offset (the generic loop spends more instructions on its counters and fragment reloads than on tiles)
template <int NT>
__device__ __forceinline__ void batch_downdate_tiles(double *S, const double2 *K, const double2 *G, int lane) {
  const int fr = lane >> 2, fk = lane & 3;
  const double *Kd = reinterpret_cast<const double *>(K), *Gd = reinterpret_cast<const double *>(G);
  double2 *t0 = reinterpret_cast<double2 *>(S) + batch_tile_chunk(0, fr, fk);
  double2 *t1 = reinterpret_cast<double2 *>(S) + batch_tile_chunk(1, fr, fk);
  double bJ[NT];
#pragma unroll
  for (int J = 0; J < NT; ++J) bJ[J] = fk < 2 ? Gd[(8 * J + fr) * 2 + fk] : 0.0;
#pragma unroll
  for (int I = 0; I < NT; ++I) {
    const double aI = fk < 2 ? -Kd[(8 * I + fr) * 2 + fk] : 0.0;
    double2 *ct = ((I & 1) ? t1 : t0) + 32 * (I * (I + 1) / 2);
#pragma unroll
    for (int J = 0; J <= I; ++J) {
      double2 c = ct[32 * J];
      dmma884_b(c.x, c.y, aI, bJ[J]);
      ct[32 * J] = c;
    }
  }
}

template <int MINB, bool ONEPASS, int DD>
__global__ void __launch_bounds__(kBatchThreads, MINB) ekf_batch_kernel(const BatchParams p) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.n, tri = p.tri;
  const int half = (n + 1) >> 1;
  const int gk = batch_gk_entries(n, DD);
  double *S = sm;                         // stored triangle: 16-byte aligned rows (DD 0, 1) or 8x8 tiles (DD 2)
  double *mu = S + batch_s_doubles(n, DD);   // state
  double2 *G = reinterpret_cast<double2 *>(mu + n + (n & 1));   // G = Sigma H^T   (n x 2); DD 1: even | odd columns
  double2 *K = G + gk;                                          // K = G S^-1      (n x 2)
  double *ZP = reinterpret_cast<double *>(K + gk);              // polar sightings (m x 2)
  auto g_at = [&](int j) { return DD == 1 ? batch_g_idx(j, half) : j; };
  const int f = p.f0 + blockIdx.x;
  const int lane = threadIdx.x;
  if (f >= p.B) return;

  // ---- load the filter (coalesced) ---------------------------------------------------------
  {
    if constexpr (DD == 2) {   // padding of the tiles and the G / K rows beyond n must stay finite zeros
      for (int e = lane; e < batch_s_doubles(n, DD); e += 32) S[e] = 0.0;
      for (int e = lane; e < 2 * gk; e += 32) G[e] = make_double2(0.0, 0.0);
      __syncwarp();
    }
    const double *gs = p.sigma + (size_t)f * tri;
#pragma unroll 4
    for (int e = lane; e < tri; e += 32) S[__ldg(p.soff + e)] = gs[e];
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
      const int r = batch_s_idx<DD>(i, 0);
      const int r2 = batch_s_idx<DD>(i, 2);
      const double s0 = S[r];
      S[r + 1] = s0 * (-dy) + S[r + 1];
      S[r2] = s0 * dx + S[r2];
    }
    if (lane == 0) {
      double P[3][3], Bm[3][3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) P[i][j] = P[j][i] = S[batch_s_idx<DD>(i, j)];
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
        for (int j = 0; j <= i; ++j) S[batch_s_idx<DD>(i, j)] = P[i][j];
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
  // shared-memory offsets that do not change from one measurement to the next (the swizzled tile layout makes each a
  // dozen integer instructions): the lane's own candidate (ONEPASS: candidate k == lane) and the pose-panel entries of
  // the rows the lane owns in the G / K sweep (rows lane and lane + 32)
  int c_a0 = 0, c_a2 = 0, c_b0 = 0, c_b2 = 0, c_aa = 0, c_ba = 0, c_bb = 0;
  if constexpr (ONEPASS) {
    const int a = 3 + 2 * min(lane, p.N - 1);
    c_a0 = batch_s_idx<DD>(a, 0); c_a2 = batch_s_idx<DD>(a, 2);
    c_b0 = batch_s_idx<DD>(a + 1, 0); c_b2 = batch_s_idx<DD>(a + 1, 2);
    c_aa = batch_s_idx<DD>(a, a); c_ba = batch_s_idx<DD>(a + 1, a); c_bb = batch_s_idx<DD>(a + 1, a + 1);
  }
  const int g_i0 = min(lane, n - 1), g_i1 = min(lane + 32, n - 1);
  const int g00 = tri_idx<DD>(g_i0, 0), g01 = tri_idx<DD>(g_i0, 1), g02 = tri_idx<DD>(g_i0, 2);
  const int g10 = tri_idx<DD>(g_i1, 0), g11 = tri_idx<DD>(g_i1, 1), g12 = tri_idx<DD>(g_i1, 2);
  for (int t = 0; t < mf; ++t) {
    const Polar z{ZP[2 * t], ZP[2 * t + 1]};
    const double th = mu[0], px = mu[1], py = mu[2];
    double pb[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) pb[i][j] = pb[j][i] = S[batch_s_idx<DD>(i, j)];

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
      for (int k = lane; k < cnt; k += 32) {                   // slam.cpp:361-404 (ONEPASS: at most one trip)
        // columns 0, 1 of a row are adjacent in both layouts; everything else is looked up
        const int a = 3 + 2 * k;
        const int ra = ONEPASS ? c_a0 : batch_s_idx<DD>(a, 0), rb = ONEPASS ? c_b0 : batch_s_idx<DD>(a + 1, 0);
        const double pa[3] = {S[ra], S[ra + 1], S[ONEPASS ? c_a2 : batch_s_idx<DD>(a, 2)]},
                     pbr[3] = {S[rb], S[rb + 1], S[ONEPASS ? c_b2 : batch_s_idx<DD>(a + 1, 2)]};
        const Lin Lk = linearize(th, px, py, mu[a], mu[a + 1]);
        const Block5 Bk = make_block5(pb, pa, pbr, S[ONEPASS ? c_aa : batch_s_idx<DD>(a, a)],
                                      S[ONEPASS ? c_ba : batch_s_idx<DD>(a + 1, a)],
                                      S[ONEPASS ? c_bb : batch_s_idx<DD>(a + 1, a + 1)]);
        const Innov I = innovation(Bk, Lk, p.r_noise);
        const double z0 = z.r - Lk.zhat_r, z1 = normalize_angle(z.phi - Lk.zhat_b);
        double i00, i01, i11;
        const double d = mahalanobis(I.s00, I.s01, I.s11, z0, z1, i00, i01, i11);
        if constexpr (ONEPASS) {
          // the only candidate of this lane: a NaN / +inf distance keeps the sentinel, as cand_min would
          if (d < INFINITY) best = Cand{d, k};
          r_xs = Lk.h0[3]; r_ys = Lk.h0[4]; r_xd = Lk.h1[4]; r_yd = -Lk.h1[3];
          r_m00 = i00; r_m01 = i01; r_m11 = i11; r_z0 = z0; r_z1 = z1;
          break;
        } else {
          const Cand nb = cand_min(best, Cand{d, k});
          if (nb.k == k) {
            r_xs = Lk.h0[3]; r_ys = Lk.h0[4]; r_xd = Lk.h1[4]; r_yd = -Lk.h1[3];
            r_m00 = i00; r_m01 = i01; r_m11 = i11; r_z0 = z0; r_z1 = z1;
          }
          best = nb;
        }
      }
      if constexpr (ONEPASS) {
        // candidate k sits in lane k, so "lowest index on ties" is "lowest lane": arg-min of the 64-bit order key of
        // the distance with two 32-bit warp reductions and a ballot instead of five shuffle rounds.  The key is the
        // usual monotone map of IEEE doubles onto unsigned integers (exact for every non-NaN value; a NaN distance
        // never became `best`, slam.cpp:418); + 0.0 folds -0.0 into +0.0, which compare equal in the reference
        const unsigned long long bits = (unsigned long long)__double_as_longlong(best.d + 0.0);
        const unsigned long long key = bits ^ ((bits >> 63) ? ~0ULL : 0x8000000000000000ULL);
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mhi = __reduce_min_sync(0xffffffffu, hi);
        const unsigned mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
        const unsigned win = __ballot_sync(0xffffffffu, hi == mhi && lo == mlo);
        const int src = __ffs(win) - 1;
        best.d = __shfl_sync(0xffffffffu, best.d, src);
        best.k = __shfl_sync(0xffffffffu, best.k, src);
      } else {
        best = warp_argmin(best);
      }
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
    const int a = 3 + 2 * j, ra = batch_s_idx<DD>(a, 0), rb = batch_s_idx<DD>(a + 1, 0);
    double lx, ly;
    if (is_new) {                                              // slam.cpp:214-215 / :351-354
      lx = px + z.r * cos(z.phi + th);
      ly = py + z.r * sin(z.phi + th);
      if (unknown) cnt += 1;
    } else { lx = mu[a]; ly = mu[a + 1]; }
    if (!have) {                                               // new landmark / known association
      L = linearize(th, px, py, lx, ly);
      const double pa[3] = {S[ra], S[ra + 1], S[batch_s_idx<DD>(a, 2)]},
                   pbr[3] = {S[rb], S[rb + 1], S[batch_s_idx<DD>(a + 1, 2)]};
      const Block5 Bk = make_block5(pb, pa, pbr, S[batch_s_idx<DD>(a, a)], S[batch_s_idx<DD>(a + 1, a)],
                                    S[batch_s_idx<DD>(a + 1, a + 1)]);
      const Innov I = innovation(Bk, L, p.r_noise);
      zd0 = z.r - L.zhat_r; zd1 = normalize_angle(z.phi - L.zhat_b);
      (void)mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, m00, m01, m11);
    }
    __syncwarp();
    if (is_new && lane == 0) { mu[a] = lx; mu[a + 1] = ly; }
    __syncwarp();
    // G = Sigma H^T, K = G S^-1, state: one row per lane and trip; the first two trips use the hoisted offsets
    auto gk_row = [&](int i, int o0, int o1, int o2) {
      const double row[5] = {S[o0], S[o1], S[o2], S[tri_idx<DD>(i, a)], S[tri_idx<DD>(i, a + 1)]};
      double g0, g1;
      row_times_ht(row, L, g0, g1);
      const double k0 = g0 * m00 + g1 * m01, k1 = g0 * m01 + g1 * m11;
      G[g_at(i)] = make_double2(g0, g1);
      K[i] = make_double2(k0, k1);
      double v = mu[i] + (k0 * zd0 + k1 * zd1);                // slam.cpp:482
      if (i == 0) v = normalize_angle(v);                      // slam.cpp:488
      mu[i] = v;
    };
    if (lane < n) gk_row(lane, g00, g01, g02);
    if (lane + 32 < n) gk_row(lane + 32, g10, g11, g12);
    for (int i = lane + 64; i < n; i += 32) gk_row(i, tri_idx<DD>(i, 0), tri_idx<DD>(i, 1), tri_idx<DD>(i, 2));
    __syncwarp();
    // Sigma -= K G^T on the stored triangle, element by element in packed order: 32 consecutive
    // elements per step keep every lane busy (a column-per-lane walk idles half of them on the triangle);
    // (row, column) of an element comes from a small table that stays in L1
    if constexpr (DD == 2) {
      // C tile (accumulator order in shared memory) += (-K rows of block-row I) x (G rows of block-column J)^T;
      // fragment element (row lane / 4, k = lane % 4): k = 0, 1 are the two columns of K / G, k = 2, 3 are zero
      const int nt = (n + 7) >> 3, fr = lane >> 2, fk = lane & 3;
      if (nt == 6) {                       // n = 41 .. 48: BASELINE config 3 (N = 20, n = 43)
        batch_downdate_tiles<6>(S, K, G, lane);
      } else {
      const double *Kd = reinterpret_cast<const double *>(K), *Gd = reinterpret_cast<const double *>(G);
      const double2 *tile = reinterpret_cast<const double2 *>(S);
      const int ch[2] = {batch_tile_chunk(0, fr, fk), batch_tile_chunk(1, fr, fk)};   // this lane's chunk, even / odd I
      for (int I = 0; I < nt; ++I) {
        const double aI = fk < 2 ? -Kd[(8 * I + fr) * 2 + fk] : 0.0;
        double2 *ct = const_cast<double2 *>(tile) + ch[I & 1];
        for (int J = 0; J <= I; ++J, ct += 32) {
          const double bJ = fk < 2 ? Gd[(8 * J + fr) * 2 + fk] : 0.0;
          double2 c = *ct;
          dmma884_b(c.x, c.y, aI, bJ);
          *ct = c;
        }
        tile += 32 * (I + 1);
      }
      }
    } else if constexpr (DD == 1) {
      // 2x2 blocks: two K rows and two G columns serve up to four elements.  K[n] / G[n] of an odd n are read
      // (they alias the next array) but their products are masked out.
      for (int q = lane; q < p.nblk; q += 32) {
        const ushort4 bm = __ldg(p.blk_map + q);
        const double2 ka = K[bm.x], kb = K[bm.x + 1], ga = G[bm.y >> 1], gb = G[half + (bm.y >> 1)];
        // rows i0 (even) and i0 + 1 are the same padded length: the block's second row is i0 + 2 doubles on
        double2 *s0 = reinterpret_cast<double2 *>(S + bm.z), *s1 = reinterpret_cast<double2 *>(S + bm.z + bm.x + 2);
        double2 a0 = *s0, a1 = *s1;
        a0.x = a0.x - (ka.x * ga.x + ka.y * ga.y);
        a0.y = a0.y - (ka.x * gb.x + ka.y * gb.y);
        a1.x = a1.x - (kb.x * ga.x + kb.y * ga.y);
        a1.y = a1.y - (kb.x * gb.x + kb.y * gb.y);
        // elements outside the stored triangle ((i0, i0+1) on the diagonal, row n of an odd n) are padding:
        // written with whatever the arithmetic gave, never read by anything else
        *s0 = a0;
        if (bm.w & 4) *s1 = a1;
      }
    } else {
      for (int e = lane; e < tri; e += 32) {
        const uchar2 ic = __ldg(p.tri_map + e);
        const double2 ki = K[ic.x], gc = G[ic.y];
        const int o = batch_row_off(ic.x) + ic.y;
        S[o] = S[o] - (ki.x * gc.x + ki.y * gc.y);
      }
    }
    __syncwarp();
  }

  // ---- store back ----------------------------------------------------------------------------------
  {
    double *gs = p.sigma + (size_t)f * tri;
#pragma unroll 4
    for (int e = lane; e < tri; e += 32) gs[e] = S[__ldg(p.soff + e)];
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
