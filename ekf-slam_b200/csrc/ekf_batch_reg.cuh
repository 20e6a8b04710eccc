// Batched mode, register-resident shape: the filter's covariance lives in the warp's REGISTERS for the whole scan, as
// the accumulator fragments of its 8x8 tiles (DMMA m8n8k4: lane l of tile (I, J <= I) holds row 8I + l/4, columns
// 8J + 2(l%4), +1; NT (NT + 1) / 2 tiles = 21 double2 per lane for n = 41 .. 48, BASELINE config 3).  The rank-2
// downdate of a measurement (slam.cpp:264-265 / :485-486) is then 21 DMMAs on registers - no tile traffic through shared
// memory at all (in ekf_batch_kernel the 21 loads + 21 stores of the tiles are 53 % of the wavefronts of a kernel whose
// shared-memory data pipe is 80 % busy).  What the rest of a measurement reads of Sigma is mirrored in small, plainly
// indexed shared-memory arrays that the warp refreshes from its registers after every downdate:
//   Pp01[i], Pp2[i]      columns 0, 1 / column 2 of row i (the pose panel; rows < 3: the pose block, symmetric)
//   D0[k], D1[k], D2[k]  landmark k's own 2x2 block (a, a), (a+1, a), (a+1, a+1), a = 3 + 2k
// and, once the landmark j of the measurement is known, its two columns (symmetric completion of the stored triangle)
//   colA[i] = Sigma(i, a_j), colB[i] = Sigma(i, a_j + 1)
// picked out of the tiles of block-row / block-column a_j / 8 with warp-uniform branches.
// Same arithmetic, operation for operation, as ekf_batch_kernel<.., true, 2> (association, gain, state, downdate), so
// the two kernels leave bit-identical filters (tests/test_gpu_batch.py).  Requirements: N <= 32 (one candidate per
// lane), 41 <= n <= 48 (NT = 6), m <= 32 (the polar sightings live one per lane); the host falls back otherwise.
#pragma once
#include "ekf_batch.cuh"

namespace ekf {

constexpr int kRegNT = 6;                                   // block-rows of 8
constexpr int kRegRows = 8 * kRegNT;                        // 48 padded rows
constexpr int kRegTiles = kRegNT * (kRegNT + 1) / 2;        // 21
constexpr int kRegMaxLm = 24;                               // landmark slots of the D arrays (a <= 47 -> k <= 22)

// shared memory of one filter (doubles): mu | Pp2 | colA | colB | D0 D1 D2 | Pp01 (double2) | G (double2) | K (double2)
constexpr int kRegSmemDoubles = 4 * kRegRows + 3 * kRegMaxLm + 3 * 2 * kRegRows;
__host__ __device__ inline size_t batch_reg_smem_bytes() { return (size_t)kRegSmemDoubles * sizeof(double); }
__host__ __device__ inline bool batch_reg_supported(int N, int n) { return N <= 32 && ((n + 7) >> 3) == kRegNT; }

__device__ __forceinline__ constexpr int reg_tile(int I, int J) { return I * (I + 1) / 2 + J; }

// Mirrors of the pose panel and of the landmarks' own blocks, from the fragments.  Every store has a fixed lane
// predicate and a compile-time offset from a per-lane base.
__device__ __forceinline__ void reg_refresh_mirrors(const double2 (&c)[kRegTiles], double2 *Pp01, double *Pp2,
                                                    double *D0, double *D1, double *D2, int fr, int fk) {
  // pose panel: tile (I, 0), lanes fk = 0 hold columns 0, 1, lanes fk = 1 hold column 2 (.x).  Rows 0 .. 2 (the pose
  // block) are mirrored symmetric: their entries above the diagonal come from the lanes that hold the stored ones
  {
    double *Pd = reinterpret_cast<double *>(Pp01);
    const double2 t = c[0];
    if (fk == 0) {
      if (fr == 0) Pd[0] = t.x;                               // (0, 0)
      else Pp01[fr] = t;                                      // (i, 0), (i, 1), i >= 1
      if (fr == 1) Pd[1] = t.x;                               // (0, 1) := (1, 0)
      if (fr == 2) { Pp2[0] = t.x; Pp2[1] = t.y; }            // (0, 2) := (2, 0), (1, 2) := (2, 1)
    }
    if (fk == 1 && fr >= 2) Pp2[fr] = t.x;
  }
#pragma unroll
  for (int I = 1; I < kRegNT; ++I) {
    if (fk == 0) Pp01[8 * I + fr] = c[reg_tile(I, 0)];
    if (fk == 1) Pp2[8 * I + fr] = c[reg_tile(I, 0)].x;
  }
  // landmark blocks, a = 3 + 2k odd: (a, a) is the .y of lane (fr = a % 8, fk = fr / 2) of the diagonal tile, (a+1, a+1)
  // the .x of lane (fr even, fk = fr / 2), (a+1, a) the .y of lane (fr even >= 2, fk = fr / 2 - 1) - or, when a % 8 == 7,
  // of lane (0, 3) of the tile below the diagonal one.  Row r = 8I + fr belongs to landmark (r - 3) / 2 (r odd, = a)
  // or (r - 4) / 2 (r even, = a + 1).
  const bool odd = fr & 1;
  const bool own_d0 = odd && fk == (fr >> 1);                // (a, a)
  const bool own_d2 = !odd && fk == (fr >> 1);               // (a+1, a+1)
  const bool own_d1 = !odd && fr >= 2 && fk == (fr >> 1) - 1;   // (a+1, a), same tile
#pragma unroll
  for (int I = 0; I < kRegNT; ++I) {
    const int r = 8 * I + fr;
    const double2 d = c[reg_tile(I, I)];
    if (own_d0 && r >= 3) D0[(r - 3) >> 1] = d.y;
    if (own_d2 && r >= 4) D2[(r - 4) >> 1] = d.x;
    if (own_d1 && r >= 4) D1[(r - 4) >> 1] = d.y;
    if (I + 1 < kRegNT) {
      if (fr == 0 && fk == 3) D1[4 * I + 2] = c[reg_tile(I + 1, I)].y;    // (8I + 8, 8I + 7): a = 8I + 7
    }
  }
}

// colA[i] = Sigma(i, a), colB[i] = Sigma(i, a + 1) for every row i (stored triangle completed by symmetry), a odd.
// Row part (i <= a: row a of block-row a / 8) and column part (i > a: column a of block-column a / 8) never write the
// same entry.  The tests on I / J are warp-uniform.
__device__ __forceinline__ void reg_extract_columns(const double2 (&c)[kRegTiles], double *colA, double *colB, int a,
                                                    int fr, int fk) {
  const int b = a + 1;
  const int ta = a >> 3, ra = a & 7, tb = b >> 3, rb = b & 7;
  const int col = 2 * fk;                                     // first column of the lane's pair inside its tile
#pragma unroll
  for (int I = 0; I < kRegNT; ++I) {
    if (I == ta) {                                            // row a: entries (a, i), i <= a
      if (fr == ra) {
#pragma unroll
        for (int J = 0; J <= I; ++J) {
          const int i0 = 8 * J + col;
          if (i0 <= a) colA[i0] = c[reg_tile(I, J)].x;
          if (i0 + 1 <= a) colA[i0 + 1] = c[reg_tile(I, J)].y;
        }
      }
    }
    if (I == tb) {                                            // row a + 1: entries (a + 1, i), i <= a + 1
      if (fr == rb) {
#pragma unroll
        for (int J = 0; J <= I; ++J) {
          const int i0 = 8 * J + col;
          if (i0 <= b) colB[i0] = c[reg_tile(I, J)].x;
          if (i0 + 1 <= b) colB[i0 + 1] = c[reg_tile(I, J)].y;
        }
      }
    }
  }
#pragma unroll
  for (int J = 0; J < kRegNT; ++J) {
    if (J == ta) {                                            // column a (odd: the .y of lanes fk = (a % 8) / 2), rows > a
      if (fk == (ra >> 1)) {
#pragma unroll
        for (int I = J; I < kRegNT; ++I)
          if (8 * I + fr > a) colA[8 * I + fr] = c[reg_tile(I, J)].y;
      }
    }
    if (J == tb) {                                            // column a + 1 (even: the .x of lanes fk = (b % 8) / 2), rows > a + 1
      if (fk == (rb >> 1)) {
#pragma unroll
        for (int I = J; I < kRegNT; ++I)
          if (8 * I + fr > b) colB[8 * I + fr] = c[reg_tile(I, J)].x;
      }
    }
  }
}

template <int MINB>
__global__ void __launch_bounds__(kBatchThreads, MINB) ekf_batch_reg_kernel(const BatchParams p) {
  extern __shared__ __align__(16) double sm[];
  const int n = p.n, tri = p.tri;
  double *mu = sm;                                   // state (48)
  double *Pp2 = mu + kRegRows;
  double *colA = Pp2 + kRegRows, *colB = colA + kRegRows;
  double *D0 = colB + kRegRows, *D1 = D0 + kRegMaxLm, *D2 = D1 + kRegMaxLm;
  double2 *Pp01 = reinterpret_cast<double2 *>(D2 + kRegMaxLm);
  double2 *G = Pp01 + kRegRows, *K = G + kRegRows;
  const int f = p.f0 + blockIdx.x;
  const int lane = threadIdx.x;
  if (f >= p.B) return;
  const int fr = lane >> 2, fk = lane & 3;

  // ---- load the filter: packed lower triangle -> accumulator fragments -----------------------------------
  double2 c[kRegTiles];
  {
    const double *gs = p.sigma + (size_t)f * tri;
#pragma unroll
    for (int I = 0; I < kRegNT; ++I) {
      const int i = 8 * I + fr;
      const double *row = gs + (size_t)i * (i + 1) / 2 + 2 * fk;
#pragma unroll
      for (int J = 0; J <= I; ++J) {
        const int c0 = 8 * J + 2 * fk;
        double2 v = make_double2(0.0, 0.0);
        if (i < n) {
          if (J < I || c0 <= i) v.x = row[8 * J];
          if (J < I || c0 + 1 <= i) v.y = row[8 * J + 1];
        }
        c[reg_tile(I, J)] = v;
      }
    }
    for (int e = lane; e < kRegRows; e += 32) {
      mu[e] = e < n ? p.mu[(size_t)f * n + e] : 0.0;
      G[e] = make_double2(0.0, 0.0); K[e] = make_double2(0.0, 0.0);
    }
  }
  int cnt = p.cnt[f];

  // ---- pose injection + prediction: slam.cpp:186-198 -------------------------------------------
  {
    const double th = normalize_angle(p.pose[3 * f]);
    const double x = p.pose[3 * f + 1], y = p.pose[3 * f + 2];
    const double dx = x - p.prev[2 * f], dy = y - p.prev[2 * f + 1];
    // the six stored entries of the pose block (tile (0, 0): lanes 0, 4, 8 and 9), before anything is changed
    const double2 t00 = c[0];
    const double p00 = __shfl_sync(0xffffffffu, t00.x, 0);
    const double p10 = __shfl_sync(0xffffffffu, t00.x, 4), p11 = __shfl_sync(0xffffffffu, t00.y, 4);
    const double p20 = __shfl_sync(0xffffffffu, t00.x, 8), p21 = __shfl_sync(0xffffffffu, t00.y, 8);
    const double p22 = __shfl_sync(0xffffffffu, t00.x, 9);
    // rows >= 3 of the pose panel: C = B A^T (columns 1 and 2 take multiples of column 0)
#pragma unroll
    for (int I = 0; I < kRegNT; ++I) {
      double2 &t = c[reg_tile(I, 0)];
      const double s0 = __shfl_sync(0xffffffffu, t.x, lane & ~3);      // Sigma(i, 0) sits in the row's lane fk = 0
      if (8 * I + fr >= 3) {
        if (fk == 0) t.y = s0 * (-dy) + t.y;
        if (fk == 1) t.x = s0 * dx + t.x;
      }
    }
    // pose block: A P A^T + Qbar on the symmetric completion, every lane the same arithmetic
    double P[3][3], Bm[3][3];
    P[0][0] = p00; P[1][0] = P[0][1] = p10; P[1][1] = p11; P[2][0] = P[0][2] = p20; P[2][1] = P[1][2] = p21; P[2][2] = p22;
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
    if (lane == 0) { c[0].x = P[0][0]; mu[0] = th; mu[1] = x; mu[2] = y; }
    if (lane == 4) { c[0].x = P[1][0]; c[0].y = P[1][1]; }
    if (lane == 8) { c[0].x = P[2][0]; c[0].y = P[2][1]; }
    if (lane == 9) { c[0].x = P[2][2]; }
  }
  reg_refresh_mirrors(c, Pp01, Pp2, D0, D1, D2, fr, fk);
  __syncwarp();

  // ---- measurement loop ----------------------------------------------------------------------
  const int mf = p.m_count ? min(p.m_count[f], p.m) : p.m;
  const bool unknown = (p.ids == nullptr);
  int err = 0;
  // polar form of every sighting of the scan, one lane each (slam.cpp:345-347 does not depend on the state)
  double z_r = 0.0, z_phi = 0.0;
  if (lane < mf) {
    const Polar z = to_polar(p.xy[((size_t)f * p.m + lane) * 2], p.xy[((size_t)f * p.m + lane) * 2 + 1]);
    z_r = z.r; z_phi = z.phi;
  }
  const int ka = 3 + 2 * min(lane, p.N - 1);          // the lane's own candidate (candidate k == lane)
  const int kk = min(lane, p.N - 1);
  const int g_i0 = min(lane, n - 1), g_i1 = min(lane + 32, n - 1);
  for (int t = 0; t < mf; ++t) {
    const Polar z{__shfl_sync(0xffffffffu, z_r, t), __shfl_sync(0xffffffffu, z_phi, t)};
    const double th = mu[0], px = mu[1], py = mu[2];
    double pb[3][3];
    {
      const double2 r0 = Pp01[0], r1 = Pp01[1], r2 = Pp01[2];
      pb[0][0] = r0.x; pb[1][0] = pb[0][1] = r1.x; pb[1][1] = r1.y;
      pb[2][0] = pb[0][2] = r2.x; pb[2][1] = pb[1][2] = r2.y; pb[2][2] = Pp2[2];
      (void)r0.y;
    }
    int j = -1, is_new = 0;
    Lin L;
    double m00 = 0.0, m01 = 0.0, m11 = 0.0, zd0 = 0.0, zd1 = 0.0;
    bool have = false;
    if (unknown) {
      Cand best{INFINITY, 0x7fffffff};
      double r_xs = 0.0, r_ys = 0.0, r_xd = 0.0, r_yd = 0.0, r_m00 = 0.0, r_m01 = 0.0, r_m11 = 0.0, r_z0 = 0.0,
             r_z1 = 0.0;
      if (lane < cnt) {                                        // slam.cpp:361-404, candidate k = lane
        const int k = lane, a = ka;
        const double2 qa = Pp01[a], qb = Pp01[a + 1];
        const double pa[3] = {qa.x, qa.y, Pp2[a]}, pbr[3] = {qb.x, qb.y, Pp2[a + 1]};
        const Lin Lk = linearize(th, px, py, mu[a], mu[a + 1]);
        const Block5 Bk = make_block5(pb, pa, pbr, D0[kk], D1[kk], D2[kk]);
        const Innov I = innovation(Bk, Lk, p.r_noise);
        const double z0 = z.r - Lk.zhat_r, z1 = normalize_angle(z.phi - Lk.zhat_b);
        double i00, i01, i11;
        const double d = mahalanobis(I.s00, I.s01, I.s11, z0, z1, i00, i01, i11);
        if (d < INFINITY) best = Cand{d, k};                   // a NaN / +inf distance keeps the sentinel
        r_xs = Lk.h0[3]; r_ys = Lk.h0[4]; r_xd = Lk.h1[4]; r_yd = -Lk.h1[3];
        r_m00 = i00; r_m01 = i01; r_m11 = i11; r_z0 = z0; r_z1 = z1;
      }
      {
        // candidate k sits in lane k: "lowest index on ties" is "lowest lane" (see ekf_batch_kernel)
        const unsigned long long bits = (unsigned long long)__double_as_longlong(best.d + 0.0);
        const unsigned long long key = bits ^ ((bits >> 63) ? ~0ULL : 0x8000000000000000ULL);
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mhi = __reduce_min_sync(0xffffffffu, hi);
        const unsigned mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
        const unsigned win = __ballot_sync(0xffffffffu, hi == mhi && lo == mlo);
        const int src = __ffs(win) - 1;
        best.d = __shfl_sync(0xffffffffu, best.d, src);
        best.k = __shfl_sync(0xffffffffu, best.k, src);
      }
      if (best.d <= p.thr && best.k < cnt) {                   // slam.cpp:406-408,418-440
        j = best.k;
        const int src = j & 31;
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
    const int a = 3 + 2 * j;
    double lx, ly;
    if (is_new) {                                              // slam.cpp:214-215 / :351-354
      lx = px + z.r * cos(z.phi + th);
      ly = py + z.r * sin(z.phi + th);
      if (unknown) cnt += 1;
    } else { lx = mu[a]; ly = mu[a + 1]; }
    if (!have) {                                               // new landmark / known association
      L = linearize(th, px, py, lx, ly);
      const double2 qa = Pp01[a], qb = Pp01[a + 1];
      const double pa[3] = {qa.x, qa.y, Pp2[a]}, pbr[3] = {qb.x, qb.y, Pp2[a + 1]};
      const Block5 Bk = make_block5(pb, pa, pbr, D0[j], D1[j], D2[j]);
      const Innov I = innovation(Bk, L, p.r_noise);
      zd0 = z.r - L.zhat_r; zd1 = normalize_angle(z.phi - L.zhat_b);
      (void)mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, m00, m01, m11);
    }
    reg_extract_columns(c, colA, colB, a, fr, fk);
    __syncwarp();
    if (is_new && lane == 0) { mu[a] = lx; mu[a + 1] = ly; }
    __syncwarp();
    // G = Sigma H^T, K = G S^-1, state: one row per lane and trip
    auto gk_row = [&](int i) {
      const double2 q = Pp01[i];
      const double row[5] = {q.x, q.y, Pp2[i], colA[i], colB[i]};
      double g0, g1;
      row_times_ht(row, L, g0, g1);
      const double k0 = g0 * m00 + g1 * m01, k1 = g0 * m01 + g1 * m11;
      G[i] = make_double2(g0, g1);
      K[i] = make_double2(k0, k1);
      double v = mu[i] + (k0 * zd0 + k1 * zd1);                // slam.cpp:482
      if (i == 0) v = normalize_angle(v);                      // slam.cpp:488
      mu[i] = v;
    };
    if (lane < n) gk_row(g_i0);
    if (lane + 32 < n) gk_row(g_i1);
    __syncwarp();
    // Sigma -= K G^T: C tile += (-K rows of block-row I) x (G rows of block-column J)^T on the fp64 tensor path;
    // fragment element (row lane / 4, k = lane % 4): k = 0, 1 are the two columns of K / G, k = 2, 3 are zero
    {
      const double *Kd = reinterpret_cast<const double *>(K), *Gd = reinterpret_cast<const double *>(G);
      double bJ[kRegNT];
#pragma unroll
      for (int J = 0; J < kRegNT; ++J) bJ[J] = fk < 2 ? Gd[(8 * J + fr) * 2 + fk] : 0.0;
#pragma unroll
      for (int I = 0; I < kRegNT; ++I) {
        const double aI = fk < 2 ? -Kd[(8 * I + fr) * 2 + fk] : 0.0;
#pragma unroll
        for (int J = 0; J <= I; ++J) dmma884_b(c[reg_tile(I, J)].x, c[reg_tile(I, J)].y, aI, bJ[J]);
      }
    }
    reg_refresh_mirrors(c, Pp01, Pp2, D0, D1, D2, fr, fk);
    __syncwarp();
  }

  // ---- store back ----------------------------------------------------------------------------------
  {
    double *gs = p.sigma + (size_t)f * tri;
#pragma unroll
    for (int I = 0; I < kRegNT; ++I) {
      const int i = 8 * I + fr;
      double *row = gs + (size_t)i * (i + 1) / 2 + 2 * fk;
#pragma unroll
      for (int J = 0; J <= I; ++J) {
        const int c0 = 8 * J + 2 * fk;
        if (i < n) {
          if (J < I || c0 <= i) row[8 * J] = c[reg_tile(I, J)].x;
          if (J < I || c0 + 1 <= i) row[8 * J + 1] = c[reg_tile(I, J)].y;
        }
      }
    }
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
