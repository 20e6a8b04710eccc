// Small utility kernels around the filter storage: initialisation (slam.cpp:128-132), block
// get/set for the accessors and fixtures, and the procedural large-map fixture of SURVEY.md
// section 8(d).  None of these is on the timed path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_layout.h"
#include "ekf_math.cuh"

namespace ekf {

// Sigma_0 = diag(0, 0, 0, prior * I_2N): slam.cpp:128-132 (storage was zero-filled before).
__global__ void init_prior_kernel(double *sigma, EkfLay lay, int n_int, double prior) {
  const int i = 4 + blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_int && ekf_lay_owns_row(lay, i)) sigma[ekf_lay_elem(lay, i, i)] = prior;
}

// out[(i-i0)*ld + (j-j0)] = Sigma[i][j] (external indices, symmetric fill); entries whose stored
// element belongs to another row shard read as 0 (the caller sums over the shards)
__global__ void get_block_kernel(const double *sigma, EkfLay lay, int i0, int i1, int j0, int j1, double *out) {
  const int64_t ld = j1 - j0;
  const int64_t total = (int64_t)(i1 - i0) * ld;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = i0 + (int)(e / ld), j = j0 + (int)(e % ld);
    const int ii = ekf_int_index(i), jj = ekf_int_index(j);
    const int r = ii >= jj ? ii : jj, c = ii >= jj ? jj : ii;
    out[e] = ekf_lay_owns_row(lay, r) ? sigma[ekf_lay_elem(lay, r, c)] : 0.0;
  }
}

// writes the entries of the block with row >= col into the stored triangle
__global__ void set_block_kernel(double *sigma, EkfLay lay, int i0, int i1, int j0, int j1, const double *in) {
  const int64_t ld = j1 - j0;
  const int64_t total = (int64_t)(i1 - i0) * ld;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = i0 + (int)(e / ld), j = j0 + (int)(e % ld);
    if (i >= j && ekf_lay_owns_row(lay, ekf_int_index(i)))
      sigma[ekf_lay_elem(lay, ekf_int_index(i), ekf_int_index(j))] = in[e];
  }
}

// external state vector <-> internal (padded) layout
__global__ void state_to_ext_kernel(const double *mu, double *out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = mu[ekf_int_index(i)];
}
__global__ void state_from_ext_kernel(double *mu, const double *in, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) mu[ekf_int_index(i)] = in[i];
}

// Mahalanobis distance of ONE body-frame sighting to every mapped landmark against the filter as it stands:
// the sweep of slam.cpp:358-404 without the tentative landmark, the decision or the update.  One thread per
// landmark; the 5x5 block {theta, x, y, lx, ly}^2 comes straight from the stored triangle.  Same device
// functions, in the same order, as phase A of ekf_seq_kernel - a diagnostic for the decision margins.
__global__ void mahalanobis_sweep_kernel(const double *sigma, EkfLay lay, const double *mu, const int *cnt, double mx,
                                         double my, double r_noise, double *out, int cap) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= *cnt || k >= cap) return;
  double pb[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) pb[i][j] = pb[j][i] = sigma[ekf_lay_elem(lay, i, j)];
  const int a = 4 + 2 * k;
  const double *ra = sigma + ekf_lay_elem(lay, a, 0), *rb = sigma + ekf_lay_elem(lay, a + 1, 0);
  const double pa3[3] = {ra[0], ra[1], ra[2]}, pb3[3] = {rb[0], rb[1], rb[2]};
  const double *dg = sigma + ekf_lay_elem(lay, a, a);
  const Polar z = to_polar(mx, my);
  const Lin L = linearize(mu[0], mu[1], mu[2], mu[a], mu[a + 1]);
  const Block5 B = make_block5(pb, pa3, pb3, dg[0], dg[EKF_STRIP], dg[EKF_STRIP + 1]);
  const Innov I = innovation(B, L, r_noise);
  const double zd0 = z.r - L.zhat_r;
  const double zd1 = normalize_angle(z.phi - L.zhat_b);
  double i00, i01, i11;
  out[k] = mahalanobis(I.s00, I.s01, I.s11, zd0, zd1, i00, i01, i11);
}

// ---- procedural fixture --------------------------------------------------------------------------
__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}
// uniform in [-1, 1), a pure function of (seed, a, b)
__host__ __device__ inline double hash_sym(uint64_t seed, uint64_t a, uint64_t b) {
  const uint64_t h = splitmix64(seed ^ splitmix64(a * 0x100000001B3ULL + b));
  return (double)(h >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}
constexpr int kProcRank = 8;
__device__ inline double proc_u(uint64_t seed, int i, int r) { return 0.01 * hash_sym(seed, (uint64_t)i, 100 + r); }

// Sigma = Dblk + U U^T on the `count` mapped landmarks and the pose; unmapped landmarks keep the
// reference prior on their diagonal and zero elsewhere.  One thread per stored entry.
__global__ void fill_procedural_sigma_kernel(double *sigma, EkfLay lay, int n_pad, int count, uint64_t seed,
                                             double prior) {
  const int64_t total = ekf_lay_elems(lay);
  const int n_active = 4 + 2 * count;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    // invert the strip layout: find strip J with strip_offset(J) <= e
    int lo = 0, hi = lay.re / EKF_STRIP - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (ekf_lay_strip_offset(lay, mid) <= e) lo = mid; else hi = mid - 1;
    }
    const int64_t rel = e - ekf_lay_strip_offset(lay, lo);
    const int i = ekf_lay_first_row(lay, lo) + (int)(rel / EKF_STRIP), j = lo * EKF_STRIP + (int)(rel % EKF_STRIP);
    double v = 0.0;
    const int n_int_all = n_pad;  // rows beyond the real dimension stay zero through the test below
    (void)n_int_all;
    if (j <= i && i != 3 && j != 3) {
      if (i < n_active) {
        for (int r = 0; r < kProcRank; ++r) v += proc_u(seed, i, r) * proc_u(seed, j, r);
        if (i == j) {
          v += 1.0e-3 * (1.5 + 0.5 * hash_sym(seed, (uint64_t)i, 7));
        } else if (i >= 4 && j == i - 1 && (i & 1)) {
          // in-landmark correlation, |rho| <= 0.3 of the smaller diagonal term
          v += 3.0e-4 * hash_sym(seed, (uint64_t)i, 9);
        }
      } else if (i == j) {
        v = prior;
      }
    }
    sigma[e] = v;
  }
}

// landmark k at column k % grid_w, row k / grid_w of a unit grid, jitter +-0.2 (true position);
// the state estimate adds an error of +-0.01.
__global__ void fill_procedural_state_kernel(double *mu, int n_pad, int N, int count, int grid_w, uint64_t seed) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= N) return;
  double x = 0.0, y = 0.0;
  if (k < count) {
    x = (double)(k % grid_w) + 0.2 * hash_sym(seed, (uint64_t)k, 1) + 0.01 * hash_sym(seed, (uint64_t)k, 3);
    y = (double)(k / grid_w) + 0.2 * hash_sym(seed, (uint64_t)k, 2) + 0.01 * hash_sym(seed, (uint64_t)k, 4);
  }
  mu[4 + 2 * k] = x;
  mu[5 + 2 * k] = y;
  (void)n_pad;
}

}  // namespace ekf
