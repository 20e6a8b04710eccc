// Storage layout of the EKF covariance in HBM (shared by host and device code).
//
// The reference keeps Sigma as a dense column-major arma::mat of order n = 3 + 2N
// (nuslam/src/slam.cpp:128-132,673) and rewrites all n^2 entries per measurement.  Here only
// the lower triangle exists, in "column strips":
//
//   * internal index:  pose theta,x,y -> 0,1,2; index 3 is a zero pad; landmark j -> rows
//     a = 4+2j, b = 5+2j.  Every landmark pair is therefore 16-byte aligned, and a 2x2
//     landmark block never straddles a 32-byte sector or a strip boundary.
//   * n_pad = n_int rounded up to a multiple of 128 (pad rows/cols are zero and stay zero).
//   * strip J (J = 0 .. n_pad/128 - 1) holds columns [128J, 128J+128) for rows [128J, n_pad),
//     row-major, 128 doubles (1 KiB) per row, rows consecutive.  Any run of rows inside a
//     strip is one contiguous byte range - the covariance pass streams it with 1-D bulk
//     (TMA) copies and no tensor map.
//   * the strictly-upper part of each strip's first 128 rows is scratch (never read by the
//     filter); it costs 64 KiB per strip (0.6 % at N = 10,000).
#ifndef EKF_SLAM_B200_LAYOUT_H
#define EKF_SLAM_B200_LAYOUT_H

#include <stdint.h>

#if defined(__CUDACC__)
#define EKF_HD __host__ __device__ __forceinline__
#else
#define EKF_HD static inline
#endif

#define EKF_STRIP 128      /* strip width (columns) */
#define EKF_PBLK 32        /* row blocking of the W / V panels */
#define EKF_CHUNK_ROWS 64  /* rows per streamed chunk in the covariance pass */

/* external (reference) index -> internal index */
EKF_HD int ekf_int_index(int ext) { return ext < 3 ? ext : ext + 1; }
EKF_HD int ekf_n_int(int n_landmarks) { return 4 + 2 * n_landmarks; }
EKF_HD int ekf_n_pad(int n_landmarks) {
  return (ekf_n_int(n_landmarks) + EKF_STRIP - 1) / EKF_STRIP * EKF_STRIP;
}
/* element offset of the first element of strip J */
EKF_HD int64_t ekf_strip_offset(int n_pad, int J) {
  return (int64_t)EKF_STRIP * ((int64_t)J * n_pad - (int64_t)EKF_STRIP * ((int64_t)J * (J - 1) / 2));
}
/* total number of stored doubles */
EKF_HD int64_t ekf_sigma_elems(int n_pad) { return ekf_strip_offset(n_pad, n_pad / EKF_STRIP); }
/* element offset of Sigma[i][j], internal indices, requires i >= (j & ~127) (i.e. inside strip) */
EKF_HD int64_t ekf_elem(int n_pad, int i, int j) {
  const int J = j / EKF_STRIP;
  return ekf_strip_offset(n_pad, J) + (int64_t)(i - J * EKF_STRIP) * EKF_STRIP + (j - J * EKF_STRIP);
}
/* offset of the symmetric entry (i,j) for any order of internal indices */
EKF_HD int64_t ekf_elem_sym(int n_pad, int i, int j) {
  return i >= j ? ekf_elem(n_pad, i, j) : ekf_elem(n_pad, j, i);
}

/* ---- row-sharded storage (one large map over several GPUs) --------------------------------------
 * A shard owns the rows [rb, re) of the lower triangle (rb, re multiples of 128, re <= n_pad) and stores
 * them in the same column strips: strip J holds columns [128J, 128J+128) for rows [max(128J, rb), re).
 * rb = 0, re = n_pad is the whole matrix and gives exactly the offsets of the functions above. */
typedef struct EkfLay { int rb, re; } EkfLay;
EKF_HD EkfLay ekf_lay_full(int n_pad) { EkfLay l; l.rb = 0; l.re = n_pad; return l; }
EKF_HD int ekf_lay_first_row(EkfLay l, int J) { return J * EKF_STRIP > l.rb ? J * EKF_STRIP : l.rb; }
EKF_HD int64_t ekf_lay_strip_offset(EkfLay l, int J) {
  const int Jb = l.rb / EKF_STRIP;
  if (J <= Jb) return (int64_t)EKF_STRIP * J * (l.re - l.rb);
  const int64_t head = (int64_t)EKF_STRIP * Jb * (l.re - l.rb);
  const int64_t cnt = J - Jb;
  const int64_t tri = (int64_t)J * (J - 1) / 2 - (int64_t)Jb * (Jb - 1) / 2;
  return head + (int64_t)EKF_STRIP * (cnt * l.re - (int64_t)EKF_STRIP * tri);
}
EKF_HD int64_t ekf_lay_elems(EkfLay l) { return ekf_lay_strip_offset(l, l.re / EKF_STRIP); }
/* requires rb <= i < re and i >= (j & ~127) */
EKF_HD int64_t ekf_lay_elem(EkfLay l, int i, int j) {
  const int J = j / EKF_STRIP;
  return ekf_lay_strip_offset(l, J) + (int64_t)(i - ekf_lay_first_row(l, J)) * EKF_STRIP + (j - J * EKF_STRIP);
}
EKF_HD int ekf_lay_owns_row(EkfLay l, int i) { return i >= l.rb && i < l.re; }
/* equal-area split of the triangle over `count` shards: boundary g at n_pad * sqrt(g / count), rounded
 * to the strip width; shard g owns [bound(g), bound(g+1)) */
EKF_HD int ekf_shard_bound(int n_pad, int g, int count) {
  if (g <= 0) return 0;
  if (g >= count) return n_pad;
  /* integer square root of n_pad^2 * g / count */
  const int64_t x = (int64_t)n_pad * n_pad * g / count;
  int64_t r = n_pad;
  while (r * r > x) r = (r + x / r) / 2;
  int b = (int)((r + EKF_STRIP / 2) / EKF_STRIP) * EKF_STRIP;
  if (b > n_pad) b = n_pad;
  return b;
}

/* W / V panels: n_pad rows x kmax columns (kmax a multiple of 4), laid out for the fp64 tensor-core
 * fragments of the covariance pass (mma.m8n8k4): [row / 32][k / 4][slot(row % 32)][k % 4].
 *   - a 32-row block of the first k2 columns is one contiguous range of k2 * 256 bytes (one bulk copy);
 *   - inside a k-step (4 columns) a row's 4 values are contiguous (32 B), and the 8 rows of one
 *     8 x 4 fragment are 8 consecutive slots: a fragment load is one conflict-free 256-byte
 *     warp-wide LDS.64 (lane l reads double l of the 8-slot group);
 *   - slot() interleaves the rows so that the four columns {4t .. 4t+3} a thread owns in two adjacent
 *     8-column tiles are contiguous in Sigma (one 256-bit load / store per thread):
 *         row = 16*jj + 4*t + 2*jb + e   ->   slot = 16*jj + 8*jb + 2*t + e.
 *     Rows (a, a+1) of one landmark (a even) stay in adjacent slots. */
EKF_HD int ekf_panel_slot(int r) { return (r & 16) | ((r & 2) << 2) | ((r >> 1) & 6) | (r & 1); }
/* inverse: row (0..31) held by a slot */
EKF_HD int ekf_panel_row_of_slot(int s) { return (s & 16) | ((s & 6) << 1) | ((s >> 2) & 2) | (s & 1); }
EKF_HD int64_t ekf_panel_elem(int kmax, int row, int k) {
  return (((int64_t)(row >> 5) * (kmax >> 2) + (k >> 2)) << 7) + (ekf_panel_slot(row & 31) << 2) + (k & 3);
}

#endif
