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
