// Landmark extraction, the step right before the EKF hot path: a batch of
// LaserScans -> clusters -> circle test -> algebraic circle fit -> (c_x, c_y, r) detections, i.e. the
// sightings `ekf_update_unknown` consumes.  One CTA per scan.
//
// Replaces, per scan,
//   nuslam/src/landmarks.cpp:58-106    getClusters: beam -> (x, y) (lidar 0.032 m ahead of the base),
//                                      split where consecutive readings are > threshold apart; the reading
//                                      that opens a gap is dropped (:80-85); the last run is merged into
//                                      the first when the scan wraps (:93-103)
//   nuslam/src/landmarks.cpp:109-156   size filter 4..39, circle test, fit, r < 0.2 && range < 2.0
//   turtlelib/src/landmark_detection.cpp:5-48    checkCircle (inscribed angles: stddev < 0.2, 1.3 < mean < 2.6)
//   turtlelib/src/landmark_detection.cpp:50-135  fitCircle: SVD of Z = [z x y 1], Y = V S V^T,
//                                      Q = Y H^-1 Y, eigenvector of the smallest positive eigenvalue,
//                                      A = Y^-1 A*   (arma::svd / eig_sym / solve in the reference)
// The dense factorizations are 4-wide, so each cluster is one thread: a one-sided (Hestenes) Jacobi
// SVD of Z (accurate for the small singular values the reference's 1e-11 test looks at) and a cyclic
// Jacobi eigen-solver for the symmetric 4x4 Q.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_math.cuh"

namespace ekf {

constexpr int kLmThreads = 128;
constexpr int kLmMaxPts = 39;        // clusters.at(i).size() < 40: landmarks.cpp:122
constexpr int kLmErrNoGap = 3;       // the reference throws at clusters.at(0): landmarks.cpp:95

struct LmParams {
  const float *ranges;     // [n_scans][n_ranges]
  int n_scans, n_ranges, max_out;
  float angle_min, angle_increment;
  double thr;              // distance_clustering_threshold (landmarks.cpp:195)
  double *out;             // [n_scans][max_out][3] (c_x, c_y, r), cluster order
  int *count;              // [n_scans] detections found (may exceed max_out; only max_out are stored)
  int *status;             // [n_scans] 0 or kLmErrNoGap
};

__host__ __device__ inline size_t lm_smem_bytes(int n_ranges) {
  return (size_t)n_ranges * (2 * sizeof(double) + 2 * sizeof(int)) + 4 * (size_t)n_ranges * sizeof(double) + 64;
}

// one-sided Jacobi SVD of Z (n x 4, column-major in z[4][n]): on exit the columns are U S, v holds V
__device__ inline void svd4(double z[4][kLmMaxPts], int n, double v[4][4], double s[4]) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 3; ++p)
      for (int q = p + 1; q < 4; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int i = 0; i < n; ++i) { al += z[p][i] * z[p][i]; be += z[q][i] * z[q][i]; ga += z[p][i] * z[q][i]; }
        if (ga == 0.0) continue;
        const double rel = fabs(ga) / sqrt(al * be);
        if (rel > off) off = rel;
        if (rel < 1.0e-16) continue;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
        for (int i = 0; i < n; ++i) {
          const double a = z[p][i], b = z[q][i];
          z[p][i] = c * a - sn * b; z[q][i] = sn * a + c * b;
        }
        for (int i = 0; i < 4; ++i) {
          const double a = v[i][p], b = v[i][q];
          v[i][p] = c * a - sn * b; v[i][q] = sn * a + c * b;
        }
      }
    if (off < 1.0e-15) break;
  }
  for (int j = 0; j < 4; ++j) {
    double a = 0.0;
    for (int i = 0; i < n; ++i) a += z[j][i] * z[j][i];
    s[j] = sqrt(a);
  }
}

// cyclic Jacobi for a symmetric 4x4: a -> diagonal (eigenvalues), e = eigenvectors (columns)
__device__ inline void eig4(double a[4][4], double e[4][4]) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) e[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < 4; ++i) { diag += a[i][i] * a[i][i]; for (int j = i + 1; j < 4; ++j) off += a[i][j] * a[i][j]; }
    if (off <= 1.0e-34 * diag || off == 0.0) break;
    for (int p = 0; p < 3; ++p)
      for (int q = p + 1; q < 4; ++q) {
        if (a[p][q] == 0.0) continue;
        const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(1.0 + theta * theta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
        for (int k = 0; k < 4; ++k) {            // A <- A J
          const double x = a[k][p], y = a[k][q];
          a[k][p] = c * x - sn * y; a[k][q] = sn * x + c * y;
        }
        for (int k = 0; k < 4; ++k) {            // A <- J^T A
          const double x = a[p][k], y = a[q][k];
          a[p][k] = c * x - sn * y; a[q][k] = sn * x + c * y;
        }
        for (int k = 0; k < 4; ++k) {
          const double x = e[k][p], y = e[k][q];
          e[k][p] = c * x - sn * y; e[k][q] = sn * x + c * y;
        }
      }
  }
}

// turtlelib/src/landmark_detection.cpp:50-135 on n points (x[i], y[i])
__device__ inline void fit_circle(const double *x, const double *y, int n, double &cx, double &cy, double &R) {
  double mx = 0.0, my = 0.0;
  for (int i = 0; i < n; ++i) { mx += x[i]; my += y[i]; }
  mx /= n; my /= n;
  double z[4][kLmMaxPts];
  double zm = 0.0;
  for (int i = 0; i < n; ++i) {
    const double xi = x[i] - mx, yi = y[i] - my;
    z[0][i] = xi * xi + yi * yi; z[1][i] = xi; z[2][i] = yi; z[3][i] = 1.0;
    zm += z[0][i];
  }
  zm /= n;
  double v[4][4], s[4];
  svd4(z, n, v, s);
  int imin = 0;
  for (int j = 1; j < 4; ++j) if (s[j] < s[imin]) imin = j;
  double A[4];
  if (s[imin] < 10.0e-12) {
    for (int i = 0; i < 4; ++i) A[i] = v[i][imin];
  } else {
    double Y[4][4], T[4][4], Q[4][4], E[4][4];
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        double a = 0.0;
        for (int k = 0; k < 4; ++k) a += v[i][k] * s[k] * v[j][k];
        Y[i][j] = a;
      }
    // H^-1 = [[0,0,0,.5],[0,1,0,0],[0,0,1,0],[.5,0,0,-2 zmean]]: T = Y H^-1
    for (int i = 0; i < 4; ++i) {
      T[i][0] = 0.5 * Y[i][3]; T[i][1] = Y[i][1]; T[i][2] = Y[i][2]; T[i][3] = 0.5 * Y[i][0] - 2.0 * zm * Y[i][3];
    }
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        double a = 0.0;
        for (int k = 0; k < 4; ++k) a += T[i][k] * Y[k][j];
        Q[i][j] = a;
      }
    for (int i = 0; i < 4; ++i)
      for (int j = i + 1; j < 4; ++j) { const double h = 0.5 * (Q[i][j] + Q[j][i]); Q[i][j] = h; Q[j][i] = h; }
    eig4(Q, E);
    int idx = -1;
    double min_val = 10.0e6;
    for (int i = 0; i < 4; ++i)
      if (Q[i][i] < min_val && Q[i][i] > 0.0) { min_val = Q[i][i]; idx = i; }
    if (idx < 0) {                               // no positive eigenvalue: the reference keeps index 0 = the smallest
      idx = 0;
      for (int i = 1; i < 4; ++i) if (Q[i][i] < Q[idx][idx]) idx = i;
    }
    // A = Y^-1 A* = V S^-1 V^T A*
    double w[4];
    for (int k = 0; k < 4; ++k) {
      double a = 0.0;
      for (int i = 0; i < 4; ++i) a += v[i][k] * E[i][idx];
      w[k] = a / s[k];
    }
    for (int i = 0; i < 4; ++i) {
      double a = 0.0;
      for (int k = 0; k < 4; ++k) a += v[i][k] * w[k];
      A[i] = a;
    }
  }
  const double a = -A[1] / 2.0 / A[0], b = -A[2] / 2.0 / A[0];
  R = sqrt((A[1] * A[1] + A[2] * A[2] - 4.0 * A[0] * A[3]) / 4.0 / A[0] / A[0]);
  cx = a + mx; cy = b + my;
}

// turtlelib/src/landmark_detection.cpp:5-48
__device__ inline bool check_circle(const double *x, const double *y, int n) {
  const double sx = x[0], sy = y[0], ex = x[n - 1], ey = y[n - 1];
  const double c = sqrt(sq_sum(sx - ex, sy - ey));
  double sum = 0.0;
  double ang[kLmMaxPts];
  for (int j = 1; j < n - 1; ++j) {
    const double a = sqrt(sq_sum(sx - x[j], sy - y[j])), b = sqrt(sq_sum(ex - x[j], ey - y[j]));
    ang[j - 1] = acos((c * c - a * a - b * b) / (-2.0 * a * b));
    sum += ang[j - 1];
  }
  const int k = n - 2;
  const double mn = sum / k;
  double var = 0.0;
  for (int j = 0; j < k; ++j) var += (ang[j] - mn) * (ang[j] - mn);
  const double stdev = k > 1 ? sqrt(var / (k - 1)) : 0.0;   // arma::stddev: N-1 normalisation
  return stdev < 0.2 && 1.3 < mn && mn < 2.6;
}

// block-wide exclusive prefix sum of one int per thread (kLmThreads threads); returns the total too
__device__ inline int block_exscan(int v, int *s_warp, int &total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  int base = 0, tot = 0;
  for (int w = 0; w < kLmThreads / 32; ++w) { if (w < warp) base += s_warp[w]; tot += s_warp[w]; }
  __syncthreads();
  total = tot;
  return base + inc - v;
}

__global__ void __launch_bounds__(kLmThreads) ekf_landmarks_kernel(const LmParams p) {
  extern __shared__ __align__(16) unsigned char lm_smem[];
  const int n = p.n_ranges;
  double *px = reinterpret_cast<double *>(lm_smem);
  double *py = px + n;
  double *res = py + n;                 // [n][4]: per cluster (cx, cy, r, keep)
  int *gpos = reinterpret_cast<int *>(res + 4 * (size_t)n);   // gap positions, in order
  int *s_warp = gpos + n;               // scan scratch (needs 4) - the tail of the buffer
  const int scan = blockIdx.x, tid = threadIdx.x;
  if (scan >= p.n_scans) return;
  const float *rg = p.ranges + (size_t)scan * n;

  // landmarks.cpp:66-69
  for (int i = tid; i < n; i += kLmThreads) {
    const double ang = normalize_angle((double)i * (double)p.angle_increment) + (double)p.angle_min;
    const double r = (double)rg[i];
    px[i] = r * cos(ang) - 0.032;
    py[i] = r * sin(ang);
  }
  __syncthreads();

  // gaps (landmarks.cpp:75-86), ordered positions by a block scan over contiguous per-thread slices
  const int per = (n + kLmThreads - 1) / kLmThreads;
  const int i0 = tid * per, i1 = min(n, i0 + per);
  int local = 0;
  for (int i = max(i0, 1); i < i1; ++i)
    if (sqrt(sq_sum(px[i] - px[i - 1], py[i] - py[i - 1])) > p.thr) ++local;
  int G = 0;
  int rank = block_exscan(local, s_warp, G);
  for (int i = max(i0, 1); i < i1; ++i)
    if (sqrt(sq_sum(px[i] - px[i - 1], py[i] - py[i - 1])) > p.thr) gpos[rank++] = i;
  __syncthreads();

  if (G == 0) {                         // clusters.at(0) throws in the reference (landmarks.cpp:95)
    if (tid == 0) { p.count[scan] = 0; p.status[scan] = kLmErrNoGap; }
    return;
  }
  const bool wrap = sqrt(sq_sum(px[0] - px[n - 1], py[0] - py[n - 1])) <= p.thr;   // landmarks.cpp:93-103
  const int n_clusters = wrap ? G : G + 1;

  // one thread per cluster: size filter, circle test, fit, final filter (landmarks.cpp:117-149)
  int keep_local = 0;
  for (int k = tid; k < n_clusters; k += kLmThreads) {
    const int start = (k == 0) ? 0 : gpos[k - 1] + 1;
    const int end = (k < G) ? gpos[k] : n;                     // exclusive
    int len = end - start;
    const int tail0 = gpos[G - 1] + 1, tail = (k == 0 && wrap) ? n - tail0 : 0;
    const int total = len + tail;
    double keep = 0.0, cx = 0.0, cy = 0.0, R = 0.0;
    if (total > 3 && total <= kLmMaxPts) {
      double x[kLmMaxPts], y[kLmMaxPts];
      for (int i = 0; i < len; ++i) { x[i] = px[start + i]; y[i] = py[start + i]; }
      for (int i = 0; i < tail; ++i) { x[len + i] = px[tail0 + i]; y[len + i] = py[tail0 + i]; }
      if (check_circle(x, y, total)) {
        fit_circle(x, y, total, cx, cy, R);
        if (R < 0.2 && sqrt(sq_sum(cx, cy)) < 2.0) keep = 1.0;
      }
    }
    res[4 * k] = cx; res[4 * k + 1] = cy; res[4 * k + 2] = R; res[4 * k + 3] = keep;
  }
  __syncthreads();
  // ordered compaction: detections come out in cluster order (marker id = running index, :137-148)
  const int perc = (n_clusters + kLmThreads - 1) / kLmThreads;
  const int c0 = tid * perc, c1 = min(n_clusters, c0 + perc);
  for (int k = c0; k < c1; ++k) keep_local += res[4 * k + 3] != 0.0;
  int total_keep = 0;
  int orank = block_exscan(keep_local, s_warp, total_keep);
  for (int k = c0; k < c1; ++k)
    if (res[4 * k + 3] != 0.0) {
      if (orank < p.max_out) {
        double *o = p.out + ((size_t)scan * p.max_out + orank) * 3;
        o[0] = res[4 * k]; o[1] = res[4 * k + 1]; o[2] = res[4 * k + 2];
      }
      ++orank;
    }
  if (tid == 0) { p.count[scan] = total_keep; p.status[scan] = 0; }
}

}  // namespace ekf
