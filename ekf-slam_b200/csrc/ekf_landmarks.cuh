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
// Jacobi eigen-solver for the symmetric 4x4 Q.  A cluster's points stay where the scan put them in shared
// memory (ClusterView maps the cluster's index to the scan's, the wrapped first cluster has two ranges) and
// the n x 4 matrix Z the SVD rotates in place lives in a shared workspace with one slot per reading - clusters
// are disjoint sets of readings, so they never collide.  (Per-thread arrays of 39 points went to local
// memory, 2.2 KB per thread, and every access of the Jacobi loops was an L2 round trip: 9.2 ms per 65,536
// scans; consecutive clusters are also dealt to different warps so that fewer fits serialise in one.)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_math.cuh"

namespace ekf {

constexpr int kLmThreadsDefault = 32;   // one warp per scan (measured against 64 and 128: see DESIGN.md section 10)
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

// A cluster's points inside the scan's shared-memory arrays: index i of the cluster is reading start + i for
// i < len, else tail0 + (i - len) (the run at the end of the scan that wraps into the first cluster).
struct ClusterView {
  int start, len, tail0;
  __device__ __forceinline__ int at(int i) const { return i < len ? start + i : tail0 + (i - len); }
};

// one-sided Jacobi SVD of Z (n x 4; column c of point i at zw[c * ld + cv.at(i)]): on exit the columns are U S,
// v holds V
__device__ inline void svd4(double *zw, int ld, const ClusterView &cv, int n, double v[4][4], double s[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double *zp = zw + p * ld, *zq = zw + q * ld;
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int i = 0; i < n; ++i) {
          const int o = cv.at(i);
          const double a = zp[o], b = zq[o];
          al += a * a; be += b * b; ga += a * b;
        }
        if (ga == 0.0) continue;
        const double rel = fabs(ga) / sqrt(al * be);
        if (rel > off) off = rel;
        if (rel < 1.0e-16) continue;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
        for (int i = 0; i < n; ++i) {
          const int o = cv.at(i);
          const double a = zp[o], b = zq[o];
          zp[o] = c * a - sn * b; zq[o] = sn * a + c * b;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const double a = v[i][p], b = v[i][q];
          v[i][p] = c * a - sn * b; v[i][q] = sn * a + c * b;
        }
      }
    if (off < 1.0e-15) break;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    double a = 0.0;
    for (int i = 0; i < n; ++i) { const double e = zw[j * ld + cv.at(i)]; a += e * e; }
    s[j] = sqrt(a);
  }
}

// cyclic Jacobi for a symmetric 4x4: a -> diagonal (eigenvalues), e = eigenvectors (columns)
__device__ inline void eig4(double a[4][4], double e[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) e[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0, diag = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      diag += a[i][i] * a[i][i];
#pragma unroll
      for (int j = i + 1; j < 4; ++j) off += a[i][j] * a[i][j];
    }
    if (off <= 1.0e-34 * diag || off == 0.0) break;
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        if (a[p][q] == 0.0) continue;
        const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(1.0 + theta * theta));
        const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
#pragma unroll
        for (int k = 0; k < 4; ++k) {            // A <- A J
          const double x = a[k][p], y = a[k][q];
          a[k][p] = c * x - sn * y; a[k][q] = sn * x + c * y;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {            // A <- J^T A
          const double x = a[p][k], y = a[q][k];
          a[p][k] = c * x - sn * y; a[q][k] = sn * x + c * y;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const double x = e[k][p], y = e[k][q];
          e[k][p] = c * x - sn * y; e[k][q] = sn * x + c * y;
        }
      }
  }
}

// turtlelib/src/landmark_detection.cpp:50-135 on the n points of a cluster; zw = Z workspace (4 columns of ld)
__device__ inline void fit_circle(const double *px, const double *py, double *zw, int ld, const ClusterView &cv, int n,
                                  double &cx, double &cy, double &R) {
  double mx = 0.0, my = 0.0;
  for (int i = 0; i < n; ++i) { const int o = cv.at(i); mx += px[o]; my += py[o]; }
  mx /= n; my /= n;
  double zm = 0.0;
  for (int i = 0; i < n; ++i) {
    const int o = cv.at(i);
    const double xi = px[o] - mx, yi = py[o] - my;
    const double z0 = xi * xi + yi * yi;
    zw[o] = z0; zw[ld + o] = xi; zw[2 * ld + o] = yi; zw[3 * ld + o] = 1.0;
    zm += z0;
  }
  zm /= n;
  double v[4][4], s[4];
  svd4(zw, ld, cv, n, v, s);
  int imin = 0;
#pragma unroll
  for (int j = 1; j < 4; ++j) if (s[j] < s[imin]) imin = j;
  double smin = s[0];
#pragma unroll
  for (int j = 1; j < 4; ++j) if (j == imin) smin = s[j];
  double A[4];
  if (smin < 10.0e-12) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      double a = v[i][0];
#pragma unroll
      for (int j = 1; j < 4; ++j) if (j == imin) a = v[i][j];
      A[i] = a;
    }
  } else {
    double Y[4][4], T[4][4], Q[4][4], E[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        double a = 0.0;
#pragma unroll
        for (int k = 0; k < 4; ++k) a += v[i][k] * s[k] * v[j][k];
        Y[i][j] = a;
      }
    // H^-1 = [[0,0,0,.5],[0,1,0,0],[0,0,1,0],[.5,0,0,-2 zmean]]: T = Y H^-1
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      T[i][0] = 0.5 * Y[i][3]; T[i][1] = Y[i][1]; T[i][2] = Y[i][2]; T[i][3] = 0.5 * Y[i][0] - 2.0 * zm * Y[i][3];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        double a = 0.0;
#pragma unroll
        for (int k = 0; k < 4; ++k) a += T[i][k] * Y[k][j];
        Q[i][j] = a;
      }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = i + 1; j < 4; ++j) { const double h = 0.5 * (Q[i][j] + Q[j][i]); Q[i][j] = h; Q[j][i] = h; }
    eig4(Q, E);
    int idx = -1;
    double min_val = 10.0e6;
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (Q[i][i] < min_val && Q[i][i] > 0.0) { min_val = Q[i][i]; idx = i; }
    if (idx < 0) {                               // no positive eigenvalue: the reference keeps index 0 = the smallest
      idx = 0;
      double best = Q[0][0];
#pragma unroll
      for (int i = 1; i < 4; ++i) if (Q[i][i] < best) { best = Q[i][i]; idx = i; }
    }
    double ev[4];                                // column idx of E
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      double a = E[i][0];
#pragma unroll
      for (int j = 1; j < 4; ++j) if (j == idx) a = E[i][j];
      ev[i] = a;
    }
    // A = Y^-1 A* = V S^-1 V^T A*
    double w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      double a = 0.0;
#pragma unroll
      for (int i = 0; i < 4; ++i) a += v[i][k] * ev[i];
      w[k] = a / s[k];
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      double a = 0.0;
#pragma unroll
      for (int k = 0; k < 4; ++k) a += v[i][k] * w[k];
      A[i] = a;
    }
  }
  const double a = -A[1] / 2.0 / A[0], b = -A[2] / 2.0 / A[0];
  R = sqrt((A[1] * A[1] + A[2] * A[2] - 4.0 * A[0] * A[3]) / 4.0 / A[0] / A[0]);
  cx = a + mx; cy = b + my;
}

// turtlelib/src/landmark_detection.cpp:5-48; ang = workspace of one double per reading (the cluster's own slots)
__device__ inline bool check_circle(const double *px, const double *py, double *ang, const ClusterView &cv, int n) {
  const int o0 = cv.at(0), oe = cv.at(n - 1);
  const double sx = px[o0], sy = py[o0], ex = px[oe], ey = py[oe];
  const double c = sqrt(sq_sum(sx - ex, sy - ey));
  double sum = 0.0;
  for (int j = 1; j < n - 1; ++j) {
    const int o = cv.at(j);
    const double a = sqrt(sq_sum(sx - px[o], sy - py[o])), b = sqrt(sq_sum(ex - px[o], ey - py[o]));
    const double an = acos((c * c - a * a - b * b) / (-2.0 * a * b));
    ang[o] = an;
    sum += an;
  }
  const int k = n - 2;
  const double mn = sum / k;
  double var = 0.0;
  for (int j = 1; j < n - 1; ++j) { const double d = ang[cv.at(j)] - mn; var += d * d; }
  const double stdev = k > 1 ? sqrt(var / (k - 1)) : 0.0;   // arma::stddev: N-1 normalisation
  return stdev < 0.2 && 1.3 < mn && mn < 2.6;
}

// block-wide exclusive prefix sum of one int per thread (kLmThreads threads); returns the total too
template <int kLmThreads>
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

template <int kLmThreads>
__global__ void __launch_bounds__(kLmThreads) ekf_landmarks_kernel(const LmParams p) {
  extern __shared__ __align__(16) unsigned char lm_smem[];
  const int n = p.n_ranges;
  double *px = reinterpret_cast<double *>(lm_smem);
  double *py = px + n;
  double *res = py + n;                 // [4][n]: workspace of the fits (Z columns / inscribed angles), then per cluster
                                        // (cx, cy, r, keep) at the cluster's first reading
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
  int rank = block_exscan<kLmThreads>(local, s_warp, G);
  for (int i = max(i0, 1); i < i1; ++i)
    if (sqrt(sq_sum(px[i] - px[i - 1], py[i] - py[i - 1])) > p.thr) gpos[rank++] = i;
  __syncthreads();

  if (G == 0) {                         // clusters.at(0) throws in the reference (landmarks.cpp:95)
    if (tid == 0) { p.count[scan] = 0; p.status[scan] = kLmErrNoGap; }
    return;
  }
  const bool wrap = sqrt(sq_sum(px[0] - px[n - 1], py[0] - py[n - 1])) <= p.thr;   // landmarks.cpp:93-103
  const int n_clusters = wrap ? G : G + 1;

  // one thread per cluster: size filter, circle test, fit, final filter (landmarks.cpp:117-149).  Cluster k goes to
  // lane k / 4 of warp k % 4; its result (cx, cy, r, keep) is left in the four workspace columns at the cluster's
  // first reading (an empty cluster - two gaps in a row, or a gap at the last reading - has no slot and is not kept)
  for (int k = (tid & 31) * (kLmThreads / 32) + (tid >> 5); k < n_clusters; k += kLmThreads) {
    ClusterView cv;
    cv.start = (k == 0) ? 0 : gpos[k - 1] + 1;
    const int end = (k < G) ? gpos[k] : n;                     // exclusive
    cv.len = end - cv.start;
    cv.tail0 = gpos[G - 1] + 1;
    const int tail = (k == 0 && wrap) ? n - cv.tail0 : 0;
    const int total = cv.len + tail;
    double keep = 0.0, cx = 0.0, cy = 0.0, R = 0.0;
    if (total > 3 && total <= kLmMaxPts) {
      if (check_circle(px, py, res, cv, total)) {
        fit_circle(px, py, res, n, cv, total, cx, cy, R);
        if (R < 0.2 && sqrt(sq_sum(cx, cy)) < 2.0) keep = 1.0;
      }
    }
    if (cv.len > 0) { res[cv.start] = cx; res[n + cv.start] = cy; res[2 * n + cv.start] = R; res[3 * n + cv.start] = keep; }
  }
  __syncthreads();
  // ordered compaction: detections come out in cluster order (marker id = running index, :137-148)
  auto slot_of = [&](int k) {                                  // the cluster's first reading, or -1 for an empty cluster
    const int start = (k == 0) ? 0 : gpos[k - 1] + 1;
    const int end = (k < G) ? gpos[k] : n;
    return end > start ? start : -1;
  };
  int keep_local = 0;
  const int perc = (n_clusters + kLmThreads - 1) / kLmThreads;
  const int c0 = tid * perc, c1 = min(n_clusters, c0 + perc);
  for (int k = c0; k < c1; ++k) { const int o = slot_of(k); keep_local += (o >= 0 && res[3 * n + o] != 0.0); }
  int total_keep = 0;
  int orank = block_exscan<kLmThreads>(keep_local, s_warp, total_keep);
  for (int k = c0; k < c1; ++k) {
    const int o = slot_of(k);
    if (o >= 0 && res[3 * n + o] != 0.0) {
      if (orank < p.max_out) {
        double *d = p.out + ((size_t)scan * p.max_out + orank) * 3;
        d[0] = res[o]; d[1] = res[n + o]; d[2] = res[2 * n + o];
      }
      ++orank;
    }
  }
  if (tid == 0) { p.count[scan] = total_keep; p.status[scan] = 0; }
}

}  // namespace ekf
