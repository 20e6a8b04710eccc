// Landmark extraction, the step right before the EKF hot path: a batch of
// LaserScans -> clusters -> circle test -> algebraic circle fit -> (c_x, c_y, r) detections, i.e. the
// sightings `ekf_update_unknown` consumes.  Three kernels per batch of scans.
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
//
// Shape (round 2, from three ncu captures of the earlier shapes - profiles/r2z_*, r2ac_*):
//   1. ekf_landmarks_segment_kernel, a warp per scan: points (the beams' (cos, sin) are the same for every scan of a
//      call: a table made once per call), gaps by a warp scan, and the clusters that pass the size filter appended to
//      the batch's task list (one atomic per scan; a scan's tasks are contiguous and in cluster order);
//   2. ekf_landmarks_fit_kernel, a THREAD per candidate cluster of the whole batch: circle test and fit.  A scan has
//      only a handful of candidates, so with the fits inside the per-scan warp 3.8 of 32 lanes were active per
//      instruction; here every lane of a warp holds a cluster (of whatever scan).  The points are re-evaluated from
//      the ranges and the table wherever they are needed (same expression as in kernel 1: LmPoints::get); Z is never
//      stored: a modified Gram-Schmidt QR evaluated on the fly (five passes over the cluster's points) leaves the
//      4 x 4 triangular factor R in registers and the one-sided (Hestenes) Jacobi SVD runs on R - same singular
//      values and right singular vectors as Z, accurate to eps * |Z| like the reference's LAPACK path; the
//      inscribed angles of the circle test are summed in one pass (shifted by the first angle);
//   3. ekf_landmarks_collect_kernel, a thread per scan: the kept detections in cluster order.
// Code size matters here (divergent threads all over a long kernel): the Jacobi rotation scalars are one
// __noinline__ copy and the loops over a cluster's points are not unrolled (109 -> 58 KB gave +32 %).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_math.cuh"

namespace ekf {

constexpr int kLmThreads = 32;       // segmentation: one warp per scan
constexpr int kLmMaxPts = 39;        // clusters.at(i).size() < 40: landmarks.cpp:122
constexpr int kLmErrNoGap = 3;       // the reference throws at clusters.at(0): landmarks.cpp:95

// One cluster that reaches the circle test (4 .. 39 readings): which scan, and its readings as a ClusterView
struct LmTask { int scan, start, len, tail0, tail; };

struct LmParams {
  const float *ranges;     // [n_scans][n_ranges]
  const double2 *beam;     // [n_ranges] (cos, sin) of the beam angles of this call (lm_beam_table_kernel)
  int n_scans, n_ranges, max_out;
  double thr;              // distance_clustering_threshold (landmarks.cpp:195)
  double *out;             // [n_scans][max_out][3] (c_x, c_y, r), cluster order
  int *count;              // [n_scans] detections found (may exceed max_out; only max_out are stored)
  int *status;             // [n_scans] 0 or kLmErrNoGap
  // between the three kernels of a call
  LmTask *tasks;           // [task_cap] the batch's candidate clusters, each scan's contiguous and in cluster order
  double4 *res;            // [task_cap] (c_x, c_y, r, keep) of each task
  int *n_tasks;            // running total (zeroed before the call)
  int *first_task;         // [n_scans] where the scan's tasks start
  int *n_task_of;          // [n_scans] how many it has
  int task_cap;
};

__host__ __device__ inline size_t lm_smem_bytes(int n_ranges) {
  return (size_t)n_ranges * (2 * sizeof(double) + sizeof(unsigned short)) + 16;
}
// a candidate cluster has at least four readings of its own
__host__ __device__ inline int lm_max_tasks_per_scan(int n_ranges) { return n_ranges / 4 + 1; }

// The readings of one scan as points in the robot frame (landmarks.cpp:66-69: the lidar sits 0.032 m ahead).  Both the
// segmentation kernel and the fit kernel evaluate a point through this one expression.
struct LmPoints {
  const float *rg;
  const double2 *beam;
  __device__ __forceinline__ void get(int o, double &x, double &y) const {
    const double r = (double)__ldg(rg + o);
    const double2 cs = __ldg(beam + o);
    x = r * cs.x - 0.032;
    y = r * cs.y;
  }
};

// landmarks.cpp:66-67: the angle of beam i, the same for every scan of a call
__global__ void lm_beam_table_kernel(double2 *beam, int n, float angle_min, float angle_increment) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double ang = normalize_angle((double)i * (double)angle_increment) + (double)angle_min;
  beam[i] = make_double2(cos(ang), sin(ang));
}

// A cluster's points inside the scan's shared-memory arrays: index i of the cluster is reading start + i for
// i < len, else tail0 + (i - len) (the run at the end of the scan that wraps into the first cluster).
struct ClusterView {
  int start, len, tail0;
  __device__ __forceinline__ int at(int i) const { return i < len ? start + i : tail0 + (i - len); }
};

// Jacobi rotation that zeroes the (p, q) coupling: (c, s) from zeta = (beta - alpha) / (2 gamma).  Deliberately ONE
// copy in the kernel (two divisions and two square roots in fp64 are ~150 instructions; inlined into the twelve unrolled
// rotations of svd4x4 and eig4 the kernel was 109 KB of code and its warps - sixteen per SM, each somewhere else in it -
// waited for instruction fetch 6.5 cycles per issued instruction, `stalled_no_instruction` in the ncu capture)
__device__ __noinline__ double2 jacobi_cs(double num, double den) {
  const double zeta = num / den;
  const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
  const double c = 1.0 / sqrt(1.0 + t * t);
  return make_double2(c, c * t);
}

// one-sided Jacobi SVD of a 4 x 4 matrix held by columns (z[c][row]): on exit the columns are U S, v holds V
__device__ __forceinline__ void svd4x4(double z[4][4], double v[4][4], double s[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
  // the cosine of the angle between two columns, |gamma| / sqrt(alpha beta), is compared in squared form (no square
  // root, no division): rotate while it is >= 1e-16, done when every pair of a sweep was below 1e-15
  for (int sweep = 0; sweep < 30; ++sweep) {
    bool done = true;
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { al += z[p][i] * z[p][i]; be += z[q][i] * z[q][i]; ga += z[p][i] * z[q][i]; }
        if (ga != 0.0) {
          const double g2 = ga * ga, ab = al * be;
          if (!(g2 < 1.0e-30 * ab)) done = false;
          if (g2 >= 1.0e-32 * ab) {
            const double2 cs = jacobi_cs(be - al, 2.0 * ga);
            const double c = cs.x, sn = cs.y;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const double a = z[p][i], b = z[q][i];
              z[p][i] = c * a - sn * b; z[q][i] = sn * a + c * b;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const double a = v[i][p], b = v[i][q];
              v[i][p] = c * a - sn * b; v[i][q] = sn * a + c * b;
            }
          }
        }
      }
    if (done) break;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    double a = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) a += z[j][i] * z[j][i];
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
    if (off <= 1.0e-30 * diag || off == 0.0) break;      // off-diagonal norm below 1e-15 of the diagonal's
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        if (a[p][q] == 0.0) continue;
        const double2 cs = jacobi_cs(a[q][q] - a[p][p], 2.0 * a[p][q]);
        const double c = cs.x, sn = cs.y;
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

// turtlelib/src/landmark_detection.cpp:50-135 on the n points of a cluster.  Z = [z x y 1] (centred points,
// z = x^2 + y^2) is reduced to its triangular factor R by modified Gram-Schmidt without being stored: pass c
// recomputes, per point, the row of Z and what the earlier columns left of it, and accumulates column c's norm
// and its products with the later columns.  The SVD of R gives Z's singular values and V.
__device__ inline void fit_circle(const LmPoints &P, const ClusterView &cv, int n, double &cx, double &cy, double &R) {
  double mx = 0.0, my = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) { double x, y; P.get(cv.at(i), x, y); mx += x; my += y; }
  mx /= n; my /= n;
  double s00 = 0.0, s01 = 0.0, s02 = 0.0, s03 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    double x, y;
    P.get(cv.at(i), x, y);
    const double xi = x - mx, yi = y - my;
    const double z0 = xi * xi + yi * yi;
    s00 += z0 * z0; s01 += z0 * xi; s02 += z0 * yi; s03 += z0;
  }
  const double zm = s03 / n;                                   // mean of z (landmark_detection.cpp:72-77)
  const double r00 = sqrt(s00), i0 = r00 > 0.0 ? 1.0 / r00 : 0.0;
  const double r01 = s01 * i0, r02 = s02 * i0, r03 = s03 * i0;
  double s11 = 0.0, s12 = 0.0, s13 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    double x, y;
    P.get(cv.at(i), x, y);
    const double xi = x - mx, yi = y - my;
    const double q0 = (xi * xi + yi * yi) * i0;
    const double a1 = xi - r01 * q0, a2 = yi - r02 * q0, a3 = 1.0 - r03 * q0;
    s11 += a1 * a1; s12 += a1 * a2; s13 += a1 * a3;
  }
  const double r11 = sqrt(s11), i1 = r11 > 0.0 ? 1.0 / r11 : 0.0;
  const double r12 = s12 * i1, r13 = s13 * i1;
  double s22 = 0.0, s23 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    double x, y;
    P.get(cv.at(i), x, y);
    const double xi = x - mx, yi = y - my;
    const double q0 = (xi * xi + yi * yi) * i0;
    const double a1 = xi - r01 * q0, a2 = yi - r02 * q0, a3 = 1.0 - r03 * q0;
    const double q1 = a1 * i1;
    const double b2 = a2 - r12 * q1, b3 = a3 - r13 * q1;
    s22 += b2 * b2; s23 += b2 * b3;
  }
  const double r22 = sqrt(s22), i2 = r22 > 0.0 ? 1.0 / r22 : 0.0;
  const double r23 = s23 * i2;
  double s33 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    double x, y;
    P.get(cv.at(i), x, y);
    const double xi = x - mx, yi = y - my;
    const double q0 = (xi * xi + yi * yi) * i0;
    const double a1 = xi - r01 * q0, a2 = yi - r02 * q0, a3 = 1.0 - r03 * q0;
    const double q1 = a1 * i1;
    const double b2 = a2 - r12 * q1, b3 = a3 - r13 * q1;
    const double c3 = b3 - r23 * (b2 * i2);
    s33 += c3 * c3;
  }
  const double r33 = sqrt(s33);
  double rc[4][4] = {{r00, 0.0, 0.0, 0.0}, {r01, r11, 0.0, 0.0}, {r02, r12, r22, 0.0}, {r03, r13, r23, r33}};   // columns of R
  double v[4][4], s[4];
  svd4x4(rc, v, s);
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

// turtlelib/src/landmark_detection.cpp:5-48.  Mean and N-1 standard deviation of the inscribed angles in one pass
// (sums of the angles' differences from the first one: no cancellation to speak of, no workspace)
__device__ inline bool check_circle(const LmPoints &P, const ClusterView &cv, int n) {
  double sx, sy, ex, ey;
  P.get(cv.at(0), sx, sy);
  P.get(cv.at(n - 1), ex, ey);
  const double c = sqrt(sq_sum(sx - ex, sy - ey));
  double first = 0.0, sd = 0.0, sd2 = 0.0;
#pragma unroll 1
  for (int j = 1; j < n - 1; ++j) {
    double x, y;
    P.get(cv.at(j), x, y);
    const double a = sqrt(sq_sum(sx - x, sy - y)), b = sqrt(sq_sum(ex - x, ey - y));
    const double an = acos((c * c - a * a - b * b) / (-2.0 * a * b));
    if (j == 1) first = an;
    const double d = an - first;
    sd += d; sd2 += d * d;
  }
  const int k = n - 2;
  const double mn = first + sd / k;
  const double var = k > 1 ? (sd2 - sd * sd / k) / (k - 1) : 0.0;
  const double stdev = sqrt(var > 0.0 ? var : 0.0);          // arma::stddev: N-1 normalisation
  return stdev < 0.2 && 1.3 < mn && mn < 2.6;
}

// ---- kernel 1 of 3: segmentation, one warp per scan -> the scan's candidate clusters as tasks --------------------
__global__ void __launch_bounds__(kLmThreads) ekf_landmarks_segment_kernel(const LmParams p) {
  extern __shared__ __align__(16) unsigned char lm_smem[];
  const int n = p.n_ranges;
  double *px = reinterpret_cast<double *>(lm_smem);
  double *py = px + n;
  unsigned short *gpos = reinterpret_cast<unsigned short *>(py + n);   // gap positions, in order (n <= 4096)
  const int scan = blockIdx.x, lane = threadIdx.x;
  if (scan >= p.n_scans) return;
  const LmPoints P{p.ranges + (size_t)scan * n, p.beam};

  for (int i = lane; i < n; i += kLmThreads) P.get(i, px[i], py[i]);
  __syncwarp();

  // gaps (landmarks.cpp:75-86), ordered positions by a warp scan over contiguous per-lane slices
  const int per = (n + kLmThreads - 1) / kLmThreads;
  const int i0 = lane * per, i1 = min(n, i0 + per);
  auto is_gap = [&](int i) { return sqrt(sq_sum(px[i] - px[i - 1], py[i] - py[i - 1])) > p.thr; };
  unsigned long long gmask = 0;                          // the slice's gaps when it is at most 64 readings long
  int local = 0;
#pragma unroll 1
  for (int i = max(i0, 1); i < i1; ++i)
    if (is_gap(i)) { ++local; if (per <= 64) gmask |= 1ULL << (i - i0); }
  int inc = local;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= off) inc += o;
  }
  const int G = __shfl_sync(0xffffffffu, inc, 31);
  int rank = inc - local;
#pragma unroll 1
  for (int i = max(i0, 1); i < i1; ++i)
    if (per <= 64 ? (int)((gmask >> (i - i0)) & 1ULL) : (int)is_gap(i)) gpos[rank++] = (unsigned short)i;
  __syncwarp();

  if (G == 0) {                         // clusters.at(0) throws in the reference (landmarks.cpp:95)
    if (lane == 0) { p.first_task[scan] = 0; p.n_task_of[scan] = 0; p.status[scan] = kLmErrNoGap; }
    return;
  }
  const bool wrap = sqrt(sq_sum(px[0] - px[n - 1], py[0] - py[n - 1])) <= p.thr;   // landmarks.cpp:93-103
  const int n_clusters = wrap ? G : G + 1;
  const int tail0 = (int)gpos[G - 1] + 1;

  // the clusters that pass the size filter (landmarks.cpp:122: more than 3, fewer than 40 readings), in cluster order:
  // counted, then written behind a base taken from the batch's running total
  auto view = [&](int k, ClusterView &cv, int &tail) {
    cv.start = (k == 0) ? 0 : (int)gpos[k - 1] + 1;
    const int end = (k < G) ? (int)gpos[k] : n;                  // exclusive
    cv.len = end - cv.start;
    cv.tail0 = tail0;
    tail = (k == 0 && wrap) ? n - tail0 : 0;
    const int total = cv.len + tail;
    return total > 3 && total <= kLmMaxPts;
  };
  int n_cand = 0;
  for (int base = 0; base < n_clusters; base += kLmThreads) {
    ClusterView cv; int tail;
    const bool cand = base + lane < n_clusters && view(base + lane, cv, tail);
    n_cand += __popc(__ballot_sync(0xffffffffu, cand));
  }
  int first = 0;
  if (lane == 0 && n_cand > 0) first = atomicAdd(p.n_tasks, n_cand);
  first = __shfl_sync(0xffffffffu, first, 0);
  if (first + n_cand > p.task_cap) n_cand = max(0, p.task_cap - first);   // (cannot happen: the capacity is the bound)
  int written = 0;
  for (int base = 0; base < n_clusters; base += kLmThreads) {
    ClusterView cv; int tail;
    const bool cand = base + lane < n_clusters && view(base + lane, cv, tail);
    const unsigned cm = __ballot_sync(0xffffffffu, cand);
    const int j = written + __popc(cm & ((1u << lane) - 1u));
    if (cand && j < n_cand) p.tasks[first + j] = LmTask{scan, cv.start, cv.len, cv.tail0, tail};
    written += __popc(cm);
  }
  if (lane == 0) { p.first_task[scan] = first; p.n_task_of[scan] = n_cand; p.status[scan] = 0; }
}

// ---- kernel 2 of 3: circle test + fit, one THREAD per candidate cluster of the whole batch ---------------------------
// (A scan has only a handful of candidates: with the fits inside the per-scan warp 3.8 of 32 lanes were active per
// instruction.  Here the lanes of a warp hold clusters of different scans and all of them work.)
constexpr int kLmFitThreads = 128;
__global__ void __launch_bounds__(kLmFitThreads, 4) ekf_landmarks_fit_kernel(const LmParams p) {
  const int total = min(*p.n_tasks, p.task_cap);
  for (int t = blockIdx.x * kLmFitThreads + threadIdx.x; t < total; t += gridDim.x * kLmFitThreads) {
    const LmTask task = p.tasks[t];
    const LmPoints P{p.ranges + (size_t)task.scan * p.n_ranges, p.beam};
    ClusterView cv;
    cv.start = task.start; cv.len = task.len; cv.tail0 = task.tail0;
    const int n_pts = task.len + task.tail;
    double cx = 0.0, cy = 0.0, R = 0.0, keep = 0.0;
    if (check_circle(P, cv, n_pts)) {                            // landmarks.cpp:124-149
      fit_circle(P, cv, n_pts, cx, cy, R);
      if (R < 0.2 && sqrt(sq_sum(cx, cy)) < 2.0) keep = 1.0;
    }
    p.res[t] = make_double4(cx, cy, R, keep);
  }
}

// ---- kernel 3 of 3: the kept detections of every scan, in cluster order (marker id = running index, :137-148) ----------
__global__ void ekf_landmarks_collect_kernel(const LmParams p) {
  const int scan = blockIdx.x * blockDim.x + threadIdx.x;
  if (scan >= p.n_scans) return;
  const int first = p.first_task[scan], cnt = p.n_task_of[scan];
  int written = 0;
  for (int j = 0; j < cnt; ++j) {
    const double4 r = p.res[first + j];
    if (r.w != 0.0) {
      if (written < p.max_out) {
        double *d = p.out + ((size_t)scan * p.max_out + written) * 3;
        d[0] = r.x; d[1] = r.y; d[2] = r.z;
      }
      ++written;
    }
  }
  p.count[scan] = written;
}

}  // namespace ekf
