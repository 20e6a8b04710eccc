// Landmark extraction, the step right before the EKF hot path: a batch of
// LaserScans -> clusters -> circle test -> algebraic circle fit -> (c_x, c_y, r) detections, i.e. the
// sightings `ekf_update_unknown` consumes.  One warp per scan.
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
// Shape (round 2; the ncu capture of the first shape - a thread per cluster rotating the n x 4 matrix Z in a
// shared workspace - showed a latency-bound kernel: 10 warps per SM, 16 % of the issue slots, 16 cycles per
// instruction, a third of the instructions in the Jacobi loops over the points):
//   * the beams' (cos, sin) are the same for every scan of a launch: a table made once per call;
//   * a cluster is still one thread (the factorizations are 4 wide), but Z is never stored: a modified
//     Gram-Schmidt QR evaluated on the fly (five passes over the cluster's points, each recomputing the point's
//     row of Z and its projections from px / py) leaves the 4 x 4 triangular factor R in registers, and the
//     one-sided (Hestenes) Jacobi SVD runs on R - same singular values and right singular vectors as Z, R
//     accurate to eps * |Z| like the reference's LAPACK path, 100 n instead of 340 n instructions per cluster;
//   * the inscribed angles of the circle test are summed in one pass (shifted by the first angle), so they
//     need no workspace either;
//   * detections go straight to the output in cluster order (ballot + popcount, no staging).
// Shared memory per scan is then px, py and the gap positions only (18 bytes per reading, 6.5 KB at 360 beams
// instead of 20 KB), and resident scans per SM are bounded by registers (16 warps), not shared memory (10).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_math.cuh"

namespace ekf {

constexpr int kLmThreads = 32;       // one warp per scan
constexpr int kLmMaxPts = 39;        // clusters.at(i).size() < 40: landmarks.cpp:122
constexpr int kLmErrNoGap = 3;       // the reference throws at clusters.at(0): landmarks.cpp:95

struct LmParams {
  const float *ranges;     // [n_scans][n_ranges]
  const double2 *beam;     // [n_ranges] (cos, sin) of the beam angles of this call (lm_beam_table_kernel)
  int n_scans, n_ranges, max_out;
  double thr;              // distance_clustering_threshold (landmarks.cpp:195)
  double *out;             // [n_scans][max_out][3] (c_x, c_y, r), cluster order
  int *count;              // [n_scans] detections found (may exceed max_out; only max_out are stored)
  int *status;             // [n_scans] 0 or kLmErrNoGap
};

__host__ __device__ inline size_t lm_smem_bytes(int n_ranges) {
  return (size_t)n_ranges * (2 * sizeof(double) + sizeof(unsigned short)) + 16;
}

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
__device__ inline void fit_circle(const double *px, const double *py, const ClusterView &cv, int n, double &cx,
                                  double &cy, double &R) {
  double mx = 0.0, my = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) { const int o = cv.at(i); mx += px[o]; my += py[o]; }
  mx /= n; my /= n;
  double s00 = 0.0, s01 = 0.0, s02 = 0.0, s03 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    const int o = cv.at(i);
    const double xi = px[o] - mx, yi = py[o] - my;
    const double z0 = xi * xi + yi * yi;
    s00 += z0 * z0; s01 += z0 * xi; s02 += z0 * yi; s03 += z0;
  }
  const double zm = s03 / n;                                   // mean of z (landmark_detection.cpp:72-77)
  const double r00 = sqrt(s00), i0 = r00 > 0.0 ? 1.0 / r00 : 0.0;
  const double r01 = s01 * i0, r02 = s02 * i0, r03 = s03 * i0;
  double s11 = 0.0, s12 = 0.0, s13 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    const int o = cv.at(i);
    const double xi = px[o] - mx, yi = py[o] - my;
    const double q0 = (xi * xi + yi * yi) * i0;
    const double a1 = xi - r01 * q0, a2 = yi - r02 * q0, a3 = 1.0 - r03 * q0;
    s11 += a1 * a1; s12 += a1 * a2; s13 += a1 * a3;
  }
  const double r11 = sqrt(s11), i1 = r11 > 0.0 ? 1.0 / r11 : 0.0;
  const double r12 = s12 * i1, r13 = s13 * i1;
  double s22 = 0.0, s23 = 0.0;
#pragma unroll 1
  for (int i = 0; i < n; ++i) {
    const int o = cv.at(i);
    const double xi = px[o] - mx, yi = py[o] - my;
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
    const int o = cv.at(i);
    const double xi = px[o] - mx, yi = py[o] - my;
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
__device__ inline bool check_circle(const double *px, const double *py, const ClusterView &cv, int n) {
  const int o0 = cv.at(0), oe = cv.at(n - 1);
  const double sx = px[o0], sy = py[o0], ex = px[oe], ey = py[oe];
  const double c = sqrt(sq_sum(sx - ex, sy - ey));
  double first = 0.0, sd = 0.0, sd2 = 0.0;
#pragma unroll 1
  for (int j = 1; j < n - 1; ++j) {
    const int o = cv.at(j);
    const double a = sqrt(sq_sum(sx - px[o], sy - py[o])), b = sqrt(sq_sum(ex - px[o], ey - py[o]));
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

// (16 resident scans per SM: 128 registers; sized for 20 or 24 it spills and is no faster - measured)
__global__ void __launch_bounds__(kLmThreads, 16) ekf_landmarks_kernel(const LmParams p) {
  extern __shared__ __align__(16) unsigned char lm_smem[];
  const int n = p.n_ranges;
  double *px = reinterpret_cast<double *>(lm_smem);
  double *py = px + n;
  unsigned short *gpos = reinterpret_cast<unsigned short *>(py + n);   // gap positions, in order (n <= 4096)
  const int scan = blockIdx.x, lane = threadIdx.x;
  if (scan >= p.n_scans) return;
  const float *rg = p.ranges + (size_t)scan * n;

  // landmarks.cpp:66-69
  for (int i = lane; i < n; i += kLmThreads) {
    const double r = (double)rg[i];
    const double2 cs = p.beam[i];
    px[i] = r * cs.x - 0.032;
    py[i] = r * cs.y;
  }
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
    if (lane == 0) { p.count[scan] = 0; p.status[scan] = kLmErrNoGap; }
    return;
  }
  const bool wrap = sqrt(sq_sum(px[0] - px[n - 1], py[0] - py[n - 1])) <= p.thr;   // landmarks.cpp:93-103
  const int n_clusters = wrap ? G : G + 1;

  // one lane per cluster: size filter, circle test, fit, final filter (landmarks.cpp:117-149); detections leave in
  // cluster order (marker id = running index, :137-148)
  int written = 0;
  for (int base = 0; base < n_clusters; base += kLmThreads) {
    const int k = base + lane;
    bool keep = false;
    double cx = 0.0, cy = 0.0, R = 0.0;
    if (k < n_clusters) {
      ClusterView cv;
      cv.start = (k == 0) ? 0 : (int)gpos[k - 1] + 1;
      const int end = (k < G) ? (int)gpos[k] : n;                // exclusive
      cv.len = end - cv.start;
      cv.tail0 = (int)gpos[G - 1] + 1;
      const int tail = (k == 0 && wrap) ? n - cv.tail0 : 0;
      const int total = cv.len + tail;
      if (total > 3 && total <= kLmMaxPts) {
        if (check_circle(px, py, cv, total)) {
          fit_circle(px, py, cv, total, cx, cy, R);
          keep = R < 0.2 && sqrt(sq_sum(cx, cy)) < 2.0;
        }
      }
    }
    const unsigned km = __ballot_sync(0xffffffffu, keep);
    const int orank = written + __popc(km & ((1u << lane) - 1u));
    if (keep && orank < p.max_out) {
      double *d = p.out + ((size_t)scan * p.max_out + orank) * 3;
      d[0] = cx; d[1] = cy; d[2] = R;
    }
    written += __popc(km);
  }
  if (lane == 0) { p.count[scan] = written; p.status[scan] = 0; }
}

}  // namespace ekf
