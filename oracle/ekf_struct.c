/* Oracle B - structured O(n^2)-per-measurement CPU restatement of the reference EKF.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/ekf_literal.py for the rule): used by tests/ as the
 * checker for large maps and by bench.py as the reported CPU baseline ("port").  Never on
 * the product path.
 *
 * PARITY UNPINNED with respect to reference outputs (the reference has no tests for this
 * path and cannot be built here); cross-validated against oracle A (ekf_literal.py, the
 * dense line-by-line form) in tests/test_oracle_cross.py.
 *
 * Same mathematics as /root/reference/nuslam/src/slam.cpp, same full (non-symmetric)
 * n x n column-major Sigma as Armadillo holds, but every product with the 2 x n Jacobian
 * uses only its five non-zero columns {0,1,2,3+2j,4+2j}:
 *     (I - K H) Sigma  ==  Sigma - K (H Sigma)          slam.cpp:264-265 / :485-486
 *     K = Sigma H^T inv(H Sigma H^T + R)                slam.cpp:252 / :476
 *     psi_k = H_k Sigma H_k^T + R  (5x5 block only)     slam.cpp:395
 *     A Sigma A^T + Qbar  (rows/cols 1,2 only)          slam.cpp:194-198 / :331-335
 * pthreads over columns / candidates (no OpenMP runtime in this image).  Compile with -ffp-contract=off so the arithmetic is
 * plain IEEE multiply/add in the written order.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#define EKFO_PI 3.14159265358979323846 /* turtlelib geometry2d.hpp:14 */

/* ---- minimal parallel-for (pthreads) -------------------------------------------------- */
typedef void (*ekfo_body)(long lo, long hi, void *ctx);
typedef struct { ekfo_body fn; void *ctx; long lo, hi; } ekfo_task;
static int g_threads = 0;
int ekfo_num_threads(void) {
  if (g_threads == 0) {
    const char *e = getenv("EKFO_THREADS");
    long t = e ? atol(e) : sysconf(_SC_NPROCESSORS_ONLN);
    if (t < 1) t = 1;
    if (t > 256) t = 256;
    g_threads = (int)t;
  }
  return g_threads;
}
void ekfo_set_threads(int t) { g_threads = t < 1 ? 1 : (t > 256 ? 256 : t); }
static void *ekfo_run(void *arg) { ekfo_task *t = (ekfo_task *)arg; t->fn(t->lo, t->hi, t->ctx); return NULL; }
static void par_for(long n, long min_per_thread, ekfo_body fn, void *ctx) {
  int nt = ekfo_num_threads();
  if (n / (min_per_thread > 0 ? min_per_thread : 1) < nt) nt = (int)(n / (min_per_thread > 0 ? min_per_thread : 1));
  if (nt <= 1) { fn(0, n, ctx); return; }
  pthread_t th[256];
  ekfo_task tk[256];
  for (int i = 0; i < nt; ++i) {
    tk[i].fn = fn; tk[i].ctx = ctx; tk[i].lo = n * i / nt; tk[i].hi = n * (i + 1) / nt;
    if (i > 0) pthread_create(&th[i], NULL, ekfo_run, &tk[i]);
  }
  ekfo_run(&tk[0]);
  for (int i = 1; i < nt; ++i) pthread_join(th[i], NULL);
}

typedef struct ekfo {
  int N;            /* landmark capacity (slam.cpp:665) */
  int n;            /* 3 + 2N */
  double q, r, thr; /* slam.cpp:666-667,671 */
  double *state;    /* n            (slam.cpp:674) */
  double *sigma;    /* n*n col-major (slam.cpp:673) */
  unsigned cnt;     /* counter_obstacles (slam.cpp:670) */
  double prev_x, prev_y; /* filter_previous_configuration translation (slam.cpp:660) */
  double *K, *HS, *dist; /* scratch: n*2, 2*n, N+1 */
} ekfo_t;

/* turtlelib/src/geometry2d.cpp:5-14 */
double ekfo_normalize_angle(double rad) {
  double diff = fmod(rad + EKFO_PI, 2 * EKFO_PI);
  if (diff <= 0.0) return diff + EKFO_PI;
  return diff - EKFO_PI;
}

ekfo_t *ekfo_create(int N, double q, double r, double thr, double prior) {
  ekfo_t *f = (ekfo_t *)calloc(1, sizeof(ekfo_t));
  if (!f) return NULL;
  f->N = N;
  f->n = 3 + 2 * N;
  f->q = q; f->r = r; f->thr = thr;
  size_t n = (size_t)f->n;
  f->state = (double *)calloc(n, sizeof(double));
  f->sigma = (double *)calloc(n * n, sizeof(double));
  f->K = (double *)calloc(n * 2, sizeof(double));
  f->HS = (double *)calloc(n * 2, sizeof(double));
  f->dist = (double *)calloc((size_t)N + 1, sizeof(double));
  if (!f->state || !f->sigma || !f->K || !f->HS || !f->dist) return NULL;
  for (size_t i = 3; i < n; ++i) f->sigma[i * n + i] = prior; /* slam.cpp:128-132 */
  return f;
}

void ekfo_destroy(ekfo_t *f) {
  if (!f) return;
  free(f->state); free(f->sigma); free(f->K); free(f->HS); free(f->dist); free(f);
}

double *ekfo_state(ekfo_t *f) { return f->state; }
double *ekfo_sigma(ekfo_t *f) { return f->sigma; }
int ekfo_n(ekfo_t *f) { return f->n; }
unsigned ekfo_count(ekfo_t *f) { return f->cnt; }
void ekfo_set_count(ekfo_t *f, unsigned c) { f->cnt = c; }
void ekfo_set_prev(ekfo_t *f, double x, double y) { f->prev_x = x; f->prev_y = y; }
void ekfo_get_prev(ekfo_t *f, double *xy) { xy[0] = f->prev_x; xy[1] = f->prev_y; }

#define S_(i, j) sig[(size_t)(j) * n + (size_t)(i)]

/* slam.cpp:186-198 (== :323-335): pose overwrite from the composed odometry pose, then
 * Sigma <- A Sigma A^T + Qbar with A = I, A(1,0) = -dy, A(2,0) = dx. */
typedef struct { double *sig; size_t n; double dx, dy; } pred_ctx;
static void pred_rows(long lo, long hi, void *vc) { /* B = A * Sigma: rows 1,2 += multiples of row 0 */
  pred_ctx *c = (pred_ctx *)vc;
  double *sig = c->sig; const size_t n = c->n;
  for (long j = lo; j < hi; ++j) {
    const double s0 = S_(0, j);
    S_(1, j) = (-c->dy) * s0 + S_(1, j);
    S_(2, j) = c->dx * s0 + S_(2, j);
  }
}
static void pred_cols(long lo, long hi, void *vc) { /* C = B * A^T: cols 1,2 += multiples of col 0 */
  pred_ctx *c = (pred_ctx *)vc;
  double *sig = c->sig; const size_t n = c->n;
  for (long i = lo; i < hi; ++i) {
    const double b0 = S_(i, 0);
    S_(i, 1) = b0 * (-c->dy) + S_(i, 1);
    S_(i, 2) = b0 * c->dx + S_(i, 2);
  }
}
void ekfo_predict(ekfo_t *f, double theta, double x, double y) {
  const size_t n = (size_t)f->n;
  double *sig = f->sigma;
  f->state[0] = ekfo_normalize_angle(theta);
  f->state[1] = x;
  f->state[2] = y;
  pred_ctx c = {sig, n, x - f->prev_x, y - f->prev_y};
  par_for((long)n, 4096, pred_rows, &c);
  par_for((long)n, 4096, pred_cols, &c);
  S_(0, 0) += f->q; S_(1, 1) += f->q; S_(2, 2) += f->q; /* slam.cpp:135-136 */
}

/* The five non-zero Jacobian columns for landmark j and their values.
 * slam.cpp:234-249 / :377-392 / :458-473.  hs is 2x5 row-major over cols {0,1,2,a,b}. */
static void jacobian5(const double *state, int j, double hs[10], int c[5]) {
  const double delta_x = state[3 + 2 * j] - state[1];
  const double delta_y = state[3 + 2 * j + 1] - state[2];
  const double d = delta_x * delta_x + delta_y * delta_y;
  const double sd = sqrt(d);
  c[0] = 0; c[1] = 1; c[2] = 2; c[3] = 3 + 2 * j; c[4] = 4 + 2 * j;
  hs[0] = 0.0;  hs[1] = -delta_x / sd; hs[2] = -delta_y / sd; hs[3] = delta_x / sd;  hs[4] = delta_y / sd;
  hs[5] = -1.0; hs[6] = delta_y / d;   hs[7] = -delta_x / d;  hs[8] = -delta_y / d;  hs[9] = delta_x / d;
}

/* slam.cpp:219-227 / :364-370 / :444-451 */
static void expected_measurement(const double *state, int j, double zhat[2]) {
  zhat[0] = sqrt(pow(state[3 + 2 * j] - state[1], 2) + pow(state[3 + 2 * j + 1] - state[2], 2));
  zhat[1] = ekfo_normalize_angle(
      atan2(state[3 + 2 * j + 1] - state[2], state[3 + 2 * j] - state[1]) - state[0]);
}

/* psi = H Sigma H^T + R restricted to the 5x5 block; evaluated as (H Sigma) H^T like the
 * reference's left-to-right three-operand product. */
static void innovation_cov(const ekfo_t *f, const double hs[10], const int c[5], double psi[4]) {
  const size_t n = (size_t)f->n;
  const double *sig = f->sigma;
  double hsb[10]; /* (H Sigma)[:, c] */
  for (int r = 0; r < 2; ++r)
    for (int q = 0; q < 5; ++q) {
      double acc = 0.0;
      for (int p = 0; p < 5; ++p) acc += hs[r * 5 + p] * S_(c[p], c[q]);
      hsb[r * 5 + q] = acc;
    }
  for (int r = 0; r < 2; ++r)
    for (int s = 0; s < 2; ++s) {
      double acc = 0.0;
      for (int q = 0; q < 5; ++q) acc += hsb[r * 5 + q] * hs[s * 5 + q];
      psi[r * 2 + s] = acc + (r == s ? f->r : 0.0);
    }
}

static void inv2(const double a[4], double out[4]) {
  const double det = a[0] * a[3] - a[1] * a[2];
  out[0] = a[3] / det; out[1] = -a[1] / det; out[2] = -a[2] / det; out[3] = a[0] / det;
}

/* slam.cpp:219-267 / :443-488 for landmark j and measurement z = (range, bearing). */
typedef struct { ekfo_t *f; const double *hs; const int *c; const double *si; } upd_ctx;
static void upd_hs(long lo, long hi, void *vc) { /* HS = H * Sigma (rows c of Sigma) */
  upd_ctx *u = (upd_ctx *)vc;
  const size_t n = (size_t)u->f->n; const double *sig = u->f->sigma; double *HS = u->f->HS;
  for (long col = lo; col < hi; ++col) {
    double a0 = 0.0, a1 = 0.0;
    for (int p = 0; p < 5; ++p) {
      const double v = S_(u->c[p], col);
      a0 += u->hs[p] * v;
      a1 += u->hs[5 + p] * v;
    }
    HS[col] = a0;
    HS[n + col] = a1;
  }
}
static void upd_gain(long lo, long hi, void *vc) { /* K = (Sigma H^T) inv(psi) (columns c of Sigma) */
  upd_ctx *u = (upd_ctx *)vc;
  const size_t n = (size_t)u->f->n; const double *sig = u->f->sigma; double *K = u->f->K;
  for (long i = lo; i < hi; ++i) {
    double g0 = 0.0, g1 = 0.0;
    for (int p = 0; p < 5; ++p) {
      const double v = S_(i, u->c[p]);
      g0 += v * u->hs[p];
      g1 += v * u->hs[5 + p];
    }
    K[i] = g0 * u->si[0] + g1 * u->si[2];
    K[n + i] = g0 * u->si[1] + g1 * u->si[3];
  }
}
static void upd_sigma(long lo, long hi, void *vc) { /* Sigma <- Sigma - K * HS, column by column */
  upd_ctx *u = (upd_ctx *)vc;
  const size_t n = (size_t)u->f->n; double *sig = u->f->sigma;
  const double *K = u->f->K, *HS = u->f->HS;
  for (long col = lo; col < hi; ++col) {
    const double h0 = HS[col], h1 = HS[n + col];
    double *scol = sig + (size_t)col * n;
    for (size_t i = 0; i < n; ++i) scol[i] = scol[i] - (K[i] * h0 + K[n + i] * h1);
  }
}
static void measurement_update(ekfo_t *f, int j, const double z[2]) {
  const size_t n = (size_t)f->n;
  double hs[10], zhat[2], psi[4], si[4];
  int c[5];
  expected_measurement(f->state, j, zhat);
  jacobian5(f->state, j, hs, c);
  double *HS = f->HS; /* 2 x n, row r at HS + r*n */
  double *K = f->K;   /* n x 2, col s at K + s*n */
  upd_ctx u = {f, hs, c, si};
  par_for((long)n, 2048, upd_hs, &u);
  for (int r = 0; r < 2; ++r)
    for (int s = 0; s < 2; ++s) {
      double acc = 0.0;
      for (int q = 0; q < 5; ++q) acc += HS[(size_t)r * n + c[q]] * hs[s * 5 + q];
      psi[r * 2 + s] = acc + (r == s ? f->r : 0.0);
    }
  inv2(psi, si);
  par_for((long)n, 2048, upd_gain, &u);
  double zd[2] = {z[0] - zhat[0], ekfo_normalize_angle(z[1] - zhat[1])};
  for (size_t i = 0; i < n; ++i) f->state[i] = f->state[i] + (K[i] * zd[0] + K[n + i] * zd[1]);
  par_for((long)n, 64, upd_sigma, &u);
  f->state[0] = ekfo_normalize_angle(f->state[0]);
}

static void end_of_scan(ekfo_t *f) { /* slam.cpp:273-274,291 / :490-491,508 */
  f->prev_x = f->state[1];
  f->prev_y = f->state[2];
}

/* slam.cpp:201-271.  action may be NULL (all ADD); value 2 == Marker::DELETE is skipped.
 * Returns 0, or -1 for an id outside [0, N). */
int ekfo_update_known(ekfo_t *f, int m, const int *ids, const double *xy, const unsigned char *action) {
  for (int i = 0; i < m; ++i) {
    if (action && action[i] == 2) continue;
    const int id = ids[i];
    if (id < 0 || id >= f->N) return -1;
    const double mx = xy[2 * i], my = xy[2 * i + 1];
    const double z[2] = {sqrt(pow(mx, 2) + pow(my, 2)), atan2(my, mx)};
    double *s = f->state;
    if (s[3 + 2 * id] == 0.0 && s[3 + 2 * id + 1] == 0.0) { /* slam.cpp:213-216 */
      s[3 + 2 * id] = s[1] + z[0] * cos(z[1] + s[0]);
      s[3 + 2 * id + 1] = s[2] + z[0] * sin(z[1] + s[0]);
    }
    measurement_update(f, id, z);
  }
  end_of_scan(f);
  return 0;
}

typedef struct { ekfo_t *f; const double *z; } assoc_ctx;
static void assoc_body(long lo, long hi, void *vc) { /* slam.cpp:361-416, existing landmarks */
  assoc_ctx *a = (assoc_ctx *)vc;
  ekfo_t *f = a->f; const double *s = f->state; const double *z = a->z;
  for (long k = lo; k < hi; ++k) {
    double hs[10], zhat[2], psi[4], pi[4];
    int c[5];
    expected_measurement(s, (int)k, zhat);
    jacobian5(s, (int)k, hs, c);
    innovation_cov(f, hs, c, psi);
    inv2(psi, pi);
    const double zd0 = z[0] - zhat[0], zd1 = ekfo_normalize_angle(z[1] - zhat[1]);
    /* (z^T inv) z, left to right: slam.cpp:401 */
    const double t0 = zd0 * pi[0] + zd1 * pi[2], t1 = zd0 * pi[1] + zd1 * pi[3];
    f->dist[k] = t0 * zd0 + t1 * zd1;
  }
}

/* slam.cpp:338-489.  assoc[i] receives the landmark index measurement i was fused with.
 * Returns 0, or -2 when the tentative landmark would not fit (the reference indexes out of
 * bounds there: slam.cpp:351 with :665) - unless the map is exactly full and the measurement
 * matches an existing landmark, which this restatement allows so that a full map stays
 * usable (the tentative slot is then simply not materialised; its distance is the
 * threshold either way, slam.cpp:406-408). */
int ekfo_update_unknown_ex(ekfo_t *f, int m, const double *xy, int *assoc, double *best_d, double *second_d);
int ekfo_update_unknown(ekfo_t *f, int m, const double *xy, int *assoc) {
  return ekfo_update_unknown_ex(f, m, xy, assoc, NULL, NULL);
}

/* The same, also reporting the decision margins: best_d[i] / second_d[i] = the smallest and
 * second-smallest Mahalanobis distance of measurement i over the EXISTING landmarks (the tentative entry excluded;
 * +inf when there are fewer).  A decision can only differ between two implementations when best_d is within their
 * arithmetic difference of the threshold, or of second_d. */
int ekfo_update_unknown_ex(ekfo_t *f, int m, const double *xy, int *assoc, double *best_d, double *second_d) {
  for (int i = 0; i < m; ++i) {
    const double mx = xy[2 * i], my = xy[2 * i + 1];
    const double z[2] = {sqrt(pow(mx, 2) + pow(my, 2)), atan2(my, mx)};
    double *s = f->state;
    const int cnt = (int)f->cnt;
    const int full = cnt >= f->N;
    if (!full) { /* slam.cpp:351-356 */
      s[3 + 2 * cnt] = s[1] + z[0] * cos(z[1] + s[0]);
      s[3 + 2 * cnt + 1] = s[2] + z[0] * sin(z[1] + s[0]);
    }
    double *dist = f->dist;
    assoc_ctx ac = {f, z};
    par_for((long)cnt, 64, assoc_body, &ac);
    dist[cnt] = f->thr; /* the tentative landmark: slam.cpp:406-408 */
    if (best_d || second_d) {
      double b1 = INFINITY, b2 = INFINITY;
      for (int k = 0; k < cnt; ++k) {
        if (dist[k] < b1) { b2 = b1; b1 = dist[k]; }
        else if (dist[k] < b2) b2 = dist[k];
      }
      if (best_d) best_d[i] = b1;
      if (second_d) second_d[i] = b2;
    }
    /* arma::index_min: strict '<' from +inf, first minimum wins, NaN never selected */
    double best = INFINITY;
    int min_index = 0;
    for (int k = 0; k <= cnt; ++k)
      if (dist[k] < best) { best = dist[k]; min_index = k; }
    if (min_index == cnt) { /* slam.cpp:421: keep the tentative landmark */
      if (full) return -2;
      f->cnt = (unsigned)cnt + 1;
    } else if (!full) { /* slam.cpp:433-440: retract it */
      s[3 + 2 * cnt] = 0.0;
      s[3 + 2 * cnt + 1] = 0.0;
    }
    assoc[i] = min_index;
    measurement_update(f, min_index, z);
  }
  end_of_scan(f);
  return 0;
}

