/* ekf_slam_b200.h - C ABI of the B200-native EKF-SLAM core.
 *
 * The reference (maxipalay/ekf-slam) has no plugin / FFI boundary for its filter: the whole
 * EKF is written inline in two ROS callbacks of class Slam.  This ABI is the seam
 * SURVEY.md section 8(b) makes explicit - it replaces
 *     nuslam/src/slam.cpp:127-139,659-676   filter members and their initialisation
 *     nuslam/src/slam.cpp:186-198 / 323-335 pose injection + prediction
 *     nuslam/src/slam.cpp:201-271           known-association measurement loop
 *     nuslam/src/slam.cpp:338-489           unknown-association (Mahalanobis) loop
 * and leaves the node shell (tf, markers, odometry; slam.cpp:273-316, 490-530, 575-634) as
 * the caller.  INTEGRATION.md shows the edit a maintainer would make in slam.cpp.
 *
 * Conventions
 *   - every function returns an ekf_status (0 = OK); nothing throws across the boundary;
 *   - plain pointers and sizes only; "host" pointers are caller-owned host memory,
 *     "device" pointers (suffix _dev) are caller-owned device memory on the filter's GPU;
 *   - state vectors use the reference's layout [theta, x, y, m1x, m1y, ...] (slam.cpp:186-188,
 *     214-215), n = 3 + 2N; Sigma blocks are row-major n x n blocks of the symmetric matrix;
 *   - one caller thread per handle (the reference's executor is single-threaded,
 *     slam.cpp:680-686); host-pointer calls are synchronous at return, *_dev calls are
 *     stream-ordered and asynchronous (errors latch in the handle, see ekf_last_status);
 *   - there is NO CPU fallback: every entry point needs a CUDA device and fails with
 *     EKF_ERR_CUDA / EKF_ERR_NO_DEVICE otherwise.
 */
#ifndef EKF_SLAM_B200_H
#define EKF_SLAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum ekf_status {
  EKF_OK = 0,
  EKF_ERR_INVALID = 1,    /* bad argument (null pointer, negative size, unknown option) */
  EKF_ERR_CAPACITY = 2,   /* a new landmark was needed but counter_obstacles == N
                             (the reference indexes out of bounds here: slam.cpp:351 with :665) */
  EKF_ERR_EMPTY_SCAN = 3, /* m == 0 (the reference throws at msg.markers.at(0): slam.cpp:281,498) */
  EKF_ERR_BAD_ID = 4,     /* known-association id outside [0, N) (Armadillo bounds error) */
  EKF_ERR_CUDA = 5,       /* a CUDA call failed; see ekf_last_cuda_error */
  EKF_ERR_NO_DEVICE = 6,  /* no usable sm_100 device */
  EKF_ERR_NOMEM = 7,
  EKF_ERR_IO = 8          /* bag file unreadable / not MCAP / unsupported encoding; see ekf_bag_last_error */
} ekf_status;

/* Filter constants.  Defaults (ekf_default_config) are the reference's hard-coded members:
 * n_landmarks{50}, q_noise{1e-2}, r_noise{1e-2}, mah_threshold = 2.0 (slam.cpp:665-671) and the
 * 10e6 landmark prior (slam.cpp:130). */
typedef struct ekf_config {
  int32_t n_landmarks;   /* N: landmark capacity; n = 3 + 2N */
  double q_noise;        /* Qbar = diag(q,q,q,0,...)        slam.cpp:135-136 */
  double r_noise;        /* R = r * I2                      slam.cpp:139 */
  double mah_threshold;  /* new-landmark threshold          slam.cpp:671,406-408 */
  double prior;          /* initial landmark variance       slam.cpp:130 */
  int32_t device;        /* CUDA device ordinal, -1 = current device */
  int32_t max_scan;      /* hint: largest m per call (0 = 64); larger scans still work */
  /* row sharding of one large map over several GPUs (one process per GPU): this handle owns
   * the covariance rows of shard `shard_rank` out of `shard_count` (equal-area split).
   * shard_count <= 1 means a whole, unsharded filter. */
  int32_t shard_rank;
  int32_t shard_count;
} ekf_config;

typedef struct ekf_filter ekf_filter;       /* one EKF (slam.cpp:659-676) */
typedef struct ekf_batch ekf_batch;         /* B independent small EKFs, one CTA each */
typedef struct ekf_landmarks ekf_landmarks; /* landmark extraction from LaserScans (the EKF's producer) */
typedef struct ekf_nusim ekf_nusim;         /* nusim's sensor models as device-side generators */

void ekf_default_config(ekf_config *cfg);
const char *ekf_status_string(int status);
const char *ekf_version(void);
/* Device self-test of the shared-reciprocal fp64 division the kernels use for the reference's quotients
 * (slam.cpp:234-249 Jacobian entries, the 2x2 inverse behind :401 / :476): `count` hashed operand pairs (all
 * exponents, specials, and the filter's own magnitudes) are divided both ways on device `device` (-1: current).
 * *mismatches = pairs inside the fast path's guard whose quotient differs in any bit from IEEE `a / b` (must be 0);
 * *outside_guard = pairs the guard sends to the plain division. */
int ekf_selftest_division(int32_t device, uint64_t seed, int32_t count, uint64_t *mismatches, uint64_t *outside_guard);
/* The same for the branch-free atan2 behind the expected bearing (slam.cpp:224-227 / :367-370): against the CUDA math
 * library's atan2 on hashed (y, x) pairs - all exponents, specials, and the filter's own range.  *mismatches must be 0;
 * *outside_guard = pairs (non-finite, both zero, extreme ratios) that go to the library call. */
int ekf_selftest_atan2(int32_t device, uint64_t seed, int32_t count, uint64_t *mismatches, uint64_t *outside_guard);

/* ---- lifetime: replaces slam.cpp:127-139 ------------------------------------------------ */
int ekf_create(const ekf_config *cfg, ekf_filter **out);
int ekf_destroy(ekf_filter *f);
/* run the handle's kernels on a caller stream (a cudaStream_t cast to void*; NULL = the
 * handle's own stream).  Lets a harness time the work with events on that stream. */
int ekf_set_stream(ekf_filter *f, void *cuda_stream);
int ekf_synchronize(ekf_filter *f);
/* latched status of asynchronous work since the last call of this function (syncs) */
int ekf_last_status(ekf_filter *f);
const char *ekf_last_cuda_error(ekf_filter *f);

/* ---- the scan update -------------------------------------------------------------------- */
/* slam.cpp:184-198: `pose` is the already composed t_map_odom * t_odom_robot as (theta, x, y)
 * (raw theta; it is wrapped inside like slam.cpp:186).  Overwrites state[0..2], then
 * Sigma <- A Sigma A^T + Qbar with dx, dy taken against the pose the filter held at the end of
 * the previous update (filter_previous_configuration, slam.cpp:190-191,291). */
int ekf_predict(ekf_filter *f, double theta, double x, double y);

/* slam.cpp:201-271: m sightings with known landmark ids.  xy = body-frame (x, y) pairs
 * (marker.pose.position), action may be NULL; entries equal to 2 (Marker::DELETE) are skipped
 * (slam.cpp:205).  A landmark whose two state slots are exactly 0.0 is initialised from the
 * sighting first (slam.cpp:213-216). */
int ekf_update_known(ekf_filter *f, int32_t m, const int32_t *ids, const double *xy,
                     const uint8_t *action);

/* slam.cpp:338-489: m sightings, Mahalanobis association, processed sequentially (measurement
 * i+1 sees the state and Sigma left by measurement i).  out_assoc[i] = landmark index the
 * sighting was fused with (the value of min_index at slam.cpp:443; a new landmark gets the old
 * counter_obstacles).  out_count = counter_obstacles afterwards.  Either may be NULL. */
int ekf_update_unknown(ekf_filter *f, int32_t m, const double *xy, int32_t *out_assoc,
                       int32_t *out_count);

/* predict + update in one call (what each sensor callback does), inputs and outputs in DEVICE
 * memory, asynchronous: pose_dev[3], xy_dev[2m], ids_dev[m] (known) / NULL (unknown),
 * action_dev[m] or NULL, assoc_dev[m] or NULL.  Used by the throughput harness; errors latch. */
int ekf_scan_dev(ekf_filter *f, const double *pose_dev, int32_t m, const double *xy_dev,
                 const int32_t *ids_dev, const uint8_t *action_dev, int32_t *assoc_dev);
/* the same with HOST buffers: copies in, runs, copies assoc / count / pose / status back.  Returns as soon
 * as those are final, i.e. after the sequential phase of the scan; the covariance pass of the scan may
 * still be running on the handle's stream.  Nothing observable depends on that: every accessor and every
 * later call is ordered behind it on the same stream (ekf_synchronize waits for it explicitly).  The
 * same holds for ekf_predict / ekf_update_known / ekf_update_unknown. */
int ekf_scan(ekf_filter *f, const double pose[3], int32_t m, const double *xy, const int32_t *ids,
             const uint8_t *action, int32_t *out_assoc, int32_t *out_count, double out_pose[3]);

/* ---- accessors (slam.cpp:273-274, 304-306, 518-520) ------------------------------------- */
int ekf_state_dim(ekf_filter *f, int32_t *n);                 /* n = 3 + 2N */
int ekf_get_pose(ekf_filter *f, double out[3]);               /* theta, x, y */
int ekf_get_state(ekf_filter *f, double *out, int32_t n);     /* first n entries of the state */
int ekf_get_count(ekf_filter *f, int32_t *count);             /* counter_obstacles */
int ekf_get_landmarks(ekf_filter *f, double *out_xy, int32_t max_landmarks, int32_t *count);
/* rows [i0,i1) x cols [j0,j1) of the symmetric Sigma, row-major into out (leading dim j1-j0) */
int ekf_get_sigma_block(ekf_filter *f, int32_t i0, int32_t i1, int32_t j0, int32_t j1, double *out);

/* ---- fixtures: checkpoint / restore, test set-up ------------------------------------------ */
int ekf_set_state(ekf_filter *f, const double *in, int32_t n);
int ekf_set_count(ekf_filter *f, int32_t count);
int ekf_set_previous_xy(ekf_filter *f, double x, double y);   /* filter_previous_configuration */
int ekf_get_previous_xy(ekf_filter *f, double out[2]);
/* writes the entries with row >= col of the given block into the stored triangle */
int ekf_set_sigma_block(ekf_filter *f, int32_t i0, int32_t i1, int32_t j0, int32_t j1, const double *in);
/* Association sweep alone (slam.cpp:358-404 without the tentative entry, the decision or the update):
 * out_d[k] = Mahalanobis distance of the body-frame sighting (x, y) to mapped landmark k, k < min(count,
 * cap), against the filter as it stands; *count = counter_obstacles.  Read-only.  This is how tests and
 * bench.py log the decision margins (min |d - threshold|, best-vs-second gap) of a scan. */
int ekf_mahalanobis(ekf_filter *f, double x, double y, double *out_d, int32_t cap, int32_t *count);
/* The same sweep for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * ekf_mahalanobis stages the distances in a per-handle buffer that grows with cudaMalloc; this variant stages them in
 * a file-scope static __device__ array of EKF_NOALLOC_MAX_LANDMARKS doubles (512 KiB, held by every GPU that loads the
 * library until the process exits) and calls no cudaMalloc / cudaFree / allocating library routine.  Same kernel,
 * bit-identical distances.  cap > EKF_NOALLOC_MAX_LANDMARKS is rejected with EKF_ERR_INVALID (ekf_last_cuda_error
 * says why).  Every call on a GPU shares the one array, which keeps the last call's contents (nothing in it has to
 * start at zero: a call reads back only what its own kernel wrote); calls that OVERLAP on one GPU - two handles on
 * different streams, two host threads - race on it and must be serialised by the caller. */
#define EKF_NOALLOC_MAX_LANDMARKS 65536
int ekf_mahalanobis_noalloc(ekf_filter *f, double x, double y, double *out_d, int32_t cap, int32_t *count);
/* Procedural large-map fixture: `count` landmarks on a jittered
 * unit grid of `grid_w` columns, state consistent with it, Sigma = block-diagonal SPD + a
 * rank-8 hash-generated cross term.  Deterministic in `seed`; generated on the device. */
int ekf_fill_procedural(ekf_filter *f, uint64_t seed, int32_t count, int32_t grid_w);

/* ---- measured facts for the harness -------------------------------------------------------- */
/* bytes the covariance pass streams per scan (stored triangle, read + write) and the stored
 * element count; used for the roofline arithmetic in bench.py / DESIGN.md */
int ekf_sigma_storage(ekf_filter *f, int64_t *stored_doubles, int64_t *n_pad);
/* number of kernels this handle has launched since creation */
int ekf_launch_count(ekf_filter *f, int64_t *launches);
/* Overlapped scans (single GPU, unsharded handle).  on = 1: from the next ekf_scan / ekf_scan_dev the covariance pass
 * of a scan (slam.cpp:264-265 / :485-486, deferred, one pass per scan) runs OUT OF PLACE on a second stream while
 * the prediction, association and state updates of the following scan (slam.cpp:323-482) already run: the two use
 * disjoint parts of an SM.  Costs a second covariance buffer (returns EKF_ERR_NOMEM and stays off when it does not
 * fit).  Results are the same filter (the columns the next scan needs are formed as "input of the pending pass
 * minus its update", which equals what the pass stores up to rounding).  Every accessor, fixture, classic-path call
 * (ekf_predict / ekf_update_*, scans with an `action` array) and ekf_set_overlap(f, 0) first drains the pipeline.
 * ekf_synchronize waits for the pending pass as well. */
int ekf_set_overlap(ekf_filter *f, int32_t on);
/* Per-scan device timing with CUDA events on the handle's stream.  ekf_enable_timing(f, r)
 * keeps the last r scans (r = 0 switches it off); ekf_read_timing synchronises and returns, for
 * the most recent scans (oldest first), the milliseconds of the sequential phase and of the
 * covariance pass.  A scan split into several passes (m > 24) counts once per pass. */
int ekf_enable_timing(ekf_filter *f, int32_t ring_scans);
int ekf_read_timing(ekf_filter *f, float *seq_ms, float *downdate_ms, int32_t max_scans, int32_t *n_scans);
/* the same ring for the begin phase of the row-sharded path (ekf_shard_scan_begin): milliseconds of the speculative
 * association kernel and of the peer-memory push kernel (0 when the round used the caller's all-reduce) */
int ekf_read_timing_shard(ekf_filter *f, float *assoc_ms, float *push_ms, int32_t max_scans, int32_t *n_scans);

/* ---- one large map row-sharded over several GPUs (BASELINE configs 4-5 at 2/4/8 GPUs) ------------
 * One process (or one handle) per GPU, created with cfg.shard_rank / cfg.shard_count: the handle
 * stores the rows [row_begin, row_end) of Sigma's lower triangle (equal-area split) and a REPLICA of
 * everything O(n) (state, pose panel, diagonal blocks, counters).  Every call below must be made on
 * every rank with identical arguments.  The library does not talk to the other ranks: each phase
 * hands back a device buffer that the caller SUM-all-reduces over the ranks (NCCL `ncclAllReduce`,
 * ncclDouble / ncclSum, on the handle's stream) before the matching finish call.
 *   ekf_shard_sync_*   rebuilds the replicated side state from the distributed Sigma; needed once
 *                      after create / ekf_fill_procedural / ekf_set_sigma_block / ekf_set_state.
 *   ekf_shard_scan_*   one round of a scan update (slam.cpp:321-508): begin = prediction (when
 *                      pose_dev != NULL; pass NULL on continuation rounds), speculative association,
 *                      this rank's part of the needed Sigma columns -> exchange buffer; finish = the
 *                      sequential measurement loop (replicated), the covariance pass over the owned
 *                      rows.  *done_dev (device int) receives the number of sightings of the round
 *                      that were applied: min(m, 24), or fewer when an association needed a landmark
 *                      the speculation had not fetched - the caller then runs another round with the
 *                      remaining sightings (xy_dev + 2 * done).  last_round = this round reaches the
 *                      end of the scan (latches filter_previous_configuration, slam.cpp:508, when all
 *                      of its sightings get applied).
 * The ordinary ekf_scan / ekf_update_* entry points return EKF_ERR_INVALID on a sharded handle;
 * ekf_get_sigma_block returns 0 for entries stored by another rank (sum over ranks = Sigma). */
int ekf_shard_info(ekf_filter *f, int32_t *row_begin, int32_t *row_end, int64_t *stored_doubles);
/* Optional: the per-scan exchange over peer memory instead of the caller's all-reduce.  Every rank exports
 * its exchange region (a CUDA IPC handle for other processes, the plain device pointer for handles of the
 * same process) and attaches every other rank's (exactly one of ipc_handle_64 / same_process_base per
 * call).  Once all shard_count - 1 peers are attached, ekf_shard_scan_begin launches the delivery itself, on the
 * handle's own side stream - every rank stores the Sigma columns it owns straight into every rank's buffer over
 * NVLink (through a relay rank when there are more than two) - and returns *xchg_doubles = 0: there is nothing left
 * to reduce.  The sequential kernel of ekf_shard_scan_finish runs concurrently and polls the elements it needs
 * (every buffer is pre-filled with a sentinel).  A peer that never delivers latches EKF_ERR_CUDA after 10 s instead
 * of hanging; *done_dev = -1 reports that the sequential kernel's own (bounded) grid barrier gave up.
 * ekf_shard_sync_* still uses the caller's all-reduce (rare). */
int ekf_shard_peer_export(ekf_filter *f, void *ipc_handle_64, void **local_base, int64_t *region_bytes);
int ekf_shard_peer_attach(ekf_filter *f, int32_t peer_rank, const void *ipc_handle_64, void *same_process_base);
int ekf_shard_sync_begin(ekf_filter *f, double **xchg_dev, int64_t *xchg_doubles);
int ekf_shard_sync_finish(ekf_filter *f);
int ekf_shard_scan_begin(ekf_filter *f, const double *pose_dev, int32_t m, const double *xy_dev,
                         const int32_t *ids_dev, int32_t last_round, double **xchg_dev, int64_t *xchg_doubles);
int ekf_shard_scan_finish(ekf_filter *f, int32_t *assoc_dev, int32_t *done_dev, double *pose_out_dev,
                          int32_t *count_dev);
/* the same; in addition, right after the sequential phase (when association, done, pose and count are
 * final) and BEFORE the covariance pass it enqueues a device-to-host copy of results_bytes bytes from
 * results_dev to results_host (pinned; 0 bytes = none) and records results_event (a cudaEvent_t, or NULL)
 * on the handle's stream: a host that waits on the event instead of the stream prepares the next scan
 * while the pass runs. */
int ekf_shard_scan_finish_ex(ekf_filter *f, int32_t *assoc_dev, int32_t *done_dev, double *pose_out_dev,
                             int32_t *count_dev, const void *results_dev, void *results_host, int64_t results_bytes,
                             void *results_event);

/* ---- batched mode: B independent small filters, one CTA each, Sigma in shared memory ---------
 * (BASELINE config 3).  Each filter runs the whole sensor callback of the reference
 * (slam.cpp:184-291 known / :321-508 unknown association) per step.  Buffers are struct-of-arrays
 * over the filters: poses[3B] = (theta, x, y) composed odometry pose per filter, xy[B*m*2] body-frame
 * sightings, ids[B*m] landmark ids (known association; a negative id = skipped sighting, the
 * Marker::DELETE rule of slam.cpp:205) or NULL (Mahalanobis association), m_count[B] = sightings
 * actually present per filter (ragged scans, <= m) or NULL.  Outputs may be NULL.
 * cfg->n_landmarks is the per-filter capacity N (3 + 2N <= 131). */
int ekf_batch_create(const ekf_config *cfg, int32_t n_filters, ekf_batch **out);
int ekf_batch_destroy(ekf_batch *b);
int ekf_batch_set_stream(ekf_batch *b, void *cuda_stream);
int ekf_batch_synchronize(ekf_batch *b);
/* host buffers: copies in, one kernel, copies out, synchronous; returns the worst per-filter status */
int ekf_batch_step(ekf_batch *b, const double *poses, int32_t m, const double *xy, const int32_t *ids,
                   const int32_t *m_count, int32_t *out_assoc, int32_t *out_count, double *out_pose);
/* device buffers, asynchronous on the handle's stream; errors latch (ekf_batch_last_status) */
int ekf_batch_step_dev(ekf_batch *b, const double *poses_dev, int32_t m, const double *xy_dev,
                       const int32_t *ids_dev, const int32_t *m_count_dev, int32_t *assoc_dev,
                       int32_t *count_dev, double *pose_dev);
int ekf_batch_last_status(ekf_batch *b);
int ekf_batch_launch_count(ekf_batch *b, int64_t *launches);
/* accessors / fixtures for filter `index` (state: n doubles; sigma: n x n row-major, symmetric) */
int ekf_batch_get_state(ekf_batch *b, int32_t index, double *out, int32_t n);
int ekf_batch_get_sigma(ekf_batch *b, int32_t index, double *out);
int ekf_batch_get_count(ekf_batch *b, int32_t index, int32_t *count);
int ekf_batch_set_state(ekf_batch *b, int32_t index, const double *in, int32_t n);
int ekf_batch_set_sigma(ekf_batch *b, int32_t index, const double *in);
int ekf_batch_set_count(ekf_batch *b, int32_t index, int32_t count);
int ekf_batch_set_previous_xy(ekf_batch *b, int32_t index, double x, double y);
/* bytes of HBM traffic one filter-step moves */
int ekf_batch_step_bytes(ekf_batch *b, int32_t m, int64_t *bytes);

/* ---- landmark extraction: the step before the filter -------------------------
 * A batch of LaserScans (float ranges, n_ranges beams each, as in sensor_msgs/LaserScan) ->
 * per scan the circle detections (c_x, c_y, r) in the lidar-corrected robot frame, in cluster order:
 * what nuslam's `landmarks` node publishes on green/detected_obstacles (nuslam/src/landmarks.cpp:58-156
 * with turtlelib checkCircle / fitCircle, turtlelib/src/landmark_detection.cpp:5-135).  n_scans <= max_scans of the handle.
 * out_xyr[n_scans][max_out][3]; out_count[s] = detections of scan s (only the first max_out are stored);
 * out_status[s] = 0, or EKF_ERR_EMPTY_SCAN when the scan has no range discontinuity at all (the
 * reference throws at clusters.at(0), landmarks.cpp:95).  cluster_threshold = 0.2 in the reference.
 * A call first writes the handle's table of beam directions for its (angle_min, angle_increment), then runs the
 * scans, both on the call's stream: calls on ONE handle are stream-ordered like every other handle here (one caller
 * thread, slam.cpp:680-686); two calls in flight on different streams need two handles. */
int ekf_landmarks_create(int32_t max_scans, int32_t n_ranges, int32_t max_out, int32_t device, ekf_landmarks **out);
int ekf_landmarks_destroy(ekf_landmarks *h);
int ekf_landmarks_detect(ekf_landmarks *h, int32_t n_scans, const float *ranges, float angle_min,
                         float angle_increment, double cluster_threshold, double *out_xyr, int32_t *out_count,
                         int32_t *out_status);
int ekf_landmarks_detect_dev(ekf_landmarks *h, int32_t n_scans, const float *ranges_dev, float angle_min,
                             float angle_increment, double cluster_threshold, double *out_xyr_dev,
                             int32_t *out_count_dev, int32_t *out_status_dev, void *cuda_stream);

/* ---- nusim's sensor models on the device ---------------------------------------
 * The world of nusim/config/basic_world.yaml: n_obstacles cylinders (x, y, r) and a rectangular arena;
 * n_robots poses (theta, x, y) in the world frame at once (the Monte-Carlo replicas).
 *   fake sensor (nusim/src/nusim.cpp:310-349): out_xy[robot][obstacle][2] = the obstacle in the robot's
 *     body frame + noise per axis; out_ids[robot][obstacle] = obstacle index when its TRUE range is
 *     <= max_range (Marker::ADD), else -(index + 1) (Marker::DELETE) - the layout and the negative-id
 *     rule ekf_batch_step takes for known association.
 *   lidar (nusim.cpp:559-710): out_ranges[robot][360] float, beam i at i degrees from the scanner at
 *     (-0.032, 0) in the body frame; cylinders first, the arena's wall lines only when no cylinder was
 *     hit; clamped to [0.11, 10]; + noise.  Feeds ekf_landmarks_detect with angle_min 0, increment 2pi/360.
 * Noise: `noise` (same shape as the output, doubles) holds the caller's draws where the reference adds
 * sensor_normal_dist(gen); with noise == NULL and noise_std > 0 a counter-based N(0, noise_std) generator
 * keyed by (seed, robot, index) is used (nusim.cpp:202 passes the "variance" parameter as the stddev). */
int ekf_nusim_create(int32_t n_obstacles, const double *obstacles_xyr, double arena_x_length, double arena_y_length,
                     int32_t device, ekf_nusim **out);
int ekf_nusim_destroy(ekf_nusim *h);
/* the two points (x3, y3, x4, y4) per wall of the lines the lidar intersects (nusim.cpp:649-653): 16 doubles */
int ekf_nusim_wall_lines(ekf_nusim *h, double *out16);
int ekf_nusim_fake_sensor(ekf_nusim *h, int32_t n_robots, const double *poses, double max_range, double noise_std,
                          uint64_t seed, const double *noise, double *out_xy, int32_t *out_ids);
int ekf_nusim_fake_sensor_dev(ekf_nusim *h, int32_t n_robots, const double *poses_dev, double max_range,
                              double noise_std, uint64_t seed, const double *noise_dev, double *out_xy_dev,
                              int32_t *out_ids_dev, void *cuda_stream);
int ekf_nusim_lidar(ekf_nusim *h, int32_t n_robots, const double *poses, double max_range, double noise_std,
                    uint64_t seed, const double *noise, float *out_ranges);
int ekf_nusim_lidar_dev(ekf_nusim *h, int32_t n_robots, const double *poses_dev, double max_range, double noise_std,
                        uint64_t seed, const double *noise_dev, float *out_ranges_dev, void *cuda_stream);

/* ---- N4: rosbag2 (MCAP container, CDR payload) reader for the node's two input topics -------------------------
 * Replaces, for offline replay, what rclcpp + rosbag2 deliver to Slam::jointStateCallback (slam.cpp:599-634,
 * sensor_msgs/msg/JointState on `joint_states`) and Slam::sensor_cb / fake_sensor_cb (slam.cpp:318, :180,
 * visualization_msgs/msg/MarkerArray on `green/detected_obstacles` / `nusim/fake_sensor`); topic names default to
 * the recorded ones of rosbags/rosbag2_2024_03_17-18_35_57/metadata.yaml:116-121,158-163.  Host-only (no GPU).
 * The whole file is parsed at open; events are the two topics merged in MCAP log-time order (ties: joint state
 * first), i.e. the order the node's single-threaded executor (slam.cpp:680-686) would have seen them. */
typedef struct ekf_bag ekf_bag;
/* joint_topic / detections_topic: NULL = "/blue/joint_states" / "/green/detected_obstacles" */
int ekf_bag_open(const char *path, const char *joint_topic, const char *detections_topic, ekf_bag **out);
int ekf_bag_close(ekf_bag *b);
/* message of the last failed ekf_bag_open on this thread ("" if none) */
const char *ekf_bag_last_error(void);
/* counts: events, joint states, detection arrays, all messages in the file (any topic), chunks */
int ekf_bag_counts(const ekf_bag *b, int64_t *events, int64_t *joint_states, int64_t *detection_arrays,
                   int64_t *messages_total, int64_t *chunks);
/* event e: kind 0 = joint state, 1 = detection array; index into that kind's list; MCAP log time (ns) */
int ekf_bag_event(const ekf_bag *b, int64_t e, int32_t *kind, int32_t *index, uint64_t *log_time_ns);
/* joint state i: header stamp (ns), number of joints, and the first min(cap, n) positions (the callback reads
 * position[0] = left and position[1] = right wheel, slam.cpp:604) */
int ekf_bag_joint_state(const ekf_bag *b, int32_t i, int64_t *stamp_ns, int32_t *n_joints, int32_t cap,
                        double *positions);
/* detection array i: number of markers and, for the first min(cap, n): pose.position (x, y) pairs, marker.id,
 * marker.action (ADD 0 / DELETE 2) - the fields slam.cpp:203-210 / :340-347 read */
int ekf_bag_detections(const ekf_bag *b, int32_t i, int32_t *n_markers, int32_t cap, double *xy, int32_t *ids,
                       int32_t *actions);

#ifdef __cplusplus
}
#endif
#endif
