// One large map row-sharded over R handles, driven from C++ straight on the C ABI (include/ekf_slam_b200.h,
// ekf_shard_*) - the executable form of INTEGRATION.md section 10 (north star: "C++ host code calls sm_100a CUDA
// through a thin C-ABI layer").  No Python, no torch.
//
//   shard_replay [--shards R] [--landmarks N] [--m M] [--scans S] [--seed X] [--allreduce]
//
// Single process, R handles; handle r lives on device r % (visible devices).  With R <= devices that is one shard
// per GPU (peer access is enabled between them, the per-scan exchange is the library's peer-memory push over
// NVLink); with more handles than devices several shards share a GPU (how a one-GPU test box exercises the
// multi-rank protocol).  The rare side-state sync (and the per-scan exchange under --allreduce) is a sum of the
// handles' buffers: ncclAllReduce when built with -DSHARD_REPLAY_NCCL and every handle has its own device, else
// peer copies + a summing kernel.  A multi-PROCESS host does the same with cudaIpc handles
// (ekf_shard_peer_export's ipc_handle_64 -> ekf_shard_peer_attach) and one ncclAllReduce per sync.
//
// The program replays S scans of the procedural map (ekf_fill_procedural), every scan sighting the m landmarks
// nearest to the robot, on the sharded handles AND on one unsharded filter (ekf_scan), and checks: association
// as expected and identical, state and sampled covariance blocks equal to 1e-12, replicas identical.  Prints ONE
// JSON line; exit code 0 only when everything agreed.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "ekf_slam_b200.h"
#ifdef SHARD_REPLAY_NCCL
#include <nccl.h>
#endif

#define CU(call)                                                                                     \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) {                                                                         \
      std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #call, cudaGetErrorString(e_));    \
      std::exit(3);                                                                                  \
    }                                                                                                \
  } while (0)
#define EKF(call)                                                                                    \
  do {                                                                                               \
    int s_ = (call);                                                                                 \
    if (s_ != EKF_OK) {                                                                              \
      std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #call, ekf_status_string(s_));      \
      std::exit(4);                                                                                  \
    }                                                                                                \
  } while (0)

__global__ void add_into(double * acc, const double * x, long long n)
{
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    acc[i] += x[i];
  }
}

struct Shard
{
  ekf_filter * f = nullptr;
  int dev = 0;
  cudaStream_t stream = nullptr;
  double * pose = nullptr, * xy = nullptr, * pose_out = nullptr;   // device copies of the scan's inputs / outputs
  int32_t * assoc = nullptr, * done = nullptr, * count = nullptr;
};

struct Exchange
{
  // sum of the R handles' buffers, delivered back into every one of them
  std::vector<Shard> * sh = nullptr;
  double * stage = nullptr, * acc = nullptr;   // on device 0
  size_t cap = 0;
#ifdef SHARD_REPLAY_NCCL
  std::vector<ncclComm_t> comms;
#endif
  void sum(const std::vector<double *> & bufs, long long n)
  {
    auto & S = *sh;
    if (n == 0) {return;}
#ifdef SHARD_REPLAY_NCCL
    if (!comms.empty()) {
      ncclGroupStart();
      for (size_t r = 0; r < S.size(); ++r) {
        ncclAllReduce(bufs[r], bufs[r], (size_t)n, ncclDouble, ncclSum, comms[r], S[r].stream);
      }
      ncclGroupEnd();
      return;
    }
#endif
    CU(cudaSetDevice(S[0].dev));
    if ((size_t)n > cap) {
      if (stage) {CU(cudaFree(stage)); CU(cudaFree(acc));}
      CU(cudaMalloc(&stage, (size_t)n * 8)); CU(cudaMalloc(&acc, (size_t)n * 8));
      cap = (size_t)n;
    }
    for (auto & s : S) {CU(cudaSetDevice(s.dev)); CU(cudaStreamSynchronize(s.stream));}
    CU(cudaSetDevice(S[0].dev));
    CU(cudaMemcpyPeer(acc, S[0].dev, bufs[0], S[0].dev, (size_t)n * 8));
    for (size_t r = 1; r < S.size(); ++r) {
      CU(cudaMemcpyPeer(stage, S[0].dev, bufs[r], S[r].dev, (size_t)n * 8));
      add_into<<<256, 256>>>(acc, stage, n);
    }
    CU(cudaDeviceSynchronize());
    for (size_t r = 0; r < S.size(); ++r) {CU(cudaMemcpyPeer(bufs[r], S[r].dev, acc, S[0].dev, (size_t)n * 8));}
    CU(cudaDeviceSynchronize());
  }
};

int main(int argc, char ** argv)
{
  int R = 2, N = 2000, m = 20, scans = 8, allreduce = 0;
  unsigned long long seed = 7;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&]() {return i + 1 < argc ? argv[++i] : "0";};
    if (a == "--shards") {R = std::atoi(next());} else if (a == "--landmarks") {N = std::atoi(next());} else if (a == "--m") {
      m = std::atoi(next());
    } else if (a == "--scans") {scans = std::atoi(next());} else if (a == "--seed") {
      seed = std::strtoull(next(), nullptr, 10);
    } else if (a == "--allreduce") {allreduce = 1;} else {
      std::fprintf(stderr, "unknown argument %s\n", a.c_str());
      return 2;
    }
  }
  int ndev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (ndev < 1 || R < 1 || R > 16 || m < 1 || m > 24) {std::fprintf(stderr, "bad configuration\n"); return 2;}
  const int used = std::min(R, ndev);
  for (int a = 0; a < used; ++a) {
    for (int b = 0; b < used; ++b) {
      if (a == b) {continue;}
      int can = 0;
      CU(cudaDeviceCanAccessPeer(&can, a, b));
      if (!can) {std::fprintf(stderr, "no peer access %d -> %d\n", a, b); return 2;}
      CU(cudaSetDevice(a));
      cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {CU(e);}
      cudaGetLastError();
    }
  }
  const int n = 3 + 2 * N;
  const int grid_w = (int)std::ceil(std::sqrt((double)N));

  // ---- the sharded handles --------------------------------------------------------------------------------
  std::vector<Shard> S((size_t)R);
  for (int r = 0; r < R; ++r) {
    Shard & s = S[(size_t)r];
    s.dev = r % ndev;
    CU(cudaSetDevice(s.dev));
    ekf_config cfg;
    ekf_default_config(&cfg);
    cfg.n_landmarks = N; cfg.device = s.dev; cfg.max_scan = m; cfg.shard_rank = r; cfg.shard_count = R;
    EKF(ekf_create(&cfg, &s.f));
    CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    EKF(ekf_set_stream(s.f, s.stream));
    CU(cudaMalloc(&s.pose, 24)); CU(cudaMalloc(&s.xy, 16 * (size_t)m)); CU(cudaMalloc(&s.pose_out, 24));
    CU(cudaMalloc(&s.assoc, 4 * (size_t)m)); CU(cudaMalloc(&s.done, 4)); CU(cudaMalloc(&s.count, 4));
    EKF(ekf_fill_procedural(s.f, seed, N, grid_w));
  }
  if (R > 1 && !allreduce) {           // peer-memory exchange: every handle maps every other handle's region
    std::vector<void *> base((size_t)R);
    for (int r = 0; r < R; ++r) {
      char ipc[64]; int64_t bytes = 0;
      CU(cudaSetDevice(S[(size_t)r].dev));
      EKF(ekf_shard_peer_export(S[(size_t)r].f, ipc, &base[(size_t)r], &bytes));
    }
    for (int r = 0; r < R; ++r) {
      for (int q = 0; q < R; ++q) {
        if (q != r) {EKF(ekf_shard_peer_attach(S[(size_t)r].f, q, nullptr, base[(size_t)q]));}
      }
    }
  }
  Exchange X;
  X.sh = &S;
#ifdef SHARD_REPLAY_NCCL
  if (R > 1 && R <= ndev) {
    std::vector<int> devs((size_t)R);
    for (int r = 0; r < R; ++r) {devs[(size_t)r] = S[(size_t)r].dev;}
    X.comms.resize((size_t)R);
    if (ncclCommInitAll(X.comms.data(), R, devs.data()) != ncclSuccess) {X.comms.clear();}
  }
#endif
  {   // side-state sync after the fixture: Pp / D replicas from the distributed Sigma
    std::vector<double *> bufs((size_t)R);
    int64_t cnt = 0;
    for (int r = 0; r < R; ++r) {CU(cudaSetDevice(S[(size_t)r].dev)); EKF(ekf_shard_sync_begin(S[(size_t)r].f, &bufs[(size_t)r], &cnt));}
    X.sum(bufs, cnt);
    for (int r = 0; r < R; ++r) {CU(cudaSetDevice(S[(size_t)r].dev)); EKF(ekf_shard_sync_finish(S[(size_t)r].f));}
  }

  // ---- the unsharded twin -----------------------------------------------------------------------------------
  CU(cudaSetDevice(0));
  ekf_filter * one = nullptr;
  {
    ekf_config cfg;
    ekf_default_config(&cfg);
    cfg.n_landmarks = N; cfg.device = 0; cfg.max_scan = m;
    EKF(ekf_create(&cfg, &one));
    EKF(ekf_fill_procedural(one, seed, N, grid_w));
  }
  std::vector<double> state((size_t)n);
  EKF(ekf_get_state(one, state.data(), n));

  // ---- scans --------------------------------------------------------------------------------------------------
  std::mt19937_64 rng(seed + 1);
  std::normal_distribution<double> noise(0.0, 1e-3);
  bool assoc_expected = true, assoc_same = true;
  int rounds_total = 0;
  double shard_s = 0.0, one_s = 0.0;
  const double cx = grid_w / 2.0, cy = grid_w / 2.0, rad = std::max(1.5, grid_w / 5.0);
  for (int sc = 0; sc < scans; ++sc) {
    const double ang = 0.3 * sc, x = cx + rad * std::cos(ang), y = cy + rad * std::sin(ang), th = ang + M_PI / 2;
    if (sc == 0) {
      for (auto & s : S) {CU(cudaSetDevice(s.dev)); EKF(ekf_set_previous_xy(s.f, x, y));}
      CU(cudaSetDevice(0));
      EKF(ekf_set_previous_xy(one, x, y));
    }
    std::vector<std::pair<double, int>> d((size_t)N);
    for (int k = 0; k < N; ++k) {
      d[(size_t)k] = {std::hypot(state[(size_t)(3 + 2 * k)] - x, state[(size_t)(4 + 2 * k)] - y), k};
    }
    std::partial_sort(d.begin(), d.begin() + m, d.end());
    std::vector<double> xy((size_t)(2 * m));
    std::vector<int> want((size_t)m);
    const double c = std::cos(th), sn = std::sin(th);
    for (int t = 0; t < m; ++t) {
      const int k = d[(size_t)t].second;
      const double dx = state[(size_t)(3 + 2 * k)] - x, dy = state[(size_t)(4 + 2 * k)] - y;
      xy[(size_t)(2 * t)] = c * dx + sn * dy + noise(rng);
      xy[(size_t)(2 * t + 1)] = -sn * dx + c * dy + noise(rng);
      want[(size_t)t] = k;
    }
    const double pose[3] = {th, x, y};
    // sharded: rounds until every sighting is applied (one round unless a speculation fails or m > 24)
    auto t0 = std::chrono::steady_clock::now();
    for (auto & s : S) {
      CU(cudaSetDevice(s.dev));
      CU(cudaMemcpyAsync(s.pose, pose, 24, cudaMemcpyHostToDevice, s.stream));
      CU(cudaMemcpyAsync(s.xy, xy.data(), 16 * (size_t)m, cudaMemcpyHostToDevice, s.stream));
    }
    int t_done = 0;
    bool first = true;
    while (t_done < m) {
      std::vector<double *> bufs((size_t)R);
      int64_t cnt = 0;
      for (auto & s : S) {
        CU(cudaSetDevice(s.dev));
        double * b = nullptr;
        EKF(ekf_shard_scan_begin(s.f, first ? s.pose : nullptr, m - t_done, s.xy + 2 * t_done, nullptr, 1, &b, &cnt));
        bufs[(size_t)(&s - &S[0])] = b;
      }
      X.sum(bufs, cnt);                               // 0 doubles when the peer-memory push delivered the columns
      for (auto & s : S) {
        CU(cudaSetDevice(s.dev));
        EKF(ekf_shard_scan_finish(s.f, s.assoc + t_done, s.done, s.pose_out, s.count));
      }
      int done = 0;
      CU(cudaSetDevice(S[0].dev));
      CU(cudaMemcpyAsync(&done, S[0].done, 4, cudaMemcpyDeviceToHost, S[0].stream));
      CU(cudaStreamSynchronize(S[0].stream));
      ++rounds_total;
      first = false;
      if (done <= 0 && rounds_total > 8 * (sc + 1) + 8) {std::fprintf(stderr, "no progress\n"); return 5;}
      t_done += done;
    }
    for (auto & s : S) {CU(cudaSetDevice(s.dev)); CU(cudaStreamSynchronize(s.stream));}
    shard_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::vector<int32_t> got((size_t)m), got1((size_t)m);
    CU(cudaSetDevice(S[0].dev));
    CU(cudaMemcpy(got.data(), S[0].assoc, 4 * (size_t)m, cudaMemcpyDeviceToHost));
    // unsharded twin
    CU(cudaSetDevice(0));
    int32_t cnt1 = 0;
    double pose1[3];
    t0 = std::chrono::steady_clock::now();
    EKF(ekf_scan(one, pose, m, xy.data(), nullptr, nullptr, got1.data(), &cnt1, pose1));
    EKF(ekf_synchronize(one));
    one_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int t = 0; t < m; ++t) {
      assoc_expected = assoc_expected && got[(size_t)t] == want[(size_t)t];
      assoc_same = assoc_same && got[(size_t)t] == got1[(size_t)t];
    }
  }

  // ---- compare the filters ------------------------------------------------------------------------------------
  std::vector<double> st1((size_t)n), str((size_t)n);
  CU(cudaSetDevice(0));
  EKF(ekf_get_state(one, st1.data(), n));
  double state_rel = 0.0, norm = 0.0;
  bool replicas_equal = true;
  for (int r = 0; r < R; ++r) {
    CU(cudaSetDevice(S[(size_t)r].dev));
    EKF(ekf_get_state(S[(size_t)r].f, str.data(), n));
    double e = 0.0;
    norm = 0.0;
    for (int i = 0; i < n; ++i) {e += (str[(size_t)i] - st1[(size_t)i]) * (str[(size_t)i] - st1[(size_t)i]); norm += st1[(size_t)i] * st1[(size_t)i];}
    state_rel = std::max(state_rel, std::sqrt(e / norm));
    if (r > 0) {
      std::vector<double> s0((size_t)n);
      CU(cudaSetDevice(S[0].dev));
      EKF(ekf_get_state(S[0].f, s0.data(), n));
      replicas_equal = replicas_equal && std::memcmp(s0.data(), str.data(), (size_t)n * 8) == 0;
    }
  }
  // sampled full-width block rows: sum over the shards (unowned entries read as 0) vs the unsharded filter
  double sigma_rel = 0.0;
  const int rows = 32;
  std::vector<double> blk1((size_t)rows * n), blkr((size_t)rows * n), acc((size_t)rows * n);
  for (int i0 : {0, n / 3, n / 2, n - rows}) {
    CU(cudaSetDevice(0));
    EKF(ekf_get_sigma_block(one, i0, i0 + rows, 0, n, blk1.data()));
    std::fill(acc.begin(), acc.end(), 0.0);
    for (auto & s : S) {
      CU(cudaSetDevice(s.dev));
      EKF(ekf_get_sigma_block(s.f, i0, i0 + rows, 0, n, blkr.data()));
      for (size_t q = 0; q < acc.size(); ++q) {acc[q] += blkr[q];}
    }
    double e = 0.0, nn = 0.0;
    for (size_t q = 0; q < acc.size(); ++q) {e += (acc[q] - blk1[q]) * (acc[q] - blk1[q]); nn += blk1[q] * blk1[q];}
    sigma_rel = std::max(sigma_rel, std::sqrt(e / nn));
  }
  int status = 0;
  for (auto & s : S) {CU(cudaSetDevice(s.dev)); status = std::max(status, ekf_last_status(s.f));}
  const bool ok = assoc_expected && assoc_same && replicas_equal && state_rel <= 1e-12 && sigma_rel <= 1e-12 && status == 0;
  std::printf("{\"ok\": %s, \"shards\": %d, \"devices_used\": %d, \"landmarks\": %d, \"m\": %d, \"scans\": %d, "
    "\"rounds\": %d, \"exchange\": \"%s\", \"association_as_expected\": %s, \"association_equals_unsharded\": %s, "
    "\"replicas_bitwise_equal\": %s, \"state_rel_vs_unsharded\": %.3e, \"sigma_rel_vs_unsharded\": %.3e, "
    "\"sharded_scans_per_s_host_loop\": %.1f, \"unsharded_scans_per_s\": %.1f}\n",
    ok ? "true" : "false", R, used, N, m, scans, rounds_total,
    (R == 1) ? "none" : (allreduce ? "sum of the handles' buffers (all-reduce)" : "peer-memory push"),
    assoc_expected ? "true" : "false", assoc_same ? "true" : "false", replicas_equal ? "true" : "false", state_rel,
    sigma_rel, scans / shard_s, scans / one_s);
  for (auto & s : S) {CU(cudaSetDevice(s.dev)); ekf_destroy(s.f);}
  CU(cudaSetDevice(0));
  ekf_destroy(one);
  return ok ? 0 : 1;
}
