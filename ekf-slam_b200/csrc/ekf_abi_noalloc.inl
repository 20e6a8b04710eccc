// Part of ekf_abi.cu (same translation unit): ONE entry point that allocates nothing.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// ekf_mahalanobis stages its distances in a per-handle scratch buffer that grows with cudaMalloc (ensure_scratch).
// ekf_mahalanobis_noalloc is the same kernel (mahalanobis_sweep_kernel, same arithmetic, same launch shape) with the
// staging in a file-scope `static __device__` array instead: EKF_NOALLOC_MAX_LANDMARKS doubles = 512 KiB, held by
// every GPU that loads the library until the process exits.
//
//   * In host code the array's NAME is a host-side stand-in, not a device address: handing it to a kernel faults (or,
//     where the GPU can reach host memory, silently uses the wrong memory).  The device address comes from
//     cudaGetSymbolAddress, once per device, with that device current (noalloc_scratch_ptr below).
//   * Every call on a device shares the one array and finds the previous call's contents in it.  Nothing in it has to
//     start at zero - a call reads back only the k elements its own kernel has just written - so nothing is cleared.
//   * Calls that OVERLAP on one GPU (two handles on different streams, two host threads) race on the array: the
//     caller serialises them.  Calls on one handle are stream-ordered and synchronous at return, like every accessor.

namespace {

constexpr int kNoAllocMaxLandmarks = EKF_NOALLOC_MAX_LANDMARKS;
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxLandmarks];

constexpr int kNoAllocMaxDevices = 64;
double *g_noalloc_ptr[kNoAllocMaxDevices] = {};   // device address of g_noalloc_scratch on device d (nullptr: not yet asked)
std::mutex g_noalloc_mu;

// the array's address on handle f's device; f->device must be current (DeviceGuard in the caller)
int noalloc_scratch_ptr(ekf_filter *f, double **out) {
  if (f->device < 0 || f->device >= kNoAllocMaxDevices) return EKF_ERR_INVALID;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[f->device]) {
    double *p = nullptr;
    CK(f, cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
    if (!p) return EKF_ERR_CUDA;
    g_noalloc_ptr[f->device] = p;
  }
  *out = g_noalloc_ptr[f->device];
  return EKF_OK;
}

}  // namespace

extern "C" {

int ekf_mahalanobis_noalloc(ekf_filter *f, double x, double y, double *out_d, int32_t cap, int32_t *count) {
  if (!f || !out_d || cap < 0) return EKF_ERR_INVALID;
  if (f->sharded) return EKF_ERR_INVALID;   // a row shard does not hold every block the sweep reads
  if (cap > kNoAllocMaxLandmarks) {
    f->cuda_err = "ekf_mahalanobis_noalloc: cap exceeds EKF_NOALLOC_MAX_LANDMARKS (" +
                  std::to_string(kNoAllocMaxLandmarks) + " distances, the size of the static scratch array); "
                  "use ekf_mahalanobis";
    return EKF_ERR_INVALID;
  }
  int32_t c = 0;
  int rc = ekf_get_count(f, &c);            // one 4-byte copy + synchronise; allocates nothing
  if (rc != EKF_OK) return rc;
  if (count) *count = c;
  const int32_t k = c < cap ? c : cap;
  if (k == 0) return EKF_OK;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }   // orders after pending passes; no allocation
  double *scratch = nullptr;
  rc = noalloc_scratch_ptr(f, &scratch);
  if (rc != EKF_OK) return rc;
  mahalanobis_sweep_kernel<<<(k + 127) / 128, 128, 0, f->stream>>>(f->sigma, f->lay, f->mu, f->cnt, x, y,
                                                                  f->cfg.r_noise, scratch, k);
  f->launches++;
  CK(f, cudaGetLastError());
  CK(f, cudaMemcpyAsync(out_d, scratch, (size_t)k * 8, cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

}  // extern "C"
