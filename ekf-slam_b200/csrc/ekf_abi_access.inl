// Part of ekf_abi.cu: accessors and fixtures.

namespace {
constexpr size_t kBlockPieceBytes = 256u << 20;  // staging granularity for big Sigma blocks
}

extern "C" {

int ekf_state_dim(ekf_filter *f, int32_t *n) {
  if (!f || !n) return EKF_ERR_INVALID;
  *n = f->n;
  return EKF_OK;
}

int ekf_get_state(ekf_filter *f, double *out, int32_t n) {
  if (!f || !out || n < 0 || n > f->n) return EKF_ERR_INVALID;
  if (n == 0) return EKF_OK;
  DeviceGuard guard(f->device);
  int rc = ensure_scratch(f, (size_t)n * 8);
  if (rc != EKF_OK) return rc;
  state_to_ext_kernel<<<(n + 255) / 256, 256, 0, f->stream>>>(f->mu, f->scratch_dev, n);
  f->launches++;
  CK(f, cudaGetLastError());
  CK(f, cudaMemcpyAsync(out, f->scratch_dev, (size_t)n * 8, cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_get_pose(ekf_filter *f, double out[3]) { return ekf_get_state(f, out, 3); }

int ekf_set_state(ekf_filter *f, const double *in, int32_t n) {
  if (!f || !in || n < 0 || n > f->n) return EKF_ERR_INVALID;
  if (n == 0) return EKF_OK;
  DeviceGuard guard(f->device);
  int rc = ensure_scratch(f, (size_t)n * 8);
  if (rc != EKF_OK) return rc;
  CK(f, cudaMemcpyAsync(f->scratch_dev, in, (size_t)n * 8, cudaMemcpyHostToDevice, f->stream));
  state_from_ext_kernel<<<(n + 255) / 256, 256, 0, f->stream>>>(f->mu, f->scratch_dev, n);
  f->launches++;
  CK(f, cudaGetLastError());
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_get_count(ekf_filter *f, int32_t *count) {
  if (!f || !count) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaMemcpyAsync(count, f->cnt, sizeof(int), cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_set_count(ekf_filter *f, int32_t count) {
  if (!f || count < 0 || count > f->N) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaMemcpyAsync(f->cnt, &count, sizeof(int), cudaMemcpyHostToDevice, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_get_landmarks(ekf_filter *f, double *out_xy, int32_t max_landmarks, int32_t *count) {
  if (!f || max_landmarks < 0 || (!out_xy && max_landmarks > 0)) return EKF_ERR_INVALID;
  int32_t c = 0;
  int rc = ekf_get_count(f, &c);
  if (rc != EKF_OK) return rc;
  if (count) *count = c;
  const int32_t k = c < max_landmarks ? c : max_landmarks;
  if (k == 0) return EKF_OK;
  DeviceGuard guard(f->device);
  // landmarks are contiguous in the internal layout starting at index 4
  CK(f, cudaMemcpyAsync(out_xy, f->mu + 4, (size_t)k * 16, cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_set_previous_xy(ekf_filter *f, double x, double y) {
  if (!f) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  const double v[2] = {x, y};
  CK(f, cudaMemcpyAsync(f->prev_xy, v, 16, cudaMemcpyHostToDevice, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_get_previous_xy(ekf_filter *f, double out[2]) {
  if (!f || !out) return EKF_ERR_INVALID;
  DeviceGuard guard(f->device);
  CK(f, cudaMemcpyAsync(out, f->prev_xy, 16, cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

static int block_args_ok(ekf_filter *f, int32_t i0, int32_t i1, int32_t j0, int32_t j1, const void *buf) {
  return f && buf && i0 >= 0 && j0 >= 0 && i1 >= i0 && j1 >= j0 && i1 <= f->n && j1 <= f->n;
}

int ekf_get_sigma_block(ekf_filter *f, int32_t i0, int32_t i1, int32_t j0, int32_t j1, double *out) {
  if (!block_args_ok(f, i0, i1, j0, j1, out)) return EKF_ERR_INVALID;
  if (i1 == i0 || j1 == j0) return EKF_OK;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }
  const size_t ld = (size_t)(j1 - j0);
  size_t rows_per = kBlockPieceBytes / (ld * 8);
  if (rows_per < 1) rows_per = 1;
  for (int32_t r = i0; r < i1; r += (int32_t)rows_per) {
    const int32_t r1 = (int32_t)((size_t)r + rows_per < (size_t)i1 ? (size_t)r + rows_per : (size_t)i1);
    const size_t bytes = (size_t)(r1 - r) * ld * 8;
    int rc = ensure_scratch(f, bytes);
    if (rc != EKF_OK) return rc;
    get_block_kernel<<<f->num_sms * 8, 256, 0, f->stream>>>(f->sigma, f->lay, r, r1, j0, j1, f->scratch_dev);
    f->launches++;
    CK(f, cudaGetLastError());
    CK(f, cudaMemcpyAsync(out + (size_t)(r - i0) * ld, f->scratch_dev, bytes, cudaMemcpyDeviceToHost, f->stream));
    CK(f, cudaStreamSynchronize(f->stream));
  }
  return EKF_OK;
}

int ekf_set_sigma_block(ekf_filter *f, int32_t i0, int32_t i1, int32_t j0, int32_t j1, const double *in) {
  if (!block_args_ok(f, i0, i1, j0, j1, in)) return EKF_ERR_INVALID;
  if (i1 == i0 || j1 == j0) return EKF_OK;
  f->side_valid = false;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }
  const size_t ld = (size_t)(j1 - j0);
  size_t rows_per = kBlockPieceBytes / (ld * 8);
  if (rows_per < 1) rows_per = 1;
  for (int32_t r = i0; r < i1; r += (int32_t)rows_per) {
    const int32_t r1 = (int32_t)((size_t)r + rows_per < (size_t)i1 ? (size_t)r + rows_per : (size_t)i1);
    const size_t bytes = (size_t)(r1 - r) * ld * 8;
    int rc = ensure_scratch(f, bytes);
    if (rc != EKF_OK) return rc;
    CK(f, cudaMemcpyAsync(f->scratch_dev, in + (size_t)(r - i0) * ld, bytes, cudaMemcpyHostToDevice, f->stream));
    set_block_kernel<<<f->num_sms * 8, 256, 0, f->stream>>>(f->sigma, f->lay, r, r1, j0, j1, f->scratch_dev);
    f->launches++;
    CK(f, cudaGetLastError());
    CK(f, cudaStreamSynchronize(f->stream));
  }
  return EKF_OK;
}

int ekf_mahalanobis(ekf_filter *f, double x, double y, double *out_d, int32_t cap, int32_t *count) {
  if (!f || !out_d || cap < 0) return EKF_ERR_INVALID;
  if (f->sharded) return EKF_ERR_INVALID;   // a row shard does not hold every block the sweep reads
  int32_t c = 0;
  int rc = ekf_get_count(f, &c);            // (synchronises the stream; drains nothing else)
  if (rc != EKF_OK) return rc;
  if (count) *count = c;
  const int32_t k = c < cap ? c : cap;
  if (k == 0) return EKF_OK;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }
  rc = ensure_scratch(f, (size_t)k * 8);
  if (rc != EKF_OK) return rc;
  mahalanobis_sweep_kernel<<<(k + 127) / 128, 128, 0, f->stream>>>(f->sigma, f->lay, f->mu, f->cnt, x, y,
                                                                  f->cfg.r_noise, f->scratch_dev, k);
  f->launches++;
  CK(f, cudaGetLastError());
  CK(f, cudaMemcpyAsync(out_d, f->scratch_dev, (size_t)k * 8, cudaMemcpyDeviceToHost, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

int ekf_fill_procedural(ekf_filter *f, uint64_t seed, int32_t count, int32_t grid_w) {
  if (!f || count < 0 || count > f->N || grid_w < 1) return EKF_ERR_INVALID;
  f->side_valid = false;
  DeviceGuard guard(f->device);
  { const int rcd = pipe_drain(f); if (rcd != EKF_OK) return rcd; }
  fill_procedural_sigma_kernel<<<f->num_sms * 16, 256, 0, f->stream>>>(f->sigma, f->lay, f->n_pad, count, seed, f->cfg.prior);
  fill_procedural_state_kernel<<<(f->N + 255) / 256, 256, 0, f->stream>>>(f->mu, f->n_pad, f->N, count, grid_w, seed);
  f->launches += 2;
  CK(f, cudaGetLastError());
  CK(f, cudaMemcpyAsync(f->cnt, &count, sizeof(int), cudaMemcpyHostToDevice, f->stream));
  CK(f, cudaStreamSynchronize(f->stream));
  return EKF_OK;
}

}  // extern "C"
