// The covariance pass: Sigma <- Sigma - V W^T over the stored lower triangle, ONE read and ONE
// write of Sigma per scan.
//
// Replaces the m executions per scan of `sigma = (I - K*Hj) * sigma` (nuslam/src/slam.cpp:
// 264-265 / :485-486; an n x n outer product plus an n^3 product each in the reference).  With
// G_t = Sigma H_t^T and K_t = G_t S_t^-1 collected by ekf_seq.cuh into the n x 2m panels
// W = [G_0 .. G_m-1], V = [K_0 .. K_m-1]:
//        Sigma[i][l] -= sum_q V[i][q] * W[l][q],   q < 2m, for every stored (i >= l) entry.
//
// HBM-bound streaming kernel (algorithmic traffic 16 B per stored entry; 4m fp64 flop per entry,
// so the fp64 FMA pipe is a close second limiter at m = 20 - see DESIGN.md).  Structure:
//   * persistent CTAs (one per SM), each walks a contiguous range of 64-row x 128-column chunks
//     of the column strips (ekf_layout.h); a chunk is one contiguous 64 KiB byte range;
//   * a dedicated I/O warp moves chunks HBM -> shared memory -> HBM with 1-D bulk async copies
//     (cp.async.bulk, the TMA engine; completion on mbarriers), two stages deep, so the eight
//     compute warps never issue a global load or store;
//   * the strip's W rows (128 x 2m) stay resident in shared memory for the whole strip walk;
//     the chunk's V rows (64 x 2m) arrive with the chunk (L2-resident panels);
//   * each compute thread owns an 8 x 4 register tile of the chunk and runs 2m rank-1 FMAs on
//     it, operands read from shared memory as conflict-free 16-byte loads (4 x 8 lane grid).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_layout.h"

namespace ekf {

constexpr int kDdComputeWarps = 8;
constexpr int kDdComputeThreads = kDdComputeWarps * 32;
constexpr int kDdThreads = kDdComputeThreads + 32;  // + one I/O warp
constexpr int kDdStages = 2;
constexpr int kChunkElems = EKF_CHUNK_ROWS * EKF_STRIP;           // 8192 doubles
constexpr int kChunkBytes = kChunkElems * 8;                      // 64 KiB

struct DdParams {
  double *sigma;
  const double *V;   // row operand  (K panels)
  const double *W;   // column operand (G panels)
  int n_pad, kmax, k2;       // k2 = 2m panel columns in use
  int row_begin, row_end;    // rows handled by this launch (multiples of 64); whole matrix: 0, n_pad
  long long total_chunks;
};

// ---- PTX helpers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst_gmem, const void *src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- chunk enumeration: strips in order, 64-row chunks inside each strip ----------------------
struct ChunkCursor {
  int J;      // strip
  int row;    // first row of the chunk (global internal row index)
};
__device__ __forceinline__ int strip_first_row(int J, int row_begin) {
  const int r = J * EKF_STRIP;
  return r > row_begin ? r : row_begin;
}
__host__ __device__ inline long long dd_total_chunks(int n_pad, int row_begin, int row_end) {
  long long total = 0;
  for (int J = 0; J * EKF_STRIP < row_end; ++J) {
    const int first = J * EKF_STRIP > row_begin ? J * EKF_STRIP : row_begin;
    total += (row_end - first) / EKF_CHUNK_ROWS;
  }
  (void)n_pad;
  return total;
}
__device__ inline ChunkCursor dd_seek(long long c, int row_begin, int row_end) {
  ChunkCursor cur{0, 0};
  for (int J = 0; J * EKF_STRIP < row_end; ++J) {
    const int first = strip_first_row(J, row_begin);
    const long long here = (row_end - first) / EKF_CHUNK_ROWS;
    if (c < here) { cur.J = J; cur.row = first + (int)c * EKF_CHUNK_ROWS; return cur; }
    c -= here;
  }
  cur.J = -1;
  return cur;
}
__device__ __forceinline__ void dd_advance(ChunkCursor &cur, int row_begin, int row_end) {
  cur.row += EKF_CHUNK_ROWS;
  if (cur.row >= row_end) { cur.J += 1; cur.row = strip_first_row(cur.J, row_begin); }
}

// Shared-memory carve-up (dynamic): [sigma stages][V-row stages][W strip rows][mbarriers]
__host__ __device__ inline size_t dd_smem_bytes(int k2) {
  return (size_t)kDdStages * kChunkBytes                       // sigma chunks
         + (size_t)kDdStages * 2 * k2 * EKF_PBLK * 8           // V rows of the chunk: 2 blocks of 32 rows
         + (size_t)4 * k2 * EKF_PBLK * 8                       // W rows of the strip: 4 blocks of 32 rows
         + 64;                                                 // mbarriers
}

__global__ void __launch_bounds__(kDdThreads, 1) ekf_downdate_kernel(const DdParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int k2 = p.k2;
  const int blk_elems = k2 * EKF_PBLK;  // doubles per 32-row panel block in use
  double *s_sig = reinterpret_cast<double *>(smem_raw);
  double *s_v = s_sig + kDdStages * kChunkElems;
  double *s_w = s_v + kDdStages * 2 * blk_elems;
  uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_w + 4 * blk_elems);
  uint64_t *full = s_bar;          // [kDdStages] chunk + V rows landed
  uint64_t *done = s_bar + 2;      // [kDdStages] compute warps finished the stage
  uint64_t *wfull = s_bar + 4;     // strip W rows landed

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const long long c_begin = p.total_chunks * blockIdx.x / gridDim.x;
  const long long c_end = p.total_chunks * (blockIdx.x + 1) / gridDim.x;
  const int n_chunks = (int)(c_end - c_begin);

  if (tid == 0) {
    for (int s = 0; s < kDdStages; ++s) { mbar_init(&full[s], 1); mbar_init(&done[s], kDdComputeWarps); }
    mbar_init(wfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (n_chunks == 0) return;

  const uint32_t blk_bytes = (uint32_t)blk_elems * 8;

  if (warp == kDdComputeWarps) {
    // ================= I/O warp: one elected lane drives the bulk-copy engine ==================
    if ((tid & 31) == 0) {
      ChunkCursor ld = dd_seek(c_begin, p.row_begin, p.row_end);   // next chunk to load
      ChunkCursor st = ld;                                         // next chunk to store
      for (int it = 0; it < n_chunks + kDdStages; ++it) {
        const int s = it % kDdStages;
        if (it >= kDdStages) {
          // stage s holds the finished chunk it-kDdStages: write it back, then the buffer is free
          mbar_wait(&done[s], ((it / kDdStages) - 1) & 1);
          bulk_s2g(p.sigma + ekf_elem(p.n_pad, st.row, st.J * EKF_STRIP), s_sig + s * kChunkElems, kChunkBytes);
          bulk_commit();
          dd_advance(st, p.row_begin, p.row_end);
          bulk_wait_read0();
        }
        if (it < n_chunks) {
          mbar_expect_tx(&full[s], kChunkBytes + 2 * blk_bytes);
          bulk_g2s(s_sig + s * kChunkElems, p.sigma + ekf_elem(p.n_pad, ld.row, ld.J * EKF_STRIP), kChunkBytes, &full[s]);
          const int vb = ld.row / EKF_PBLK;
          bulk_g2s(s_v + (s * 2 + 0) * blk_elems, p.V + (int64_t)vb * p.kmax * EKF_PBLK, blk_bytes, &full[s]);
          bulk_g2s(s_v + (s * 2 + 1) * blk_elems, p.V + (int64_t)(vb + 1) * p.kmax * EKF_PBLK, blk_bytes, &full[s]);
          dd_advance(ld, p.row_begin, p.row_end);
        }
      }
      bulk_wait_all0();
    }
    return;
  }

  // ======================= compute warps ============================================================
  const int lane = tid & 31;
  const int wr = warp >> 2;        // 32-row block of the chunk (0..1)
  const int wc = warp & 3;         // 32-column block of the strip (0..3)
  const int rg = lane >> 3;        // 0..3
  const int g = lane & 7;          // 0..7
  // thread rows: 32*wr + 8*pp + 2*rg + e  (pp 0..3, e 0..1); cols: 32*wc + 16*c + 2*g + f (c,f 0..1)
  const double *vbase0 = s_v + wr * blk_elems + 2 * rg;
  const double *wbase = s_w + wc * blk_elems + 2 * g;

  ChunkCursor cur = dd_seek(c_begin, p.row_begin, p.row_end);
  int loadedJ = -1;
  uint32_t wphase = 0;
  for (int it = 0; it < n_chunks; ++it) {
    const int s = it % kDdStages;
    if (cur.J != loadedJ) {
      // new strip: all compute warps must be done with the old W rows before they are replaced
      named_bar_sync(1, kDdComputeThreads);
      if (tid == 0) {
        mbar_expect_tx(wfull, 4 * blk_bytes);
        for (int b = 0; b < 4; ++b)
          bulk_g2s(s_w + b * blk_elems, p.W + (int64_t)(cur.J * 4 + b) * p.kmax * EKF_PBLK, blk_bytes, wfull);
      }
      mbar_wait(wfull, wphase);
      wphase ^= 1;
      loadedJ = cur.J;
    }
    mbar_wait(&full[s], (it / kDdStages) & 1);

    double *tile = s_sig + s * kChunkElems + (32 * wr + 2 * rg) * EKF_STRIP + 32 * wc + 2 * g;
    double2 acc[4][2][2];  // [pp][e][c] -> (f = x, y)
#pragma unroll
    for (int pp = 0; pp < 4; ++pp)
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int c = 0; c < 2; ++c)
          acc[pp][e][c] = *reinterpret_cast<const double2 *>(tile + (8 * pp + e) * EKF_STRIP + 16 * c);

    const double *vb = vbase0 + s * 2 * blk_elems;
#pragma unroll 2
    for (int q = 0; q < k2; ++q) {
      double2 v[4], w[2];
#pragma unroll
      for (int pp = 0; pp < 4; ++pp) v[pp] = *reinterpret_cast<const double2 *>(vb + q * EKF_PBLK + 8 * pp);
#pragma unroll
      for (int c = 0; c < 2; ++c) w[c] = *reinterpret_cast<const double2 *>(wbase + q * EKF_PBLK + 16 * c);
#pragma unroll
      for (int pp = 0; pp < 4; ++pp)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          acc[pp][0][c].x = fma(-v[pp].x, w[c].x, acc[pp][0][c].x);
          acc[pp][0][c].y = fma(-v[pp].x, w[c].y, acc[pp][0][c].y);
          acc[pp][1][c].x = fma(-v[pp].y, w[c].x, acc[pp][1][c].x);
          acc[pp][1][c].y = fma(-v[pp].y, w[c].y, acc[pp][1][c].y);
        }
    }

#pragma unroll
    for (int pp = 0; pp < 4; ++pp)
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int c = 0; c < 2; ++c)
          *reinterpret_cast<double2 *>(tile + (8 * pp + e) * EKF_STRIP + 16 * c) = acc[pp][e][c];

    // make the generic-proxy writes visible to the bulk-copy engine, then hand the stage over
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(&done[s]);
    dd_advance(cur, p.row_begin, p.row_end);
  }
}

}  // namespace ekf
