// The covariance pass: Sigma <- Sigma - V W^T over the stored lower triangle, ONE read and ONE
// write of Sigma per scan.
//
// Replaces the m executions per scan of `sigma = (I - K*Hj) * sigma` (nuslam/src/slam.cpp:
// 264-265 / :485-486; an n x n outer product plus an n^3 product each in the reference).  With
// G_t = Sigma H_t^T and K_t = G_t S_t^-1 collected by ekf_seq.cuh into the n x 2m panels
// W = [G_0 .. G_m-1], V = [K_0 .. K_m-1]:
//        Sigma[i][l] -= sum_q V[i][q] * W[l][q],   q < 2m, for every stored (i >= l) entry.
//
// Roofline: 16 B of HBM traffic and 4m fp64 flop per stored entry.  At m = 20 that is 3.2 GB and
// 16.3 GFLOP per pass at N = 10,000: 0.49 ms at the measured 6.5 TB/s against 0.44 ms at the
// fp64 peak (37 TFLOP/s), so BOTH limits bind and neither the FMA issue slots nor the shared-memory
// operand traffic of a scalar DFMA loop can be afforded (first version of round 1, ncu capture in
// profiles/earlier/r1_downdate_ncu_full_summary.csv: the DFMA loop ran at 58 % fp64-pipe and 72 %
// shared-memory wavefront utilisation, 0.78 ms).
// Structure of this version:
//   * persistent CTAs (one per SM), each walks a contiguous range of 64-row x 128-column chunks
//     of the column strips (ekf_layout.h); a chunk is one contiguous 64 KiB byte range;
//   * the rank-2m product runs on the fp64 tensor path: mma.sync.m8n8k4.f64 (DMMA.8x8x4), 16
//     accumulator tiles (32 x 32 of Sigma) per warp, 8 warps per chunk.  Per k-step a warp issues
//     8 conflict-free LDS.64 fragment loads for 16 DMMAs (the DFMA loop needed 6 LDS.128 per
//     32 DFMA per thread);
//   * Sigma goes HBM -> registers -> HBM directly in accumulator-fragment order with 256-bit
//     accesses (SASS LDG.E.NA.ENL2.256 / STG.E.ENL2.256, profiles/sass_excerpt.txt): the row/column interleave of ekf_panel_slot() makes every
//     thread own 4 contiguous doubles, every warp access 8 full 128-byte lines; the next chunk's
//     tile is prefetched into a second register set while the current one is updated;
//   * the strip's W rows (128 x 2m) stay resident in shared memory for the whole strip walk, double
//     buffered: the producer warp requests the next strip's rows as soon as its ring reaches that
//     strip, so the compute warps change strips without draining (the short strips at the tail of the
//     triangle made the last CTAs 14 % slower than the rest before); the chunk's V rows (64 x 2m,
//     L2-resident panels) arrive through a 3-stage bulk-copy (cp.async.bulk + mbarrier) ring driven
//     by the same producer warp.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ekf_layout.h"

namespace ekf {

constexpr int kDdVStages = 3;
constexpr int kDdWBufs = 2;   // W rows of the current strip + the next one (loaded ahead by the producer)
// Two shapes of the same kernel (template parameters CR = rows per chunk, PF = register prefetch):
//   <64, true >  8 compute warps + producer, 1 CTA / SM, next chunk prefetched into registers
//   <32, false>  4 compute warps + producer, 3 CTAs / SM, no register double buffer: the load latency
//                of one CTA is covered by the DMMA phases of the other two (they drift apart freely)
constexpr int dd_threads(int chunk_rows) { return (chunk_rows / 8) * 32 + 32; }

struct DdParams {
  double *sigma_out;   // nullptr: in place; else the updated triangle goes here (same layout), `sigma` stays intact
  double *sigma;
  const double *V;   // row operand  (K panels)
  const double *W;   // column operand (G panels)
  EkfLay lay;                // rows stored by this handle (whole matrix or one row shard)
  int n_pad, kmax, k2;       // k2 = panel columns in use, a multiple of 4
  int row_begin, row_end;    // rows handled by this launch (multiples of 64); whole matrix: 0, n_pad
  long long total_chunks;
  int prefetch_ahead;        // L2 prefetch distance of the producer warp, in chunks (0 = off)
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
// L2 prefetch of a contiguous byte range (SASS: UBLKPF): turns the HBM latency of the compute warps'
// 256-bit Sigma loads into an L2 hit a few chunks later
__device__ __forceinline__ void bulk_prefetch_l2(const void *src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
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
  int J;          // strip
  int row;        // first row of the chunk (global internal row index)
  long long off;  // element offset of the chunk in the handle's Sigma storage (kept incrementally)
};
__device__ __forceinline__ int strip_first_row(int J, int row_begin) {
  const int r = J * EKF_STRIP;
  return r > row_begin ? r : row_begin;
}
__host__ __device__ inline long long dd_total_chunks(int chunk_rows, int row_begin, int row_end) {
  long long total = 0;
  for (int J = 0; J * EKF_STRIP < row_end; ++J) {
    const int first = J * EKF_STRIP > row_begin ? J * EKF_STRIP : row_begin;
    total += (row_end - first) / chunk_rows;
  }
  return total;
}
__device__ inline ChunkCursor dd_seek(long long c, int chunk_rows, EkfLay lay) {
  ChunkCursor cur{0, 0, 0};
  for (int J = 0; J * EKF_STRIP < lay.re; ++J) {
    const int first = ekf_lay_first_row(lay, J);
    const long long here = (lay.re - first) / chunk_rows;
    if (c < here) {
      cur.J = J; cur.row = first + (int)c * chunk_rows;
      cur.off = ekf_lay_elem(lay, cur.row, J * EKF_STRIP);
      return cur;
    }
    c -= here;
  }
  cur.J = -1;
  return cur;
}
__device__ __forceinline__ void dd_advance(ChunkCursor &cur, int chunk_rows, EkfLay lay) {
  cur.row += chunk_rows;
  cur.off += (long long)chunk_rows * EKF_STRIP;          // next rows of the same strip are contiguous
  if (cur.row >= lay.re) {
    cur.J += 1; cur.row = ekf_lay_first_row(lay, cur.J);
    cur.off = ekf_lay_strip_offset(lay, cur.J);
  }
}

// Shared-memory carve-up (dynamic): [V-row stages][W strip rows][mbarriers]
__host__ __device__ inline size_t dd_smem_bytes(int chunk_rows, int k2) {
  return (size_t)kDdVStages * (chunk_rows / EKF_PBLK) * k2 * EKF_PBLK * 8   // V rows of the chunk
         + (size_t)kDdWBufs * 4 * k2 * EKF_PBLK * 8            // W rows of a strip: 4 blocks of 32 rows, x2
         + 128;                                                // mbarriers
}

__device__ __forceinline__ void ld256(const double *p, double &a, double &b, double &c, double &d) {
  asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}
__device__ __forceinline__ void st256(double *p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
// D = A (8x4, row) * B (4x8, col) + C on the fp64 tensor path
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// k2 must be a multiple of 4 (the sequential kernel zero-fills the pad columns of an odd scan).
template <int CR, int PF>
__global__ void __launch_bounds__(dd_threads(CR), PF ? (CR == 32 ? 2 : 1) : 3) ekf_downdate_kernel(const DdParams p) {
  constexpr int kComputeWarps = CR / 8;            // each warp owns a 32 x 32 tile of the chunk
  constexpr int kVBlocks = CR / EKF_PBLK;          // 32-row panel blocks per chunk
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int k2 = p.k2;
  const int blk_elems = k2 * EKF_PBLK;  // doubles per 32-row panel block in use
  double *s_v = reinterpret_cast<double *>(smem_raw);
  double *s_w = s_v + kDdVStages * kVBlocks * blk_elems;
  uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_w + kDdWBufs * 4 * blk_elems);
  uint64_t *full = s_bar;                  // [kDdVStages] V rows of the chunk landed
  uint64_t *done = s_bar + kDdVStages;     // [kDdVStages] compute warps finished with the stage
  uint64_t *wfull = s_bar + 2 * kDdVStages;             // [kDdWBufs] strip W rows landed
  uint64_t *wdone = s_bar + 2 * kDdVStages + kDdWBufs;  // [kDdWBufs] compute warps left the strip

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const long long c_begin = p.total_chunks * blockIdx.x / gridDim.x;
  const long long c_end = p.total_chunks * (blockIdx.x + 1) / gridDim.x;
  const int n_chunks = (int)(c_end - c_begin);

  if (tid == 0) {
    for (int s = 0; s < kDdVStages; ++s) { mbar_init(&full[s], 1); mbar_init(&done[s], kComputeWarps); }
    for (int b = 0; b < kDdWBufs; ++b) { mbar_init(&wfull[b], 1); mbar_init(&wdone[b], kComputeWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (n_chunks == 0) return;

  const uint32_t blk_bytes = (uint32_t)blk_elems * 8;

  if (warp == kComputeWarps) {
    // ================= producer warp: one elected lane feeds the V-row ring ====================
    if ((tid & 31) == 0) {
      ChunkCursor ld = dd_seek(c_begin, CR, p.lay);
      ChunkCursor pf = ld;       // Sigma chunk to prefetch into L2: `prefetch_ahead` chunks past `ld`
      constexpr uint32_t kChunkBytes = CR * EKF_STRIP * 8;
      int pf_it = 0;
      int wJ = -1, wcount = 0;   // strip whose W rows were requested last; strips requested so far
      for (int it = 0; it < n_chunks; ++it) {
        const int s = it % kDdVStages;
        if (it >= kDdVStages) mbar_wait(&done[s], ((it / kDdVStages) - 1) & 1);
        if (ld.J != wJ) {
          // the ring runs up to kDdVStages chunks ahead of the compute warps: the next strip's W rows are
          // in shared memory before they get there, so a strip change costs them no drain.  The buffer
          // was last read for the strip before the previous one.
          const int b = wcount % kDdWBufs;
          if (wcount >= kDdWBufs) mbar_wait(&wdone[b], ((wcount / kDdWBufs) - 1) & 1);
          mbar_expect_tx(&wfull[b], 4 * blk_bytes);
          for (int q = 0; q < 4; ++q)
            bulk_g2s(s_w + (b * 4 + q) * blk_elems, p.W + (int64_t)(ld.J * 4 + q) * p.kmax * EKF_PBLK, blk_bytes,
                     &wfull[b]);
          wJ = ld.J; ++wcount;
        }
        for (; p.prefetch_ahead > 0 && pf_it <= it + p.prefetch_ahead && pf_it < n_chunks; ++pf_it) {
          if (pf_it >= 2) {      // chunks 0 and 1 are fetched by the compute warps right away
            const double *src = p.sigma + pf.off;
            bulk_prefetch_l2(src, kChunkBytes);
          }
          dd_advance(pf, CR, p.lay);
        }
        mbar_expect_tx(&full[s], kVBlocks * blk_bytes);
        const int vb = ld.row / EKF_PBLK;
#pragma unroll
        for (int b = 0; b < kVBlocks; ++b)
          bulk_g2s(s_v + (s * kVBlocks + b) * blk_elems, p.V + (int64_t)(vb + b) * p.kmax * EKF_PBLK, blk_bytes,
                   &full[s]);
        dd_advance(ld, CR, p.lay);
      }
    }
    return;
  }

  // ======================= compute warps ============================================================
  const int lane = tid & 31;
  const int wr = warp >> 2;        // 32-row block of the chunk
  const int wc = warp & 3;         // 32-column block of the strip (0..3)
  const int g = lane >> 2;         // fragment row / column group (0..7)
  const int t4 = lane & 3;         // fragment k index / column pair (0..3)
  // accumulator tile (i, j), i, j = 0..3: row = 32 wr + row_of_slot(8 i + g), columns
  // 32 wc + 16 (j >> 1) + 4 t4 + 2 (j & 1) + {0, 1}: the pair (j, j + 1) is 4 contiguous doubles.
  int toff[4];                     // element offset of (tile row i, column pair 0) inside the chunk
#pragma unroll
  for (int i = 0; i < 4; ++i)
    toff[i] = (32 * wr + ekf_panel_row_of_slot(8 * i + g)) * EKF_STRIP + 32 * wc + 4 * t4;

  ChunkCursor cur = dd_seek(c_begin, CR, p.lay);
  double c[4][4][2], cn[PF ? 4 : 1][4][2];
  if (PF) {
    const double *base = p.sigma + cur.off;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
        ld256(base + toff[i] + 16 * jj, c[i][2 * jj][0], c[i][2 * jj][1], c[i][2 * jj + 1][0], c[i][2 * jj + 1][1]);
  }
  int loadedJ = -1, wcount = 0;
  const double *wb = nullptr;
  const int ksteps = k2 >> 2;
  for (int it = 0; it < n_chunks; ++it) {
    const int s = it % kDdVStages;
    double *out = (p.sigma_out ? p.sigma_out : p.sigma) + cur.off;
    const int curJ = cur.J;
    dd_advance(cur, CR, p.lay);
    if (PF == 1) {
      if (it + 1 < n_chunks) {       // prefetch the next chunk's tile into the second register set
        const double *base = p.sigma + cur.off;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int jj = 0; jj < 2; ++jj)
            ld256(base + toff[i] + 16 * jj, cn[i][2 * jj][0], cn[i][2 * jj][1], cn[i][2 * jj + 1][0],
                  cn[i][2 * jj + 1][1]);
      }
    } else if (PF == 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
          ld256(out + toff[i] + 16 * jj, c[i][2 * jj][0], c[i][2 * jj][1], c[i][2 * jj + 1][0], c[i][2 * jj + 1][1]);
    }
    if (curJ != loadedJ) {
      // new strip: hand the old W buffer back to the producer, switch to the one it filled ahead
      if (loadedJ >= 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&wdone[(wcount - 1) % kDdWBufs]);
      }
      const int b = wcount % kDdWBufs;
      mbar_wait(&wfull[b], (wcount / kDdWBufs) & 1);
      wb = s_w + (b * 4 + wc) * blk_elems + lane;
      ++wcount;
      loadedJ = curJ;
    }
    mbar_wait(&full[s], (it / kDdVStages) & 1);

    const double *va = s_v + (s * kVBlocks + wr) * blk_elems + lane;
    int ks0 = 0;
    if (PF == 2) {
      // first k-step, THEN the register prefetch of the next chunk: every accumulator's load has been
      // consumed before new loads are issued, so the new loads cannot hold back the DMMAs of this chunk
      {
        double a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = -va[i * 32];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = wb[j * 32];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma884(c[i][j][0], c[i][j][1], a[i], b[j]);
      }
      ks0 = 1;
      if (it + 1 < n_chunks) {
        const double *base = p.sigma + cur.off;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int jj = 0; jj < 2; ++jj)
            ld256(base + toff[i] + 16 * jj, cn[i][2 * jj][0], cn[i][2 * jj][1], cn[i][2 * jj + 1][0],
                  cn[i][2 * jj + 1][1]);
      }
    }
#pragma unroll 2
    for (int ks = ks0; ks < ksteps; ++ks) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = -va[ks * 128 + i * 32];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = wb[ks * 128 + j * 32];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(c[i][j][0], c[i][j][1], a[i], b[j]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&done[s]);

#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
        st256(out + toff[i] + 16 * jj, c[i][2 * jj][0], c[i][2 * jj][1], c[i][2 * jj + 1][0], c[i][2 * jj + 1][1]);
    if (PF) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { c[i][j][0] = cn[i][j][0]; c[i][j][1] = cn[i][j][1]; }
    }
  }
}

}  // namespace ekf
