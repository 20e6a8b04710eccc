// Probe: cost of the warp arg-min on (double distance, int index) pairs, as in ekf_seq.cuh phase C.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/argmin_probe tools/argmin_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <math.h>
struct Cand { double d; int k; };
__device__ __forceinline__ Cand cand_min(Cand a, Cand b) { return (b.d < a.d || (b.d == a.d && b.k < a.k)) ? b : a; }
__global__ void probe(const double *in, long long *out, int mode, int nwarps_active) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp >= nwarps_active) return;
  Cand c{in[threadIdx.x], (int)threadIdx.x};
  int bstar = lane;
  long long t0 = clock64();
  for (int rep = 0; rep < 16; ++rep) {
    if (mode == 0) {
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        Cand o;
        o.d = __shfl_xor_sync(0xffffffffu, c.d, off);
        o.k = __shfl_xor_sync(0xffffffffu, c.k, off);
        const int ob = __shfl_xor_sync(0xffffffffu, bstar, off);
        const Cand n2 = cand_min(c, o);
        if (n2.k != c.k) bstar = ob;
        c = n2;
      }
    } else {
      unsigned long long key = (unsigned long long)__double_as_longlong(c.d);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, off);
        const int okk = __shfl_xor_sync(0xffffffffu, c.k, off);
        const int ob = __shfl_xor_sync(0xffffffffu, bstar, off);
        const bool take = ok < key || (ok == key && okk < c.k);
        if (take) { key = ok; c.k = okk; bstar = ob; }
      }
      c.d = __longlong_as_double((long long)key);
    }
    c.d += 1.0 + lane * (rep & 1);
  }
  long long t1 = clock64();
  if (lane == 0) { out[warp * 2] = t1 - t0; out[warp * 2 + 1] = (long long)c.k + bstar + (long long)c.d; }
}
int main() {
  double h[128]; for (int i = 0; i < 128; ++i) h[i] = (i % 7 == 3) ? INFINITY : 1.0 + (i * 37 % 11);
  double *din; long long *dout, hout[8];
  cudaMalloc(&din, sizeof h); cudaMalloc(&dout, sizeof hout); cudaMemcpy(din, h, sizeof h, cudaMemcpyHostToDevice);
  for (int mode = 0; mode < 2; ++mode)
    for (int nw = 1; nw <= 4; nw += 3) {
      for (int r = 0; r < 3; ++r) { probe<<<1, 128>>>(din, dout, mode, nw); cudaDeviceSynchronize(); }
      cudaMemcpy(hout, dout, sizeof hout, cudaMemcpyDeviceToHost);
      printf("mode %d (%s) warps %d: %lld cycles per 5-round arg-min\n", mode, mode ? "u64 key" : "double compare", nw, hout[0] / 16);
    }
  return 0;
}
