// Microbenchmarks that bound the EKF kernels on this box (not part of the product):
//   1. fp64 FMA pipe peak (DFMA)                 - co-limiter of the covariance pass at m = 20
//   2. fp64 tensor-core peak (DMMA m8n8k4)       - alternative inner product for that pass
//   3. both interleaved                          - do the two pipes overlap?
//   4. streaming read+write bandwidth            - the roofline denominator, as a kernel sees it
//   5. grid-barrier latency                      - cost of the one barrier per measurement
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fp64_peak tools/fp64_peak.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__global__ void dfma_kernel(double *out, int iters, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-9 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  if (s == 123.456) out[0] = s;
}

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void dmma_kernel(double *out, int iters, double a, double b) {
  double c[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) c[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma(c[2 * i], c[2 * i + 1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i];
  if (s == 123.456) out[0] = s;
}

__global__ void mixed_kernel(double *out, int iters, double a, double b) {
  double c[16], acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { c[i] = i; acc[i] = i + threadIdx.x; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      dmma(c[2 * i], c[2 * i + 1], a, b);
      acc[2 * i] = fma(acc[2 * i], a, b);
      acc[2 * i + 1] = fma(acc[2 * i + 1], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i] + acc[i];
  if (s == 123.456) out[0] = s;
}

__global__ void stream_kernel(const double2 *__restrict__ in, double2 *__restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    double2 v0 = in[i], v1 = in[i + stride], v2 = in[i + 2 * stride], v3 = in[i + 3 * stride];
    v0.x += 1.0; v1.x += 1.0; v2.x += 1.0; v3.x += 1.0;
    out[i] = v0; out[i + stride] = v1; out[i + 2 * stride] = v2; out[i + 3 * stride] = v3;
  }
  for (; i < n; i += stride) out[i] = in[i];
}

__global__ void barrier_kernel(unsigned long long *bar, int rounds) {
  for (int r = 0; r < rounds; ++r) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(bar, 1ULL);
      const unsigned long long target = (unsigned long long)(r + 1) * gridDim.x;
      unsigned long long v;
      do { asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(bar) : "memory"); } while (v < target);
      __threadfence();
    }
    __syncthreads();
  }
}

template <typename F>
static float time_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  launch(); launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  printf("device %s, %d SMs, clock %.0f MHz\n", prop.name, sms, prop.clockRate / 1000.0);
  double *out; CK(cudaMalloc(&out, 64));
  const int iters = 20000, blocks = sms * 8, threads = 256;
  {
    float ms = time_ms([&] { dfma_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double flops = 2.0 * 16 * iters * (double)blocks * threads;
    printf("DFMA   : %8.3f ms  %7.2f TFLOP/s fp64\n", ms, flops / ms * 1e-9);
  }
  {
    float ms = time_ms([&] { dmma_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double flops = 2.0 * 256 * 8 * iters * (double)blocks * (threads / 32);
    printf("DMMA   : %8.3f ms  %7.2f TFLOP/s fp64 (m8n8k4)\n", ms, flops / ms * 1e-9);
  }
  {
    float ms = time_ms([&] { mixed_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 5);
    double flops = 2.0 * 256 * 8 * iters * (double)blocks * (threads / 32) + 2.0 * 16 * iters * (double)blocks * threads;
    printf("MIXED  : %8.3f ms  %7.2f TFLOP/s fp64 (DMMA + DFMA interleaved)\n", ms, flops / ms * 1e-9);
  }
  {
    const size_t n = (size_t)1 << 28;  // 4 GiB in, 4 GiB out
    double2 *a, *b; CK(cudaMalloc(&a, n * 16)); CK(cudaMalloc(&b, n * 16));
    CK(cudaMemset(a, 0, n * 16));
    for (int bl : {sms * 4, sms * 8, sms * 16, sms * 32}) {
      float ms = time_ms([&] { stream_kernel<<<bl, 512>>>(a, b, n); }, 5);
      printf("STREAM : %8.3f ms  %7.1f GB/s read+write (%d blocks x 512)\n", ms, 2.0 * n * 16 / ms * 1e-6, bl);
    }
    float ms = time_ms([&] { CK(cudaMemcpyAsync(b, a, n * 16, cudaMemcpyDeviceToDevice)); }, 5);
    printf("MEMCPY : %8.3f ms  %7.1f GB/s read+write (cudaMemcpy D2D)\n", ms, 2.0 * n * 16 / ms * 1e-6);
    CK(cudaFree(a)); CK(cudaFree(b));
  }
  {
    unsigned long long *bar; CK(cudaMalloc(&bar, 8));
    const int rounds = 2000;
    for (int nb : {37, 74, 148}) {
      for (int th : {128, 256}) {
        float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
          CK(cudaMemset(bar, 0, 8));
          cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
          int r = rounds; void *args[] = {&bar, &r};
          CK(cudaEventRecord(e0));
          CK(cudaLaunchCooperativeKernel((const void *)barrier_kernel, dim3(nb), dim3(th), args, 0, 0));
          CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
          float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
          if (ms < best) best = ms;
        }
        printf("BARRIER: %d blocks x %d threads: %.3f us per grid barrier\n", nb, th, best * 1000.0 / rounds);
      }
    }
  }
  return 0;
}
