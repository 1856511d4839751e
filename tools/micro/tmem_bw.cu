// Microbenchmark: TMEM -> register read throughput per SM (tcgen05.ld 32x32b.x32 / .x16).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bw tmem_bw.cu && ./tmem_bw
#include <cstdio>
#include <cuda_runtime.h>
#include "../../fleetx_b200/csrc/tc5.cuh"
using namespace tc5;

template <int X>
__global__ void k(int iters, long long* out, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  float acc = 0.f;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (X == 32) {
      uint32_t r[32];
      tmem_ld_32x32(base + ((i * 32) & 255) + (warp >> 2) * 0, r);
      tmem_ld_wait();
      acc += __uint_as_float(r[0]) + __uint_as_float(r[31]);
    } else {
      uint32_t r[16];
      tmem_ld_32x16(base + ((i * 16) & 255), r);
      tmem_ld_wait();
      acc += __uint_as_float(r[0]) + __uint_as_float(r[15]);
    }
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(slot);
}

int main() {
  long long* d; float* s; cudaMalloc(&d, 8 * 148); cudaMalloc(&s, 4 * 148 * 1024);
  for (int warps : {4, 8, 16, 32}) {
    for (int x : {32, 16}) {
      const int iters = 2000;
      if (x == 32) k<32><<<148, warps * 32>>>(iters, d, s); else k<16><<<148, warps * 32>>>(iters, d, s);
      cudaError_t e = cudaDeviceSynchronize();
      long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      double bytes = (double)iters * warps * 32 * x * 4;
      printf("warps %2d x%-2d: %lld cycles, %.1f B/cycle/SM, %.1f cycles per warp-load (%s)\n", warps, x, h, bytes / h,
             (double)h / iters, cudaGetErrorString(e));
    }
  }
  return 0;
}
