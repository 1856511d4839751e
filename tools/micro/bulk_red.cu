// Throughput of cp.reduce.async.bulk (.add.f32, shared -> global) as a dQ accumulation path:
// every CTA adds CHUNK-byte tiles from shared memory into a global fp32 buffer that `overlap`
// CTAs hit concurrently. Build: nvcc -gencode arch=compute_100a,code=sm_100a -o bulk_red bulk_red.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

template <int CHUNK>
__global__ void __launch_bounds__(128) red_kernel(float* dst, size_t dst_tiles, int iters, int mode) {
  extern __shared__ __align__(128) uint8_t smem[];
  float* s = reinterpret_cast<float*>(smem);
  for (int i = threadIdx.x; i < CHUNK / 4; i += blockDim.x) s[i] = 1.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t saddr = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
    for (int it = 0; it < iters; ++it) {
      // mode 0: CTAs walk the tiles in the same order, shifted by blockIdx (several CTAs per tile over time)
      // mode 1: each CTA owns a private stripe (no sharing) — the plain-store-like ceiling
      size_t tile = mode == 0 ? (static_cast<size_t>(it) + blockIdx.x / 16) % dst_tiles
                              : (static_cast<size_t>(blockIdx.x) * iters + it) % dst_tiles;
      float* g = dst + tile * (CHUNK / 4);
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(g), "r"(saddr),
                   "r"(CHUNK)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  constexpr int CHUNK = 16384;
  const size_t dst_bytes = 134217728;  // fp32 dQ of the bench workload
  const size_t tiles = dst_bytes / CHUNK;
  float* d;
  cudaMalloc(&d, dst_bytes);
  cudaMemset(d, 0, dst_bytes);
  cudaFuncSetAttribute(red_kernel<CHUNK>, cudaFuncAttributeMaxDynamicSharedMemorySize, CHUNK);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode)
    for (int grid : {1, 16, 148, 592}) {
      const int iters = grid < 148 ? 2000 : static_cast<int>((1140ull << 20) / CHUNK / grid);  // ~1.14 GB total at full grids
      red_kernel<CHUNK><<<grid, 128, CHUNK>>>(d, tiles, iters, mode);
      cudaDeviceSynchronize();
      cudaEventRecord(e0);
      red_kernel<CHUNK><<<grid, 128, CHUNK>>>(d, tiles, iters, mode);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double gb = static_cast<double>(iters) * grid * CHUNK / 1e9;
      printf("mode %d grid %d: %.3f ms for %.2f GB reduced -> %.1f GB/s (%s)\n", mode, grid, ms, gb, gb / ms * 1e3,
             cudaGetErrorString(cudaGetLastError()));
    }
  float h[4];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("sample %f\n", h[0]);
  return 0;
}
