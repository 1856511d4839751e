// Throughput of red.global.add.v4.f32 issued from registers (coalesced: a warp covers 512 contiguous
// bytes per instruction) as an alternative dQ accumulation path to the TMA bulk reduce.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void __launch_bounds__(512) red_kernel(float* dst, size_t dst_blocks, int iters) {
  // one "block" = 128 rows x 64 floats = 32 KB, written by 512 threads x 4 v4 ops
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quarter = warp & 3, part = warp >> 2, row = quarter * 32 + lane;
  for (int it = 0; it < iters; ++it) {
    const size_t blk = (static_cast<size_t>(it) + blockIdx.x / 16) % dst_blocks;
    float* base = dst + blk * 8192;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      float* a = base + ((part * 4 + g) * 128 + row) * 4;
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "f"(1.f), "f"(1.f), "f"(1.f), "f"(1.f) : "memory");
    }
  }
}

int main() {
  const size_t dst_bytes = 134217728;
  const size_t blocks = dst_bytes / 32768;
  float* d;
  cudaMalloc(&d, dst_bytes);
  cudaMemset(d, 0, dst_bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int grid : {1, 16, 148, 296}) {
    const int iters = grid < 148 ? 500 : static_cast<int>((1140ull << 20) / 32768 / grid);
    red_kernel<<<grid, 512>>>(d, blocks, iters);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    red_kernel<<<grid, 512>>>(d, blocks, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double gb = static_cast<double>(iters) * grid * 32768 / 1e9;
    printf("grid %d: %.3f ms for %.2f GB -> %.1f GB/s (%s)\n", grid, ms, gb, gb / ms * 1e3, cudaGetErrorString(cudaGetLastError()));
  }
  float h[4];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("sample %f\n", h[0]);
  return 0;
}
