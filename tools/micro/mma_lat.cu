// Microbenchmark: tcgen05.mma issue/execute rate for dependent vs independent accumulators,
// with a minimal-instruction issue loop (precomputed descriptor words, unrolled).
#include <cstdio>
#include <cuda_runtime.h>
#include "../../fleetx_b200/csrc/tc5.cuh"
using namespace tc5;

template <int N, int NACC, int TS>
__global__ void k(int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    if (lane == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncwarp();
    tmem_alloc<512>(&slot);
  }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 0 && lane == 0) {
    const uint32_t a_lo = desc_lo(smem_u32(smem), 0), b_lo = desc_lo(smem_u32(smem) + 32768, 0);
    const uint32_t idesc = instr_desc_bf16(128, N, 0, 0);
    long long t0 = clock64();
    for (int i = 0; i < iters; i += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t d = tm + (u % NACC) * N;
        if (TS) umma_ts_lo<1>(d, tm + 448, b_lo + u * 2, idesc);
        else umma_ss_lo<1>(d, a_lo + u * 2, b_lo + u * 2, idesc);
      }
    }
    long long t1 = clock64();
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    out[blockIdx.x * 2] = t1 - t0;
    out[blockIdx.x * 2 + 1] = t2 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

template <int N, int NACC, int TS>
void run(long long* d) {
  const int iters = 1024;
  cudaFuncSetAttribute(k<N, NACC, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304);
  k<N, NACC, TS><<<148, 64, 98304>>>(iters, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("%s N=%3d nacc=%d: issue %.1f cyc/mma, complete %.1f cyc/mma (math floor %d) %s\n", TS ? "TS" : "SS", N, NACC,
         (double)h[0] / iters, (double)h[1] / iters, N / 2, cudaGetErrorString(e));
}

int main() {
  long long* d; cudaMalloc(&d, 16 * 148);
  run<64, 1, 0>(d); run<64, 2, 0>(d); run<64, 4, 0>(d);
  run<128, 1, 0>(d); run<128, 2, 0>(d);
  run<256, 1, 0>(d);
  run<64, 1, 1>(d); run<64, 2, 1>(d); run<64, 4, 1>(d);
  run<128, 1, 1>(d); run<128, 2, 1>(d);
  run<256, 1, 1>(d);
  return 0;
}
