// ag_pipeline.cu — PROTOTYPE of the signalling a fused all-gather -> GEMM needs, as a
// stand-alone multi-GPU microbenchmark. NOT part of the library and not yet run on hardware.
//
// Today (csrc/comm.cu, mode 0) a gather ends with a barrier over all blocks of all ranks and only then may the
// column-parallel GEMM start. The plan is to let the GEMM's TMA producer wait per SOURCE RANK instead: rows that
// came from rank r are usable as soon as rank r's blocks have signalled, and the consumer walks the sources in
// the order they arrive — its own shard first, then rank+1, rank+2, ... This program checks that protocol and
// measures what it buys, with a checksum kernel standing in for the GEMM:
//
//   pusher (rank r, few CTAs, own stream)    round k = 0..n-1: copy the local shard into rank (r-k)'s gather buffer
//                                            with plain peer stores, __threadfence_system, then one
//                                            red.release.sys.add per CTA on that rank's arrive[r] counter.
//                                            Before touching a peer's buffer it waits until that peer has consumed
//                                            the previous epoch (consumed[dst], read over NVLink).
//   consumer (rank q, the other CTAs)        round k = 0..n-1: src = (q+k) % n; one thread per CTA spins with
//                                            ld.acquire.sys until arrive[src] has reached epoch * pusher CTAs, then
//                                            the CTA reads that shard's rows from the LOCAL buffer.
//
// One process drives all GPUs (peer access enabled), so no torch / NCCL is involved. Spins carry a watchdog
// (report + trap) so a protocol error cannot hang a GPU. Output: per-rank checksums against the host's, and the
// time of gather+consume run back to back versus overlapped.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/micro/ag_pipeline tools/micro/ag_pipeline.cu
//   tools/micro/ag_pipeline [ngpus=2] [shard_MiB=8] [push_ctas=16] [epochs=10]
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e__ = (x);                                                         \
    if (e__ != cudaSuccess) {                                                      \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      exit(2);                                                                     \
    }                                                                              \
  } while (0)

constexpr int MAXR = 8;
constexpr int THREADS = 256;

struct Rank {
  uint4* shard;                 // local shard, `vecs` 16-byte vectors
  uint4* full;                  // gather buffer: n * vecs vectors, chunk r = rank r's shard
  uint32_t* arrive;             // [MAXR] counters: arrive[r] += 1 per pusher CTA of rank r per epoch
  uint32_t* consumed;           // [1]    counter: += 1 per consumer CTA per epoch
  float* sums;                  // [MAXR] checksum of each chunk as consumed
};
struct World {
  Rank r[MAXR];                 // device pointers, valid on every GPU (peer access)
  int n;
  size_t vecs;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t now_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void spin_until(const uint32_t* p, uint32_t target, int tag) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (static_cast<int32_t>(ld_acquire_sys(p) - target) < 0) {
    if ((++spins & 1023u) == 0) {
      const uint64_t t = now_ns();
      if (t0 == 0) t0 = t;
      if (t - t0 > 3000000000ull) {
        printf("ag_pipeline watchdog: block %d tag %d target %u have %u\n", blockIdx.x, tag, target, ld_acquire_sys(p));
        __trap();
      }
    }
  }
}

// rank `me` publishes its shard to every rank, nearest consumer-order first
__global__ void __launch_bounds__(THREADS)
push_kernel(const World w, int me, uint32_t epoch, uint32_t consumer_ctas, int paced) {
  const size_t stride = (size_t)gridDim.x * THREADS;
  for (int k = 0; k < w.n; ++k) {
    const int dst = (me - k + w.n) % w.n;
    // the destination must have consumed the previous contents of its buffer (epoch - 1) before we overwrite them
    if (threadIdx.x == 0 && epoch > 1) spin_until(w.r[dst].consumed, (epoch - 1) * consumer_ctas, 10 + k);
    __syncthreads();
    uint4* out = w.r[dst].full + (size_t)me * w.vecs;
    for (size_t v = (size_t)blockIdx.x * THREADS + threadIdx.x; v < w.vecs; v += stride) out[v] = w.r[me].shard[v];
    if (paced) {
      __threadfence_system();
      __syncthreads();
      if (threadIdx.x == 0) red_release_sys_add(w.r[dst].arrive + me, 1u);
    }
  }
  if (!paced) {  // reference behaviour: everything is signalled at the very end
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < w.n) red_release_sys_add(w.r[threadIdx.x].arrive + me, 1u);
  }
}

// rank `me` consumes the gathered tensor chunk by chunk, in arrival order; the checksum stands in for the GEMM
__global__ void __launch_bounds__(THREADS)
consume_kernel(const World w, int me, uint32_t epoch, uint32_t pusher_ctas) {
  __shared__ float red[THREADS / 32];
  const size_t stride = (size_t)gridDim.x * THREADS;
  for (int k = 0; k < w.n; ++k) {
    const int src = (me + k) % w.n;
    if (threadIdx.x == 0) spin_until(w.r[me].arrive + src, epoch * pusher_ctas, 20 + k);
    __syncthreads();
    const uint4* in = w.r[me].full + (size_t)src * w.vecs;
    float acc = 0.f;
    for (size_t v = (size_t)blockIdx.x * THREADS + threadIdx.x; v < w.vecs; v += stride) {
      uint4 u;  // weak load that bypasses L1: the data was published by the acquire above
      asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(in + v) : "memory");
      const uint32_t ws[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) acc += __uint_as_float(ws[j] << 16) + __uint_as_float(ws[j] & 0xFFFF0000u);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int i = 0; i < THREADS / 32; ++i) s += red[i];
      atomicAdd(w.r[me].sums + src, s);
    }
    __syncthreads();
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) red_release_sys_add(w.r[me].consumed, 1u);
}

int main(int argc, char** argv) {
  int n = argc > 1 ? atoi(argv[1]) : 2;
  const size_t mib = argc > 2 ? atoi(argv[2]) : 8;
  const int push_ctas = argc > 3 ? atoi(argv[3]) : 16;
  const int epochs = argc > 4 ? atoi(argv[4]) : 10;
  int have = 0;
  CK(cudaGetDeviceCount(&have));
  if (n > have || n > MAXR || n < 1) {
    fprintf(stderr, "need %d GPUs, have %d\n", n, have);
    return 1;
  }
  World w{};
  w.n = n;
  w.vecs = mib * (1u << 20) / 16;
  for (int a = 0; a < n; ++a) {
    CK(cudaSetDevice(a));
    for (int b = 0; b < n; ++b)
      if (a != b) {
        int ok = 0;
        CK(cudaDeviceCanAccessPeer(&ok, a, b));
        if (!ok) {
          fprintf(stderr, "no peer access %d -> %d\n", a, b);
          return 1;
        }
        cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
        (void)cudaGetLastError();
      }
  }
  std::vector<std::vector<__nv_bfloat16>> host(n);
  std::vector<double> want(n, 0.0);
  uint32_t lcg = 777u;
  for (int a = 0; a < n; ++a) {
    host[a].resize(w.vecs * 8);
    for (auto& e : host[a]) {
      lcg = lcg * 1664525u + 1013904223u;
      e = __float2bfloat16(((lcg >> 9) & 0xFF) / 256.0f);
      want[a] += __bfloat162float(e);
    }
    CK(cudaSetDevice(a));
    CK(cudaMalloc(&w.r[a].shard, w.vecs * 16));
    CK(cudaMalloc(&w.r[a].full, (size_t)n * w.vecs * 16));
    CK(cudaMalloc(&w.r[a].arrive, MAXR * 4));
    CK(cudaMalloc(&w.r[a].consumed, 4));
    CK(cudaMalloc(&w.r[a].sums, MAXR * 4));
    CK(cudaMemset(w.r[a].arrive, 0, MAXR * 4));
    CK(cudaMemset(w.r[a].consumed, 0, 4));
    CK(cudaMemcpy(w.r[a].shard, host[a].data(), w.vecs * 16, cudaMemcpyHostToDevice));
  }
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const int cons_ctas = sms - push_ctas;  // both kernels of a rank must be resident together: one CTA per SM in total
  std::vector<cudaStream_t> sp(n), sc(n);
  std::vector<cudaEvent_t> e0(n), e1(n);
  int lo_prio, hi_prio;
  CK(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
  for (int a = 0; a < n; ++a) {
    CK(cudaSetDevice(a));
    CK(cudaStreamCreateWithPriority(&sp[a], cudaStreamNonBlocking, hi_prio));
    CK(cudaStreamCreateWithPriority(&sc[a], cudaStreamNonBlocking, lo_prio));
    CK(cudaEventCreate(&e0[a]));
    CK(cudaEventCreate(&e1[a]));
  }

  uint32_t epoch = 0;
  for (int mode = 0; mode < 2; ++mode) {  // 0: signal at the end (today's behaviour), 1: signal per source (the plan)
    double worst_ms = 0;
    for (int it = 0; it < epochs; ++it) {
      ++epoch;
      for (int a = 0; a < n; ++a) {
        CK(cudaSetDevice(a));
        CK(cudaMemsetAsync(w.r[a].sums, 0, MAXR * 4, sc[a]));
        CK(cudaEventRecord(e0[a], sc[a]));
        CK(cudaStreamWaitEvent(sp[a], e0[a], 0));
      }
      for (int a = 0; a < n; ++a) {  // pushers first: a consumer must never occupy the SMs its own pusher needs
        CK(cudaSetDevice(a));
        push_kernel<<<push_ctas, THREADS, 0, sp[a]>>>(w, a, epoch, (uint32_t)cons_ctas, mode);
      }
      for (int a = 0; a < n; ++a) {
        CK(cudaSetDevice(a));
        consume_kernel<<<cons_ctas, THREADS, 0, sc[a]>>>(w, a, epoch, (uint32_t)push_ctas);
        CK(cudaEventRecord(e1[a], sc[a]));
      }
      for (int a = 0; a < n; ++a) {
        CK(cudaSetDevice(a));
        CK(cudaStreamSynchronize(sc[a]));
        CK(cudaStreamSynchronize(sp[a]));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0[a], e1[a]));
        if (it > 0 && ms > worst_ms) worst_ms = ms;
      }
    }
    bool ok = true;
    for (int a = 0; a < n; ++a) {
      float got[MAXR];
      CK(cudaSetDevice(a));
      CK(cudaMemcpy(got, w.r[a].sums, MAXR * 4, cudaMemcpyDeviceToHost));
      for (int s = 0; s < n; ++s)
        if (fabs(got[s] - want[s]) > 1e-3 * want[s]) {
          ok = false;
          printf("rank %d chunk %d: checksum %.1f, expected %.1f\n", a, s, got[s], want[s]);
        }
    }
    const double gb = (double)(n - 1) * w.vecs * 16 / 1e9;
    printf("%s: %d GPUs, shard %zu MiB, %d pusher CTAs: %s, slowest gather+consume %.3f ms (%.0f GB/s into each GPU)\n",
           mode ? "signal per source" : "signal at the end", n, mib, push_ctas, ok ? "checksums MATCH" : "MISMATCH",
           worst_ms, gb / (worst_ms * 1e-3));
  }
  return 0;
}
