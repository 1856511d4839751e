// common.cuh — error reporting, launch checks and TMA tensor-map encoding shared by all
// translation units of libfleetx_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/fleetx_b200.h"

namespace fx {

// Thread-local error text returned by fx_last_error(). Defined in runtime.cu.
void set_error(const char* fmt, ...);

#define FX_CHECK_ARG(cond, ...)                 \
  do {                                          \
    if (!(cond)) {                              \
      fx::set_error(__VA_ARGS__);               \
      return FX_ERR_INVALID_ARGUMENT;           \
    }                                           \
  } while (0)

#define FX_CUDA(call)                                                                     \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess) {                                                             \
      fx::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__,    \
                    __LINE__);                                                            \
      return FX_ERR_CUDA;                                                                 \
    }                                                                                     \
  } while (0)

#define FX_LAUNCH_CHECK()                                                                 \
  do {                                                                                    \
    cudaError_t e__ = cudaGetLastError();                                                 \
    if (e__ != cudaSuccess) {                                                             \
      fx::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, \
                    __LINE__);                                                            \
      return FX_ERR_CUDA;                                                                 \
    }                                                                                     \
    fx::count_launch();                                                                   \
  } while (0)

void count_launch();
int num_sms();
// CTAs the persistent kernels launched by THIS thread may use: #SMs, or less after fx_set_sm_budget()
int sm_budget();

// FX_* environment tuning knobs, read ONCE (thread-safe) the first time any entry point asks; later changes of
// the environment are ignored, and no launch path calls getenv.
struct Knobs {
  int gemm_bn;         // FX_GEMM_BN: 128 / 256 forces the N tile; 0 = heuristic
  int gemm_group;      // FX_GEMM_GROUP: M-tiles per raster group; 0 = default
  int comm_variant;    // FX_COMM_VARIANT: 0 / 1 / 2 forces a tp_allreduce_kernel variant where it applies; -1 = default
  bool attn_bwd_split; // FX_ATTN_BWD=split: the deterministic two-kernel attention backward at head_dim 64
};
const Knobs& knobs();

// Encode a bf16 tensor map with 128-byte swizzle. dims/strides innermost first; strides in
// bytes for dims 1.. (dim 0 is contiguous). Returns 0 or FX_ERR_*.
int make_tma_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box);

inline int make_tma_bf16_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer,
                            uint64_t outer_stride_bytes, uint32_t box_inner, uint32_t box_outer) {
  uint64_t dims[2] = {inner, outer};
  uint64_t strides[1] = {outer_stride_bytes};
  uint32_t box[2] = {box_inner, box_outer};
  return make_tma_bf16(map, base, 2, dims, strides, box);
}

// uint32 2-D map without swizzle (dropout bit masks)
int make_tma_u32_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t outer_stride_bytes,
                    uint32_t box_inner, uint32_t box_outer);

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// A kernel without shared memory of its own can only share an SM with a resident GEMM CTA (whose dynamic window
// takes the whole 228 KB carve-out) if it is launched for the same shared-memory / L1 split: ask for the
// maximum-shared carve-out once per device. FLAGS: a static std::atomic<bool>[64] owned by the caller.
#define FX_PREFER_MAX_SHARED(kernel, flags)                                                              \
  do {                                                                                                   \
    int dev__ = 0;                                                                                       \
    if (cudaGetDevice(&dev__) == cudaSuccess && dev__ >= 0 && dev__ < 64 && !flags[dev__]) {             \
      cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,                       \
                           cudaSharedmemCarveoutMaxShared);                                              \
      flags[dev__] = true;                                                                               \
    }                                                                                                    \
  } while (0)

}  // namespace fx
