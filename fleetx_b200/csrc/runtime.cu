// runtime.cu — process-wide bits of libfleetx_b200.so: error text, launch counter, device
// properties and the cuTensorMapEncodeTiled entry point (fetched from the driver at run time so
// the library links against nothing but the CUDA runtime).
#include <stdlib.h>

#include <atomic>
#include <mutex>

#include "common.cuh"

namespace fx {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int num_sms() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

static thread_local int g_sm_budget = 0;
int sm_budget() {
  const int n = num_sms();
  return (g_sm_budget > 0 && g_sm_budget < n) ? g_sm_budget : n;
}

const Knobs& knobs() {
  static Knobs k;
  static std::once_flag once;
  std::call_once(once, [] {
    k.gemm_bn = 0; k.gemm_group = 0; k.comm_variant = -1; k.attn_bwd_split = false;
    if (const char* e = getenv("FX_GEMM_BN")) k.gemm_bn = atoi(e) == 128 ? 128 : 256;
    if (const char* e = getenv("FX_GEMM_GROUP")) k.gemm_group = atoi(e) > 0 ? atoi(e) : 0;
    if (const char* e = getenv("FX_COMM_VARIANT")) k.comm_variant = atoi(e);
    if (const char* e = getenv("FX_ATTN_BWD")) k.attn_bwd_split = e[0] == 's';
  });
  return k;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tma_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return FX_ERR_CUDA;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gdim, gstr,
                  bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d): base %p rank %d dims [%llu,%llu] "
              "stride1 %llu box [%u,%u]",
              (int)r, base, rank, (unsigned long long)dims[0],
              (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0], rank > 1 ? box[1] : 0);
    return FX_ERR_CUDA;
  }
  return FX_OK;
}

int make_tma_u32_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t outer_stride_bytes,
                    uint32_t box_inner, uint32_t box_outer) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
    return FX_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstr[1] = {outer_stride_bytes};
  cuuint32_t bx[2] = {box_inner, box_outer};
  cuuint32_t es[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), gdim, gstr, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (u32) failed (CUresult %d)", (int)r);
    return FX_ERR_CUDA;
  }
  return FX_OK;
}

}  // namespace fx

extern "C" {

const char* fx_last_error(void) { return fx::g_err; }
int fx_version(void) { return 100; }
uint64_t fx_launch_count(void) { return fx::g_launches.load(std::memory_order_relaxed); }
int fx_set_sm_budget(int ctas) {
  const int old = fx::g_sm_budget;
  fx::g_sm_budget = ctas > 0 ? ctas : 0;
  return old;
}

}  // extern "C"
