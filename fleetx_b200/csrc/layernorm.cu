// layernorm.cu — K1: LayerNorm forward / backward (nn.LayerNorm(d_model, epsilon=1e-5),
// hybrid_model.py:513-514; applied at :534 and :554). HBM-bound: one CTA (128..1024 threads, by h) per
// row, the row lives in registers (16-byte loads), statistics by warp shuffle + a 4-entry
// shared-memory exchange. Algorithmic bytes per token (bf16): forward 4h (+8 stats),
// backward 6h (8h with the fused skip-path gradient).
#include <mutex>

#include "common.cuh"
#include "tc5.cuh"

namespace fx {
namespace ln {

using tc5::bf16_hi;
using tc5::bf16_lo;
using tc5::pack_bf16;

constexpr int MAX_PARTIAL_ROWS = 592;  // 148 SMs x 4: rows of the dgamma/dbeta partial buffer

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the CTA of NT threads. `red` holds NT/32 entries; safe to call back to back.
template <int NT>
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // previous readers of red[] are done
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int i = 1; i < NT / 32; ++i) r += red[i];
  return r;
}
template <int NT>
__device__ __forceinline__ float2 block_sum2(float a, float b, float2* red) {
  a = warp_sum(a);
  b = warp_sum(b);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = make_float2(a, b);
  __syncthreads();
  float2 r = red[0];
#pragma unroll
  for (int i = 1; i < NT / 32; ++i) {
    r.x += red[i].x;
    r.y += red[i].y;
  }
  return r;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                    pack_bf16(f[6], f[7]));
}

// VPT = 16-byte vectors per thread, NT = threads per CTA; NT * VPT * 8 >= h.
template <int VPT, int NT>
__global__ void __launch_bounds__(NT)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
              const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
              float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int h,
              float eps) {
  __shared__ float red[NT / 32];
  const int nvec = h >> 3;
  const float inv_h = 1.0f / static_cast<float>(h);

  // gamma/beta are row-invariant: keep them in registers across the row loop
  float g[VPT][8], bt[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * NT;
    if (v < nvec) {
      const float4* gp = reinterpret_cast<const float4*>(gamma + v * 8);
      const float4* bp = reinterpret_cast<const float4*>(beta + v * 8);
      const float4 g0 = __ldg(gp), g1 = __ldg(gp + 1), b0 = __ldg(bp), b1 = __ldg(bp + 1);
      g[i][0] = g0.x; g[i][1] = g0.y; g[i][2] = g0.z; g[i][3] = g0.w;
      g[i][4] = g1.x; g[i][5] = g1.y; g[i][6] = g1.z; g[i][7] = g1.w;
      bt[i][0] = b0.x; bt[i][1] = b0.y; bt[i][2] = b0.z; bt[i][3] = b0.w;
      bt[i][4] = b1.x; bt[i][5] = b1.y; bt[i][6] = b1.z; bt[i][7] = b1.w;
    }
  }

  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const uint4* xp = reinterpret_cast<const uint4*>(x + (size_t)row * h);
    float f[VPT][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        unpack8(__ldg(xp + v), f[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += f[i][j];
      }
    }
    const float mean = block_sum<NT>(s, red) * inv_h;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = f[i][j] - mean;
          ss += d * d;
        }
      }
    }
    const float var = block_sum<NT>(ss, red) * inv_h;
    const float rstd = rsqrtf(var + eps);
    if (threadIdx.x == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
    uint4* yp = reinterpret_cast<uint4*>(y + (size_t)row * h);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (f[i][j] - mean) * rstd * g[i][j] + bt[i][j];
        yp[v] = pack8(o);
      }
    }
  }
}

// Backward. Each CTA walks rows blockIdx.x, +gridDim.x, ... and keeps its share of the
// dgamma/dbeta sums in registers; they land in partial[blockIdx.x][{0,1}][h] at the end.
template <int VPT, int NT>
__global__ void __launch_bounds__(NT)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
              const float* __restrict__ gamma, const float* __restrict__ mean,
              const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres,
              __nv_bfloat16* __restrict__ dx, float* __restrict__ partial, int rows, int h) {
  __shared__ float2 red[NT / 32];
  const int nvec = h >> 3;
  const float inv_h = 1.0f / static_cast<float>(h);

  float g[VPT][8], dg[VPT][8], db[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * NT;
#pragma unroll
    for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = 0.f;
    if (v < nvec) {
      const float4* gp = reinterpret_cast<const float4*>(gamma + v * 8);
      const float4 g0 = __ldg(gp), g1 = __ldg(gp + 1);
      g[i][0] = g0.x; g[i][1] = g0.y; g[i][2] = g0.z; g[i][3] = g0.w;
      g[i][4] = g1.x; g[i][5] = g1.y; g[i][6] = g1.z; g[i][7] = g1.w;
    }
  }

  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const uint4* xp = reinterpret_cast<const uint4*>(x + (size_t)row * h);
    const uint4* dyp = reinterpret_cast<const uint4*>(dy + (size_t)row * h);
    const float mu = __ldg(mean + row), rs = __ldg(rstd + row);
    float xh[VPT][8], gy[VPT][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        float xv[8], dv[8];
        unpack8(__ldg(xp + v), xv);
        unpack8(__ldg(dyp + v), dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mu) * rs;
          gy[i][j] = dv[j] * g[i][j];
          s1 += gy[i][j];
          s2 += gy[i][j] * xh[i][j];
          dg[i][j] += dv[j] * xh[i][j];
          db[i][j] += dv[j];
        }
      }
    }
    const float2 s = block_sum2<NT>(s1, s2, red);
    const float c1 = s.x * inv_h, c2 = s.y * inv_h;
    uint4* dxp = reinterpret_cast<uint4*>(dx + (size_t)row * h);
    const uint4* drp = reinterpret_cast<const uint4*>(dres ? dres + (size_t)row * h : nullptr);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (gy[i][j] - c1 - xh[i][j] * c2);
        if (dres != nullptr) {
          float r[8];
          unpack8(__ldg(drp + v), r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxp[v] = pack8(o);
      }
    }
  }

  float* pg = partial + (size_t)blockIdx.x * 2 * h;
  float* pb = pg + h;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * NT;
    if (v < nvec) {
      float4* a = reinterpret_cast<float4*>(pg + v * 8);
      float4* b = reinterpret_cast<float4*>(pb + v * 8);
      a[0] = make_float4(dg[i][0], dg[i][1], dg[i][2], dg[i][3]);
      a[1] = make_float4(dg[i][4], dg[i][5], dg[i][6], dg[i][7]);
      b[0] = make_float4(db[i][0], db[i][1], db[i][2], db[i][3]);
      b[1] = make_float4(db[i][4], db[i][5], db[i][6], db[i][7]);
    }
  }
}

// out[c] = sum_r partial[r][c]; `width` = 2h covers dgamma and dbeta in one pass.
// 32 x 8 threads: x walks columns (coalesced), y splits the partial rows.
__global__ void __launch_bounds__(256)
ln_bwd_finish_kernel(const float* __restrict__ partial, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, int nparts, int h) {
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int width = 2 * h;
  float s = 0.f;
  if (c < width)
    for (int r = threadIdx.y; r < nparts; r += 8) s += partial[(size_t)r * width + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < width) {
#pragma unroll
    for (int i = 1; i < 8; ++i) s += red[i][threadIdx.x];
    if (c < h) dgamma[c] = s;
    else dbeta[c - h] = s;
  }
}

// CTA width grows with h so that a thread never holds more than two 16-byte vectors of a row
// (the backward keeps 5 row-sized arrays live in registers).
static int pick_threads(int h) {
  int nt = 128;
  while (nt < 1024 && nt * 16 < h) nt <<= 1;
  return nt;
}

#define FX_LN_DISPATCH(KERNEL, ...)                                                        \
  do {                                                                                     \
    const int nt__ = pick_threads(h);                                                      \
    const int vpt__ = ceil_div(h / 8, nt__);                                               \
    if (nt__ == 128 && vpt__ == 1) KERNEL<1, 128><<<grid, 128, 0, st>>>(__VA_ARGS__);      \
    else if (nt__ == 128) KERNEL<2, 128><<<grid, 128, 0, st>>>(__VA_ARGS__);               \
    else if (nt__ == 256) KERNEL<2, 256><<<grid, 256, 0, st>>>(__VA_ARGS__);               \
    else if (nt__ == 512) KERNEL<2, 512><<<grid, 512, 0, st>>>(__VA_ARGS__);               \
    else KERNEL<2, 1024><<<grid, 1024, 0, st>>>(__VA_ARGS__);                              \
  } while (0)
static int bwd_grid(int rows) { return rows < MAX_PARTIAL_ROWS ? rows : MAX_PARTIAL_ROWS; }

}  // namespace ln
}  // namespace fx

extern "C" int fx_layernorm_fwd(void* stream, const void* x, const float* gamma, const float* beta,
                                void* y, float* mean, float* rstd, int rows, int h, float eps) {
  using namespace fx;
  FX_CHECK_ARG(rows > 0 && h > 0, "fx_layernorm_fwd: empty input (rows %d, h %d)", rows, h);
  FX_CHECK_ARG(h % 8 == 0 && h <= 16384, "fx_layernorm_fwd: h must be a multiple of 8 and <= 16384 (got %d)", h);
  FX_CHECK_ARG(x && gamma && beta && y && mean && rstd, "fx_layernorm_fwd: null pointer");
  FX_CHECK_ARG(((uintptr_t)x | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta) % 16 == 0,
               "fx_layernorm_fwd: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = rows < num_sms() * 16 ? rows : num_sms() * 16;
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto Y = static_cast<__nv_bfloat16*>(y);
  using namespace fx::ln;
  FX_LN_DISPATCH(ln_fwd_kernel, X, gamma, beta, Y, mean, rstd, rows, h, eps);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" size_t fx_layernorm_bwd_workspace_bytes(int rows, int h) {
  if (rows <= 0 || h <= 0) return 0;
  return (size_t)fx::ln::bwd_grid(rows) * 2 * (size_t)h * sizeof(float);
}

extern "C" int fx_layernorm_bwd(void* stream, const void* dy, const void* x, const float* gamma,
                                const float* mean, const float* rstd, const void* dresidual,
                                void* dx, float* dgamma, float* dbeta, void* workspace, int rows,
                                int h) {
  using namespace fx;
  FX_CHECK_ARG(rows > 0 && h > 0, "fx_layernorm_bwd: empty input (rows %d, h %d)", rows, h);
  FX_CHECK_ARG(h % 8 == 0 && h <= 16384, "fx_layernorm_bwd: h must be a multiple of 8 and <= 16384 (got %d)", h);
  FX_CHECK_ARG(dy && x && gamma && mean && rstd && dx && dgamma && dbeta && workspace,
               "fx_layernorm_bwd: null pointer");
  FX_CHECK_ARG(((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dx | (uintptr_t)dresidual | (uintptr_t)gamma |
                (uintptr_t)workspace) % 16 == 0,
               "fx_layernorm_bwd: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = ln::bwd_grid(rows);
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto DR = static_cast<const __nv_bfloat16*>(dresidual);
  auto DX = static_cast<__nv_bfloat16*>(dx);
  float* part = static_cast<float*>(workspace);
  using namespace fx::ln;
  FX_LN_DISPATCH(ln_bwd_kernel, DY, X, gamma, mean, rstd, DR, DX, part, rows, h);
  FX_LAUNCH_CHECK();
  ln::ln_bwd_finish_kernel<<<ceil_div(2 * h, 32), dim3(32, 8), 0, st>>>(part, dgamma, dbeta, grid, h);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

// ------------------------------------------------------------------------------------------
// LayerNorm backward for callers that may not allocate device memory.
//
// Who needs this: hosts that own all GPU memory up front and forbid allocation afterwards — a
// serving or training runtime whose executor planned every buffer before the step started, a
// Paddle/static-graph custom op that is handed exactly its declared inputs and outputs and no
// allocator, code that runs where cudaMalloc would stall the whole device (it synchronises) or
// is simply not permitted, e.g. inside a latency-critical loop. fx_layernorm_bwd needs a partial-
// sum workspace from such a caller; this variant brings its own instead: a fixed array in the
// library's device image. Every GPU that loads the library carries it until the process exits.
//
// Limits: h <= FX_LN_NOALLOC_MAX_H (4096); any number of rows. The array is 148 x 2 x 4096 fp32
// = 4,849,664 bytes. The backward kernel runs on at most 148 CTAs here (the normal entry point
// uses up to 592), so it is somewhat slower; results are the same up to fp32 summation order.
namespace fx {
namespace ln {

constexpr int NOALLOC_PARTS = 148;
constexpr int NOALLOC_MAX_H = 4096;

// File-scope device array. In HOST code this name is only a handle for the runtime, not a device
// address: it must go through cudaGetSymbolAddress before a kernel can be given it.
static __device__ __align__(256) float g_ln_bwd_noalloc_scratch[NOALLOC_PARTS * 2 * NOALLOC_MAX_H];

static std::mutex g_noalloc_mu;
static float* g_noalloc_ptr[64] = {nullptr};  // per device, resolved once with that device current

// Returns the device address of the scratch array on the CURRENT device (nullptr + error text on failure).
static float* noalloc_scratch() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("fx_layernorm_bwd_noalloc: cannot identify the current CUDA device");
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_ln_bwd_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("fx_layernorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                cudaGetErrorString(e));
      return nullptr;
    }
    g_noalloc_ptr[dev] = ptr;
  }
  return g_noalloc_ptr[dev];
}

}  // namespace ln
}  // namespace fx

extern "C" int fx_layernorm_bwd_noalloc(void* stream, const void* dy, const void* x, const float* gamma,
                                        const float* mean, const float* rstd, const void* dresidual,
                                        void* dx, float* dgamma, float* dbeta, int rows, int h) {
  using namespace fx;
  FX_CHECK_ARG(rows > 0 && h > 0, "fx_layernorm_bwd_noalloc: empty input (rows %d, h %d)", rows, h);
  FX_CHECK_ARG(h % 8 == 0, "fx_layernorm_bwd_noalloc: h must be a multiple of 8 (got %d)", h);
  FX_CHECK_ARG(h <= ln::NOALLOC_MAX_H,
               "fx_layernorm_bwd_noalloc: h = %d exceeds the built-in scratch (max h %d); use fx_layernorm_bwd "
               "with a caller-provided workspace", h, ln::NOALLOC_MAX_H);
  FX_CHECK_ARG(dy && x && gamma && mean && rstd && dx && dgamma && dbeta, "fx_layernorm_bwd_noalloc: null pointer");
  FX_CHECK_ARG(((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dx | (uintptr_t)dresidual | (uintptr_t)gamma) % 16 == 0,
               "fx_layernorm_bwd_noalloc: pointers must be 16-byte aligned");
  float* part = ln::noalloc_scratch();  // first call on a device resolves the address (no allocation)
  if (part == nullptr) return FX_ERR_CUDA;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = rows < ln::NOALLOC_PARTS ? rows : ln::NOALLOC_PARTS;
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto DR = static_cast<const __nv_bfloat16*>(dresidual);
  auto DX = static_cast<__nv_bfloat16*>(dx);
  using namespace fx::ln;
  // Nothing in the scratch has to start at zero: the first kernel overwrites partial[0..grid) x 2h in
  // full and the second reads exactly that range, so whatever an earlier call left behind is never seen.
  FX_LN_DISPATCH(ln_bwd_kernel, DY, X, gamma, mean, rstd, DR, DX, part, rows, h);
  FX_LAUNCH_CHECK();
  ln::ln_bwd_finish_kernel<<<ceil_div(2 * h, 32), dim3(32, 8), 0, st>>>(part, dgamma, dbeta, grid, h);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
