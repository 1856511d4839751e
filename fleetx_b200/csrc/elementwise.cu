// elementwise.cu — K5: bias + dropout + residual forward / backward
// (hybrid_model.py:546-547 dropout1, :556-561 dropout2, mode "upscale_in_train") and the
// column sums that produce the bias gradients of the four linears. All HBM-bound, 16-byte
// accesses, one 8-element group per thread per step; the dropout mask is regenerated from
// Philox in backward instead of being stored. Algorithmic bytes per token (bf16): forward 6h,
// backward 4h, column sum 2n.
#include <atomic>

#include "common.cuh"
#include "philox.cuh"
#include "tc5.cuh"

namespace fx {
namespace ew {

using tc5::bf16_hi;
using tc5::bf16_lo;
using tc5::pack_bf16;

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                    pack_bf16(f[6], f[7]));
}

// keep[j] for the 8 elements of group `grp`
__device__ __forceinline__ void keep_flags(bool (&keep)[8], size_t grp, float p, uint32_t thresh,
                                           uint64_t seed, uint64_t offset,
                                           const uint8_t* __restrict__ mask_in) {
  if (mask_in != nullptr) {
    const uint2 m = __ldg(reinterpret_cast<const uint2*>(mask_in) + grp);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      keep[j] = ((m.x >> (8 * j)) & 0xFFu) != 0;
      keep[4 + j] = ((m.y >> (8 * j)) & 0xFFu) != 0;
    }
  } else if (p > 0.f) {
    const Philox8 r = philox4x32_10(seed, offset, grp);
#pragma unroll
    for (int j = 0; j < 8; ++j) keep[j] = philox_keep(r, j, thresh);
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) keep[j] = true;
  }
}

__global__ void __launch_bounds__(256)
bdr_fwd_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ bias,
               const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ out,
               size_t ngroups, int hvec, float p, float scale, uint32_t thresh, uint64_t seed,
               uint64_t offset, const uint8_t* __restrict__ mask_in, uint8_t* __restrict__ mask_out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t grp = (size_t)blockIdx.x * blockDim.x + threadIdx.x; grp < ngroups; grp += stride) {
    float v[8], r[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(y) + grp), v);
    unpack8(__ldg(reinterpret_cast<const uint4*>(residual) + grp), r);
    if (bias != nullptr) {
      float b[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(bias) + (grp % hvec)), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += b[j];
    }
    bool keep[8];
    keep_flags(keep, grp, p, thresh, seed, offset, mask_in);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = r[j] + (keep[j] ? v[j] * scale : 0.f);
    reinterpret_cast<uint4*>(out)[grp] = pack8(o);
    if (mask_out != nullptr) {
      uint2 m = make_uint2(0, 0);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        m.x |= (keep[j] ? 1u : 0u) << (8 * j);
        m.y |= (keep[4 + j] ? 1u : 0u) << (8 * j);
      }
      reinterpret_cast<uint2*>(mask_out)[grp] = m;
    }
  }
}

__global__ void __launch_bounds__(256)
bdr_bwd_kernel(const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dy,
               size_t ngroups, float p, float scale, uint32_t thresh, uint64_t seed,
               uint64_t offset, const uint8_t* __restrict__ mask_in) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t grp = (size_t)blockIdx.x * blockDim.x + threadIdx.x; grp < ngroups; grp += stride) {
    float v[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(dout) + grp), v);
    bool keep[8];
    keep_flags(keep, grp, p, thresh, seed, offset, mask_in);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = keep[j] ? v[j] * scale : 0.f;
    reinterpret_cast<uint4*>(dy)[grp] = pack8(o);
  }
}

// ------------------------------------------------------------------ column sum
// grid (ceil(n/256), splits); block 32 x 8. Thread (tx,ty) owns 8 columns and rows
// ty + 8*(blockIdx.y + k*splits). Partial sums go to part[blockIdx.y][n] (fp32).
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int ldx, float* __restrict__ part,
                      int rows, int n) {
  // block (32, 8): thread (tx, ty) owns 8 columns and every (8 * gridDim.y)-th row from blockIdx.y * 8 + ty on, four
  // 16-byte loads in flight; each thread stores its own partial row (no shared memory: a block fits beside a
  // resident GEMM CTA when the host runs the column sums next to the weight-gradient GEMM that reads the same matrix)
  const int col = (blockIdx.x * 32 + threadIdx.x) * 8;
  if (col >= n) return;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  const int step = gridDim.y * 8;
  int r = blockIdx.y * 8 + threadIdx.y;
  for (; r + 3 * step < rows; r += 4 * step) {
    uint4 q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) q[u] = __ldg(reinterpret_cast<const uint4*>(x + (size_t)(r + u * step) * ldx + col));
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float v[8];
      unpack8(q[u], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
  }
  for (; r < rows; r += step) {
    float v[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(x + (size_t)r * ldx + col)), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += v[j];
  }
  float4* o = reinterpret_cast<float4*>(part + (size_t)(blockIdx.y * 8 + threadIdx.y) * n + col);
  o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
  o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

__global__ void __launch_bounds__(256)
colsum_finish_kernel(const float* __restrict__ part, void* __restrict__ out, int out_f32,
                     int splits, int n) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  float s = 0.f;
  for (int i = 0; i < splits; ++i) s += part[(size_t)i * n + c];
  if (out_f32) static_cast<float*>(out)[c] = s;
  else static_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16_rn(s);
}

static int colsum_splits(int rows, int n) {
  const int col_blocks = ceil_div(n, 256);
  int splits = ceil_div(num_sms() * 4, col_blocks);
  const int max_splits = ceil_div(rows, 8);
  if (splits > max_splits) splits = max_splits;
  if (splits > 256) splits = 256;
  if (splits < 1) splits = 1;
  return splits;
}

}  // namespace ew
}  // namespace fx

extern "C" int fx_bias_dropout_residual_fwd(void* stream, const void* y, const void* bias,
                                            const void* residual, void* out, int rows, int h,
                                            float p, uint64_t seed, uint64_t offset,
                                            const uint8_t* mask_in, uint8_t* mask_out) {
  using namespace fx;
  FX_CHECK_ARG(rows > 0 && h > 0, "fx_bias_dropout_residual_fwd: empty input (rows %d, h %d)", rows, h);
  FX_CHECK_ARG(h % 8 == 0, "fx_bias_dropout_residual_fwd: h must be a multiple of 8 (got %d)", h);
  FX_CHECK_ARG(p >= 0.f && p < 1.f, "fx_bias_dropout_residual_fwd: dropout p must be in [0,1) (got %f)", p);
  FX_CHECK_ARG(y && residual && out, "fx_bias_dropout_residual_fwd: null pointer");
  FX_CHECK_ARG(((uintptr_t)y | (uintptr_t)bias | (uintptr_t)residual | (uintptr_t)out) % 16 == 0 &&
                   ((uintptr_t)mask_in | (uintptr_t)mask_out) % 8 == 0,
               "fx_bias_dropout_residual_fwd: misaligned pointer");
  const size_t ngroups = (size_t)rows * (h / 8);
  const size_t want = (ngroups + 255) / 256;
  const int grid = (int)(want < (size_t)num_sms() * 8 ? want : (size_t)num_sms() * 8);
  ew::bdr_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(bias),
      static_cast<const __nv_bfloat16*>(residual), static_cast<__nv_bfloat16*>(out), ngroups, h / 8, p,
      1.0f / (1.0f - p), dropout_threshold(p), seed, offset, mask_in, mask_out);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_bias_dropout_residual_bwd(void* stream, const void* dout, void* dy, int rows,
                                            int h, float p, uint64_t seed, uint64_t offset,
                                            const uint8_t* mask_in) {
  using namespace fx;
  FX_CHECK_ARG(rows > 0 && h > 0, "fx_bias_dropout_residual_bwd: empty input (rows %d, h %d)", rows, h);
  FX_CHECK_ARG(h % 8 == 0, "fx_bias_dropout_residual_bwd: h must be a multiple of 8 (got %d)", h);
  FX_CHECK_ARG(p >= 0.f && p < 1.f, "fx_bias_dropout_residual_bwd: dropout p must be in [0,1) (got %f)", p);
  FX_CHECK_ARG(dout && dy, "fx_bias_dropout_residual_bwd: null pointer");
  FX_CHECK_ARG(((uintptr_t)dout | (uintptr_t)dy) % 16 == 0 && (uintptr_t)mask_in % 8 == 0,
               "fx_bias_dropout_residual_bwd: misaligned pointer");
  const size_t ngroups = (size_t)rows * (h / 8);
  const size_t want = (ngroups + 255) / 256;
  const int grid = (int)(want < (size_t)num_sms() * 8 ? want : (size_t)num_sms() * 8);
  ew::bdr_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dout), static_cast<__nv_bfloat16*>(dy), ngroups, p,
      1.0f / (1.0f - p), dropout_threshold(p), seed, offset, mask_in);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" size_t fx_colsum_workspace_bytes(int rows, int n) {
  if (rows <= 0 || n <= 0) return 0;
  return (size_t)fx::ew::colsum_splits(rows, n) * 8 * (size_t)n * sizeof(float);
}

extern "C" int fx_colsum_bf16(void* stream, const void* x, int ldx, void* out, int out_dtype,
                              void* workspace, int rows, int n) {
  using namespace fx;
  FX_CHECK_ARG(rows > 0 && n > 0, "fx_colsum_bf16: empty input (rows %d, n %d)", rows, n);
  FX_CHECK_ARG(n % 8 == 0 && ldx % 8 == 0 && ldx >= n, "fx_colsum_bf16: n, ldx must be multiples of 8, ldx >= n");
  FX_CHECK_ARG(out_dtype == FX_BF16 || out_dtype == FX_F32, "fx_colsum_bf16: bad out_dtype %d", out_dtype);
  FX_CHECK_ARG(x && out && workspace, "fx_colsum_bf16: null pointer");
  FX_CHECK_ARG(((uintptr_t)x | (uintptr_t)workspace) % 16 == 0, "fx_colsum_bf16: misaligned pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int splits = ew::colsum_splits(rows, n);
  float* part = static_cast<float*>(workspace);
  static std::atomic<bool> carve1[64], carve2[64];
  FX_PREFER_MAX_SHARED(ew::colsum_partial_kernel, carve1);
  FX_PREFER_MAX_SHARED(ew::colsum_finish_kernel, carve2);
  ew::colsum_partial_kernel<<<dim3(ceil_div(n, 256), splits), dim3(32, 8), 0, st>>>(
      static_cast<const __nv_bfloat16*>(x), ldx, part, rows, n);
  FX_LAUNCH_CHECK();
  ew::colsum_finish_kernel<<<ceil_div(n, 256), 256, 0, st>>>(part, out, out_dtype == FX_F32, splits * 8, n);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
