// kvcache.cu — incremental (KV-cache) self-attention of the decoder layer in inference:
// MultiHeadAttention with use_cache=True, hybrid_model.py:165-186 (_fuse_prepare_qkv: k, v of the new tokens
// are concatenated behind cache.k / cache.v, [b, nh, L, hd]) and :268-298 (core_attn in eval: no dropout).
//
// Two kernels. kv_append copies the k | v thirds of the packed qkv_proj output of the t new tokens into the
// cache buffers at positions [L, L + t). attn_decode then computes, for each new token i, softmax over cache
// positions [0, L + i] — HBM-bound (the whole K and V of the head stream past once per new token): one CTA per
// (head, batch, new token); each thread walks its share of the cache rows with an online softmax and a private
// hd-wide accumulator, then the 128 partial states merge through shared memory.
#include "common.cuh"
#include "tc5.cuh"

namespace fx {
namespace kvc {

using tc5::bf16_hi;
using tc5::bf16_lo;
using tc5::pack_bf16;

// qkv [b, t, nh, 3, hd] -> cache_k / cache_v [b, nh, cap, hd] rows L .. L+t-1
__global__ void __launch_bounds__(256)
kv_append_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ ck, __nv_bfloat16* __restrict__ cv, int b,
                 int t, int nh, int hd, int cap, int L) {
  const int hv = hd >> 3;
  const size_t total = static_cast<size_t>(b) * t * nh * hv;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < total;
       v += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(v % hv);
    size_t r = v / hv;
    const int head = static_cast<int>(r % nh);
    r /= nh;
    const int tok = static_cast<int>(r % t), batch = static_cast<int>(r / t);
    const uint4* src = reinterpret_cast<const uint4*>(qkv + ((static_cast<size_t>(batch) * t + tok) * nh + head) * 3 * hd);
    const size_t dst = ((static_cast<size_t>(batch) * nh + head) * cap + L + tok) * hv + c;
    reinterpret_cast<uint4*>(ck)[dst] = __ldg(src + hv + c);
    reinterpret_cast<uint4*>(cv)[dst] = __ldg(src + 2 * hv + c);
  }
}

template <int HD>
__global__ void __launch_bounds__(128)
attn_decode_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ ck,
                   const __nv_bfloat16* __restrict__ cv, __nv_bfloat16* __restrict__ out, int t, int nh, int cap, int L,
                   float scale_log2e) {
  constexpr int HV = HD / 8;
  __shared__ float q_s[HD];
  __shared__ float m_s[128], l_s[128];
  __shared__ float acc_s[32][HD + 1];
  const int head = blockIdx.x, batch = blockIdx.y, tok = blockIdx.z;
  const int tid = threadIdx.x;
  const int n_kv = L + tok + 1;  // the new token attends to the cache and to the new tokens up to itself
  const __nv_bfloat16* qp = qkv + ((static_cast<size_t>(batch) * t + tok) * nh + head) * 3 * HD;
  if (tid < HD) q_s[tid] = __bfloat162float(qp[tid]) * scale_log2e;
  __syncthreads();
  const uint4* kp = reinterpret_cast<const uint4*>(ck + (static_cast<size_t>(batch) * nh + head) * cap * HD);
  const uint4* vp = reinterpret_cast<const uint4*>(cv + (static_cast<size_t>(batch) * nh + head) * cap * HD);
  float m = -INFINITY, l = 0.f;
  float acc[HD];
#pragma unroll
  for (int i = 0; i < HD; ++i) acc[i] = 0.f;
  for (int j = tid; j < n_kv; j += 128) {
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < HV; ++c) {
      const uint4 u = __ldg(kp + static_cast<size_t>(j) * HV + c);
      s += q_s[8 * c + 0] * bf16_lo(u.x) + q_s[8 * c + 1] * bf16_hi(u.x) + q_s[8 * c + 2] * bf16_lo(u.y) +
           q_s[8 * c + 3] * bf16_hi(u.y) + q_s[8 * c + 4] * bf16_lo(u.z) + q_s[8 * c + 5] * bf16_hi(u.z) +
           q_s[8 * c + 6] * bf16_lo(u.w) + q_s[8 * c + 7] * bf16_hi(u.w);
    }
    const float mn = fmaxf(m, s);
    const float alpha = exp2f(m - mn), p = exp2f(s - mn);
    l = l * alpha + p;
#pragma unroll
    for (int c = 0; c < HV; ++c) {
      const uint4 u = __ldg(vp + static_cast<size_t>(j) * HV + c);
      acc[8 * c + 0] = acc[8 * c + 0] * alpha + p * bf16_lo(u.x);
      acc[8 * c + 1] = acc[8 * c + 1] * alpha + p * bf16_hi(u.x);
      acc[8 * c + 2] = acc[8 * c + 2] * alpha + p * bf16_lo(u.y);
      acc[8 * c + 3] = acc[8 * c + 3] * alpha + p * bf16_hi(u.y);
      acc[8 * c + 4] = acc[8 * c + 4] * alpha + p * bf16_lo(u.z);
      acc[8 * c + 5] = acc[8 * c + 5] * alpha + p * bf16_hi(u.z);
      acc[8 * c + 6] = acc[8 * c + 6] * alpha + p * bf16_lo(u.w);
      acc[8 * c + 7] = acc[8 * c + 7] * alpha + p * bf16_hi(u.w);
    }
    m = mn;
  }
  // merge the 128 (m, l, acc) states: global max first, then rescaled sums, 32 threads' accumulators at a time
  m_s[tid] = m;
  __syncthreads();
  float gm = -INFINITY;
  for (int i = 0; i < 128; ++i) gm = fmaxf(gm, m_s[i]);
  const float w = m == -INFINITY ? 0.f : exp2f(m - gm);
  l_s[tid] = l * w;
  float total[(HD + 127) / 128];
#pragma unroll
  for (int k = 0; k < (HD + 127) / 128; ++k) total[k] = 0.f;
  for (int g = 0; g < 4; ++g) {
    __syncthreads();
    if ((tid >> 5) == g) {
#pragma unroll
      for (int i = 0; i < HD; ++i) acc_s[tid & 31][i] = acc[i] * w;
    }
    __syncthreads();
    for (int col = tid; col < HD; col += 128) {
      float sum = 0.f;
#pragma unroll 8
      for (int r = 0; r < 32; ++r) sum += acc_s[r][col];
      total[(col - tid) / 128] += sum;
    }
  }
  __syncthreads();
  float lsum = 0.f;
  for (int i = 0; i < 128; ++i) lsum += l_s[i];
  __nv_bfloat16* op = out + (static_cast<size_t>(batch) * t + tok) * nh * HD + static_cast<size_t>(head) * HD;
  for (int col = tid; col < HD; col += 128) op[col] = __float2bfloat16_rn(total[(col - tid) / 128] / lsum);
}

}  // namespace kvc
}  // namespace fx

extern "C" int fx_kv_cache_append(void* stream, const void* qkv, void* cache_k, void* cache_v, int b, int t, int nh, int hd,
                                  int cap, int L) {
  using namespace fx;
  FX_CHECK_ARG(b > 0 && t > 0 && nh > 0 && (hd == 64 || hd == 128), "fx_kv_cache_append: bad shape (b %d, t %d, nh %d, hd %d)", b, t, nh, hd);
  FX_CHECK_ARG(L >= 0 && L + t <= cap, "fx_kv_cache_append: %d cached + %d new tokens exceed the capacity %d", L, t, cap);
  FX_CHECK_ARG(qkv && cache_k && cache_v, "fx_kv_cache_append: null pointer");
  FX_CHECK_ARG(((uintptr_t)qkv | (uintptr_t)cache_k | (uintptr_t)cache_v) % 16 == 0, "fx_kv_cache_append: pointers must be 16-byte aligned");
  const size_t total = static_cast<size_t>(b) * t * nh * (hd / 8);
  const int grid = static_cast<int>((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  kvc::kv_append_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(qkv), static_cast<__nv_bfloat16*>(cache_k), static_cast<__nv_bfloat16*>(cache_v), b, t, nh, hd, cap, L);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_attn_decode(void* stream, const void* qkv, const void* cache_k, const void* cache_v, void* out, int b, int t,
                              int nh, int hd, int cap, int L, float scale) {
  using namespace fx;
  FX_CHECK_ARG(b > 0 && t > 0 && nh > 0 && (hd == 64 || hd == 128), "fx_attn_decode: bad shape (b %d, t %d, nh %d, hd %d)", b, t, nh, hd);
  FX_CHECK_ARG(L >= 0 && L + t <= cap, "fx_attn_decode: %d cached + %d new tokens exceed the capacity %d", L, t, cap);
  FX_CHECK_ARG(b <= 65535 && t <= 65535, "fx_attn_decode: b and t must be <= 65535");
  FX_CHECK_ARG(qkv && cache_k && cache_v && out && scale > 0.f, "fx_attn_decode: bad arguments");
  FX_CHECK_ARG(((uintptr_t)qkv | (uintptr_t)cache_k | (uintptr_t)cache_v | (uintptr_t)out) % 16 == 0, "fx_attn_decode: pointers must be 16-byte aligned");
  const dim3 grid(nh, b, t);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float sl2 = scale * 1.4426950408889634f;
  if (hd == 64)
    kvc::attn_decode_kernel<64><<<grid, 128, 0, st>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(cache_k),
                                                      static_cast<const __nv_bfloat16*>(cache_v), static_cast<__nv_bfloat16*>(out), t, nh, cap, L, sl2);
  else
    kvc::attn_decode_kernel<128><<<grid, 128, 0, st>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(cache_k),
                                                       static_cast<const __nv_bfloat16*>(cache_v), static_cast<__nv_bfloat16*>(out), t, nh, cap, L, sl2);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
