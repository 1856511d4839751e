// kvcache.cu — incremental (KV-cache) self-attention of the decoder layer in inference:
// MultiHeadAttention with use_cache=True, hybrid_model.py:165-186 (_fuse_prepare_qkv: k, v of the new tokens
// are concatenated behind cache.k / cache.v, [b, nh, L, hd]) and :268-298 (core_attn in eval: no dropout).
//
// Two kernels. kv_append copies the k | v thirds of the packed qkv_proj output of the t new tokens into the
// cache buffers at positions [L, L + t). attn_decode then computes, for each new token i, softmax over cache
// positions [0, L + i] — HBM-bound (the whole K and V of the head stream past once per new token); see the
// kernel for how the rows are spread over lanes.
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

// One CTA (8 warps) per (head, batch, new token). A cache row (hd bf16) is read by hd/8 adjacent lanes, 16
// bytes each, so a warp's load covers 32/(hd/8) whole rows, contiguous in memory; the row's score is summed
// over those lanes with shuffles, and each lane keeps the online-softmax state (m, l) of its row slot and an
// 8-wide slice of the output accumulator — few registers, so many warps and many loads are in flight. UNR
// rows per slot are loaded before the first use. At the end the row slots merge inside the warp (shuffles),
// the warps through shared memory.
template <int HD>
__global__ void __launch_bounds__(256)
attn_decode_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ ck,
                   const __nv_bfloat16* __restrict__ cv, __nv_bfloat16* __restrict__ out, int t, int nh, int cap, int L,
                   float scale_log2e) {
  constexpr int HV = HD / 8;        // lanes per row
  constexpr int RPW = 32 / HV;      // rows per warp per load
  constexpr int NW = 8, UNR = 4;
  __shared__ float m_s[NW], l_s[NW];
  __shared__ float acc_s[NW][HD];
  const int head = blockIdx.x, batch = blockIdx.y, tok = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane % HV, slot = lane / HV;
  const int n_kv = L + tok + 1;  // the new token attends to the cache and to the new tokens up to itself
  const __nv_bfloat16* qp = qkv + ((static_cast<size_t>(batch) * t + tok) * nh + head) * 3 * HD + sub * 8;
  float q[8];
  {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(qp));
    q[0] = bf16_lo(u.x); q[1] = bf16_hi(u.x); q[2] = bf16_lo(u.y); q[3] = bf16_hi(u.y);
    q[4] = bf16_lo(u.z); q[5] = bf16_hi(u.z); q[6] = bf16_lo(u.w); q[7] = bf16_hi(u.w);
#pragma unroll
    for (int i = 0; i < 8; ++i) q[i] *= scale_log2e;
  }
  const uint4* kp = reinterpret_cast<const uint4*>(ck + (static_cast<size_t>(batch) * nh + head) * cap * HD) + sub;
  const uint4* vp = reinterpret_cast<const uint4*>(cv + (static_cast<size_t>(batch) * nh + head) * cap * HD) + sub;
  float m = -INFINITY, l = 0.f;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int step = NW * RPW;  // rows per CTA per load
  // the trip count is the same for all lanes of a warp (the shuffles below need every lane): row slots that
  // run past the end load a clamped row and give it the score -inf
  for (int jb = warp * RPW; jb < n_kv; jb += step * UNR) {
    const int j0 = jb + slot;
    uint4 kr[UNR], vr[UNR];
#pragma unroll
    for (int r = 0; r < UNR; ++r) {
      const int j = min(j0 + r * step, n_kv - 1);
      kr[r] = __ldg(kp + static_cast<size_t>(j) * HV);
      vr[r] = __ldg(vp + static_cast<size_t>(j) * HV);
    }
    float sc[UNR];
    float mn = m;
#pragma unroll
    for (int r = 0; r < UNR; ++r) {
      const uint4 u = kr[r];
      float d = q[0] * bf16_lo(u.x) + q[1] * bf16_hi(u.x) + q[2] * bf16_lo(u.y) + q[3] * bf16_hi(u.y) +
                q[4] * bf16_lo(u.z) + q[5] * bf16_hi(u.z) + q[6] * bf16_lo(u.w) + q[7] * bf16_hi(u.w);
#pragma unroll
      for (int o = 1; o < HV; o <<= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
      sc[r] = (j0 + r * step < n_kv) ? d : -INFINITY;
      mn = fmaxf(mn, sc[r]);
    }
    const float alpha = mn == -INFINITY ? 1.f : exp2f(m - mn);  // a slot may have seen no row yet
    l *= alpha;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= alpha;
#pragma unroll
    for (int r = 0; r < UNR; ++r) {
      const float p = sc[r] == -INFINITY ? 0.f : exp2f(sc[r] - mn);  // 0 for rows past the end
      const uint4 u = vr[r];
      l += p;
      acc[0] += p * bf16_lo(u.x); acc[1] += p * bf16_hi(u.x); acc[2] += p * bf16_lo(u.y); acc[3] += p * bf16_hi(u.y);
      acc[4] += p * bf16_lo(u.z); acc[5] += p * bf16_hi(u.z); acc[6] += p * bf16_lo(u.w); acc[7] += p * bf16_hi(u.w);
    }
    m = mn;
  }
  // merge the row slots of the warp (lanes with the same `sub`), then the warps
#pragma unroll
  for (int o = HV; o < 32; o <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), l2 = __shfl_xor_sync(0xffffffffu, l, o);
    const float mn = fmaxf(m, m2);
    const float w1 = m == -INFINITY ? 0.f : exp2f(m - mn), w2 = m2 == -INFINITY ? 0.f : exp2f(m2 - mn);
    l = l * w1 + l2 * w2;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = acc[i] * w1 + __shfl_xor_sync(0xffffffffu, acc[i], o) * w2;
    m = mn;
  }
  if (slot == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) acc_s[warp][sub * 8 + i] = acc[i];
    if (sub == 0) { m_s[warp] = m; l_s[warp] = l; }
  }
  __syncthreads();
  if (threadIdx.x < HD) {
    float gm = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w) gm = fmaxf(gm, m_s[w]);
    float tot = 0.f, lsum = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float sw = m_s[w] == -INFINITY ? 0.f : exp2f(m_s[w] - gm);
      tot += acc_s[w][threadIdx.x] * sw;
      lsum += l_s[w] * sw;
    }
    out[(static_cast<size_t>(batch) * t + tok) * nh * HD + static_cast<size_t>(head) * HD + threadIdx.x] =
        __float2bfloat16_rn(tot / lsum);
  }
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
    kvc::attn_decode_kernel<64><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(cache_k),
                                                      static_cast<const __nv_bfloat16*>(cache_v), static_cast<__nv_bfloat16*>(out), t, nh, cap, L, sl2);
  else
    kvc::attn_decode_kernel<128><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(cache_k),
                                                       static_cast<const __nv_bfloat16*>(cache_v), static_cast<__nv_bfloat16*>(out), t, nh, cap, L, sl2);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
