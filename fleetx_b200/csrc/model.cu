// model.cu — the pieces of GPTModelHybrid / GPTForPretrainingHybrid around the decoder stack
// (SURVEY §8f #2): GPTEmbeddings (hybrid_model.py:574-622: vocab-parallel word embedding + position
// embedding + dropout), and the vocab-parallel softmax cross entropy of GPTPretrainingCriterionHybird
// (:791-832: fleet.meta_parallel.ParallelCrossEntropy / CrossEntropyLoss(reduction="none"), then the
// loss-mask weighted mean). The logits themselves are a plain fx_gemm_bf16 against the word embedding
// (parallel_matmul, :45-66). All HBM-bound, 16-byte accesses; one token row per CTA / warp group.
#include "common.cuh"
#include "philox.cuh"
#include "tc5.cuh"

namespace fx {
namespace model {

using tc5::bf16_hi;
using tc5::bf16_lo;
using tc5::pack_bf16;

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// out[t, :] = dropout(word[t, :] + wpe[pos[t], :]),  word[t, :] = word_in[t, :] if given (the allreduced
// partial lookups of the mp ranks), else wte[ids[t] - vocab_start, :] (zeros when the id belongs to another
// rank: VocabParallelEmbedding). wpe == null: no position term (the first pass of the mp > 1 path).
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ pos_ids,
                     const __nv_bfloat16* __restrict__ wte, int64_t vocab_start, int64_t vocab_rows,
                     const __nv_bfloat16* __restrict__ word_in, const __nv_bfloat16* __restrict__ wpe,
                     __nv_bfloat16* __restrict__ out, size_t T, int s, int hvec, float scale, uint32_t thresh,
                     uint64_t seed, uint64_t offset) {
  const size_t total = T * hvec;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < total; v += stride) {
    const size_t t = v / hvec;
    const int c = static_cast<int>(v - t * hvec);
    float x[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (word_in != nullptr) {
      unpack8(__ldg(reinterpret_cast<const uint4*>(word_in) + v), x);
    } else {
      const int64_t row = __ldg(ids + t) - vocab_start;
      if (row >= 0 && row < vocab_rows) unpack8(__ldg(reinterpret_cast<const uint4*>(wte) + row * hvec + c), x);
    }
    if (wpe != nullptr) {
      const int64_t pr = pos_ids != nullptr ? __ldg(pos_ids + t) : static_cast<int64_t>(t % s);
      float pe[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(wpe) + pr * hvec + c), pe);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] += pe[j];
    }
    if (thresh != 0) {
      const Philox8 r = philox4x32_10(seed, offset, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = philox_keep(r, j, thresh) ? x[j] * scale : 0.f;
    }
    reinterpret_cast<uint4*>(out)[v] = pack8(x);
  }
}

// d = dropout'(dout); dwte[ids[t] - vocab_start, :] += d[t, :] (this rank's rows only), dwpe[pos[t], :] += d[t, :].
// fp32 accumulators, vector reductions (red.global.add.v4.f32): rows repeat (a token id many times, every
// position once per batch entry), and the word accumulator already holds the logits GEMM's share of the
// tied weight's gradient.
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const __nv_bfloat16* __restrict__ dout, const int64_t* __restrict__ ids,
                     const int64_t* __restrict__ pos_ids, float* __restrict__ dwte, int64_t vocab_start,
                     int64_t vocab_rows, float* __restrict__ dwpe, size_t T, int s, int hvec, float scale,
                     uint32_t thresh, uint64_t seed, uint64_t offset) {
  const size_t total = T * hvec;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < total; v += stride) {
    const size_t t = v / hvec;
    const int c = static_cast<int>(v - t * hvec);
    float d[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(dout) + v), d);
    if (thresh != 0) {
      const Philox8 r = philox4x32_10(seed, offset, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = philox_keep(r, j, thresh) ? d[j] * scale : 0.f;
    }
    const int64_t row = __ldg(ids + t) - vocab_start;
    if (dwte != nullptr && row >= 0 && row < vocab_rows) {
      float* a = dwte + (row * hvec + c) * 8;
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "f"(d[0]), "f"(d[1]), "f"(d[2]), "f"(d[3]) : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a + 4), "f"(d[4]), "f"(d[5]), "f"(d[6]), "f"(d[7]) : "memory");
    }
    if (dwpe != nullptr) {
      const int64_t pr = pos_ids != nullptr ? __ldg(pos_ids + t) : static_cast<int64_t>(t % s);
      float* a = dwpe + (pr * hvec + c) * 8;
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "f"(d[0]), "f"(d[1]), "f"(d[2]), "f"(d[3]) : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a + 4), "f"(d[4]), "f"(d[5]), "f"(d[6]), "f"(d[7]) : "memory");
    }
  }
}

__global__ void __launch_bounds__(256)
f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t nvec) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * v), b = __ldg(reinterpret_cast<const float4*>(src) + 2 * v + 1);
    reinterpret_cast<uint4*>(dst)[v] = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
  }
}

static int grid_for(size_t items, int per_block) {
  const size_t want = (items + per_block - 1) / per_block;
  const size_t cap = static_cast<size_t>(num_sms()) * 16;
  return static_cast<int>(want < cap ? (want ? want : 1) : cap);
}

// ---------------------------------------------------------------------------- softmax cross entropy
// logits bf16 [T, Vl]: this rank's vocabulary slice [vocab_start, vocab_start + Vl) (Vl = V on one rank).
// Pass 1, one CTA per token row, one sweep with an online (max, sum) per thread:
//   lmax[t] = max_c l[t,c],  lsum[t] = sum_c exp(l[t,c] - lmax[t]),  target[t] = l[t, label - vocab_start] or 0
// With mp > 1 the host takes gmax = allreduce_max(lmax), rescales lsum by exp(lmax - gmax) (ce_rescale) and
// allreduce-sums lsum and target — ParallelCrossEntropy's three collectives on [T] vectors.
__global__ void __launch_bounds__(256)
ce_stats_kernel(const __nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels, int64_t vocab_start,
                int vl, float* __restrict__ lmax, float* __restrict__ lsum, float* __restrict__ target) {
  __shared__ float sm[8], ss[8];
  const size_t t = blockIdx.x;
  const uint4* row = reinterpret_cast<const uint4*>(logits + t * vl);
  const int nvec = vl >> 3;
  float m = -INFINITY, s = 0.f;
  for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
    float x[8];
    unpack8(__ldg(row + v), x);
    float mx = x[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) mx = fmaxf(mx, x[j]);
    const float mn = fmaxf(m, mx);
    float acc = s * __expf(m - mn);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += __expf(x[j] - mn);
    m = mn;
    s = acc;
  }
  // combine (m, s) pairs: warp, then block
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    const float mn = fmaxf(m, m2);
    s = (m == -INFINITY ? 0.f : s * __expf(m - mn)) + (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mn));
    m = mn;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { sm[warp] = m; ss[warp] = s; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float M = sm[0], S = ss[0];
    for (int w = 1; w < static_cast<int>(blockDim.x >> 5); ++w) {
      const float mn = fmaxf(M, sm[w]);
      S = (M == -INFINITY ? 0.f : S * __expf(M - mn)) + (sm[w] == -INFINITY ? 0.f : ss[w] * __expf(sm[w] - mn));
      M = mn;
    }
    lmax[t] = M;
    lsum[t] = S;
    const int64_t col = __ldg(labels + t) - vocab_start;
    target[t] = (col >= 0 && col < vl) ? __bfloat162float(logits[t * vl + col]) : 0.f;
  }
}

__global__ void ce_rescale_kernel(const float* __restrict__ lmax, const float* __restrict__ gmax, float* __restrict__ lsum,
                                  size_t T) {
  const size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t < T) lsum[t] *= __expf(lmax[t] - gmax[t]);
}

// loss[t] = log(gsum[t]) + gmax[t] - target[t]   (CrossEntropyLoss(reduction="none"), hybrid_model.py:798-799)
// out[0] = sum_t loss[t] * mask[t] / sum_t mask[t]  (:826-828), out[1] = sum_t mask[t]. One block.
__global__ void __launch_bounds__(1024)
ce_loss_kernel(const float* __restrict__ gmax, const float* __restrict__ gsum, const float* __restrict__ target,
               const float* __restrict__ mask, float* __restrict__ loss_tok, float* __restrict__ out, size_t T) {
  __shared__ float sa[32], sb[32];
  float num = 0.f, den = 0.f;
  for (size_t t = threadIdx.x; t < T; t += blockDim.x) {
    const float l = __logf(gsum[t]) + gmax[t] - target[t];
    if (loss_tok != nullptr) loss_tok[t] = l;
    const float w = mask != nullptr ? mask[t] : 1.f;
    num += l * w;
    den += w;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    num += __shfl_xor_sync(0xffffffffu, num, o);
    den += __shfl_xor_sync(0xffffffffu, den, o);
  }
  if ((threadIdx.x & 31) == 0) { sa[threadIdx.x >> 5] = num; sb[threadIdx.x >> 5] = den; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
    for (int w = 0; w < 32; ++w) { a += sa[w]; b += sb[w]; }
    out[0] = a / b;
    out[1] = b;
  }
}

// dlogits[t,c] = (exp(l[t,c] - gmax[t]) / gsum[t] - [c == label]) * dloss * mask[t] / sum(mask)
__global__ void __launch_bounds__(256)
ce_bwd_kernel(const __nv_bfloat16* logits, __nv_bfloat16* dlogits,  // may alias (in-place): no __restrict__, no __ldg
              const int64_t* __restrict__ labels, int64_t vocab_start, int vl, const float* __restrict__ gmax,
              const float* __restrict__ gsum, const float* __restrict__ mask, const float* __restrict__ loss_out,
              const float* __restrict__ dloss) {
  const size_t t = blockIdx.x;
  const float w = (mask != nullptr ? mask[t] : 1.f) * (dloss != nullptr ? dloss[0] : 1.f) / loss_out[1];
  const float m = gmax[t], inv = w / gsum[t];
  const int64_t col = __ldg(labels + t) - vocab_start;
  const uint4* row = reinterpret_cast<const uint4*>(logits + t * vl);
  uint4* drow = reinterpret_cast<uint4*>(dlogits + t * vl);
  const int nvec = vl >> 3;
  for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
    float x[8];
    unpack8(row[v], x);  // plain load: each vector is read, then overwritten, by the same thread
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = __expf(x[j] - m) * inv - ((v * 8 + j) == col ? w : 0.f);
    drow[v] = pack8(x);
  }
}

}  // namespace model
}  // namespace fx

extern "C" int fx_embedding_fwd(void* stream, const int64_t* ids, const int64_t* pos_ids, const void* wte,
                                int64_t vocab_start, int64_t vocab_rows, const void* word_in, const void* wpe, void* out,
                                int64_t T, int s, int h, float p, uint64_t seed, uint64_t offset) {
  using namespace fx;
  FX_CHECK_ARG(T > 0 && s > 0 && h > 0 && h % 8 == 0, "fx_embedding_fwd: T %lld, s %d, h %d (h must be a multiple of 8)",
               (long long)T, s, h);
  FX_CHECK_ARG(out && (word_in || (ids && wte && vocab_rows > 0)), "fx_embedding_fwd: needs word_in, or ids + wte");
  FX_CHECK_ARG(p >= 0.f && p < 1.f, "fx_embedding_fwd: dropout p must be in [0,1) (got %f)", p);
  FX_CHECK_ARG(((uintptr_t)wte | (uintptr_t)word_in | (uintptr_t)wpe | (uintptr_t)out) % 16 == 0,
               "fx_embedding_fwd: pointers must be 16-byte aligned");
  const size_t total = static_cast<size_t>(T) * (h / 8);
  model::embedding_fwd_kernel<<<model::grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      ids, pos_ids, static_cast<const __nv_bfloat16*>(wte), vocab_start, vocab_rows,
      static_cast<const __nv_bfloat16*>(word_in), static_cast<const __nv_bfloat16*>(wpe),
      static_cast<__nv_bfloat16*>(out), static_cast<size_t>(T), s, h / 8, 1.0f / (1.0f - p),
      p > 0.f ? dropout_threshold(p) : 0u, seed, offset);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_embedding_bwd(void* stream, const void* dout, const int64_t* ids, const int64_t* pos_ids, float* dwte,
                                int64_t vocab_start, int64_t vocab_rows, float* dwpe, int64_t T, int s, int h, float p,
                                uint64_t seed, uint64_t offset) {
  using namespace fx;
  FX_CHECK_ARG(T > 0 && s > 0 && h > 0 && h % 8 == 0, "fx_embedding_bwd: T %lld, s %d, h %d (h must be a multiple of 8)",
               (long long)T, s, h);
  FX_CHECK_ARG(dout && ids && (dwte || dwpe), "fx_embedding_bwd: null pointer");
  FX_CHECK_ARG(p >= 0.f && p < 1.f, "fx_embedding_bwd: dropout p must be in [0,1) (got %f)", p);
  FX_CHECK_ARG(((uintptr_t)dout | (uintptr_t)dwte | (uintptr_t)dwpe) % 16 == 0, "fx_embedding_bwd: pointers must be 16-byte aligned");
  const size_t total = static_cast<size_t>(T) * (h / 8);
  model::embedding_bwd_kernel<<<model::grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dout), ids, pos_ids, dwte, vocab_start, vocab_rows, dwpe, static_cast<size_t>(T), s,
      h / 8, 1.0f / (1.0f - p), p > 0.f ? dropout_threshold(p) : 0u, seed, offset);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_f32_to_bf16(void* stream, const float* src, void* dst, int64_t n) {
  using namespace fx;
  FX_CHECK_ARG(n > 0 && n % 8 == 0 && src && dst, "fx_f32_to_bf16: n must be a positive multiple of 8 (got %lld)", (long long)n);
  FX_CHECK_ARG(((uintptr_t)src | (uintptr_t)dst) % 16 == 0, "fx_f32_to_bf16: pointers must be 16-byte aligned");
  model::f32_to_bf16_kernel<<<model::grid_for(n / 8, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), static_cast<size_t>(n / 8));
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_cross_entropy_stats(void* stream, const void* logits, const int64_t* labels, int64_t vocab_start,
                                      int64_t T, int vl, float* lmax, float* lsum, float* target) {
  using namespace fx;
  FX_CHECK_ARG(T > 0 && vl > 0 && vl % 8 == 0, "fx_cross_entropy_stats: T %lld, local vocabulary %d (must be a multiple of 8)",
               (long long)T, vl);
  FX_CHECK_ARG(logits && labels && lmax && lsum && target, "fx_cross_entropy_stats: null pointer");
  FX_CHECK_ARG((uintptr_t)logits % 16 == 0, "fx_cross_entropy_stats: logits must be 16-byte aligned");
  model::ce_stats_kernel<<<static_cast<unsigned>(T), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(logits), labels, vocab_start, vl, lmax, lsum, target);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_cross_entropy_rescale(void* stream, const float* lmax, const float* gmax, float* lsum, int64_t T) {
  using namespace fx;
  FX_CHECK_ARG(T > 0 && lmax && gmax && lsum, "fx_cross_entropy_rescale: bad arguments");
  model::ce_rescale_kernel<<<static_cast<unsigned>((T + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      lmax, gmax, lsum, static_cast<size_t>(T));
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_cross_entropy_loss(void* stream, const float* gmax, const float* gsum, const float* target,
                                     const float* loss_mask, float* loss_tok, float* out2, int64_t T) {
  using namespace fx;
  FX_CHECK_ARG(T > 0 && gmax && gsum && target && out2, "fx_cross_entropy_loss: bad arguments");
  model::ce_loss_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(gmax, gsum, target, loss_mask, loss_tok, out2,
                                                                          static_cast<size_t>(T));
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_cross_entropy_bwd(void* stream, const void* logits, void* dlogits, const int64_t* labels,
                                    int64_t vocab_start, int64_t T, int vl, const float* gmax, const float* gsum,
                                    const float* loss_mask, const float* out2, const float* dloss) {
  using namespace fx;
  FX_CHECK_ARG(T > 0 && vl > 0 && vl % 8 == 0, "fx_cross_entropy_bwd: T %lld, local vocabulary %d (must be a multiple of 8)",
               (long long)T, vl);
  FX_CHECK_ARG(logits && dlogits && labels && gmax && gsum && out2, "fx_cross_entropy_bwd: null pointer");
  FX_CHECK_ARG(((uintptr_t)logits | (uintptr_t)dlogits) % 16 == 0, "fx_cross_entropy_bwd: logits must be 16-byte aligned");
  model::ce_bwd_kernel<<<static_cast<unsigned>(T), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(logits), static_cast<__nv_bfloat16*>(dlogits), labels, vocab_start, vl, gmax, gsum,
      loss_mask, out2, dloss);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
