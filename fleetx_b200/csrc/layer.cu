// layer.cu — fx_decoder_layer_fwd / fx_decoder_layer_bwd: the whole TransformerDecoderLayer
// (hybrid_model.py:525-566 with MultiHeadAttention.forward :300-347) as one native call.
//
// No kernel lives here: the two entry points carve the caller's `saved` and `workspace` buffers and run
// the library's own entry points in the order of the reference forward, resp. its hand-derived adjoint
// (Paddle autograd in the reference), all on one stream. Tensor parallelism follows
// fleet.meta_parallel.{Column,Row}ParallelLinear: forward allreduce after the out_proj / linear2 local
// matmuls (fused with bias + dropout + residual), backward allreduce of dX entering qkv_proj / linear1.
#include <math.h>

#include "common.cuh"

namespace fx {
namespace layer {

constexpr size_t ALIGN = 256;
static size_t up(size_t n) { return (n + ALIGN - 1) / ALIGN * ALIGN; }

// Offsets (bytes) of the activations kept between forward and backward.
struct SavedPlan {
  size_t ln1, mean1, rstd1, qkv, ctx, lse, mask, x1, ln2, mean2, rstd2, z, act, r1, r2, mean3, rstd3, total;
};
// Offsets of scratch buffers. Forward and backward plans share one workspace (max of the two).
struct WorkPlan {
  size_t part;                                     // [T,h] partial sums of a row-parallel GEMM (no TP buffer given)
  size_t d_y, d_z, d_in, d_x1, d_ctx, d_qkv, d_red, aux;  // backward
  size_t total;
};

struct Dims {
  size_t T, h, q3, hl, f;
};
static Dims dims(const fx_layer_desc* d) {
  Dims m;
  m.T = static_cast<size_t>(d->b) * d->s;
  m.h = d->h;
  m.hl = static_cast<size_t>(d->nh_local) * d->hd;
  m.q3 = 3 * m.hl;
  m.f = d->f_local;
  return m;
}

static SavedPlan plan_saved(const fx_layer_desc* d) {
  const Dims m = dims(d);
  SavedPlan p{};
  size_t o = 0;
  auto take = [&](size_t bytes) {
    const size_t at = o;
    o += up(bytes);
    return at;
  };
  const bool pre = d->normalize_before != 0;
  p.ln1 = pre ? take(m.T * m.h * 2) : 0;
  p.mean1 = take(m.T * 4);
  p.rstd1 = take(m.T * 4);
  p.qkv = take(m.T * m.q3 * 2);
  p.ctx = take(m.T * m.hl * 2);
  p.lse = take(static_cast<size_t>(d->b) * d->nh_local * d->s * 4);
  p.mask = d->p_attn > 0.f ? take(fx_attn_dropout_mask_bytes(d->b, d->s, d->nh_local)) : 0;
  p.x1 = take(m.T * m.h * 2);
  p.ln2 = pre ? take(m.T * m.h * 2) : 0;
  p.mean2 = pre ? take(m.T * 4) : 0;
  p.rstd2 = pre ? take(m.T * 4) : 0;
  p.z = take(m.T * m.f * 2);
  p.act = take(m.T * m.f * 2);
  p.r1 = pre ? 0 : take(m.T * m.h * 2);  // post-LN: input of LN1
  p.r2 = pre ? 0 : take(m.T * m.h * 2);  // post-LN: input of LN2
  p.mean3 = pre ? 0 : take(m.T * 4);
  p.rstd3 = pre ? 0 : take(m.T * 4);
  p.total = o;
  return p;
}

static WorkPlan plan_work(const fx_layer_desc* d) {
  const Dims m = dims(d);
  WorkPlan p{};
  size_t o = 0;
  auto take = [&](size_t bytes) {
    const size_t at = o;
    o += up(bytes);
    return at;
  };
  p.part = take(m.T * m.h * 2);
  p.d_y = take(m.T * m.h * 2);
  p.d_z = take(m.T * m.f * 2);
  p.d_in = take(m.T * m.h * 2);
  p.d_x1 = take(m.T * m.h * 2);
  p.d_ctx = take(m.T * m.hl * 2);
  p.d_qkv = take(m.T * m.q3 * 2);
  p.d_red = take(m.T * m.h * 2);
  size_t aux = fx_layernorm_bwd_workspace_bytes(static_cast<int>(m.T), d->h);
  const size_t widest = m.f > m.q3 ? m.f : m.q3;
  const size_t cs = fx_colsum_workspace_bytes(static_cast<int>(m.T), static_cast<int>(widest > m.h ? widest : m.h));
  const size_t ab = fx_attn_causal_bwd_workspace_bytes(d->b, d->s, d->nh_local, d->hd);
  if (cs > aux) aux = cs;
  if (ab > aux) aux = ab;
  p.aux = take(aux);
  p.total = o;
  return p;
}

static int check_desc(const fx_layer_desc* d, const char* who) {
  FX_CHECK_ARG(d != nullptr, "%s: null descriptor", who);
  FX_CHECK_ARG(d->b > 0 && d->s > 0 && d->h > 0 && d->nh_local > 0 && d->f_local > 0,
               "%s: empty shape (b %d, s %d, h %d, nh_local %d, f_local %d)", who, d->b, d->s, d->h, d->nh_local, d->f_local);
  FX_CHECK_ARG(d->hd == 64 || d->hd == 128, "%s: head_dim must be 64 or 128 (got %d)", who, d->hd);
  FX_CHECK_ARG(d->h % 8 == 0 && d->f_local % 8 == 0, "%s: h and f_local must be multiples of 8", who);
  FX_CHECK_ARG((static_cast<size_t>(d->b) * d->s) % 8 == 0, "%s: b*s must be a multiple of 8 (got %d x %d)", who, d->b, d->s);
  FX_CHECK_ARG(d->norm1_w && d->norm1_b && d->norm2_w && d->norm2_b && d->qkv_w && d->qkv_b && d->out_w && d->out_b &&
                   d->lin1_w && d->lin1_b && d->lin2_w && d->lin2_b,
               "%s: null parameter pointer", who);
  if (d->tp != nullptr) {
    const fx_tp_comm* c = d->tp;
    FX_CHECK_ARG(c->world >= 1 && c->rank >= 0 && c->rank < c->world, "%s: bad tp rank %d of %d", who, c->rank, c->world);
    FX_CHECK_ARG(c->world == 1 || (c->peer_bufs && c->peer_signal_pads && c->local_buf && c->epoch && c->num_blocks > 0),
                 "%s: incomplete fx_tp_comm", who);
  }
  return FX_OK;
}

static bool tp_on(const fx_layer_desc* d) { return d->tp != nullptr && d->tp->world > 1; }

// Plain allreduce of dX (backward of _c_identity): where the sum ends up. More than two ranks: the switch
// writes it into every rank's copy of the symmetric buffer (in place). Two ranks: the one-shot peer-load
// variant is faster but the peer still reads this rank's copy, so the sum goes to scratch.
static void* reduced_dx(const fx_layer_desc* d, void* part, void* scratch) {
  return (tp_on(d) && d->tp->world == 2) ? scratch : part;
}

// sum over ranks of `part` (+ bias, dropout, residual when out != null; else the sum stays in part)
static int reduce_bdr(void* stream, const fx_layer_desc* d, void* part, void* out, const void* bias,
                      const void* residual, float p, uint64_t seed, uint64_t offset) {
  const Dims m = dims(d);
  if (!tp_on(d)) {
    if (out == nullptr) return FX_OK;  // plain allreduce over one rank
    return fx_bias_dropout_residual_fwd(stream, part, bias, residual, out, static_cast<int>(m.T), d->h, p, seed, offset,
                                        nullptr, nullptr);
  }
  const fx_tp_comm* c = d->tp;
  const uint32_t epoch = ++*c->epoch;
  const uint64_t off = (static_cast<const char*>(part) - static_cast<const char*>(c->local_buf)) / 2;
  return fx_tp_allreduce_bdr(stream, c->peer_bufs, c->peer_signal_pads, c->multicast_ptr, c->rank, c->world, epoch,
                             c->num_blocks, out, static_cast<int>(m.T), d->h, bias, residual, p, seed, offset, nullptr, off, 0);
}

}  // namespace layer
}  // namespace fx

#define FX_TRY(call)            \
  do {                          \
    const int rc__ = (call);    \
    if (rc__ != FX_OK) return rc__; \
  } while (0)

extern "C" size_t fx_decoder_layer_saved_bytes(const fx_layer_desc* d) {
  if (fx::layer::check_desc(d, "fx_decoder_layer_saved_bytes") != FX_OK) return 0;
  return fx::layer::plan_saved(d).total;
}

extern "C" size_t fx_decoder_layer_workspace_bytes(const fx_layer_desc* d) {
  if (fx::layer::check_desc(d, "fx_decoder_layer_workspace_bytes") != FX_OK) return 0;
  return fx::layer::plan_work(d).total;
}

extern "C" int fx_decoder_layer_fwd(void* stream, const fx_layer_desc* d, const void* x, void* out, void* saved,
                                    void* workspace) {
  using namespace fx::layer;
  FX_TRY(check_desc(d, "fx_decoder_layer_fwd"));
  FX_CHECK_ARG(x && out && saved && workspace, "fx_decoder_layer_fwd: null buffer");
  FX_CHECK_ARG((reinterpret_cast<uintptr_t>(saved) | reinterpret_cast<uintptr_t>(workspace)) % ALIGN == 0,
               "fx_decoder_layer_fwd: saved / workspace must be 256-byte aligned");
  const Dims m = dims(d);
  const SavedPlan S = plan_saved(d);
  const WorkPlan W = plan_work(d);
  char* sv = static_cast<char*>(saved);
  char* ws = static_cast<char*>(workspace);
  const int T = static_cast<int>(m.T), h = d->h, q3 = static_cast<int>(m.q3), hl = static_cast<int>(m.hl), f = d->f_local;
  const bool pre = d->normalize_before != 0;
  const float scale = static_cast<float>(1.0 / sqrt(static_cast<double>(d->hd)));  // hybrid_model.py:270 head_dim**-0.5
  // partial sums of the row-parallel GEMMs land in the symmetric buffer when there is one
  void* part = tp_on(d) ? d->tp->local_buf : static_cast<void*>(ws + W.part);
  float* mean1 = reinterpret_cast<float*>(sv + S.mean1);
  float* rstd1 = reinterpret_cast<float*>(sv + S.rstd1);
  void* x1 = sv + S.x1;

  const void* attn_in = x;
  if (pre) {  // :534
    FX_TRY(fx_layernorm_fwd(stream, x, d->norm1_w, d->norm1_b, sv + S.ln1, mean1, rstd1, T, h, d->eps));
    attn_in = sv + S.ln1;
  }
  // qkv_proj (:166) with its bias in the epilogue
  FX_TRY(fx_gemm_bf16(stream, 0, 0, T, q3, h, attn_in, h, d->qkv_w, q3, sv + S.qkv, q3, FX_BF16, 0, FX_EPI_BIAS, d->qkv_b,
                      nullptr, 0));
  // :167-173 head split + :268-298 core attention
  FX_TRY(fx_attn_causal_fwd(stream, sv + S.qkv, sv + S.ctx, reinterpret_cast<float*>(sv + S.lse),
                            d->p_attn > 0.f ? sv + S.mask : nullptr, d->b, d->s, d->nh_local, d->hd, scale, d->p_attn,
                            d->seed_attn, d->offset_attn));
  // out_proj local matmul (:340), then _mp_allreduce + bias + dropout1 + residual (:546-547)
  FX_TRY(fx_gemm_bf16(stream, 0, 0, T, h, hl, sv + S.ctx, hl, d->out_w, h, part, h, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  void* r1 = pre ? x1 : static_cast<void*>(sv + S.r1);
  FX_TRY(reduce_bdr(stream, d, part, r1, d->out_b, x, d->p_hidden1, d->seed_h1, d->offset_h1));
  const void* ffn_in;
  if (pre) {  // :554
    FX_TRY(fx_layernorm_fwd(stream, x1, d->norm2_w, d->norm2_b, sv + S.ln2, reinterpret_cast<float*>(sv + S.mean2),
                            reinterpret_cast<float*>(sv + S.rstd2), T, h, d->eps));
    ffn_in = sv + S.ln2;
  } else {  // :550
    FX_TRY(fx_layernorm_fwd(stream, r1, d->norm1_w, d->norm1_b, x1, mean1, rstd1, T, h, d->eps));
    ffn_in = x1;
  }
  // linear1 + GELU (:558-559), pre-activation kept for backward; linear2 local matmul
  FX_TRY(fx_gemm_bf16(stream, 0, 0, T, f, h, ffn_in, h, d->lin1_w, f, sv + S.act, f, FX_BF16, 0, FX_EPI_BIAS_GELU, d->lin1_b,
                      sv + S.z, f));
  FX_TRY(fx_gemm_bf16(stream, 0, 0, T, h, f, sv + S.act, f, d->lin2_w, h, part, h, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  // _mp_allreduce + bias + dropout2 + residual (:556-561)
  void* r2 = pre ? out : static_cast<void*>(sv + S.r2);
  FX_TRY(reduce_bdr(stream, d, part, r2, d->lin2_b, x1, d->p_hidden2, d->seed_h2, d->offset_h2));
  if (!pre)  // :564
    FX_TRY(fx_layernorm_fwd(stream, r2, d->norm2_w, d->norm2_b, out, reinterpret_cast<float*>(sv + S.mean3),
                            reinterpret_cast<float*>(sv + S.rstd3), T, h, d->eps));
  return FX_OK;
}

extern "C" int fx_decoder_layer_bwd(void* stream, const fx_layer_desc* d, const void* dout, const void* x,
                                    const void* saved, void* dx, const fx_layer_grads* g, void* workspace) {
  using namespace fx::layer;
  FX_TRY(check_desc(d, "fx_decoder_layer_bwd"));
  FX_CHECK_ARG(dout && x && saved && dx && g && workspace, "fx_decoder_layer_bwd: null buffer");
  FX_CHECK_ARG(g->norm1_w && g->norm1_b && g->norm2_w && g->norm2_b && g->qkv_w && g->qkv_b && g->out_w && g->out_b &&
                   g->lin1_w && g->lin1_b && g->lin2_w && g->lin2_b,
               "fx_decoder_layer_bwd: null gradient pointer");
  FX_CHECK_ARG((reinterpret_cast<uintptr_t>(saved) | reinterpret_cast<uintptr_t>(workspace)) % ALIGN == 0,
               "fx_decoder_layer_bwd: saved / workspace must be 256-byte aligned");
  const Dims m = dims(d);
  const SavedPlan S = plan_saved(d);
  const WorkPlan W = plan_work(d);
  const char* sv = static_cast<const char*>(saved);
  char* ws = static_cast<char*>(workspace);
  const int T = static_cast<int>(m.T), h = d->h, q3 = static_cast<int>(m.q3), hl = static_cast<int>(m.hl), f = d->f_local;
  const bool pre = d->normalize_before != 0;
  const float scale = static_cast<float>(1.0 / sqrt(static_cast<double>(d->hd)));  // hybrid_model.py:270 head_dim**-0.5
  void* part = tp_on(d) ? d->tp->local_buf : static_cast<void*>(ws + W.part);
  void* aux = ws + W.aux;
  const float* mean1 = reinterpret_cast<const float*>(sv + S.mean1);
  const float* rstd1 = reinterpret_cast<const float*>(sv + S.rstd1);
  const void* x1 = sv + S.x1;

  const void* dy = dout;  // gradient arriving at r2
  if (!pre) {             // out = LN2(r2), :564
    FX_TRY(fx_layernorm_bwd(stream, dout, sv + S.r2, d->norm2_w, reinterpret_cast<const float*>(sv + S.mean3),
                            reinterpret_cast<const float*>(sv + S.rstd3), nullptr, ws + W.d_x1, g->norm2_w, g->norm2_b, aux, T, h));
    dy = ws + W.d_x1;
  }
  // r2 = x1 + dropout2(y2 + b2): gradient of the dropout branch; the skip share is folded in below
  void* d_y2 = ws + W.d_y;
  FX_TRY(fx_bias_dropout_residual_bwd(stream, dy, d_y2, T, h, d->p_hidden2, d->seed_h2, d->offset_h2, nullptr));
  // (dY · W2^T) ∘ gelu'(z), then dX of linear1 into the (symmetric) partial buffer
  FX_TRY(fx_gemm_bf16(stream, 0, 1, T, f, h, d_y2, h, d->lin2_w, h, ws + W.d_z, f, FX_BF16, 0, FX_EPI_DGELU, nullptr,
                      const_cast<char*>(sv + S.z), f));
  FX_TRY(fx_gemm_bf16(stream, 0, 1, T, h, f, ws + W.d_z, f, d->lin1_w, f, part, h, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  void* red = reduced_dx(d, part, ws + W.d_red);
  FX_TRY(reduce_bdr(stream, d, part, red == part ? nullptr : red, nullptr, nullptr, 0.f, 0, 0));  // backward of _c_identity
  FX_TRY(fx_colsum_bf16(stream, d_y2, h, g->lin2_b, FX_BF16, aux, T, h));
  FX_TRY(fx_gemm_bf16(stream, 1, 0, f, h, T, sv + S.act, f, d_y2, h, g->lin2_w, h, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  FX_TRY(fx_colsum_bf16(stream, ws + W.d_z, f, g->lin1_b, FX_BF16, aux, T, f));
  const void* ffn_in = pre ? sv + S.ln2 : x1;
  FX_TRY(fx_gemm_bf16(stream, 1, 0, h, f, T, ffn_in, h, ws + W.d_z, f, g->lin1_w, f, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));

  const void* d_r1;
  void* d_ao;
  if (pre) {
    // x1 feeds LN2 and the skip: d(x1) = LN2'(d_ffn_in) + dY
    FX_TRY(fx_layernorm_bwd(stream, red, x1, d->norm2_w, reinterpret_cast<const float*>(sv + S.mean2),
                            reinterpret_cast<const float*>(sv + S.rstd2), dy, ws + W.d_x1, g->norm2_w, g->norm2_b, aux, T, h));
    d_r1 = ws + W.d_x1;
    d_ao = ws + W.d_y;
  } else {
    // x1 = LN1(r1) feeds linear1 and the skip: d(x1) = d_ffn_in + dY (dY lives in d_x1: add in place is not
    // possible with distinct in/out kernels, so the sum goes to d_in), then through LN1
    FX_TRY(fx_bias_dropout_residual_fwd(stream, red, nullptr, dy, ws + W.d_in, T, h, 0.f, 0, 0, nullptr, nullptr));
    FX_TRY(fx_layernorm_bwd(stream, ws + W.d_in, sv + S.r1, d->norm1_w, mean1, rstd1, nullptr, ws + W.d_y, g->norm1_w,
                            g->norm1_b, aux, T, h));
    d_r1 = ws + W.d_y;
    d_ao = ws + W.d_in;
  }
  // r1 = x + dropout1(attn_out + bo)
  FX_TRY(fx_bias_dropout_residual_bwd(stream, d_r1, d_ao, T, h, d->p_hidden1, d->seed_h1, d->offset_h1, nullptr));
  FX_TRY(fx_gemm_bf16(stream, 0, 1, T, hl, h, d_ao, h, d->out_w, h, ws + W.d_ctx, hl, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  FX_TRY(fx_attn_causal_bwd(stream, ws + W.d_ctx, sv + S.qkv, sv + S.ctx, reinterpret_cast<const float*>(sv + S.lse),
                            d->p_attn > 0.f ? sv + S.mask : nullptr, ws + W.d_qkv, aux, d->b, d->s, d->nh_local, d->hd, scale,
                            d->p_attn));
  FX_TRY(fx_gemm_bf16(stream, 0, 1, T, h, q3, ws + W.d_qkv, q3, d->qkv_w, q3, part, h, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  FX_TRY(reduce_bdr(stream, d, part, red == part ? nullptr : red, nullptr, nullptr, 0.f, 0, 0));  // backward of _c_identity
  FX_TRY(fx_colsum_bf16(stream, d_ao, h, g->out_b, FX_BF16, aux, T, h));
  FX_TRY(fx_gemm_bf16(stream, 1, 0, hl, h, T, sv + S.ctx, hl, d_ao, h, g->out_w, h, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  FX_TRY(fx_colsum_bf16(stream, ws + W.d_qkv, q3, g->qkv_b, FX_BF16, aux, T, q3));
  const void* attn_in = pre ? sv + S.ln1 : x;
  FX_TRY(fx_gemm_bf16(stream, 1, 0, h, q3, T, attn_in, h, ws + W.d_qkv, q3, g->qkv_w, q3, FX_BF16, 0, FX_EPI_NONE, nullptr, nullptr, 0));
  if (pre) {  // x feeds LN1 and the skip
    FX_TRY(fx_layernorm_bwd(stream, red, x, d->norm1_w, mean1, rstd1, d_r1, dx, g->norm1_w, g->norm1_b, aux, T, h));
  } else {
    FX_TRY(fx_bias_dropout_residual_fwd(stream, red, nullptr, d_r1, dx, T, h, 0.f, 0, 0, nullptr, nullptr));
  }
  return FX_OK;
}
