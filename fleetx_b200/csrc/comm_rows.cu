// comm_rows.cu — the tensor-parallel allreduce with the OWNER of each row doing the row-wise work.
//
// Reference semantics: fleet.meta_parallel.RowParallelLinear's `_mp_allreduce` + bias (hybrid_model.py:157-163,
// 505-511), then dropout + residual (:546-547, :556-561) and the LayerNorm that follows (:553-554); in
// backward the allreduce that is `_c_identity`'s adjoint in front of ColumnParallelLinear, then LayerNorm
// backward, the skip-path add and the dropout backward (Paddle autograd of :531-554).
//
// In the reference (and in fx_tp_allreduce_bdr) every mp rank repeats that row-wise work on the whole
// replicated [T, h] tensor: it does not shrink with the mp degree. Here rank r owns rows
// [r*T/n, (r+1)*T/n). One kernel
//   1. waits until every rank's partial sums are in place (flag handshake),
//   2. pulls ITS rows, summed by the switch (multimem.ld_reduce, fp32 accumulation),
//   3. applies the row-wise chain to them — a row lives in the registers of 1, 2 or 4 warps —,
//   4. keeps what only the owner needs later (x1, LayerNorm statistics, d_x1; parameter-gradient partial
//      sums go to fp32 accumulators) and multicasts the tensor every rank needs next (LayerNorm output /
//      layer output / d(attention out) / dx) into every rank's copy of a second symmetric buffer
//      (multimem.st), which the next GEMM reads from ITS OWN GPU's copy,
//   5. passes one rank barrier.
// Both NVLink directions are busy at once (rows being pulled while finished rows are pushed), no row-wise
// kernel runs on replicated data, and the bytes over NVLink are those of the plain allreduce.
#include <atomic>

#include "comm.cuh"
#include "philox.cuh"

namespace fx {
namespace comm {

constexpr int ROWS_THREADS = 512;
constexpr int ROWS_WARPS = ROWS_THREADS / 32;

struct RowsParams {
  uint64_t sig[MAX_RANKS];
  int rank, world;
  uint32_t epoch;
  uint64_t mc_in, mc_out;
  int rows, h, r0, r1;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;   // FWD_LN: [rows, h] (global row); FWD_OUT: [own, h] (local row)
  float p_drop, scale;
  uint32_t thresh;
  uint64_t seed, offset;
  uint64_t drop_vec0;              // index, in the whole tensor the mask is defined on, of this call's first 16-byte vector
  const uint8_t* mask_in;          // [rows, h] keep flags (test hook), global row
  const float* gamma;
  const float* beta;
  float eps;
  __nv_bfloat16* x_own;            // FWD_LN: out x1 [own, h]; BWD_LN_BDR: in x1 [own, h]
  const __nv_bfloat16* x_full;     // BWD_LN: in x [rows, h]
  float* mean;                     // FWD_LN out / BWD_LN_BDR in: [own]; BWD_LN in: [rows]
  float* rstd;
  const __nv_bfloat16* dres;       // BWD_LN_BDR: [rows, h] (global row); BWD_LN: [own, h] (local row)
  __nv_bfloat16* dx_own;           // BWD_LN_BDR: out d_x1 [own, h]
  float* acc;                      // BWD: fp32 [3, h] dgamma | dbeta | dbias accumulators (atomics; host zeroes)
  uint64_t mc_sum;                 // BWD_LN, optional: multicast address of a small fp32 vector to sum over the ranks ...
  float* sum_out;                  // ... into this local vector of sum_n floats once every rank's rows are done
  int sum_n;
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum of (a, b) over the W warps that share a row. red: [ROWS_WARPS] float2 in shared memory.
template <int W>
__device__ __forceinline__ float2 row_sum2(float a, float b, float2* red, int warp, int lane) {
  a = warp_sum(a);
  b = warp_sum(b);
  if (W == 1) return make_float2(a, b);
  const int g = warp / W;
  if (lane == 0) red[warp] = make_float2(a, b);
  asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "n"(32 * W) : "memory");
  float2 r = red[g * W];
#pragma unroll
  for (int i = 1; i < W; ++i) {
    r.x += red[g * W + i].x;
    r.y += red[g * W + i].y;
  }
  asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "n"(32 * W) : "memory");  // red[] may be rewritten
  return r;
}

__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// keep flags of the 8 elements of vector `gv` (index of the 16-byte vector in the whole [rows, h] tensor)
__device__ __forceinline__ uint32_t keep8(const RowsParams& p, size_t gv) {
  if (p.mask_in != nullptr) {
    const uint2 m = __ldg(reinterpret_cast<const uint2*>(p.mask_in) + gv);
    uint32_t k = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      k |= ((m.x >> (8 * j)) & 0xFFu) ? (1u << j) : 0u;
      k |= ((m.y >> (8 * j)) & 0xFFu) ? (1u << (4 + j)) : 0u;
    }
    return k;
  }
  if (p.p_drop > 0.f) {
    const Philox8 rnd = philox4x32_10(p.seed, p.offset, p.drop_vec0 + gv);
    uint32_t k = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) k |= philox_keep(rnd, j, p.thresh) ? (1u << j) : 0u;
    return k;
  }
  return 0xFFu;
}

enum { MODE_FWD_LN = 0, MODE_FWD_OUT = 1, MODE_BWD_LN_BDR = 2, MODE_BWD_LN = 3 };

// W = warps per row, VPL = 16-byte vectors per lane: 32 * W * VPL * 8 >= h.
template <int MODE, int W, int VPL>
__global__ void __launch_bounds__(ROWS_THREADS, 1)
tp_rows_kernel(const RowsParams p) {
  extern __shared__ float acc_s[];  // BWD modes: [3][8][hv] column accumulators, layout (e * hv + v): bank = lane
  __shared__ float2 red[ROWS_WARPS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int G = ROWS_WARPS / W;  // rows in flight per block
  const int g = warp / W, w = warp % W;
  const int hv = p.h >> 3;
  const float inv_h = 1.0f / static_cast<float>(p.h);
  constexpr bool BWD = MODE == MODE_BWD_LN_BDR || MODE == MODE_BWD_LN;

  if (BWD) {
    for (int i = tid; i < 3 * p.h; i += ROWS_THREADS) acc_s[i] = 0.f;
  }
  ranks_ready(p);  // ends with __syncthreads()

  for (int row = p.r0 + blockIdx.x * G + g; row < p.r1; row += gridDim.x * G) {
    const size_t gbase = static_cast<size_t>(row) * hv;          // vector index of the row in [rows, h]
    const size_t lbase = static_cast<size_t>(row - p.r0) * hv;   // ... in an [own, h] tensor
    uint4 raw[VPL];
    // ---- the row, summed over the ranks by the switch; everything else the row needs is requested now too
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
      const int v = w * 32 + lane + j * 32 * W;
      if (v < hv) raw[j] = multimem_ld_reduce_bf16x8(p.mc_in + (gbase + v) * 16);
    }

    if (MODE == MODE_FWD_LN || MODE == MODE_FWD_OUT) {
      uint4 res[VPL];
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = w * 32 + lane + j * 32 * W;
        if (v < hv) res[j] = __ldg(reinterpret_cast<const uint4*>(p.residual) + (MODE == MODE_FWD_LN ? gbase : lbase) + v);
      }
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = w * 32 + lane + j * 32 * W;
        if (v >= hv) continue;
        float y[8], r[8], o[8];
        unpack8(raw[j], y);
        unpack8(res[j], r);
        if (p.bias != nullptr) {
          float b[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(p.bias) + v), b);
#pragma unroll
          for (int e = 0; e < 8; ++e) y[e] += b[e];
        }
        const uint32_t k = keep8(p, gbase + v);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = r[e] + (((k >> e) & 1u) ? y[e] * p.scale : 0.f);
        raw[j] = pack8(o);  // x1 / out as stored (bf16)
        if (MODE == MODE_FWD_OUT) {
          multimem_st_v4(p.mc_out + (gbase + v) * 16, raw[j]);
        } else {
          reinterpret_cast<uint4*>(p.x_own)[lbase + v] = raw[j];
          unpack8(raw[j], o);  // LayerNorm sees the stored values, as the separate kernel would
#pragma unroll
          for (int e = 0; e < 8; ++e) s += o[e];
        }
      }
      if (MODE == MODE_FWD_LN) {
        const float mean = row_sum2<W>(s, 0.f, red, warp, lane).x * inv_h;
        float ss = 0.f;
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
          const int v = w * 32 + lane + j * 32 * W;
          if (v >= hv) continue;
          float o[8];
          unpack8(raw[j], o);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float d = o[e] - mean;
            ss += d * d;
          }
        }
        const float var = row_sum2<W>(ss, 0.f, red, warp, lane).x * inv_h;
        const float rstd = rsqrtf(var + p.eps);
        if (w == 0 && lane == 0) {
          p.mean[row - p.r0] = mean;
          p.rstd[row - p.r0] = rstd;
        }
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
          const int v = w * 32 + lane + j * 32 * W;
          if (v >= hv) continue;
          float o[8], gm[8], bt[8];
          unpack8(raw[j], o);
          load8f(p.gamma + v * 8, gm);
          load8f(p.beta + v * 8, bt);
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = (o[e] - mean) * rstd * gm[e] + bt[e];
          multimem_st_v4(p.mc_out + (gbase + v) * 16, pack8(o));
        }
      }
    } else {
      // ---- LayerNorm backward on the summed gradient (+ skip-path gradient, + dropout backward)
      uint4 xb[VPL];
      const bool own_x = MODE == MODE_BWD_LN_BDR;
      const float mean = p.mean[own_x ? row - p.r0 : row], rstd = p.rstd[own_x ? row - p.r0 : row];
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = w * 32 + lane + j * 32 * W;
        if (v < hv)
          xb[j] = own_x ? __ldg(reinterpret_cast<const uint4*>(p.x_own) + lbase + v)
                        : __ldg(reinterpret_cast<const uint4*>(p.x_full) + gbase + v);
      }
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = w * 32 + lane + j * 32 * W;
        if (v >= hv) continue;
        float dy[8], x[8], gm[8];
        unpack8(raw[j], dy);
        unpack8(xb[j], x);
        load8f(p.gamma + v * 8, gm);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xhat = (x[e] - mean) * rstd;
          const float gd = gm[e] * dy[e];
          s1 += gd;
          s2 += gd * xhat;
          atomicAdd(&acc_s[e * hv + v], dy[e] * xhat);                // dgamma
          atomicAdd(&acc_s[(8 + e) * hv + v], dy[e]);                 // dbeta
        }
      }
      const float2 c = row_sum2<W>(s1, s2, red, warp, lane);
      const float c1 = c.x * inv_h, c2 = c.y * inv_h;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        const int v = w * 32 + lane + j * 32 * W;
        if (v >= hv) continue;
        float dy[8], x[8], gm[8], dr[8], o[8];
        unpack8(raw[j], dy);
        unpack8(xb[j], x);
        load8f(p.gamma + v * 8, gm);
        unpack8(__ldg(reinterpret_cast<const uint4*>(p.dres) + (MODE == MODE_BWD_LN_BDR ? gbase : lbase) + v), dr);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xhat = (x[e] - mean) * rstd;
          o[e] = rstd * (gm[e] * dy[e] - c1 - xhat * c2) + dr[e];
        }
        uint4 dxv = pack8(o);
        if (MODE == MODE_BWD_LN) {
          multimem_st_v4(p.mc_out + (gbase + v) * 16, dxv);           // dx, for every rank
        } else {
          reinterpret_cast<uint4*>(p.dx_own)[lbase + v] = dxv;        // d_x1: only the owner needs it again
          unpack8(dxv, o);                                            // dropout backward on the stored values
          const uint32_t k = keep8(p, gbase + v);
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = ((k >> e) & 1u) ? o[e] * p.scale : 0.f;
          dxv = pack8(o);
          multimem_st_v4(p.mc_out + (gbase + v) * 16, dxv);           // d(attention out + bias), for every rank
          unpack8(dxv, o);
#pragma unroll
          for (int e = 0; e < 8; ++e) atomicAdd(&acc_s[(16 + e) * hv + v], o[e]);  // its column sums: d(out_proj.bias)
        }
      }
    }
  }

  if (BWD) {
    __syncthreads();
    const int n_acc = MODE == MODE_BWD_LN_BDR ? 3 : 2;
    for (int i = tid; i < n_acc * p.h; i += ROWS_THREADS) {
      const int which = i / p.h, rem = i - which * p.h;   // rem = e * hv + v
      const int e = rem / hv, v = rem - e * hv;
      const float val = acc_s[i];
      if (val != 0.f) atomicAdd(p.acc + which * p.h + v * 8 + e, val);
    }
  }
  // every rank's rows have landed in every copy of the output buffer; nobody still reads the partial sums
  rank_barrier(p, SLOT_RS_DONE, SLOT_TICKET_A, true, true, 1);
  const bool sum_tail = MODE == MODE_BWD_LN && p.sum_out != nullptr;
  if (sum_tail) {
    // ... and every rank's parameter-gradient partial sums are final: each rank pulls the whole (small) vector
    // through the switch. Nobody may touch its copy again while a peer still reads it: barrier B waits.
    for (int i = (blockIdx.x * ROWS_THREADS + tid) * 4; i < p.sum_n; i += gridDim.x * ROWS_THREADS * 4)
      *reinterpret_cast<float4*>(p.sum_out + i) = multimem_ld_reduce_f32x4(p.mc_sum + static_cast<uint64_t>(i) * 4);
  }
  rank_barrier(p, SLOT_AG_DONE, SLOT_TICKET_B, sum_tail, false, 2);
}

// Sum of a small fp32 vector over the ranks (the parameter-gradient partial sums each rank made over its own
// rows): every rank pulls the whole vector through the switch and writes it to ordinary local memory.
struct SmallParams {
  uint64_t sig[MAX_RANKS];
  int rank, world;
  uint32_t epoch;
  uint64_t mc;     // multicast address of the fp32 vector
  float* out;
  int n;           // multiple of 4
};

__global__ void __launch_bounds__(ROWS_THREADS)
tp_allreduce_f32_kernel(const SmallParams p) {
  ranks_ready(p);
  for (int i = (blockIdx.x * ROWS_THREADS + threadIdx.x) * 4; i < p.n; i += gridDim.x * ROWS_THREADS * 4)
    *reinterpret_cast<float4*>(p.out + i) = multimem_ld_reduce_f32x4(p.mc + static_cast<uint64_t>(i) * 4);
  rank_barrier(p, SLOT_RS_DONE, SLOT_TICKET_A, false, false, 1);
  // nobody may touch its copy of the vector again while a peer's reduction still reads it
  rank_barrier(p, SLOT_AG_DONE, SLOT_TICKET_B, true, false, 2);
}

template <int MODE, int W, int VPL>
static int launch_rows_wv(cudaStream_t st, int num_blocks, const RowsParams& q) {
  const size_t smem = (MODE == MODE_BWD_LN_BDR || MODE == MODE_BWD_LN) ? static_cast<size_t>(3) * q.h * sizeof(float) : 0;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  FX_CUDA(cudaGetDevice(&dev));
  if (smem > 0 && dev < 64 && !attr_set[dev]) {
    FX_CUDA(cudaFuncSetAttribute(tp_rows_kernel<MODE, W, VPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 8192 * 4));
    attr_set[dev] = true;
  }
  tp_rows_kernel<MODE, W, VPL><<<num_blocks, ROWS_THREADS, smem, st>>>(q);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

template <int MODE>
static int launch_rows(cudaStream_t st, int num_blocks, const RowsParams& q) {
  if (q.h <= 512) return launch_rows_wv<MODE, 1, 2>(st, num_blocks, q);
  if (q.h <= 2048) return launch_rows_wv<MODE, 1, 8>(st, num_blocks, q);
  if (q.h <= 4096) return launch_rows_wv<MODE, 2, 8>(st, num_blocks, q);
  return launch_rows_wv<MODE, 4, 8>(st, num_blocks, q);
}

}  // namespace comm
}  // namespace fx

extern "C" int fx_tp_rows_max_h(void) { return 8192; }

extern "C" int fx_tp_rows_fused(void* stream, const fx_tp_rows_desc* d) {
  using namespace fx;
  FX_CHECK_ARG(d != nullptr, "fx_tp_rows_fused: null descriptor");
  FX_CHECK_ARG(d->world >= 2 && d->world <= comm::MAX_RANKS && d->rank >= 0 && d->rank < d->world,
               "fx_tp_rows_fused: bad rank/world (%d/%d)", d->rank, d->world);
  FX_CHECK_ARG(d->rows > 0 && d->h > 0 && d->h % 8 == 0 && d->h <= 8192,
               "fx_tp_rows_fused: rows %d, h %d (h must be a multiple of 8, at most 8192)", d->rows, d->h);
  FX_CHECK_ARG(d->peer_signal_pads && d->multicast_in && d->multicast_out, "fx_tp_rows_fused: null pointer");
  FX_CHECK_ARG(d->epoch > 0, "fx_tp_rows_fused: epoch must start at 1");
  FX_CHECK_ARG(d->num_blocks > 0 && d->num_blocks <= num_sms(), "fx_tp_rows_fused: num_blocks must be in [1, #SMs] "
               "(all blocks of all ranks wait on each other, so they must be co-resident)");
  FX_CHECK_ARG(d->p >= 0.f && d->p < 1.f, "fx_tp_rows_fused: dropout p must be in [0,1) (got %f)", d->p);
  FX_CHECK_ARG(d->mode >= 0 && d->mode <= 3, "fx_tp_rows_fused: unknown mode %d", d->mode);
  comm::RowsParams q{};
  for (int i = 0; i < d->world; ++i) q.sig[i] = d->peer_signal_pads[i];
  q.rank = d->rank; q.world = d->world; q.epoch = d->epoch;
  q.mc_in = d->multicast_in; q.mc_out = d->multicast_out;
  q.rows = d->rows; q.h = d->h;
  const int per = (d->rows + d->world - 1) / d->world;
  q.r0 = d->rank * per < d->rows ? d->rank * per : d->rows;
  q.r1 = q.r0 + per < d->rows ? q.r0 + per : d->rows;
  q.bias = static_cast<const __nv_bfloat16*>(d->bias);
  q.residual = static_cast<const __nv_bfloat16*>(d->residual);
  q.p_drop = d->p; q.scale = 1.0f / (1.0f - d->p);
  q.thresh = dropout_threshold(d->p);
  q.seed = d->seed; q.offset = d->offset;
  FX_CHECK_ARG(d->dropout_elem_offset % 8 == 0, "fx_tp_rows_fused: dropout_elem_offset must be a multiple of 8");
  q.drop_vec0 = d->dropout_elem_offset / 8;
  q.mask_in = d->mask_in;
  q.gamma = d->gamma; q.beta = d->beta; q.eps = d->eps;
  q.x_own = static_cast<__nv_bfloat16*>(d->x_own);
  q.x_full = static_cast<const __nv_bfloat16*>(d->x_full);
  q.mean = d->mean; q.rstd = d->rstd;
  q.dres = static_cast<const __nv_bfloat16*>(d->dres);
  q.dx_own = static_cast<__nv_bfloat16*>(d->dx_own);
  q.acc = d->acc;
  q.mc_sum = d->multicast_sum; q.sum_out = d->sum_out; q.sum_n = d->sum_n;
  FX_CHECK_ARG(q.sum_out == nullptr || (d->mode == comm::MODE_BWD_LN && q.mc_sum && q.sum_n > 0 && q.sum_n % 4 == 0 &&
                                        (uintptr_t)q.sum_out % 16 == 0),
               "fx_tp_rows_fused: the trailing sum needs mode BWD_LN, a multicast address and a multiple of 4 floats");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (d->mode) {
    case comm::MODE_FWD_LN:
      FX_CHECK_ARG(q.residual && q.gamma && q.beta && q.x_own && q.mean && q.rstd, "fx_tp_rows_fused(FWD_LN): null pointer");
      return comm::launch_rows<comm::MODE_FWD_LN>(st, d->num_blocks, q);
    case comm::MODE_FWD_OUT:
      FX_CHECK_ARG(q.residual, "fx_tp_rows_fused(FWD_OUT): null pointer");
      return comm::launch_rows<comm::MODE_FWD_OUT>(st, d->num_blocks, q);
    case comm::MODE_BWD_LN_BDR:
      FX_CHECK_ARG(q.x_own && q.mean && q.rstd && q.gamma && q.dres && q.dx_own && q.acc,
                   "fx_tp_rows_fused(BWD_LN_BDR): null pointer");
      if (!d->acc_accumulate) FX_CUDA(cudaMemsetAsync(q.acc, 0, sizeof(float) * 3 * q.h, st));
      return comm::launch_rows<comm::MODE_BWD_LN_BDR>(st, d->num_blocks, q);
    default:
      FX_CHECK_ARG(q.x_full && q.mean && q.rstd && q.gamma && q.dres && q.acc, "fx_tp_rows_fused(BWD_LN): null pointer");
      if (!d->acc_accumulate) FX_CUDA(cudaMemsetAsync(q.acc, 0, sizeof(float) * 2 * q.h, st));
      return comm::launch_rows<comm::MODE_BWD_LN>(st, d->num_blocks, q);
  }
}

extern "C" int fx_tp_allreduce_f32(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank,
                                   int world, uint32_t epoch, float* out, int n) {
  using namespace fx;
  FX_CHECK_ARG(world >= 2 && world <= comm::MAX_RANKS && rank >= 0 && rank < world,
               "fx_tp_allreduce_f32: bad rank/world (%d/%d)", rank, world);
  FX_CHECK_ARG(peer_signal_pads && multicast_ptr && out && n > 0 && n % 4 == 0 && (uintptr_t)out % 16 == 0,
               "fx_tp_allreduce_f32: null / misaligned pointer or n %d not a positive multiple of 4", n);
  FX_CHECK_ARG(epoch > 0, "fx_tp_allreduce_f32: epoch must start at 1");
  comm::SmallParams q{};
  for (int i = 0; i < world; ++i) q.sig[i] = peer_signal_pads[i];
  q.rank = rank; q.world = world; q.epoch = epoch;
  q.mc = multicast_ptr; q.out = out; q.n = n;
  const int blocks = (n / 4 + comm::ROWS_THREADS - 1) / comm::ROWS_THREADS;
  comm::tp_allreduce_f32_kernel<<<blocks < 16 ? blocks : 16, comm::ROWS_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(q);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
