// gemm.cu — K2/K6: persistent, warp-specialised bf16 GEMM on tcgen05 for the four linears of
// the decoder layer and their gradients (hybrid_model.py:125-131,157-163,497-511).
//
//   warp 8      TMA producer: one lane streams 128xBK A tiles and BNxBK B tiles into a
//               4..6-stage shared-memory ring (128-byte swizzle), signalling `full` mbarriers
//   warp 9      MMA issuer: one lane issues tcgen05.mma (M=128, N=BN, K=16) into one of two
//               TMEM accumulators (2 x BN fp32 columns), tcgen05.commit frees ring slots
//   warps 0..7  epilogue: tcgen05.ld the finished accumulator (one row per thread), transpose
//               32x32 blocks through shared memory so global accesses are sector-coalesced,
//               apply bias / GELU / GELU' / accumulate, store bf16 or fp32 — overlapped with
//               the next tile's main loop through the second accumulator
//
// Operands may be K-major or MN-major in memory (Paddle's [in,out] weights make the forward
// B operand MN-major and the weight gradient MN-major on both sides); the shared-memory
// descriptors carry the difference, so no transposed copies are ever made.
#include <stdlib.h>

#include <atomic>

#include "common.cuh"
#include "tc5.cuh"

namespace fx {
namespace gemm {

using namespace tc5;

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int NUM_EPI_WARPS = 8;  // two per TMEM lane quarter, interleaved over 32-column chunks
constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
// The single-thread TMA / MMA issuers sit in the HIGHEST warp ids: the SM's warp arbiter favours
// higher ids, and an MMA issuer starved of issue slots by the epilogue warps stalls the tensor pipe.
constexpr int TMA_WARP = NUM_EPI_WARPS, MMA_WARP = NUM_EPI_WARPS + 1;
constexpr int A_STAGE_BYTES = BM * BK * 2;
constexpr int CHUNK_BYTES = BK * 128;  // one 64-wide MN chunk of an MN-major operand

struct Params {
  int M, N, K;
  int a_mn, b_mn;
  void* C;
  int ldc, c_f32, accumulate;
  int epi;
  const __nv_bfloat16* bias;
  __nv_bfloat16* aux;
  int ldaux;
  int num_m, num_n, num_tiles, num_kb;
  int group;  // M-blocks per rasterisation group
  // split-K: a work item is (tile, split); split s covers k-blocks [s * kb_per_split, ...) and writes its fp32
  // partial tile to C + s * split_stride (a [splits, M, ldc] workspace that a second kernel sums)
  int splits, kb_per_split;
  size_t split_stride;
};

template <int BN>
struct Cfg {
  static constexpr int B_STAGE_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = BN == 256 ? 4 : 6;
  static constexpr int BAR_BYTES = 256;
  static constexpr int STG_BYTES = NUM_EPI_WARPS * 32 * 33 * 4;  // per-epilogue-warp 32x32 fp32 transpose tile (+1 pad)
  // no alignment slack: the dynamic window starts 1024-aligned (checked in the kernel), and the 1 KB this saves is
  // what lets a small kernel with no shared memory of its own (dropout keep bits, dropout backward, ...) share the
  // SM with a resident GEMM CTA when the host runs it on another stream
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + STG_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
};

__device__ __forceinline__ void tile_coords(const Params& p, int tile, int& m_blk, int& n_blk) {
  const int GROUP = p.group;  // M-blocks that share each B tile while it is hot in L2
  const int per_group = GROUP * p.num_n;
  const int g = tile / per_group;
  const int first_m = g * GROUP;
  const int gsize = min(GROUP, p.num_m - first_m);
  const int r = tile - g * per_group;
  m_blk = first_m + r % gsize;
  n_blk = r / gsize;
}

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));  // one MUFU op, rel. error 2^-11 (bf16 output: 2^-8)
  return y;
}
__device__ __forceinline__ float gelu_tanh(float z) {
  // 0.5 z (1 + tanh(u)),  u = sqrt(2/pi) (z + 0.044715 z^3)   (F.gelu(approximate=True), hybrid_model.py:558)
  const float u = z * (0.7978845608028654f + 0.035677408136300125f * z * z);
  const float hz = 0.5f * z;
  return fmaf(hz, tanh_approx(u), hz);
}
__device__ __forceinline__ float gelu_tanh_grad(float z) {
  const float z2 = z * z;
  const float u = z * (0.7978845608028654f + 0.035677408136300125f * z2);
  const float t = tanh_approx(u);
  const float du = 0.7978845608028654f + 0.10703222440890037f * z2;  // du/dz
  // d/dz [0.5 z (1 + t)] = 0.5 (1 + t) + 0.5 z (1 - t^2) du
  return 0.5f * (1.0f + t) + 0.5f * z * (1.0f - t * t) * du;
}

// Epilogue math + store for one 32x32 fp32 block that a warp has just transposed through its
// shared-memory staging tile: lane l owns columns [cl, cl+4) (cl = 4*(l%8)) of rows l/8 + 4*i,
// so every global access of the warp covers whole 32-byte sectors of 4 consecutive rows.
__device__ __forceinline__ void epilogue_block(const Params& p, const float* stg, int row0, int col0,
                                               int lane, size_t c_off) {
  const int cl = (lane & 7) * 4;
  const int col = col0 + cl;
  if (col >= p.N) return;  // N % 8 == 0, so a 4-column group is entirely in or out
  float bias[4] = {0.f, 0.f, 0.f, 0.f};
  if (p.epi == FX_EPI_BIAS || p.epi == FX_EPI_BIAS_GELU) {
    const uint2 b = __ldg(reinterpret_cast<const uint2*>(p.bias + col));
    bias[0] = bf16_lo(b.x); bias[1] = bf16_hi(b.x); bias[2] = bf16_lo(b.y); bias[3] = bf16_hi(b.y);
  }
  // Loads the epilogue depends on (saved pre-activation for GELU', old C for accumulate) are all
  // issued up front: one exposed HBM round trip per 32x32 block instead of one per row group.
  uint2 zpre[8];
  if (p.epi == FX_EPI_DGELU) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int row = row0 + i * 4 + (lane >> 3);
      zpre[i] = row < p.M ? __ldg(reinterpret_cast<const uint2*>(p.aux + (size_t)row * p.ldaux + col)) : make_uint2(0, 0);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int rl = i * 4 + (lane >> 3);
    const int row = row0 + rl;
    if (row >= p.M) continue;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = stg[rl * 33 + cl + j] + bias[j];

    if (p.epi == FX_EPI_BIAS_GELU) {
      if (p.aux != nullptr)
        *reinterpret_cast<uint2*>(p.aux + (size_t)row * p.ldaux + col) =
            make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
      // the activation sees the bf16-rounded pre-activation, exactly what a non-fused
      // linear -> gelu pair (hybrid_model.py:558-559) would see
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = gelu_tanh(__bfloat162float(__float2bfloat16_rn(v[j])));
    } else if (p.epi == FX_EPI_DGELU) {
      const uint2 z = zpre[i];
      v[0] *= gelu_tanh_grad(bf16_lo(z.x)); v[1] *= gelu_tanh_grad(bf16_hi(z.x));
      v[2] *= gelu_tanh_grad(bf16_lo(z.y)); v[3] *= gelu_tanh_grad(bf16_hi(z.y));
    }

    if (p.c_f32) {
      float4* cp = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.C) + c_off + (size_t)row * p.ldc + col);
      float4 o = make_float4(v[0], v[1], v[2], v[3]);
      if (p.accumulate) {
        const float4 old = *cp;
        o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
      }
      *cp = o;
    } else {
      uint2* cp = reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col);
      if (p.accumulate) {
        const uint2 old = *cp;
        v[0] += bf16_lo(old.x); v[1] += bf16_hi(old.x); v[2] += bf16_lo(old.y); v[3] += bf16_hi(old.y);
      }
      *cp = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
    }
  }
}

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
            const Params p) {
  using C = Cfg<BN>;
  extern __shared__ __align__(1024) uint8_t smem[];
  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) {
    printf("fleetx_b200: GEMM shared memory is not 1024-byte aligned\n");
    __trap();
  }
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
  uint64_t* empty = full + C::STAGES;
  uint64_t* tfull = empty + C::STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  float* staging = reinterpret_cast<float*>(smem + C::STAGES * C::STAGE_BYTES + C::BAR_BYTES);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == TMA_WARP && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < C::STAGES; ++i) {
        mbar_init(&full[i], 1);
        mbar_init(&empty[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull[i], 1);
        mbar_init(&tempty[i], NUM_EPI_WARPS);
      }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<C::TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == TMA_WARP) {
    // ---------------------------------------------------------------- TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < p.num_tiles * p.splits; item += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(p, item / p.splits, m_blk, n_blk);
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        const int kb0 = (item % p.splits) * p.kb_per_split;
        const int kbn = min(p.kb_per_split, p.num_kb - kb0);
        for (int kb = 0; kb < kbn; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          mbar_expect_tx(&full[stage], C::STAGE_BYTES);
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          uint8_t* sb = sa + A_STAGE_BYTES;
          const int k0 = (kb0 + kb) * BK;
          if (!p.a_mn) {
            tma_load_2d(sa, &tma_a, &full[stage], k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * CHUNK_BYTES, &tma_a, &full[stage], m0 + j * 64, k0);
          }
          if (!p.b_mn) {
            tma_load_2d(sb, &tma_b, &full[stage], k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * CHUNK_BYTES, &tma_b, &full[stage], n0 + j * 64, k0);
          }
          if (++stage == C::STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ---------------------------------------------------------------- MMA issuer
    const uint32_t idesc = instr_desc_bf16(BM, BN, (uint32_t)p.a_mn, (uint32_t)p.b_mn);
    // Descriptor low words: everything but the stage base and the k-step offset is loop-invariant,
    // and the issuing thread runs ~5 cycles per dependent instruction against 128 cycles per MMA.
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t a_lo0 = desc_lo(smem0, p.a_mn ? CHUNK_BYTES : 0);
    const uint32_t b_lo0 = desc_lo(smem0 + A_STAGE_BYTES, p.b_mn ? CHUNK_BYTES : 0);
    const uint32_t a_k = (p.a_mn ? 16 * 128 : 32) >> 4, b_k = (p.b_mn ? 16 * 128 : 32) >> 4;  // per 16-wide k-step
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < p.num_tiles * p.splits; item += gridDim.x) {
      mbar_wait(&tempty[acc], acc_phase ^ 1, 2);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      const int kbn = min(p.kb_per_split, p.num_kb - (item % p.splits) * p.kb_per_split);
      for (int kb = 0; kb < kbn; ++kb) {
        mbar_wait(&full[stage], phase, 3);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_lo = a_lo0 + stage * (C::STAGE_BYTES >> 4), b_lo = b_lo0 + stage * (C::STAGE_BYTES >> 4);
          umma_ss_lo<2>(tmem_d, a_lo, b_lo, idesc, kb != 0);
          umma_ss_lo<1>(tmem_d, a_lo + a_k, b_lo + b_k, idesc);
          umma_ss_lo<1>(tmem_d, a_lo + 2 * a_k, b_lo + 2 * b_k, idesc);
          umma_ss_lo<1>(tmem_d, a_lo + 3 * a_k, b_lo + 3 * b_k, idesc);
          umma_commit(&empty[stage]);
          if (kb == kbn - 1) umma_commit(&tfull[acc]);
        }
        __syncwarp();
        if (++stage == C::STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ---------------------------------------------------------------- epilogue
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    float* stg = staging + warp * (32 * 33);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < p.num_tiles * p.splits; item += gridDim.x) {
      int m_blk, n_blk;
      tile_coords(p, item / p.splits, m_blk, n_blk);
      const size_t c_off = static_cast<size_t>(item % p.splits) * p.split_stride;
      const int row0 = m_blk * BM + q * 32;
      const int n0 = n_blk * BN;
      mbar_wait(&tfull[acc], acc_phase, 4);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = warp >> 2; c < BN / 32; c += NUM_EPI_WARPS / 4) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        // thread = accumulator row -> staging row; (lane*33 + i) % 32 banks are conflict-free
#pragma unroll
        for (int i = 0; i < 32; ++i) stg[lane * 33 + i] = __uint_as_float(r[i]);
        __syncwarp();
        epilogue_block(p, stg, row0, n0 + c * 32, lane, c_off);
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

template <int BN>
static int launch(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb, Params& p) {
  using C = Cfg<BN>;
  static std::atomic<bool> attr_set[64];  // idempotent set-up: an atomic flag per device is enough
  int dev = 0;
  FX_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    FX_CUDA(cudaFuncSetAttribute(gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 C::SMEM_BYTES));
    attr_set[dev] = true;
  }
  p.num_m = ceil_div(p.M, BM);
  p.num_n = ceil_div(p.N, BN);
  p.num_tiles = p.num_m * p.num_n;
  p.num_kb = ceil_div(p.K, BK);
  if (p.splits < 1) p.splits = 1;
  p.kb_per_split = ceil_div(p.num_kb, p.splits);
  p.splits = ceil_div(p.num_kb, p.kb_per_split);  // no empty split
  const int items = p.num_tiles * p.splits;
  const int grid = items < sm_budget() ? items : sm_budget();
  gemm_kernel<BN><<<grid, NUM_THREADS, C::SMEM_BYTES, stream>>>(ta, tb, p);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

}  // namespace gemm
}  // namespace fx

namespace fx {
namespace gemm {

// out[i] = sum over splits of ws[s][i] (fp32 partial tiles of a split-K GEMM), as fp32 or bf16
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ ws, void* __restrict__ out, int out_f32, size_t n4, size_t stride4,
                     int splits) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < n4; i += static_cast<size_t>(gridDim.x) * 256) {
    float4 a = __ldg(reinterpret_cast<const float4*>(ws) + i);
    for (int sp = 1; sp < splits; ++sp) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(ws) + sp * stride4 + i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    if (out_f32) reinterpret_cast<float4*>(out)[i] = a;
    else reinterpret_cast<uint2*>(out)[i] = make_uint2(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w));
  }
}

// How many k-ranges a GEMM with few output tiles should be cut into so that its work items fill the machine
// (the weight gradients at a high mp degree: e.g. 256 x 2048 x 16384 is 16 tiles for 148 SMs).
static int auto_splits(int M, int N, int K) {
  const int ctas = sm_budget();
  const long tiles = (long)ceil_div(M, BM) * ceil_div(N, 256);
  if (tiles >= ctas || K < 4096) return 1;
  // time ~ waves of work items / splits (+ a little per split for the fp32 partial tiles and their sum);
  // keep >= 16 k-blocks per split so that the drain of a tile stays small next to its main loop
  int best = 1;
  float best_cost = 1e30f;
  for (int sp = 1; sp <= 8 && K / sp >= 1024; ++sp) {
    const long waves = (tiles * sp + ctas - 1) / ctas;
    const float cost = static_cast<float>(waves) / sp + 0.03f * (sp - 1);
    if (cost < best_cost - 1e-6f) {
      best_cost = cost;
      best = sp;
    }
  }
  return best;
}

}  // namespace gemm
}  // namespace fx

static int gemm_entry(void* stream, int trans_a, int trans_b, int M, int N, int K, const void* A, int lda,
                      const void* B, int ldb, void* C, int ldc, int c_dtype, int accumulate, int epilogue,
                      const void* bias, void* aux, int ldaux, int splits, void* split_ws) {
  using namespace fx;
  FX_CHECK_ARG(M > 0 && N > 0 && K > 0, "fx_gemm_bf16: M, N, K must be positive (got %d %d %d)", M, N, K);
  // N, K (and M when A is stored [K, M]) are contiguous dimensions of 16-byte vector / TMA accesses; a plain
  // row count M may be anything (decoding runs the linears on a handful of token rows)
  FX_CHECK_ARG(N % 8 == 0 && K % 8 == 0 && (!trans_a || M % 8 == 0),
               "fx_gemm_bf16: N, K (and M with trans_a) must be multiples of 8 (got M %d N %d K %d)", M, N, K);
  FX_CHECK_ARG(A && B && C, "fx_gemm_bf16: null operand");
  FX_CHECK_ARG(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "fx_gemm_bf16: leading dims must be multiples of 8");
  FX_CHECK_ARG(lda >= (trans_a ? M : K) && ldb >= (trans_b ? K : N) && ldc >= N,
               "fx_gemm_bf16: leading dimension smaller than the row length");
  FX_CHECK_ARG(c_dtype == FX_BF16 || c_dtype == FX_F32, "fx_gemm_bf16: bad c_dtype %d", c_dtype);
  FX_CHECK_ARG(epilogue >= FX_EPI_NONE && epilogue <= FX_EPI_DGELU, "fx_gemm_bf16: bad epilogue %d", epilogue);
  FX_CHECK_ARG(!(epilogue == FX_EPI_BIAS || epilogue == FX_EPI_BIAS_GELU) || bias,
               "fx_gemm_bf16: epilogue needs a bias");
  FX_CHECK_ARG(epilogue != FX_EPI_DGELU || aux, "fx_gemm_bf16: FX_EPI_DGELU needs aux");
  FX_CHECK_ARG(!aux || (ldaux % 8 == 0 && ldaux >= N), "fx_gemm_bf16: bad ldaux %d", ldaux);
  FX_CHECK_ARG(((uintptr_t)A | (uintptr_t)B | (uintptr_t)C | (uintptr_t)bias | (uintptr_t)aux) % 16 == 0,
               "fx_gemm_bf16: pointers must be 16-byte aligned");

  gemm::Params p{};
  const int sms = sm_budget();
  // Tile time is bound by L2->SM operand traffic per k-block: (128 + BN) * 64 * 2 bytes, i.e.
  // 48 KB for BN = 256 and 32 KB for BN = 128 (for half the flops) — weigh waves by that.
  auto cost = [&](int bn) {
    const long tiles = (long)ceil_div(M, gemm::BM) * ceil_div(N, bn);
    return ((tiles + sms - 1) / sms) * (bn == 256 ? 48 : 32);
  };
  int BN = (N > 128 && cost(256) <= cost(128)) ? 256 : 128;
  if (splits > 1 && N > 128) BN = 256;  // the split count was chosen for 256-wide tiles
  p.group = 8;
  if (knobs().gemm_bn) BN = knobs().gemm_bn;            // tuning knobs (read once, runtime.cu)
  if (knobs().gemm_group) p.group = knobs().gemm_group;

  CUtensorMap ta, tb;
  int rc;
  if (!trans_a) rc = make_tma_bf16_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, 64, gemm::BM);
  else rc = make_tma_bf16_2d(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 64, gemm::BK);
  if (rc) return rc;
  if (trans_b) rc = make_tma_bf16_2d(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, 64, BN);
  else rc = make_tma_bf16_2d(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, gemm::BK);
  if (rc) return rc;

  p.M = M; p.N = N; p.K = K;
  p.a_mn = trans_a ? 1 : 0;
  p.b_mn = trans_b ? 0 : 1;
  p.C = C; p.ldc = ldc; p.c_f32 = (c_dtype == FX_F32); p.accumulate = accumulate ? 1 : 0;
  p.epi = epilogue;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.aux = static_cast<__nv_bfloat16*>(aux);
  p.ldaux = ldaux;
  p.splits = 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (splits > 1) {
    // every split writes its fp32 partial tiles into its own [M, N] slab of the workspace; a second kernel sums them
    p.C = split_ws; p.ldc = N; p.c_f32 = 1; p.accumulate = 0;
    p.splits = splits;
    p.split_stride = static_cast<size_t>(M) * N;
    rc = BN == 256 ? gemm::launch<256>(st, ta, tb, p) : gemm::launch<128>(st, ta, tb, p);
    if (rc) return rc;
    const size_t n4 = static_cast<size_t>(M) * N / 4;
    const int blocks = static_cast<int>(n4 / 256 + 1 < (size_t)(4 * num_sms()) ? n4 / 256 + 1 : 4 * num_sms());
    gemm::splitk_reduce_kernel<<<blocks, 256, 0, st>>>(static_cast<const float*>(split_ws), C, c_dtype == FX_F32, n4, n4,
                                                       p.splits);
    FX_LAUNCH_CHECK();
    return FX_OK;
  }
  return BN == 256 ? gemm::launch<256>(st, ta, tb, p) : gemm::launch<128>(st, ta, tb, p);
}

extern "C" int fx_gemm_bf16(void* stream, int trans_a, int trans_b, int M, int N, int K,
                            const void* A, int lda, const void* B, int ldb, void* C, int ldc,
                            int c_dtype, int accumulate, int epilogue, const void* bias, void* aux,
                            int ldaux) {
  return gemm_entry(stream, trans_a, trans_b, M, N, K, A, lda, B, ldb, C, ldc, c_dtype, accumulate, epilogue, bias, aux,
                    ldaux, 1, nullptr);
}

extern "C" int fx_gemm_splitk_splits(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 1;
  return fx::gemm::auto_splits(M, N, K);
}

extern "C" size_t fx_gemm_splitk_workspace_bytes(int M, int N, int splits) {
  if (M <= 0 || N <= 0 || splits <= 1) return 0;
  return static_cast<size_t>(splits) * M * N * sizeof(float);
}

extern "C" int fx_gemm_bf16_splitk(void* stream, int trans_a, int trans_b, int M, int N, int K, const void* A, int lda,
                                   const void* B, int ldb, void* C, int c_dtype, int splits, void* workspace) {
  using namespace fx;
  FX_CHECK_ARG(splits >= 1 && splits <= 16, "fx_gemm_bf16_splitk: splits must be in [1, 16] (got %d)", splits);
  FX_CHECK_ARG(splits == 1 || (workspace != nullptr && (uintptr_t)workspace % 16 == 0),
               "fx_gemm_bf16_splitk: null or misaligned workspace");
  FX_CHECK_ARG(M > 0 && N > 0 && ((long)M * N) % 4 == 0, "fx_gemm_bf16_splitk: M * N must be a positive multiple of 4");
  return gemm_entry(stream, trans_a, trans_b, M, N, K, A, lda, B, ldb, C, N, c_dtype, 0, FX_EPI_NONE, nullptr, nullptr, 0,
                    splits, workspace);
}
