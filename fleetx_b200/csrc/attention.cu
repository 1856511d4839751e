// attention.cu — K3/K4: fused causal self-attention, forward and backward, on tcgen05.
//
// Replaces hybrid_model.py:165-186 (_fuse_prepare_qkv: reshape/transpose/split of the packed
// qkv_proj output) and :268-298 (core_attn: scaled QK^T, causal softmax, dropout under the
// 'local_seed' stream, PV, head merge). The s x s score / probability tensors the reference
// writes three times never leave the SM here.
//
// Data layout: qkv [b, s, nh, 3, hd] bf16 exactly as qkv_proj produced it — Q, K and V tiles
// are cut out of it by ONE 4-D TMA tensor map (dims hd | 3·nh | s | b), so the head split costs
// no copy; out / dout are [b, s, nh·hd] (heads merged), dqkv has the qkv layout again.
//
// Forward (one CTA = 128 query rows of one head; 320 threads):
//   warp 8     TMA: Q once, then K_j / V_j through a 2-deep ring
//   warp 9     MMA: S_j = Q·K_j^T (128x128, K = hd) and PV_j = P_j·V_j (128 x hd, K = 128) into
//              TMEM; V is consumed as an MN-major operand straight from the same TMA box
//   warps 0-7  softmax: two threads per query row (64 kv columns each; the row maximum is exchanged
//              through shared memory); two sweeps over S in TMEM (max, then exp2),
//              P_j (with dropout folded in) goes back to shared memory as bf16 in the
//              128B-swizzled K-major layout the PV MMA reads; O is kept in registers and updated
//              as O = alpha·O + PV_j one step behind, so the tensor core always has work queued
// Backward (two kernels, deterministic, no atomics):
//   dKV kernel: CTA owns a 128-row KV tile, walks the query tiles at or below the diagonal:
//              S^T = K·Q^T, dP^T = V·dO^T, dV += P^T·dO, dK += dS^T·Q
//   dQ  kernel: CTA owns a 128-row Q tile, walks KV tiles: S = Q·K^T, dP = dO·V^T, dQ += dS·K
//   with delta = rowsum(dO ∘ O) from a small pre-pass (workspace, fp32 [b, nh, s]).
#include <stdlib.h>
#include <type_traits>

#include <atomic>

#include "common.cuh"
#include "philox.cuh"
#include "tc5.cuh"

namespace fx {
namespace attn {

using namespace tc5;

constexpr int FWD_SM_WARPS = 8;  // softmax warps: two per TMEM lane quarter, 64 of a tile's 128 kv columns each
constexpr int NUM_THREADS = 64 + FWD_SM_WARPS * 32;
// Warp roles. The single-thread TMA / MMA issuers sit in the HIGHEST warp ids: the SM's warp arbiter
// favours higher ids, and an MMA issuer starved of issue slots by the math warps stalls the tensor pipe.
constexpr int FWD_TMA_WARP = FWD_SM_WARPS, FWD_MMA_WARP = FWD_SM_WARPS + 1;  // warps 0..7: softmax
constexpr float LOG2E = 1.4426950408889634f;

// byte offset of 16-byte chunk `c16` of row `r` inside a [rows x 64 bf16] 128B-swizzled tile
__device__ __forceinline__ uint32_t sw128_off(int r, int c16) {
  return static_cast<uint32_t>(r * 128 + ((c16 ^ (r & 7)) << 4));
}

__device__ __forceinline__ void named_bar_sync_fwd(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

struct FwdParams {
  __nv_bfloat16* out;
  float* lse;
  int b, s, nh;
  float scale, p_drop, inv_keep, log2_inv_keep;
  const uint32_t* mask;     // dropout keep bits [b, nh, s, wpr] (see attn_dropout_mask_kernel), or null
  int wpr;                  // 32-bit mask words per query row = ceil(s / 128) * 4
};

// ---------------------------------------------------------------------------- dropout keep bits
// The keep bits of attention dropout are drawn ONCE per layer call by this kernel and read by the
// forward and both backward kernels (Philox inside those kernels cost more issue slots than the
// softmax itself). Definition unchanged (philox.cuh): query row q, 8 consecutive kv per Philox call,
// group index ((batch·nh + head)·s + q)·ceil(s/8) + kv/8.
//
// Layout: word w of row q covers kv in [32w, 32w+32). Element e = 4k + 2u + t of the word (k < 8,
// u, t < 2) sits at bit 8·(2u + t) + 7 − k: after `word << k` the four elements 4k..4k+3 are the sign
// bits of bytes 0..3, so one PRMT (sign-replicate mode) turns them into bf16x2 or fp32 AND-masks.
struct MaskParams {
  uint32_t* mask;
  int b, s, nh, wpr;
  uint32_t thresh2, off_lo, off_hi;
  uint64_t bh0;  // (batch, head) index of this call's first pair inside the batch the mask is defined on
  PhiloxKeys keys;
};

constexpr int MASK_THREADS = 512;  // one such block fits beside a resident GEMM CTA (no shared memory, 32 registers)
__global__ void __launch_bounds__(MASK_THREADS)
attn_dropout_mask_kernel(const MaskParams p) {
  // CTA = 32 query rows (lane = row) of one (batch, head); its 16 warps walk the kv words 16 at a time up
  // to the end of the diagonal 128-tile (tiles beyond it are never read: causal).
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q = blockIdx.x * 32 + lane;
  if (q >= p.s) return;
  const int n_words = min(p.wpr, (blockIdx.x * 32 / 128 + 1) * 4);
  const uint64_t bh = blockIdx.y;
  const uint64_t s8 = static_cast<uint64_t>((p.s + 7) / 8);
  const uint64_t grp_row = ((p.bh0 + bh) * p.s + q) * s8;
  uint32_t* out_row = p.mask + (bh * p.s + q) * p.wpr;
  for (int w = warp; w < n_words; w += MASK_THREADS / 32) {
    uint32_t word = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const Philox8 rnd = philox4x32_10_rk(p.keys, p.off_lo, p.off_hi, grp_row + static_cast<uint64_t>(w * 4 + c));
#pragma unroll
      for (int half = 0; half < 2; ++half) {  // k = 2c + half: lanes[2·half] holds u = 0, lanes[2·half + 1] holds u = 1
        const uint32_t f0 = keep_flags2(rnd.lanes[2 * half], p.thresh2);
        const uint32_t f1 = keep_flags2(rnd.lanes[2 * half + 1], p.thresh2);
        // bytes [f0.b1, f0.b3, f1.b1, f1.b3]: their sign bits are the keep flags of elements 4k .. 4k+3
        const uint32_t t = prmt(f0, f1, 0x7531u) & 0x80808080u;
        word |= t >> (2 * c + half);
      }
    }
    out_row[w] = word;
  }
}

template <int HD>
struct FwdCfg {
  static constexpr int TILE_BYTES = 128 * HD * 2;  // one Q / K / V tile
  static constexpr int SUB_BYTES = 128 * 128;      // one [128 rows x 64 cols] swizzled sub-tile
  static constexpr int P_BYTES = 128 * 128 * 2;
  static constexpr int NS = HD == 64 ? 1 : 2;      // S accumulators in TMEM
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_K = TILE_BYTES;
  static constexpr int OFF_V = 3 * TILE_BYTES;
  static constexpr int OFF_P = 5 * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_P + P_BYTES;   // 19 mbarriers + TMEM slot: 160 bytes
  static constexpr int OFF_XCH = OFF_BAR + 160;     // [2 halves][128 rows] bf16: row-max exchange between the
                                                    // two threads of a row (bf16 keeps 2 CTAs/SM within smem)
  static constexpr int SMEM_BYTES = OFF_XCH + 512;
  static constexpr int TMEM_PV = NS * 128;         // column of the first PV accumulator
  static constexpr int TMEM_COLS = HD == 64 ? 256 : 512;
  static constexpr int MIN_CTAS = HD == 64 ? 2 : 1;
};

template <int HD>
__global__ void __launch_bounds__(NUM_THREADS, FwdCfg<HD>::MIN_CTAS)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tma_qkv, const FwdParams p) {
  using C = FwdCfg<HD>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // 2
  uint64_t* k_empty = bars + 3;       // 2
  uint64_t* v_full = bars + 5;        // 2
  uint64_t* v_empty = bars + 7;       // 2
  uint64_t* s_full = bars + 9;        // 2
  uint64_t* s_empty = bars + 11;      // 2
  uint64_t* p_full = bars + 13;       // 1
  uint64_t* p_empty = bars + 14;      // 1
  uint64_t* pv_full = bars + 15;      // 2
  uint64_t* pv_empty = bars + 17;     // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_qt = (p.s + 127) / 128;
  const int qt = num_qt - 1 - blockIdx.x;  // heaviest (longest KV walk) tiles first
  const int head = blockIdx.y;
  const int batch = blockIdx.z;
  const int n_kv = qt + 1;

  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) {
    printf("fleetx_b200: attention shared memory is not 1024-byte aligned\n");
    __trap();
  }
  if (warp == FWD_TMA_WARP && lane == 0) tma_prefetch_desc(&tma_qkv);
  if (warp == FWD_MMA_WARP) {
    if (lane == 0) {
      mbar_init(q_full, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&k_full[i], 1);
        mbar_init(&k_empty[i], 1);
        mbar_init(&v_full[i], 1);
        mbar_init(&v_empty[i], 1);
        mbar_init(&s_full[i], 1);
        mbar_init(&s_empty[i], FWD_SM_WARPS);
        mbar_init(&pv_full[i], 1);
        mbar_init(&pv_empty[i], FWD_SM_WARPS);
      }
      mbar_init(p_full, FWD_SM_WARPS);
      mbar_init(p_empty, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<C::TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == FWD_TMA_WARP) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      mbar_expect_tx(q_full, C::TILE_BYTES);
#pragma unroll
      for (int h2 = 0; h2 < HD / 64; ++h2)
        tma_load_4d(smem + C::OFF_Q + h2 * C::SUB_BYTES, &tma_qkv, q_full, h2 * 64, head * 3 + 0, qt * 128, batch);
      for (int j = 0; j < n_kv; ++j) {
        const int buf = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&k_empty[buf], ph ^ 1, 10);
        mbar_expect_tx(&k_full[buf], C::TILE_BYTES);
#pragma unroll
        for (int h2 = 0; h2 < HD / 64; ++h2)
          tma_load_4d(smem + C::OFF_K + buf * C::TILE_BYTES + h2 * C::SUB_BYTES, &tma_qkv, &k_full[buf], h2 * 64,
                      head * 3 + 1, j * 128, batch);
        mbar_wait(&v_empty[buf], ph ^ 1, 11);
        mbar_expect_tx(&v_full[buf], C::TILE_BYTES);
#pragma unroll
        for (int h2 = 0; h2 < HD / 64; ++h2)
          tma_load_4d(smem + C::OFF_V + buf * C::TILE_BYTES + h2 * C::SUB_BYTES, &tma_qkv, &v_full[buf], h2 * 64,
                      head * 3 + 2, j * 128, batch);
      }
    }
  } else if (warp == FWD_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t idesc_s = instr_desc_bf16(128, 128, 0, 0);
    const uint32_t idesc_pv = instr_desc_bf16(128, HD, 0, 1);
    const uint32_t q_addr = smem_u32(smem + C::OFF_Q);
    const uint32_t p_addr = smem_u32(smem + C::OFF_P);

    // descriptor low words (see tc5.cuh: the issuing thread must spend only a few instructions per MMA)
    const uint32_t q_lo = desc_lo(q_addr, 0), p_lo = desc_lo(p_addr, 0);
    const uint32_t k_lo0 = desc_lo(smem_u32(smem + C::OFF_K), 0);
    const uint32_t v_lo0 = desc_lo(smem_u32(smem + C::OFF_V), C::SUB_BYTES);

    auto issue_s = [&](int j) {
      const int buf = j & 1, sb = j % C::NS;
      mbar_wait(&k_full[buf], (j >> 1) & 1, 20);
      mbar_wait(&s_empty[sb], ((j / C::NS) & 1) ^ 1, 21);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t k_lo = k_lo0 + buf * (C::TILE_BYTES >> 4);
        const uint32_t d = tmem_base + sb * 128;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t off = ((kk >> 2) * C::SUB_BYTES + (kk & 3) * 32) >> 4;
          if (kk == 0) umma_ss_lo<0>(d, q_lo + off, k_lo + off, idesc_s);
          else umma_ss_lo<1>(d, q_lo + off, k_lo + off, idesc_s);
        }
        umma_commit(&s_full[sb]);
        umma_commit(&k_empty[buf]);
      }
      __syncwarp();
    };

    mbar_wait(q_full, 0, 22);
    issue_s(0);
    for (int j = 0; j < n_kv; ++j) {
      if (j + 1 < n_kv) issue_s(j + 1);
      const int buf = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      mbar_wait(p_full, j & 1, 23);
      mbar_wait(&v_full[buf], ph, 24);
      mbar_wait(&pv_empty[buf], ph ^ 1, 25);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t v_lo = v_lo0 + buf * (C::TILE_BYTES >> 4);
        const uint32_t d = tmem_base + C::TMEM_PV + buf * HD;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t offp = ((kk >> 2) * C::SUB_BYTES + (kk & 3) * 32) >> 4;
          if (kk == 0) umma_ss_lo<0>(d, p_lo + offp, v_lo + kk * (2048 >> 4), idesc_pv);
          else umma_ss_lo<1>(d, p_lo + offp, v_lo + kk * (2048 >> 4), idesc_pv);
        }
        umma_commit(&pv_full[buf]);
        umma_commit(&v_empty[buf]);
        umma_commit(p_empty);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------ softmax / output
    // Thread pair (warp w, warp w+4; same lane) shares a query row: each takes 64 of the tile's 128 kv
    // columns and half of the row's output columns.
    const int quarter = warp & 3, half = warp >> 2;
    const int row_local = quarter * 32 + lane;
    const int row = qt * 128 + row_local;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const uint32_t p_addr = smem_u32(smem + C::OFF_P) + half * C::SUB_BYTES;  // sub-tile `half` = my 64 columns
    const float sl2 = p.scale * LOG2E;
    const bool drop = p.p_drop > 0.f;
    // this query row's dropout keep bits: 4 words per 128-wide kv tile, two of them mine
    const uint2* mask_row = reinterpret_cast<const uint2*>(
        p.mask + (static_cast<size_t>(batch * p.nh + head) * p.s + min(row, p.s - 1)) * p.wpr) + half;
    __nv_bfloat16* xch = reinterpret_cast<__nv_bfloat16*>(smem + C::OFF_XCH);
    constexpr int OC = HD / 2;  // output columns per thread

    uint64_t o2[OC / 2];  // packed fp32 pairs (FFMA2): o2[i] = output columns 2i, 2i+1 of my half
#pragma unroll
    for (int i = 0; i < OC / 2; ++i) o2[i] = 0ull;
    float m = -INFINITY, l = 0.f, alpha_prev = 0.f;

    auto o_update = [&](int j, float alpha) {
      const int buf = j & 1;
      mbar_wait(&pv_full[buf], (j >> 1) & 1, 30);
      tc_fence_after();
      const uint64_t alpha2 = f2_pack(alpha, alpha);
#pragma unroll
      for (int c = 0; c < OC / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(lane_base + C::TMEM_PV + buf * HD + half * OC + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) o2[c * 16 + i] = fma2(o2[c * 16 + i], alpha2, f2_pack_u(r[2 * i], r[2 * i + 1]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&pv_empty[buf]);
    };

    // Dropout's 1/(1-p) is folded into the exponent (p' = 2^(s·sl2 - m·sl2 + log2 inv_keep)), so the
    // row sum below runs over p' too and is scaled back once at the end.
    const float log2_inv_keep = p.log2_inv_keep;
    const uint64_t sl2_2 = f2_pack(sl2, sl2);

    // One kv tile. DIAG (only the diagonal tile needs the causal predicate) is a compile-time flag: as a
    // run-time one the compiler if-converted it into a compare + select per element of EVERY tile.
    auto tile = [&](int j, auto diag_c) {
      constexpr bool DIAG = decltype(diag_c)::value;
      const int sb = j % C::NS;
      uint2 mw = make_uint2(0, 0);
      if (drop) mw = __ldg(mask_row + 2 * j);  // issued before the wait: the load overlaps sweep 1
      mbar_wait(&s_full[sb], (j / C::NS) & 1, 31);
      tc_fence_after();
      // sweep 1: maximum of my 64 raw scores
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(lane_base + sb * 128 + half * 64 + c * 32, r);
        tmem_ld_wait();
        if (DIAG) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            mx = fmaxf(mx, (half * 64 + c * 32 + i > row_local) ? -INFINITY : __uint_as_float(r[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; i += 2) mx = fmaxf(mx, fmaxf(__uint_as_float(r[i]), __uint_as_float(r[i + 1])));
        }
      }
      // Both threads of the row must scale with the SAME maximum. Each publishes its own, rounded UP to
      // bf16 (any bound >= the true maximum is a valid stabiliser), and both take the max of the two.
      const __nv_bfloat16 mine_up = __float2bfloat16_ru(mx);
      xch[half * 128 + row_local] = mine_up;
      named_bar_sync_fwd(1 + quarter, 64);
      const float other = __bfloat162float(xch[(half ^ 1) * 128 + row_local]);
      named_bar_sync_fwd(1 + quarter, 64);  // partner has read before the next tile overwrites
      const float m_new = fmaxf(m, fmaxf(__bfloat162float(mine_up), other));
      const float alpha = ex2_approx((m - m_new) * sl2);
      const float neg_m = log2_inv_keep - m_new * sl2;
      const uint64_t neg_m2 = f2_pack(neg_m, neg_m);
      // P buffer must have been consumed by PV_{j-1}
      mbar_wait(p_empty, (j & 1) ^ 1, 32);
      // sweep 2: probabilities -> bf16 P tile (dropout folded in), row sum of the undropped p'
      uint64_t rowsum2 = 0ull;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(lane_base + sb * 128 + half * 64 + c * 32, r);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const uint64_t a = fma2(f2_pack_u(r[i], r[i + 1]), sl2_2, neg_m2);
          float e0 = ex2_approx(f2_lo(a));
          float e1 = ex2_approx(f2_hi(a));
          if (DIAG) {
            e0 = (half * 64 + c * 32 + i > row_local) ? 0.f : e0;
            e1 = (half * 64 + c * 32 + i + 1 > row_local) ? 0.f : e1;
          }
          rowsum2 = add2(rowsum2, f2_pack(e0, e1));
          pk[i >> 1] = pack_bf16(e0, e1);
        }
        if (drop) {
          const uint32_t word = c == 0 ? mw.x : mw.y;
#pragma unroll
          for (int k = 0; k < 8; ++k) {  // sign bits of bytes 0..3 of (word << k) = elements 4k .. 4k+3
            const uint32_t x = word << k;
            pk[2 * k] &= prmt(x, 0u, 0x9988u);
            pk[2 * k + 1] &= prmt(x, 0u, 0xBBAAu);
          }
        }
        // my columns [c*32, c*32+32) of sub-tile `half`: 16-byte chunks c*4 .. c*4+3
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(p_addr + sw128_off(row_local, c * 4 + g), pk[g * 4 + 0], pk[g * 4 + 1], pk[g * 4 + 2],
                       pk[g * 4 + 3]);
      }
      tc_fence_before();
      fence_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&s_empty[sb]);
        mbar_arrive(p_full);
      }
      l = l * alpha + (f2_lo(rowsum2) + f2_hi(rowsum2));
      m = m_new;
      if (j > 0) o_update(j - 1, alpha_prev);
      alpha_prev = alpha;
    };

    for (int j = 0; j < n_kv - 1; ++j) tile(j, std::false_type{});
    tile(n_kv - 1, std::true_type{});  // n_kv - 1 == qt: the diagonal tile
    o_update(n_kv - 1, alpha_prev);

    // the two partial row sums share every rescale, so they simply add; exchange them in fp32 through
    // the P tile, which nothing reads any more (PV of the last tile has completed: o_update waited)
    float* lx = reinterpret_cast<float*>(smem + C::OFF_P);
    lx[half * 128 + row_local] = l;
    named_bar_sync_fwd(1 + quarter, 64);
    l += lx[(half ^ 1) * 128 + row_local];

    if (row < p.s) {
      const float inv_l = p.inv_keep / l;  // l sums p' = p * inv_keep
      uint4* op = reinterpret_cast<uint4*>(p.out + (static_cast<size_t>(batch) * p.s + row) * (p.nh * HD) + head * HD +
                                           half * OC);
#pragma unroll
      for (int g = 0; g < OC / 8; ++g)
        op[g] = make_uint4(pack_bf16(f2_lo(o2[g * 4 + 0]) * inv_l, f2_hi(o2[g * 4 + 0]) * inv_l),
                           pack_bf16(f2_lo(o2[g * 4 + 1]) * inv_l, f2_hi(o2[g * 4 + 1]) * inv_l),
                           pack_bf16(f2_lo(o2[g * 4 + 2]) * inv_l, f2_hi(o2[g * 4 + 2]) * inv_l),
                           pack_bf16(f2_lo(o2[g * 4 + 3]) * inv_l, f2_hi(o2[g * 4 + 3]) * inv_l));
      if (half == 0)
        p.lse[(static_cast<size_t>(batch) * p.nh + head) * p.s + row] = m * p.scale + __logf(l) - __logf(p.inv_keep);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == FWD_MMA_WARP) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

static int make_qkv_map(CUtensorMap* map, const void* qkv, int b, int s, int nh, int hd) {
  const uint64_t dims[4] = {(uint64_t)hd, (uint64_t)3 * nh, (uint64_t)s, (uint64_t)b};
  const uint64_t strides[3] = {(uint64_t)hd * 2, (uint64_t)nh * 3 * hd * 2, (uint64_t)s * nh * 3 * hd * 2};
  const uint32_t box[4] = {64, 1, 128, 1};
  return make_tma_bf16(map, qkv, 4, dims, strides, box);
}

template <int HD>
static int launch_fwd(cudaStream_t st, const CUtensorMap& map, const FwdParams& p) {
  using C = FwdCfg<HD>;
  static std::atomic<bool> attr_set[64];  // idempotent set-up: an atomic flag per device is enough
  int dev = 0;
  FX_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    FX_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set[dev] = true;
  }
  dim3 grid((p.s + 127) / 128, p.nh, p.b);
  attn_fwd_kernel<HD><<<grid, NUM_THREADS, C::SMEM_BYTES, st>>>(map, p);
  FX_LAUNCH_CHECK();
  return FX_OK;
}


// =====================================================================================
// Backward
// =====================================================================================
constexpr int BWD_EW_WARPS = 16;  // elementwise warps: 4 per TMEM lane quarter, 16 of a tile's 64 columns each
constexpr int BWD_THREADS = 64 + BWD_EW_WARPS * 32;  // warps 0..15 elementwise, then TMA, then MMA (see FWD_MMA_WARP)
constexpr int BWD_TMA_WARP = BWD_EW_WARPS, BWD_MMA_WARP = BWD_EW_WARPS + 1;

#ifdef FX_TRACE
// Debug-only phase timing (tools/micro/trace_attn.py builds a private copy of this file with
// -DFX_TRACE). g_trace[i] accumulates clock64 deltas of phase i for one elementwise warp per CTA.
__device__ unsigned long long g_trace[16];
#define FX_T0() long long t_trace = clock64()
#define FX_T(i)                                                                       \
  do {                                                                                \
    if (warp == 0 && lane == 0) {                                                     \
      const long long n_ = clock64();                                                 \
      atomicAdd(&g_trace[i], static_cast<unsigned long long>(n_ - t_trace));          \
      t_trace = n_;                                                                   \
    }                                                                                 \
  } while (0)
#define FX_TM(i)                                                                      \
  do {                                                                                \
    if (lane == 0) {                                                                  \
      const long long n_ = clock64();                                                 \
      atomicAdd(&g_trace[i], static_cast<unsigned long long>(n_ - t_trace));          \
      t_trace = n_;                                                                   \
    }                                                                                 \
  } while (0)
#else
#define FX_T0() do {} while (0)
#define FX_T(i) do {} while (0)
#define FX_TM(i) do {} while (0)
#endif

struct BwdParams {
  // both statistics are stored NEGATED: they enter the element-wise math as addends of packed FMA / ADD
  const float* lse2;    // [b, nh, s_pad]  −(lse·log2(e) − log2(inv_keep)); −inf in the padding
  const float* delta;   // [b, nh, s_pad]  −rowsum(dO ∘ O) / inv_keep;       0 in the padding
  __nv_bfloat16* dqkv;  // [b, s, nh, 3, hd]
  int b, s, nh, s_pad;  // s_pad = s rounded up to 128
  float scale, p_drop, inv_keep;
  const uint32_t* mask;  // dropout keep bits [b, nh, s, wpr], or null
  int wpr;
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Row statistics the two backward kernels read per query row, padded so tiles never need a bound
// check: with dropout folded in as p' = p·inv_keep,
//   dS = p'·(keep·dP − delta'),  p' = 2^(s·scale·log2e − lse2),
//   lse2 = lse·log2e − log2(inv_keep),  delta' = rowsum(dO ∘ O) / inv_keep.
// Both are stored negated (they are addends of a packed FMA / ADD in the kernels). Padding rows get
// −lse2 = −inf (p' = 0) and delta' = 0.
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ out,
                  const float* __restrict__ lse, float* __restrict__ lse2, float* __restrict__ delta, int b, int s,
                  int nh, int hd, int s_pad, float log2_inv_keep, float keep) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // over [b, s_pad, nh]
  const size_t total = (size_t)b * s_pad * nh;
  if (idx >= total) return;
  const int head = (int)(idx % nh);
  const size_t br = idx / nh;
  const int row = (int)(br % s_pad);
  const int batch = (int)(br / s_pad);
  const size_t o_idx = ((size_t)batch * nh + head) * s_pad + row;
  if (row >= s) {
    lse2[o_idx] = -INFINITY;
    delta[o_idx] = 0.f;
    return;
  }
  const size_t i_idx = (((size_t)batch * s + row) * nh + head) * hd;
  const uint4* a = reinterpret_cast<const uint4*>(dout + i_idx);
  const uint4* o = reinterpret_cast<const uint4*>(out + i_idx);
  float acc = 0.f;
  for (int v = 0; v < hd / 8; ++v) {
    const uint4 x = __ldg(a + v), y = __ldg(o + v);
    const uint32_t xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) acc += bf16_lo(xs[j]) * bf16_lo(ys[j]) + bf16_hi(xs[j]) * bf16_hi(ys[j]);
  }
  delta[o_idx] = -acc * keep;
  lse2[o_idx] = log2_inv_keep - lse[((size_t)batch * nh + head) * s + row] * LOG2E;
}

// TMEM accumulator columns [taddr, taddr + NCOLS) of this thread's row -> bf16, times `mul`, to dst.
template <int NCOLS>
__device__ __forceinline__ void drain_scaled(uint32_t taddr, __nv_bfloat16* dst, float mul, bool valid) {
  static_assert(NCOLS == 16 || NCOLS == 32, "");
  uint32_t r[NCOLS];
  if constexpr (NCOLS == 16) tmem_ld_32x16(taddr, r);
  else tmem_ld_32x32(taddr, r);
  tmem_ld_wait();
  if (valid) {
    uint4* o = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int g = 0; g < NCOLS / 8; ++g)
      o[g] = make_uint4(pack_bf16(__uint_as_float(r[g * 8 + 0]) * mul, __uint_as_float(r[g * 8 + 1]) * mul),
                        pack_bf16(__uint_as_float(r[g * 8 + 2]) * mul, __uint_as_float(r[g * 8 + 3]) * mul),
                        pack_bf16(__uint_as_float(r[g * 8 + 4]) * mul, __uint_as_float(r[g * 8 + 5]) * mul),
                        pack_bf16(__uint_as_float(r[g * 8 + 6]) * mul, __uint_as_float(r[g * 8 + 7]) * mul));
  }
}

// ---------------------------------------------------------------------------- dQ kernel
template <int HD>
struct DqCfg {
  static constexpr int QT_BYTES = 128 * HD * 2;   // Q / dO tile (128 rows)
  static constexpr int KV_BYTES = 64 * HD * 2;    // K / V tile (64 rows)
  static constexpr int KV_SUB = 64 * 128;         // one 64-col sub-tile of a 64-row tile
  // K/V ring depth: a slot is only released when the dQ MMA that reads K retires, so the ring must
  // be deep enough for the next loads to be in flight a few tiles ahead of their use
  static constexpr int NKV = HD == 64 ? 4 : 3;
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_DO = QT_BYTES;
  static constexpr int OFF_K = 2 * QT_BYTES;
  static constexpr int OFF_V = OFF_K + NKV * KV_BYTES;
  static constexpr int OFF_BAR = OFF_V + NKV * KV_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  // TMEM: S[NB] | dP[NB] | dQ. dS (bf16) is written back over the dP columns it came from and is
  // the A operand of the dQ MMA straight from TMEM. Three S/dP buffers let the MMA warp run two
  // tiles ahead of the elementwise warps.
  static constexpr int NB = 3;
  static constexpr int TM_S = 0, TM_DP = NB * 64, TM_DQ = 2 * NB * 64;
  static constexpr int TMEM_COLS = 512;
};

template <int HD>
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tma_q,   // qkv, box rows 128
                   const __grid_constant__ CUtensorMap tma_kv,  // qkv, box rows 64
                   const __grid_constant__ CUtensorMap tma_do,  // dout, box rows 128
                   const BwdParams p) {
  using C = DqCfg<HD>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  uint64_t* qdo_full = bars;        // 1
  uint64_t* kv_full = bars + 1;     // NKV (<= 4)
  uint64_t* kv_empty = bars + 5;    // NKV
  uint64_t* s_full = bars + 9;      // NB (<= 4)
  uint64_t* ds_full = bars + 13;    // NB
  uint64_t* dq_full = bars + 17;    // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_qt = (p.s + 127) / 128;
  const int qt = num_qt - 1 - blockIdx.x;
  const int head = blockIdx.y, batch = blockIdx.z;
  const int n_kv = min((p.s + 63) / 64, 2 * qt + 2);

  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) __trap();
  if (warp == BWD_TMA_WARP && lane == 0) {
    tma_prefetch_desc(&tma_q);
    tma_prefetch_desc(&tma_kv);
    tma_prefetch_desc(&tma_do);
  }
  if (warp == BWD_MMA_WARP) {
    if (lane == 0) {
      mbar_init(qdo_full, 1);
      mbar_init(dq_full, 1);
      for (int i = 0; i < C::NKV; ++i) {
        mbar_init(&kv_full[i], 1);
        mbar_init(&kv_empty[i], 1);
      }
      for (int i = 0; i < C::NB; ++i) {
        mbar_init(&s_full[i], 1);
        mbar_init(&ds_full[i], BWD_EW_WARPS);
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

  if (warp == BWD_TMA_WARP) {
    if (elect_one()) {
      mbar_expect_tx(qdo_full, 2 * C::QT_BYTES);
#pragma unroll
      for (int h2 = 0; h2 < HD / 64; ++h2) {
        tma_load_4d(smem + C::OFF_Q + h2 * 16384, &tma_q, qdo_full, h2 * 64, head * 3 + 0, qt * 128, batch);
        tma_load_4d(smem + C::OFF_DO + h2 * 16384, &tma_do, qdo_full, h2 * 64, head, qt * 128, batch);
      }
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % C::NKV;
        mbar_wait(&kv_empty[st], ((j / C::NKV) & 1) ^ 1, 40);
        mbar_expect_tx(&kv_full[st], 2 * C::KV_BYTES);
#pragma unroll
        for (int h2 = 0; h2 < HD / 64; ++h2) {
          tma_load_4d(smem + C::OFF_K + st * C::KV_BYTES + h2 * C::KV_SUB, &tma_kv, &kv_full[st], h2 * 64,
                      head * 3 + 1, j * 64, batch);
          tma_load_4d(smem + C::OFF_V + st * C::KV_BYTES + h2 * C::KV_SUB, &tma_kv, &kv_full[st], h2 * 64,
                      head * 3 + 2, j * 64, batch);
        }
      }
    }
  } else if (warp == BWD_MMA_WARP) {
    const uint32_t idesc_s = instr_desc_bf16(128, 64, 0, 0);
    const uint32_t idesc_dq = instr_desc_bf16(128, HD, 0, 1);
    const uint32_t q_lo = desc_lo(smem_u32(smem + C::OFF_Q), 0), do_lo = desc_lo(smem_u32(smem + C::OFF_DO), 0);
    const uint32_t k_lo0 = desc_lo(smem_u32(smem + C::OFF_K), 0), v_lo0 = desc_lo(smem_u32(smem + C::OFF_V), 0);
    const uint32_t kmn_lo0 = desc_lo(smem_u32(smem + C::OFF_K), C::KV_SUB);  // K as the MN-major B operand of dQ

    auto issue_sdp = [&](int j) {
      // S/dP buffer j % NB was last used by tile j-NB, whose dQ MMA has been issued already (the
      // tensor pipe runs MMAs in issue order), after every elementwise warp had finished with it.
      const int buf = j % C::NB, st = j % C::NKV;
      mbar_wait(&kv_full[st], (j / C::NKV) & 1, 41);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t k_lo = k_lo0 + st * (C::KV_BYTES >> 4), v_lo = v_lo0 + st * (C::KV_BYTES >> 4);
        const uint32_t d_s = tmem_base + C::TM_S + buf * 64, d_dp = tmem_base + C::TM_DP + buf * 64;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t offa = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4, offb = ((kk >> 2) * C::KV_SUB + (kk & 3) * 32) >> 4;
          if (kk == 0) {
            umma_ss_lo<0>(d_s, q_lo + offa, k_lo + offb, idesc_s);
            umma_ss_lo<0>(d_dp, do_lo + offa, v_lo + offb, idesc_s);
          } else {
            umma_ss_lo<1>(d_s, q_lo + offa, k_lo + offb, idesc_s);
            umma_ss_lo<1>(d_dp, do_lo + offa, v_lo + offb, idesc_s);
          }
        }
        umma_commit(&s_full[buf]);
      }
      __syncwarp();
    };

    mbar_wait(qdo_full, 0, 43);
    for (int j = 0; j < C::NB - 1 && j < n_kv; ++j) issue_sdp(j);
    for (int j = 0; j < n_kv; ++j) {
      if (j + C::NB - 1 < n_kv) issue_sdp(j + C::NB - 1);
      const int buf = j % C::NB, st = j % C::NKV;
      mbar_wait(&ds_full[buf], (j / C::NB) & 1, 44);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t k_lo = kmn_lo0 + st * (C::KV_BYTES >> 4);
        const uint32_t a = tmem_base + C::TM_DP + buf * 64;
        // k-step kk = the 16 kv columns of elementwise part kk
        umma_ts_lo<2>(tmem_base + C::TM_DQ, a, k_lo, idesc_dq, j != 0);
        umma_ts_lo<1>(tmem_base + C::TM_DQ, a + 16, k_lo + (2048 >> 4), idesc_dq);
        umma_ts_lo<1>(tmem_base + C::TM_DQ, a + 32, k_lo + 2 * (2048 >> 4), idesc_dq);
        umma_ts_lo<1>(tmem_base + C::TM_DQ, a + 48, k_lo + 3 * (2048 >> 4), idesc_dq);
        umma_commit(&kv_empty[st]);
        if (j == n_kv - 1) umma_commit(dq_full);
      }
      __syncwarp();
    }
  } else {
    const int quarter = warp & 3, part = warp >> 2;  // part: which 16 of the tile's 64 kv columns
    const int row_local = quarter * 32 + lane;
    const int row = qt * 128 + row_local;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sl2 = p.scale * LOG2E;
    const bool drop = p.p_drop > 0.f;
    // keep bits of this query row; this thread's 16 kv columns are half of word 2j + part/2 of tile j
    const uint32_t* mask_row = p.mask + (static_cast<size_t>(batch * p.nh + head) * p.s + min(row, p.s - 1)) * p.wpr + (part >> 1);
    const int mshift = (part & 1) * 4;
    const size_t stat_idx = (static_cast<size_t>(batch) * p.nh + head) * p.s_pad + row;  // row < s_pad always
    const float neg_lse2 = p.lse2[stat_idx];
    const uint64_t sl2_2 = f2_pack(sl2, sl2), nl2_2 = f2_pack(neg_lse2, neg_lse2);
    const uint64_t nd_2 = f2_pack(p.delta[stat_idx], p.delta[stat_idx]);

    // One 128 x 64 tile: p' = 2^(s·sl2 − lse2), dS = p'·(keep·dP − delta'). DIAG / DROP are compile-time
    // so that the common tile (below the diagonal) carries no predicate or select per element.
    auto tile = [&](int j, auto diag_c, auto drop_c) {
      constexpr bool DIAG = decltype(diag_c)::value, DROP = decltype(drop_c)::value;
      const int buf = j % C::NB;
      uint32_t mword = 0;
      if constexpr (DROP) mword = __ldg(mask_row + 2 * j);
      mbar_wait(&s_full[buf], (j / C::NB) & 1, 45);
      tc_fence_after();
      uint32_t rs[16], rp[16];
      tmem_ld_32x16(lane_base + C::TM_S + buf * 64 + part * 16, rs);
      tmem_ld_32x16(lane_base + C::TM_DP + buf * 64 + part * 16, rp);
      tmem_ld_wait();
      const int kv0 = j * 64 + part * 16;
      if constexpr (DROP) {
        // zero dP where the forward dropped P: AND the fp32 bits with an all-ones / all-zeros mask
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // elements 4k .. 4k+3 of this thread's 16 = sign bits of (word << (mshift + k))
          const uint32_t x = mword << (mshift + k);
          rp[4 * k + 0] &= prmt(x, 0u, 0x8888u);
          rp[4 * k + 1] &= prmt(x, 0u, 0x9999u);
          rp[4 * k + 2] &= prmt(x, 0u, 0xAAAAu);
          rp[4 * k + 3] &= prmt(x, 0u, 0xBBBBu);
        }
      }
      uint32_t pk[8];
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        const uint64_t a = fma2(f2_pack_u(rs[i], rs[i + 1]), sl2_2, nl2_2);
        float p0 = ex2_approx(f2_lo(a));
        float p1 = ex2_approx(f2_hi(a));
        if constexpr (DIAG) {
          p0 = (kv0 + i > row) ? 0.f : p0;
          p1 = (kv0 + i + 1 > row) ? 0.f : p1;
        }
        const uint64_t ds = mul2(f2_pack(p0, p1), add2(f2_pack_u(rp[i], rp[i + 1]), nd_2));
        pk[i >> 1] = pack_bf16(f2_lo(ds), f2_hi(ds));
      }
      // dS (bf16 pairs) goes back into the first 8 of this thread's own 16 dP columns
      tmem_st_32x8(lane_base + C::TM_DP + buf * 64 + part * 16, pk);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ds_full[buf]);
    };

    for (int j = 0; j < n_kv; ++j) {
      const bool diag = (j >= 2 * qt);  // the last two 64-wide KV tiles straddle the diagonal
      if (drop) {
        if (diag) tile(j, std::true_type{}, std::true_type{});
        else tile(j, std::false_type{}, std::true_type{});
      } else {
        if (diag) tile(j, std::true_type{}, std::false_type{});
        else tile(j, std::false_type{}, std::false_type{});
      }
    }

    // dQ = scale * (dS · K): this warp drains columns [part*HD/4, +HD/4) of its 32 rows
    mbar_wait(dq_full, 0, 47);
    tc_fence_after();
    __nv_bfloat16* dst = p.dqkv + ((static_cast<size_t>(batch) * p.s + row) * p.nh + head) * (3 * HD) + part * (HD / 4);
    drain_scaled<HD / 4>(lane_base + C::TM_DQ + part * (HD / 4), dst, p.scale, row < p.s);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == BWD_MMA_WARP) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------- dK/dV kernel
template <int HD>
struct DkvCfg {
  static constexpr int KV_BYTES = 128 * HD * 2;  // K / V tile (128 rows), persistent
  static constexpr int Q_BYTES = 64 * HD * 2;    // Q / dO tile (64 rows), 2-deep ring
  static constexpr int Q_SUB = 64 * 128;
  // Q/dO ring depth: a slot is only released when the dV/dK MMAs that read it retire, so the ring
  // must be deep enough for the next loads to be in flight a few tiles ahead of their use
  static constexpr int NQ = 4;
  static constexpr int OFF_K = 0;
  static constexpr int OFF_V = KV_BYTES;
  static constexpr int OFF_Q = 2 * KV_BYTES;
  static constexpr int OFF_DO = OFF_Q + NQ * Q_BYTES;
  static constexpr int OFF_MASK = OFF_DO + NQ * Q_BYTES;   // [NQ][64 queries][4 words]: dropout keep bits of the tile
  static constexpr int OFF_STAT = OFF_MASK + NQ * 1024;    // [NQ][lse2 | delta'][64 queries] fp32
  static constexpr int OFF_BAR = OFF_STAT + NQ * 512;
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  // TMEM: S^T[NB] | dP^T[NB] | dV | dK. P^T and dS^T (bf16) are written back over the S^T / dP^T
  // columns they came from and feed the dV / dK MMAs as A operands straight from TMEM. With three
  // buffers (head_dim 64; 128 has room for two) the MMA warp runs two tiles ahead.
  static constexpr int NB = HD == 64 ? 3 : 2;
  static constexpr int TM_ST = 0, TM_DPT = NB * 64, TM_DV = 2 * NB * 64, TM_DK = TM_DV + HD;
  static constexpr int TMEM_COLS = 512;
};

template <int HD>
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tma_kv,  // qkv, box rows 128
                    const __grid_constant__ CUtensorMap tma_q,   // qkv, box rows 64
                    const __grid_constant__ CUtensorMap tma_do,  // dout, box rows 64
                    const __grid_constant__ CUtensorMap tma_mask,  // keep bits as u32 [b·nh·s, wpr], box 4 x 64
                    const BwdParams p) {
  using C = DkvCfg<HD>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  uint64_t* kv_full = bars;          // 1
  uint64_t* q_full = bars + 1;       // NQ (<= 4)
  uint64_t* q_empty = bars + 5;      // NQ
  uint64_t* st_full = bars + 9;      // NB (<= 4)
  uint64_t* pt_full = bars + 13;     // NB
  uint64_t* acc_full = bars + 17;    // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jt = blockIdx.x;  // KV tile; low tiles walk the most query tiles and start first
  const int head = blockIdx.y, batch = blockIdx.z;
  const int i0 = 2 * jt;                  // first 64-row query tile at or below the diagonal
  const int n_q = (p.s + 63) / 64 - i0;   // >= 1 because jt*128 < s

  if (threadIdx.x == 0 && (smem_u32(smem) & 1023u) != 0) __trap();
  if (warp == BWD_TMA_WARP && lane == 0) {
    tma_prefetch_desc(&tma_kv);
    tma_prefetch_desc(&tma_q);
    tma_prefetch_desc(&tma_do);
  }
  if (warp == BWD_MMA_WARP) {
    if (lane == 0) {
      mbar_init(kv_full, 1);
      mbar_init(acc_full, 1);
      for (int i = 0; i < C::NQ; ++i) {
        mbar_init(&q_full[i], 1);
        mbar_init(&q_empty[i], 1);
      }
      for (int i = 0; i < C::NB; ++i) {
        mbar_init(&st_full[i], 1);
        mbar_init(&pt_full[i], BWD_EW_WARPS);
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

  if (warp == BWD_TMA_WARP) {
    if (elect_one()) {
      mbar_expect_tx(kv_full, 2 * C::KV_BYTES);
#pragma unroll
      for (int h2 = 0; h2 < HD / 64; ++h2) {
        tma_load_4d(smem + C::OFF_K + h2 * 16384, &tma_kv, kv_full, h2 * 64, head * 3 + 1, jt * 128, batch);
        tma_load_4d(smem + C::OFF_V + h2 * 16384, &tma_kv, kv_full, h2 * 64, head * 3 + 2, jt * 128, batch);
      }
      for (int t = 0; t < n_q; ++t) {
        const int st = t % C::NQ;
        mbar_wait(&q_empty[st], ((t / C::NQ) & 1) ^ 1, 50);
        const bool drop = p.p_drop > 0.f;
        mbar_expect_tx(&q_full[st], 2 * C::Q_BYTES + 512 + (drop ? 1024 : 0));
        if (drop)  // keep bits of query rows q0 .. q0+63 for this CTA's 128 kv columns (4 words per row)
          tma_load_2d(smem + C::OFF_MASK + st * 1024, &tma_mask, &q_full[st], 4 * jt,
                      (batch * p.nh + head) * p.s + (i0 + t) * 64);
        const size_t stat_off = (static_cast<size_t>(batch) * p.nh + head) * p.s_pad + (i0 + t) * 64;
        bulk_copy_g2s(smem + C::OFF_STAT + st * 512, p.lse2 + stat_off, 256, &q_full[st]);
        bulk_copy_g2s(smem + C::OFF_STAT + st * 512 + 256, p.delta + stat_off, 256, &q_full[st]);
#pragma unroll
        for (int h2 = 0; h2 < HD / 64; ++h2) {
          tma_load_4d(smem + C::OFF_Q + st * C::Q_BYTES + h2 * C::Q_SUB, &tma_q, &q_full[st], h2 * 64, head * 3 + 0,
                      (i0 + t) * 64, batch);
          tma_load_4d(smem + C::OFF_DO + st * C::Q_BYTES + h2 * C::Q_SUB, &tma_do, &q_full[st], h2 * 64, head,
                      (i0 + t) * 64, batch);
        }
      }
    }
  } else if (warp == BWD_MMA_WARP) {
    FX_T0();
    const uint32_t idesc_st = instr_desc_bf16(128, 64, 0, 0);
    const uint32_t idesc_acc = instr_desc_bf16(128, HD, 0, 1);
    const uint32_t k_lo = desc_lo(smem_u32(smem + C::OFF_K), 0), v_lo = desc_lo(smem_u32(smem + C::OFF_V), 0);
    const uint32_t q_lo0 = desc_lo(smem_u32(smem + C::OFF_Q), 0), do_lo0 = desc_lo(smem_u32(smem + C::OFF_DO), 0);
    // Q / dO again as MN-major B operands of the dK / dV MMAs
    const uint32_t qmn_lo0 = desc_lo(smem_u32(smem + C::OFF_Q), C::Q_SUB), domn_lo0 = desc_lo(smem_u32(smem + C::OFF_DO), C::Q_SUB);

    auto issue_st = [&](int t) {
      // S^T/dP^T buffer t % NB was last used by tile t-NB, whose dV/dK MMAs have been issued already
      // (the tensor pipe runs MMAs in issue order), after every elementwise warp had finished with it.
      const int buf = t % C::NB, st = t % C::NQ;
      FX_TM(12);
      mbar_wait(&q_full[st], (t / C::NQ) & 1, 51);
      FX_TM(8);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t q_lo = q_lo0 + st * (C::Q_BYTES >> 4), do_lo = do_lo0 + st * (C::Q_BYTES >> 4);
        const uint32_t d_st = tmem_base + C::TM_ST + buf * 64, d_dpt = tmem_base + C::TM_DPT + buf * 64;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t offa = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4, offb = ((kk >> 2) * C::Q_SUB + (kk & 3) * 32) >> 4;
          if (kk == 0) {
            umma_ss_lo<0>(d_st, k_lo + offa, q_lo + offb, idesc_st);
            umma_ss_lo<0>(d_dpt, v_lo + offa, do_lo + offb, idesc_st);
          } else {
            umma_ss_lo<1>(d_st, k_lo + offa, q_lo + offb, idesc_st);
            umma_ss_lo<1>(d_dpt, v_lo + offa, do_lo + offb, idesc_st);
          }
        }
        umma_commit(&st_full[buf]);
      }
      __syncwarp();
      FX_TM(9);
    };

    mbar_wait(kv_full, 0, 53);
    for (int t = 0; t < C::NB - 1 && t < n_q; ++t) issue_st(t);
    for (int t = 0; t < n_q; ++t) {
      if (t + C::NB - 1 < n_q) issue_st(t + C::NB - 1);
      const int buf = t % C::NB, st = t % C::NQ;
      mbar_wait(&pt_full[buf], (t / C::NB) & 1, 54);
      FX_TM(10);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t q_lo = qmn_lo0 + st * (C::Q_BYTES >> 4), do_lo = domn_lo0 + st * (C::Q_BYTES >> 4);
        const uint32_t a_p = tmem_base + C::TM_ST + buf * 64, a_ds = tmem_base + C::TM_DPT + buf * 64;
        const uint32_t d_dv = tmem_base + C::TM_DV, d_dk = tmem_base + C::TM_DK;
        // k-step kk = the 16 query columns of elementwise part kk; dV and dK interleaved
        umma_ts_lo<2>(d_dv, a_p, do_lo, idesc_acc, t != 0);
        umma_ts_lo<2>(d_dk, a_ds, q_lo, idesc_acc, t != 0);
#pragma unroll
        for (int kk = 1; kk < 4; ++kk) {
          umma_ts_lo<1>(d_dv, a_p + kk * 16, do_lo + kk * (2048 >> 4), idesc_acc);
          umma_ts_lo<1>(d_dk, a_ds + kk * 16, q_lo + kk * (2048 >> 4), idesc_acc);
        }
        umma_commit(&q_empty[st]);
        if (t == n_q - 1) umma_commit(acc_full);
      }
      __syncwarp();
      FX_TM(11);
    }
  } else {
    const int quarter = warp & 3, part = warp >> 2;  // part: which 16 of the tile's 64 query columns
    const int r_local = quarter * 32 + lane;                // kv row inside the tile
    const int kv = jt * 128 + r_local;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sl2 = p.scale * LOG2E;
    const bool drop = p.p_drop > 0.f;
    // This kv row's position inside a keep-bit word (layout: attn_dropout_mask_kernel): word r/32 of the
    // query's 4 words, element e = r % 32 at bit 8·(e & 3) + 7 − e/4; mask_sh moves that bit to bit 31.
    const int mask_word = r_local >> 5;
    const int mask_sh = 31 - (8 * (r_local & 3) + 7 - ((r_local & 31) >> 2));
    const uint64_t sl2_2 = f2_pack(sl2, sl2);

    FX_T0();
    // One 128 x 64 tile: p' = 2^(s·sl2 − lse2), P^T = keep·p', dS^T = p'·(keep·dP^T − delta'). DIAG / DROP
    // are compile-time so that the common tile carries no predicate or select per element.
    auto tile = [&](int t, auto diag_c, auto drop_c) {
      constexpr bool DIAG = decltype(diag_c)::value, DROP = decltype(drop_c)::value;
      const int buf = t % C::NB;
      const int q0 = (i0 + t) * 64;
      FX_T(0);
      const int qst = t % C::NQ;
      mbar_wait(&q_full[qst], (t / C::NQ) & 1, 58);  // the bulk-copied row statistics of this tile have landed
      mbar_wait(&st_full[buf], (t / C::NB) & 1, 55);  // S^T / dP^T are in TMEM
      FX_T(1);
      tc_fence_after();
      uint32_t rs[16], rp[16];
      tmem_ld_32x16(lane_base + C::TM_ST + buf * 64 + part * 16, rs);
      tmem_ld_32x16(lane_base + C::TM_DPT + buf * 64 + part * 16, rp);
      tmem_ld_wait();
      FX_T(2);

      const int qh = q0 + part * 16;  // first query column of this thread's 16
      const float* stat = reinterpret_cast<const float*>(smem + C::OFF_STAT + qst * 512) + part * 16;
      // keep bits of the tile as the producer's TMA left them: [64 queries][4 words]; all lanes of a warp
      // share mask_word, so every read below is a broadcast
      const uint32_t* mtile = reinterpret_cast<const uint32_t*>(smem + C::OFF_MASK + qst * 1024) + part * 16 * 4 + mask_word;
      uint32_t pk_p[8], pk_ds[8];
#pragma unroll
      for (int i4 = 0; i4 < 16; i4 += 4) {
        // (−lse2, −delta') of queries i4 .. i4+3 as two packed pairs each; warp-uniform: broadcast
        const ulonglong2 l2 = *reinterpret_cast<const ulonglong2*>(stat + i4);
        const ulonglong2 dl = *reinterpret_cast<const ulonglong2*>(stat + 64 + i4);
#pragma unroll
        for (int u = 0; u < 4; u += 2) {
          const int i = i4 + u;
          const uint64_t a = fma2(f2_pack_u(rs[i], rs[i + 1]), sl2_2, u == 0 ? l2.x : l2.y);
          float p0 = ex2_approx(f2_lo(a));
          float p1 = ex2_approx(f2_hi(a));
          if constexpr (DIAG) {
            p0 = (qh + i < kv) ? 0.f : p0;
            p1 = (qh + i + 1 < kv) ? 0.f : p1;
          }
          uint32_t d0 = rp[i], d1 = rp[i + 1], pp = pack_bf16(p0, p1);
          if constexpr (DROP) {
            const uint32_t m0 = static_cast<uint32_t>(static_cast<int32_t>(mtile[i * 4] << mask_sh) >> 31);
            const uint32_t m1 = static_cast<uint32_t>(static_cast<int32_t>(mtile[(i + 1) * 4] << mask_sh) >> 31);
            d0 &= m0;
            d1 &= m1;
            pp &= prmt(m0, m1, 0x5410u);
          }
          pk_p[i >> 1] = pp;
          const uint64_t ds = mul2(f2_pack(p0, p1), add2(f2_pack_u(d0, d1), u == 0 ? dl.x : dl.y));
          pk_ds[i >> 1] = pack_bf16(f2_lo(ds), f2_hi(ds));
        }
      }
      FX_T(3);
      // P^T and dS^T (bf16 pairs) go back into the first 8 of this thread's own 16 S^T / dP^T columns
      tmem_st_32x8(lane_base + C::TM_ST + buf * 64 + part * 16, pk_p);
      tmem_st_32x8(lane_base + C::TM_DPT + buf * 64 + part * 16, pk_ds);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&pt_full[buf]);
      FX_T(5);
    };

    for (int t = 0; t < n_q; ++t) {
      const bool diag = (t < 2);  // the first two 64-row query tiles straddle the diagonal
      if (drop) {
        if (diag) tile(t, std::true_type{}, std::true_type{});
        else tile(t, std::false_type{}, std::true_type{});
      } else {
        if (diag) tile(t, std::true_type{}, std::false_type{});
        else tile(t, std::false_type{}, std::false_type{});
      }
    }

    // dV = P_drop^T · dO, dK = scale * dS^T · Q; this warp drains columns [part*HD/4, +HD/4)
    mbar_wait(acc_full, 0, 57);
    tc_fence_after();
    __nv_bfloat16* base = p.dqkv + ((static_cast<size_t>(batch) * p.s + kv) * p.nh + head) * (3 * HD) + part * (HD / 4);
    drain_scaled<HD / 4>(lane_base + C::TM_DK + part * (HD / 4), base + HD, p.scale, kv < p.s);
    drain_scaled<HD / 4>(lane_base + C::TM_DV + part * (HD / 4), base + 2 * HD, 1.0f, kv < p.s);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == BWD_MMA_WARP) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

static int make_qkv_map_rows(CUtensorMap* map, const void* qkv, int b, int s, int nh, int hd, uint32_t box_rows) {
  const uint64_t dims[4] = {(uint64_t)hd, (uint64_t)3 * nh, (uint64_t)s, (uint64_t)b};
  const uint64_t strides[3] = {(uint64_t)hd * 2, (uint64_t)nh * 3 * hd * 2, (uint64_t)s * nh * 3 * hd * 2};
  const uint32_t box[4] = {64, 1, box_rows, 1};
  return make_tma_bf16(map, qkv, 4, dims, strides, box);
}
static int make_out_map_rows(CUtensorMap* map, const void* out, int b, int s, int nh, int hd, uint32_t box_rows) {
  const uint64_t dims[4] = {(uint64_t)hd, (uint64_t)nh, (uint64_t)s, (uint64_t)b};
  const uint64_t strides[3] = {(uint64_t)hd * 2, (uint64_t)nh * hd * 2, (uint64_t)s * nh * hd * 2};
  const uint32_t box[4] = {64, 1, box_rows, 1};
  return make_tma_bf16(map, out, 4, dims, strides, box);
}

#include "attention_bwd_fused.cuh"

// head_dim 64 runs the single-kernel backward unless FX_ATTN_BWD=split asks for the deterministic pair
static bool use_fused_bwd(int hd) {
  return hd == 64 && !knobs().attn_bwd_split;
}

template <int HD>
static int launch_bwd(cudaStream_t st, const void* dout, const void* qkv, const void* out, const float* lse,
                      const BwdParams& p, float* lse2, float* delta, float* dq_acc) {
  static std::atomic<bool> attr_set[64];  // idempotent set-up: an atomic flag per device is enough
  int dev = 0;
  FX_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    FX_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, DqCfg<HD>::SMEM_BYTES));
    FX_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, DkvCfg<HD>::SMEM_BYTES));
    if (HD == 64)
      FX_CUDA(cudaFuncSetAttribute(attn_bwd_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FusedCfg::SMEM_BYTES));
    attr_set[dev] = true;
  }
  CUtensorMap qkv128, qkv64, do128, do64, mmap;
  int rc;
  if (p.p_drop > 0.f) {
    if ((rc = make_tma_u32_2d(&mmap, p.mask, (uint64_t)p.wpr, (uint64_t)p.b * p.nh * p.s, (uint64_t)p.wpr * 4, 4, 64))) return rc;
  } else {
    mmap = CUtensorMap{};
  }
  if ((rc = make_qkv_map_rows(&qkv128, qkv, p.b, p.s, p.nh, HD, 128))) return rc;
  if ((rc = make_qkv_map_rows(&qkv64, qkv, p.b, p.s, p.nh, HD, 64))) return rc;
  if ((rc = make_out_map_rows(&do128, dout, p.b, p.s, p.nh, HD, 128))) return rc;
  if ((rc = make_out_map_rows(&do64, dout, p.b, p.s, p.nh, HD, 64))) return rc;

  const size_t total = (size_t)p.b * p.s_pad * p.nh;
  attn_delta_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
      static_cast<const __nv_bfloat16*>(dout), static_cast<const __nv_bfloat16*>(out), lse, lse2, delta, p.b, p.s,
      p.nh, HD, p.s_pad, p.p_drop > 0.f ? log2f(p.inv_keep) : 0.f, 1.0f - p.p_drop);
  FX_LAUNCH_CHECK();
  dim3 grid((p.s + 127) / 128, p.nh, p.b);
  if (HD == 64 && use_fused_bwd(HD)) {
    const size_t acc_floats = (size_t)p.b * p.nh * p.s_pad * 64;
    FX_CUDA(cudaMemsetAsync(dq_acc, 0, acc_floats * sizeof(float), st));
    attn_bwd_fused_kernel<<<grid, BWD_THREADS, FusedCfg::SMEM_BYTES, st>>>(qkv128, qkv64, do64, mmap, p, dq_acc);
    FX_LAUNCH_CHECK();
    const size_t n = (size_t)p.b * p.nh * p.s_pad * 8;
    attn_dq_convert_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dq_acc, p.dqkv, p.b, p.s, p.nh, p.s_pad, p.scale);
    FX_LAUNCH_CHECK();
    return FX_OK;
  }
  attn_bwd_dkv_kernel<HD><<<grid, BWD_THREADS, DkvCfg<HD>::SMEM_BYTES, st>>>(qkv128, qkv64, do64, mmap, p);
  FX_LAUNCH_CHECK();
  attn_bwd_dq_kernel<HD><<<grid, BWD_THREADS, DqCfg<HD>::SMEM_BYTES, st>>>(qkv128, qkv64, do128, p);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

}  // namespace attn
}  // namespace fx

extern "C" size_t fx_attn_dropout_mask_bytes(int b, int s, int nh) {
  if (b <= 0 || s <= 0 || nh <= 0) return 0;
  const size_t wpr = ((size_t)s + 127) / 128 * 4;
  return (size_t)b * nh * s * wpr * sizeof(uint32_t);
}

extern "C" int fx_attn_causal_fwd(void* stream, const void* qkv, void* out, float* lse, void* dropout_mask, int b,
                                  int s, int nh, int hd, float scale, float p_drop, uint64_t seed, uint64_t offset) {
  return fx_attn_causal_fwd_at(stream, qkv, out, lse, dropout_mask, b, s, nh, hd, scale, p_drop, seed, offset, 0, 0);
}

extern "C" int fx_attn_dropout_mask(void* stream, void* dropout_mask, int b, int s, int nh, float p_drop, uint64_t seed,
                                    uint64_t offset, int batch0) {
  using namespace fx;
  FX_CHECK_ARG(b > 0 && s > 0 && nh > 0 && batch0 >= 0, "fx_attn_dropout_mask: empty input (b %d, s %d, nh %d)", b, s, nh);
  FX_CHECK_ARG(nh <= 65535 && b * nh <= 65535, "fx_attn_dropout_mask: nh and b*nh must be <= 65535");
  FX_CHECK_ARG(p_drop > 0.f && p_drop < 1.f, "fx_attn_dropout_mask: dropout p must be in (0,1) (got %f)", p_drop);
  FX_CHECK_ARG(dropout_mask && (uintptr_t)dropout_mask % 16 == 0, "fx_attn_dropout_mask: null or misaligned buffer");
  attn::MaskParams m{};
  m.mask = static_cast<uint32_t*>(dropout_mask);
  m.b = b; m.s = s; m.nh = nh; m.wpr = (s + 127) / 128 * 4;
  m.thresh2 = dropout_threshold(p_drop) * 0x10001u;
  m.off_lo = static_cast<uint32_t>(offset); m.off_hi = static_cast<uint32_t>(offset >> 32);
  m.keys = philox_round_keys(seed);
  m.bh0 = static_cast<uint64_t>(batch0) * nh;
  dim3 grid((s + 31) / 32, b * nh);
  static std::atomic<bool> carve[64];
  FX_PREFER_MAX_SHARED(attn::attn_dropout_mask_kernel, carve);
  attn::attn_dropout_mask_kernel<<<grid, attn::MASK_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(m);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_attn_causal_fwd_at(void* stream, const void* qkv, void* out, float* lse, void* dropout_mask, int b,
                                     int s, int nh, int hd, float scale, float p_drop, uint64_t seed, uint64_t offset,
                                     int batch0, int mask_ready) {
  using namespace fx;
  FX_CHECK_ARG(batch0 >= 0, "fx_attn_causal_fwd: negative batch offset");
  FX_CHECK_ARG(b > 0 && s > 0 && nh > 0, "fx_attn_causal_fwd: empty input (b %d, s %d, nh %d)", b, s, nh);
  FX_CHECK_ARG(hd == 64 || hd == 128, "fx_attn_causal_fwd: head_dim must be 64 or 128 (got %d)", hd);
  FX_CHECK_ARG(nh <= 65535 && b * nh <= 65535, "fx_attn_causal_fwd: nh and b*nh must be <= 65535");
  FX_CHECK_ARG(p_drop >= 0.f && p_drop < 1.f, "fx_attn_causal_fwd: dropout p must be in [0,1) (got %f)", p_drop);
  FX_CHECK_ARG(scale > 0.f, "fx_attn_causal_fwd: scale must be positive");
  FX_CHECK_ARG(qkv && out && lse, "fx_attn_causal_fwd: null pointer");
  FX_CHECK_ARG(p_drop == 0.f || dropout_mask, "fx_attn_causal_fwd: dropout needs the keep-bit buffer (fx_attn_dropout_mask_bytes)");
  FX_CHECK_ARG(((uintptr_t)qkv | (uintptr_t)out | (uintptr_t)dropout_mask) % 16 == 0,
               "fx_attn_causal_fwd: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int wpr = (s + 127) / 128 * 4;
  if (p_drop > 0.f && !mask_ready) {
    const int rcm = fx_attn_dropout_mask(stream, dropout_mask, b, s, nh, p_drop, seed, offset, batch0);
    if (rcm) return rcm;
  }
  CUtensorMap map;
  int rc = attn::make_qkv_map(&map, qkv, b, s, nh, hd);
  if (rc) return rc;
  attn::FwdParams p{};
  p.out = static_cast<__nv_bfloat16*>(out);
  p.lse = lse;
  p.b = b; p.s = s; p.nh = nh;
  p.scale = scale; p.p_drop = p_drop; p.inv_keep = 1.0f / (1.0f - p_drop);
  p.log2_inv_keep = p_drop > 0.f ? log2f(p.inv_keep) : 0.f;
  p.mask = static_cast<const uint32_t*>(dropout_mask);
  p.wpr = wpr;
  return hd == 64 ? attn::launch_fwd<64>(st, map, p) : attn::launch_fwd<128>(st, map, p);
}

#ifdef FX_TRACE
extern "C" int fx_debug_trace(unsigned long long* host16, int reset) {
  if (reset) {
    unsigned long long z[16] = {0};
    return cudaMemcpyToSymbol(fx::attn::g_trace, z, sizeof(z)) == cudaSuccess ? 0 : -2;
  }
  return cudaMemcpyFromSymbol(host16, fx::attn::g_trace, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
}
#endif

extern "C" size_t fx_attn_causal_bwd_workspace_bytes(int b, int s, int nh, int hd) {
  if (b <= 0 || s <= 0 || nh <= 0) return 0;
  const size_t s_pad = ((size_t)s + 127) / 128 * 128;
  // −lse2 and −delta', padded per (batch, head) row; head_dim 64 adds the fp32 dQ accumulator of the
  // single-kernel backward ([b, nh, s_pad, 64])
  return (size_t)b * nh * s_pad * sizeof(float) * (hd == 64 ? 2 + 64 : 2);
}

extern "C" int fx_attn_causal_bwd(void* stream, const void* dout, const void* qkv, const void* out,
                                  const float* lse, const void* dropout_mask, void* dqkv, void* workspace, int b,
                                  int s, int nh, int hd, float scale, float p_drop) {
  using namespace fx;
  FX_CHECK_ARG(b > 0 && s > 0 && nh > 0, "fx_attn_causal_bwd: empty input (b %d, s %d, nh %d)", b, s, nh);
  FX_CHECK_ARG(hd == 64 || hd == 128, "fx_attn_causal_bwd: head_dim must be 64 or 128 (got %d)", hd);
  FX_CHECK_ARG(nh <= 65535 && b <= 65535, "fx_attn_causal_bwd: nh and b must be <= 65535");
  FX_CHECK_ARG(p_drop == 0.f || dropout_mask, "fx_attn_causal_bwd: dropout needs the keep bits the forward call wrote");
  FX_CHECK_ARG(p_drop >= 0.f && p_drop < 1.f, "fx_attn_causal_bwd: dropout p must be in [0,1) (got %f)", p_drop);
  FX_CHECK_ARG(scale > 0.f, "fx_attn_causal_bwd: scale must be positive");
  FX_CHECK_ARG(dout && qkv && out && lse && dqkv && workspace, "fx_attn_causal_bwd: null pointer");
  FX_CHECK_ARG(((uintptr_t)dout | (uintptr_t)qkv | (uintptr_t)out | (uintptr_t)dqkv | (uintptr_t)workspace) % 16 == 0,
               "fx_attn_causal_bwd: pointers must be 16-byte aligned");
  attn::BwdParams p{};
  p.s_pad = (s + 127) / 128 * 128;
  float* lse2 = static_cast<float*>(workspace);
  float* delta = lse2 + (size_t)b * nh * p.s_pad;
  p.lse2 = lse2;
  p.delta = delta;
  p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
  p.b = b; p.s = s; p.nh = nh;
  p.scale = scale; p.p_drop = p_drop; p.inv_keep = 1.0f / (1.0f - p_drop);
  p.mask = static_cast<const uint32_t*>(dropout_mask);
  p.wpr = (s + 127) / 128 * 4;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* dq_acc = delta + (size_t)b * nh * p.s_pad;  // head_dim 64 only (see fx_attn_causal_bwd_workspace_bytes)
  return hd == 64 ? attn::launch_bwd<64>(st, dout, qkv, out, lse, p, lse2, delta, dq_acc)
                  : attn::launch_bwd<128>(st, dout, qkv, out, lse, p, lse2, delta, nullptr);
}
