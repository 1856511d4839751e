// attention_bwd_fused.cuh — single-kernel attention backward for head_dim 64 (included by attention.cu,
// inside namespace fx::attn, after the split dK/dV and dQ kernels whose helpers it shares).
//
// The split pair recomputes S and dP (and their exponentials) twice: once per KV tile for dK/dV, once per
// Q tile for dQ. Here the CTA that owns a 128-row KV tile also produces the dQ contribution of that tile:
//   once per CTA:             K, V rows -> TMEM (bf16): A operands of the S^T / dP^T MMAs from then on
//   per 64-row query tile t:  S^T = K·Q^T, dP^T = V·dO^T           (tcgen05, into TMEM)
//                             P^T, dS^T -> TMEM (bf16); dS^T also -> shared memory (bf16, 128B-swizzled)
//                             dV += P^T·dO, dK += dS^T·Q           (A from TMEM)
//   per pair of query tiles:  dQ_partial[128 q x 64] = dS·K — the shared-memory bytes of dS^T read as an
//                             MN-major A operand — into a TMEM accumulator, drained by the element-wise
//                             warps and added into a fp32 global accumulator with red.global.add.v4.f32.
// The L2 atomic path takes ~26 B/clk per SM (5.6 TB/s chip-wide; tools/micro/bulk_red.cu, red_v4.cu) for
// the 1.1 GB of dQ partials of a bench step. A small kernel then scales the accumulator and writes bf16 dQ
// into the packed dqkv tensor. fp32 adds from different KV-tile CTAs arrive in no fixed order: dQ can
// differ in the last bf16 bit from run to run (dK, dV are exact). FX_ATTN_BWD=split selects the
// deterministic two-kernel path.

struct FusedCfg {
  static constexpr int HD = 64;
  static constexpr int KV_BYTES = 128 * HD * 2;  // K / V tile, resident
  static constexpr int Q_BYTES = 64 * HD * 2;    // Q / dO tile (64 rows)
  static constexpr int NQ = 6;                   // Q / dO ring depth
  static constexpr int DS_BYTES = 128 * 128;     // dS^T tile: [128 kv rows][64 q] bf16
  static constexpr int OFF_K = 0;
  static constexpr int OFF_V = KV_BYTES;
  static constexpr int OFF_Q = 2 * KV_BYTES;
  static constexpr int OFF_DO = OFF_Q + NQ * Q_BYTES;
  static constexpr int OFF_MASK = OFF_DO + NQ * Q_BYTES;  // [NQ][64 queries][4 words]
  static constexpr int OFF_STAT = OFF_MASK + NQ * 1024;   // [NQ][-lse2 | -delta'][64 queries] fp32
  static constexpr int OFF_DS = OFF_STAT + NQ * 512;      // 4 tiles = 2 pair buffers of 2 x 16 KB
  static constexpr int OFF_BAR = OFF_DS + 4 * DS_BYTES;
  static_assert(OFF_BAR + 256 <= 227 * 1024, "");
  static constexpr int SMEM_BYTES = OFF_BAR + 256;
  static_assert(OFF_DS % 1024 == 0, "");
  static constexpr int NB = 2;  // S^T / dP^T buffers
  // TMEM: S^T[2] | dP^T[2] | dV | dK | dQ | K | V. K and V (bf16, one kv row per lane) are copied into TMEM
  // once and feed the S^T / dP^T MMAs as A operands from there: shared-memory bandwidth (MMA operand
  // reads + element-wise loads/stores + TMA + the bulk reduce) is what bounds this kernel, and the two
  // resident 16 KB tiles would otherwise be re-read for every query tile.
  static constexpr int TM_ST = 0, TM_DPT = NB * 64, TM_DV = 2 * NB * 64, TM_DK = TM_DV + HD, TM_DQ = TM_DK + HD;
  static constexpr int TM_K = TM_DQ + HD, TM_V = TM_K + 32;
  static constexpr int TMEM_COLS = 512;  // 128 + 128 + 64 + 64 + 64 + 32 + 32
};

// floats per (batch, head, 128-query block) in the dQ accumulator
constexpr size_t DQACC_BLOCK = 128 * 64;

__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_fused_kernel(const __grid_constant__ CUtensorMap tma_kv,    // qkv, box rows 128
                      const __grid_constant__ CUtensorMap tma_q,     // qkv, box rows 64
                      const __grid_constant__ CUtensorMap tma_do,    // dout, box rows 64
                      const __grid_constant__ CUtensorMap tma_mask,  // keep bits as u32 [b·nh·s, wpr], box 4 x 64
                      const BwdParams p, float* __restrict__ dq_acc) {
  using C = FusedCfg;
  constexpr int HD = C::HD;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  uint64_t* kv_full = bars;         // 1
  uint64_t* q_full = bars + 1;                 // NQ
  uint64_t* q_empty = q_full + C::NQ;          // NQ
  uint64_t* st_full = q_empty + C::NQ;         // NB
  uint64_t* pt_full = st_full + 2;             // NB
  uint64_t* acc_full = pt_full + 2;            // 1
  uint64_t* ds_empty = acc_full + 1;           // 2 (pair buffers of dS^T in shared memory)
  uint64_t* dq_full = ds_empty + 2;            // 1 (dQ accumulator in TMEM)
  uint64_t* dq_empty = dq_full + 1;            // 1
  uint64_t* kvt_full = dq_empty + 1;           // 1 (K, V copied into TMEM)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(kvt_full + 1);
  static_assert((1 + 2 * C::NQ + 2 + 2 + 1 + 2 + 1 + 1 + 1) * 8 + 4 <= 256, "barrier area");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jt = blockIdx.x;  // KV tile; low tiles walk the most query tiles and start first
  const int head = blockIdx.y, batch = blockIdx.z;
  const int i0 = 2 * jt;                 // first 64-row query tile at or below the diagonal
  const int n_q = (p.s + 63) / 64 - i0;  // >= 1 because jt*128 < s

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
      mbar_init(dq_full, 1);
      mbar_init(dq_empty, BWD_EW_WARPS);
      mbar_init(kvt_full, 8);
      for (int i = 0; i < C::NQ; ++i) {
        mbar_init(&q_full[i], 1);
        mbar_init(&q_empty[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&st_full[i], 1);
        mbar_init(&pt_full[i], BWD_EW_WARPS);
        mbar_init(&ds_empty[i], 1);
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
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      const bool drop = p.p_drop > 0.f;
      mbar_expect_tx(kv_full, 2 * C::KV_BYTES);
      tma_load_4d(smem + C::OFF_K, &tma_kv, kv_full, 0, head * 3 + 1, jt * 128, batch);
      tma_load_4d(smem + C::OFF_V, &tma_kv, kv_full, 0, head * 3 + 2, jt * 128, batch);
      FX_T0();
      for (int t = 0; t < n_q; ++t) {
        const int st = t % C::NQ;
        FX_TM(15);
        mbar_wait(&q_empty[st], ((t / C::NQ) & 1) ^ 1, 70);
        FX_TM(14);
        mbar_expect_tx(&q_full[st], 2 * C::Q_BYTES + 512 + (drop ? 1024 : 0));
        if (drop)  // keep bits of query rows q0 .. q0+63 for this CTA's 128 kv columns (4 words per row)
          tma_load_2d(smem + C::OFF_MASK + st * 1024, &tma_mask, &q_full[st], 4 * jt,
                      (batch * p.nh + head) * p.s + (i0 + t) * 64);
        const size_t stat_off = (static_cast<size_t>(batch) * p.nh + head) * p.s_pad + (i0 + t) * 64;
        bulk_copy_g2s(smem + C::OFF_STAT + st * 512, p.lse2 + stat_off, 256, &q_full[st]);
        bulk_copy_g2s(smem + C::OFF_STAT + st * 512 + 256, p.delta + stat_off, 256, &q_full[st]);
        tma_load_4d(smem + C::OFF_Q + st * C::Q_BYTES, &tma_q, &q_full[st], 0, head * 3 + 0, (i0 + t) * 64, batch);
        tma_load_4d(smem + C::OFF_DO + st * C::Q_BYTES, &tma_do, &q_full[st], 0, head, (i0 + t) * 64, batch);
      }
    }
  } else if (warp == BWD_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer
    FX_T0();
    const uint32_t idesc_st = instr_desc_bf16(128, 64, 0, 0);   // S^T, dP^T: A (K / V) from TMEM, B (Q / dO) K-major
    const uint32_t idesc_acc = instr_desc_bf16(128, HD, 0, 1);  // dV, dK: B (dO / Q) MN-major
    const uint32_t idesc_dq = instr_desc_bf16(128, HD, 1, 1);   // dQ: A (dS) and B (K) MN-major
    const uint32_t kmn_lo = desc_lo(smem_u32(smem + C::OFF_K), 0);  // one 64-wide chunk: LBO unused
    const uint32_t q_lo0 = desc_lo(smem_u32(smem + C::OFF_Q), 0), do_lo0 = desc_lo(smem_u32(smem + C::OFF_DO), 0);
    const uint32_t dsmn_lo0 = desc_lo(smem_u32(smem + C::OFF_DS), C::DS_BYTES);      // MN-major A of dQ: two q chunks

    auto issue_st = [&](int t) {
      // S^T/dP^T buffer t % NB was last used by tile t-NB, whose dV MMA has been issued already (the
      // tensor pipe runs MMAs in issue order), after every element-wise warp had finished with it.
      const int buf = t % C::NB, st = t % C::NQ;
      FX_TM(12);
      mbar_wait(&q_full[st], (t / C::NQ) & 1, 71);
      FX_TM(t < 2 ? 13 : 8);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t q_lo = q_lo0 + st * (C::Q_BYTES >> 4), do_lo = do_lo0 + st * (C::Q_BYTES >> 4);
        const uint32_t d_st = tmem_base + C::TM_ST + buf * 64, d_dpt = tmem_base + C::TM_DPT + buf * 64;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t off = (kk * 32) >> 4;  // k-step = 16 hd columns: 32 bytes of Q / dO, 8 TMEM columns of K / V
          if (kk == 0) {
            umma_ts_lo<0>(d_st, tmem_base + C::TM_K + kk * 8, q_lo + off, idesc_st);
            umma_ts_lo<0>(d_dpt, tmem_base + C::TM_V + kk * 8, do_lo + off, idesc_st);
          } else {
            umma_ts_lo<1>(d_st, tmem_base + C::TM_K + kk * 8, q_lo + off, idesc_st);
            umma_ts_lo<1>(d_dpt, tmem_base + C::TM_V + kk * 8, do_lo + off, idesc_st);
          }
        }
        umma_commit(&st_full[buf]);
      }
      __syncwarp();
      FX_TM(9);
    };

    mbar_wait(kvt_full, 0, 72);
    issue_st(0);
    for (int t = 0; t < n_q; ++t) {
      if (t + 1 < n_q) issue_st(t + 1);
      const int buf = t % C::NB, st = t % C::NQ;
      const bool pair_done = (t & 1) || t == n_q - 1;
      const int pair = t >> 1;
      mbar_wait(&pt_full[buf], (t / C::NB) & 1, 73);
      FX_TM(10);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t q_lo = q_lo0 + st * (C::Q_BYTES >> 4), do_lo = do_lo0 + st * (C::Q_BYTES >> 4);
        const uint32_t a_p = tmem_base + C::TM_ST + buf * 64, a_ds = tmem_base + C::TM_DPT + buf * 64;
        const uint32_t d_dv = tmem_base + C::TM_DV, d_dk = tmem_base + C::TM_DK;
        // k-step kk = the 16 query columns of element-wise part kk; dV and dK interleaved.
        // Q / dO tiles are the MN-major B operands here (64 hd contiguous per query row).
        umma_ts_lo<2>(d_dv, a_p, do_lo, idesc_acc, t != 0);
        umma_ts_lo<2>(d_dk, a_ds, q_lo, idesc_acc, t != 0);
#pragma unroll
        for (int kk = 1; kk < 4; ++kk) {
          umma_ts_lo<1>(d_dv, a_p + kk * 16, do_lo + kk * (2048 >> 4), idesc_acc);
          umma_ts_lo<1>(d_dk, a_ds + kk * 16, q_lo + kk * (2048 >> 4), idesc_acc);
        }
        umma_commit(&q_empty[st]);
      }
      __syncwarp();
      if (pair_done) {
        // one dQ accumulator: pair-1 is drained right after the element-wise pass of tile 2·pair+1, whose
        // pt_full was waited for above; the dV / dK MMAs just issued cover that short gap
        mbar_wait(dq_empty, (pair & 1) ^ 1, 74);
        tc_fence_after();
      }
      if (elect_one()) {
        if (pair_done) {
          // dQ_partial[q pair (128) x hd] = dS[q x kv] · K[kv x hd]: A = the pair's two dS^T tiles read
          // MN-major (64 q contiguous per kv row), k-step = 16 kv rows = 2048 bytes in both operands
          const uint32_t a_mn = dsmn_lo0 + (pair & 1) * (2 * C::DS_BYTES >> 4);
          const uint32_t d_dq = tmem_base + C::TM_DQ;
          umma_ss_lo<0>(d_dq, a_mn, kmn_lo, idesc_dq);
#pragma unroll
          for (int kk = 1; kk < 8; ++kk) umma_ss_lo<1>(d_dq, a_mn + kk * (2048 >> 4), kmn_lo + kk * (2048 >> 4), idesc_dq);
          umma_commit(dq_full);
          umma_commit(&ds_empty[pair & 1]);
        }
        if (t == n_q - 1) umma_commit(acc_full);
      }
      __syncwarp();
      FX_TM(11);
    }
  } else {
    // ------------------------------------------------------------------ element-wise warps
    const int quarter = warp & 3, part = warp >> 2;  // part: which 16 of the tile's 64 query columns
    const int r_local = quarter * 32 + lane;         // kv row inside the tile (= TMEM lane)
    const int kv = jt * 128 + r_local;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const float sl2 = p.scale * LOG2E;
    const bool drop = p.p_drop > 0.f;
    // This kv row's position inside a keep-bit word (layout: attn_dropout_mask_kernel): word r/32 of the
    // query's 4 words, element e = r % 32 at bit 8·(e & 3) + 7 − e/4; mask_sh moves that bit to bit 31.
    const int mask_word = r_local >> 5;
    const int mask_sh = 31 - (8 * (r_local & 3) + 7 - ((r_local & 31) >> 2));
    const uint64_t sl2_2 = f2_pack(sl2, sl2);
    const uint32_t ds_row = smem_u32(smem + C::OFF_DS);  // my 32 bytes of a dS^T tile: chunks 2·part, 2·part+1 of row r_local
    const uint32_t ds_c0 = sw128_off(r_local, 2 * part), ds_c1 = sw128_off(r_local, 2 * part + 1);
    // K (parts 0) and V (parts 1) rows -> TMEM as bf16 A operands: lane = kv row, 32 columns of packed pairs
    if (part < 2) {
      mbar_wait(kv_full, 0, 85);
      const uint32_t src = smem_u32(smem + (part == 0 ? C::OFF_K : C::OFF_V));
      uint32_t r[32];
#pragma unroll
      for (int c = 0; c < 8; ++c)
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r[4 * c]), "=r"(r[4 * c + 1]), "=r"(r[4 * c + 2]), "=r"(r[4 * c + 3])
                     : "r"(src + sw128_off(r_local, c)));
      uint32_t lo[16], hi[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) { lo[i] = r[i]; hi[i] = r[16 + i]; }
      const uint32_t dst = lane_base + (part == 0 ? C::TM_K : C::TM_V);
      tmem_st_32x16(dst, lo);
      tmem_st_32x16(dst + 16, hi);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(kvt_full);
    }

    FX_T0();
    // Drain the dQ accumulator of query pair P (128-row block jt + P) and add it into the global fp32
    // accumulator with vector reductions straight from registers (red.global.add.v4.f32). The L2 atomic path
    // takes ~26 B/clk per SM whichever unit feeds it (tools/micro/bulk_red.cu, red_v4.cu); a TMA bulk reduce
    // of the block was tried first and queues in front of the tile loads for ~1.3k cycles, which starves the
    // MMA warp; four dedicated reducer warps were tried too and cost more in registers than they hide. The
    // accumulator block is laid out [16 column groups][128 rows][4]: a warp's 32 rows make one contiguous
    // 512-byte access.
    float* dq_block0 = dq_acc + ((static_cast<size_t>(batch) * p.nh + head) * (p.s_pad / 128) + jt) * DQACC_BLOCK +
                       (part * 4 * 128 + r_local) * 4;
    auto drain_dq = [&](int P, bool both_tiles) {
      mbar_wait(dq_full, P & 1, 80);
      tc_fence_after();
      uint32_t r[16];
      tmem_ld_32x16(lane_base + C::TM_DQ + part * 16, r);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_empty);  // the accumulator is free again before the reductions go out
      if (both_tiles || r_local < 64) {      // rows 64..127 are stale when the pair has a single tile
        float* dst = dq_block0 + static_cast<size_t>(P) * DQACC_BLOCK;
#pragma unroll
        for (int g = 0; g < 4; ++g)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + g * 128 * 4), "r"(r[4 * g]),
                       "r"(r[4 * g + 1]), "r"(r[4 * g + 2]), "r"(r[4 * g + 3])
                       : "memory");
      }
    };

    // One 128 x 64 tile: p' = 2^(s·sl2 − lse2), P^T = keep·p', dS^T = p'·(keep·dP^T − delta'). DIAG / DROP
    // are compile-time so that the common tile carries no predicate or select per element.
    auto tile = [&](int t, auto diag_c, auto drop_c) {
      constexpr bool DIAG = decltype(diag_c)::value, DROP = decltype(drop_c)::value;
      const int buf = t % C::NB;
      const int q0 = (i0 + t) * 64;
      const int qst = t % C::NQ;
      FX_T(0);
      // dS^T slot t % 4 belongs to pair buffer (t / 2) % 2: free once the dQ MMA two pairs back has retired
      if ((t & 1) == 0) mbar_wait(&ds_empty[(t >> 1) & 1], ((t >> 2) & 1) ^ 1, 81);
      mbar_wait(&q_full[qst], (t / C::NQ) & 1, 82);   // row statistics and keep bits of this tile have landed
      mbar_wait(&st_full[buf], (t / C::NB) & 1, 83);  // S^T / dP^T are in TMEM
      FX_T(t == 0 ? 6 : 1);
      tc_fence_after();
      uint32_t rs[16], rp[16];
      tmem_ld_32x16(lane_base + C::TM_ST + buf * 64 + part * 16, rs);
      tmem_ld_32x16(lane_base + C::TM_DPT + buf * 64 + part * 16, rp);
      tmem_ld_wait();
      FX_T(2);

      const int qh = q0 + part * 16;  // first query column of this thread's 16
      const float* stat = reinterpret_cast<const float*>(smem + C::OFF_STAT + qst * 512) + part * 16;
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
      // P^T / dS^T (bf16 pairs) go back into the first 8 of this thread's own 16 S^T / dP^T columns (A
      // operands of the dV / dK MMAs); dS^T also goes to shared memory, where the dQ MMA reads it MN-major
      tmem_st_32x8(lane_base + C::TM_ST + buf * 64 + part * 16, pk_p);
      tmem_st_32x8(lane_base + C::TM_DPT + buf * 64 + part * 16, pk_ds);
      const uint32_t ds_tile = ds_row + (t & 3) * C::DS_BYTES;
      st_shared_v4(ds_tile + ds_c0, pk_ds[0], pk_ds[1], pk_ds[2], pk_ds[3]);
      st_shared_v4(ds_tile + ds_c1, pk_ds[4], pk_ds[5], pk_ds[6], pk_ds[7]);
      tmem_st_wait();
      tc_fence_before();
      fence_async_smem();
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
      // The dQ block of pair P is drained after tile 2P+3: its MMAs were issued after tile 2P+1, two tiles
      // of element-wise work ago, so the drain does not wait; and it comes right AFTER this tile's release
      // (mbarrier arrive = membar), so the reductions have a whole tile to land before the next membar
      // would have to wait for them.
      if ((t & 1) == 1 && t >= 3) drain_dq((t >> 1) - 1, true);
    }
    // what the loop has not drained: the pair before the last when n_q is odd (its slot would have been
    // tile n_q), and the last pair — a single tile when n_q is odd
    if ((n_q & 1) == 1 && n_q >= 3) drain_dq((n_q >> 1) - 1, true);
    drain_dq((n_q - 1) >> 1, (n_q & 1) == 0);

    // dV = P_drop^T · dO, dK = scale * dS^T · Q; this warp drains columns [part*16, +16)
    mbar_wait(acc_full, 0, 84);
    tc_fence_after();
    __nv_bfloat16* base = p.dqkv + ((static_cast<size_t>(batch) * p.s + kv) * p.nh + head) * (3 * HD) + part * (HD / 4);
    drain_scaled<HD / 4>(lane_base + C::TM_DK + part * (HD / 4), base + HD, p.scale, kv < p.s);
    drain_scaled<HD / 4>(lane_base + C::TM_DV + part * (HD / 4), base + 2 * HD, 1.0f, kv < p.s);
    FX_T(7);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == BWD_MMA_WARP) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// dQ accumulator (fp32, per 128-row block [16 column groups][128 rows][4]) -> scale -> bf16 into the q
// slot of the packed dqkv tensor. One thread per 8 hd columns.
__global__ void __launch_bounds__(256)
attn_dq_convert_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, int b, int s, int nh,
                       int s_pad, float scale) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;  // over [b, nh, s_pad, 8]
  const size_t total = static_cast<size_t>(b) * nh * s_pad * 8;
  if (idx >= total) return;
  const int c8 = static_cast<int>(idx & 7);
  const size_t rowi = idx >> 3;
  const int row = static_cast<int>(rowi % s_pad);
  const size_t bh = rowi / s_pad;
  if (row >= s) return;
  const int head = static_cast<int>(bh % nh), batch = static_cast<int>(bh / nh);
  const float4* blk = reinterpret_cast<const float4*>(dq_acc + (bh * (s_pad / 128) + row / 128) * DQACC_BLOCK);
  const float4 x = __ldg(blk + (2 * c8) * 128 + (row & 127)), y = __ldg(blk + (2 * c8 + 1) * 128 + (row & 127));
  uint4 o = make_uint4(pack_bf16(x.x * scale, x.y * scale), pack_bf16(x.z * scale, x.w * scale),
                       pack_bf16(y.x * scale, y.y * scale), pack_bf16(y.z * scale, y.w * scale));
  *reinterpret_cast<uint4*>(dqkv + ((static_cast<size_t>(batch) * s + row) * nh + head) * 192 + c8 * 8) = o;
}
