// comm.cu — K7: the model-parallel allreduce of the row-parallel linears, as our own kernel over
// NVLink 5 / NVSwitch peer memory, fused with the elementwise op that follows it.
//
// Reference semantics (fleet.meta_parallel.RowParallelLinear, call sites hybrid_model.py:157-163,
// 505-511): y = allreduce_sum_mp(x_shard · W_shard), then + bias ONCE, then (in the decoder layer,
// :546-547 / :556-561) dropout under the shared 'global_seed' stream and the residual add. The
// backward of ColumnParallelLinear's `_c_identity` is the same allreduce on dX.
//
// The GEMM writes its partial result straight into a SYMMETRIC buffer (same virtual layout on every
// rank, mapped into every peer and into one NVSwitch multicast address). One kernel then does
//
//   phase 0  flag handshake: every rank's partial is complete (stream order on each GPU)
//   phase 1  reduce-scatter + all-gather through the switch: rank r owns rows [r·T/n, (r+1)·T/n);
//            it reads them with multimem.ld_reduce (the switch sums the n replicas, fp32
//            accumulation) and writes the sum back with multimem.st, which the switch replicates
//            into every rank's copy of the buffer
//   phase 2  counter barrier across all blocks of all ranks
//   phase 3  epilogue: every rank reads the now complete sum from its own copy and applies
//            out = residual + dropout(sum + bias)  (or just copies, for the backward allreduces)
//            into an ordinary local tensor
//   phase 4  counter barrier: nobody overwrites the symmetric buffer while a peer still reads it
//
// so each byte crosses NVLink once in each direction, and the bias/dropout/residual pass costs no
// extra trip through HBM. Flags and counters are monotonic (epoch-stamped), never reset.
#include "comm.cuh"
#include "philox.cuh"

namespace fx {
namespace comm {

constexpr int THREADS = 512;

struct Params {
  uint64_t sig[MAX_RANKS];   // signal pad of every rank, as mapped on this GPU
  uint64_t buf[MAX_RANKS];   // symmetric buffer of every rank, as mapped on this GPU
  uint64_t mc;               // multicast address of the symmetric buffer
  int rank, world;
  int rows, h;               // logical [rows, h] bf16
  uint32_t epoch;            // 1, 2, 3, ... identical on all ranks for the same call
  __nv_bfloat16* out;        // local result [rows, h]
  // epilogue (null bias / residual => plain allreduce)
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  float p_drop, scale;
  uint32_t thresh;
  uint64_t seed, offset;
  uint64_t group0;           // dropout group index of vector 0 (row slices of a larger [T, h] tensor)
  const uint8_t* mask_in;
  int variant;               // 0: reduced slices are multicast to every rank (multimem.st), phase 3 reads locally
                             // 1: reduced slices stay with their owner, phase 3 pulls them with peer loads
                             // 2: one shot (2 ranks): phase 3 adds the peer copy to the local one; no phases 1-2
};

__global__ void __launch_bounds__(THREADS)
tp_allreduce_kernel(const Params p) {
  const int tid = threadIdx.x;

  // ---- phase 0: everyone's partial result is in place
  ranks_ready(p);

  const int hv = p.h >> 3;                                   // 16-byte vectors per row
  const size_t stride = static_cast<size_t>(gridDim.x) * THREADS;
  if (p.variant != 2) {
  // ---- phase 1: reduce-scatter my row slice through the switch, result into my own copy
  const int rows_per = (p.rows + p.world - 1) / p.world;
  const int r0 = p.rank * rows_per;
  const int r1 = min(p.rows, r0 + rows_per);
  const size_t v0 = static_cast<size_t>(r0) * hv, v1 = static_cast<size_t>(max(r1, r0)) * hv;
  uint4* mine = reinterpret_cast<uint4*>(p.buf[p.rank]);
  for (size_t v = v0 + static_cast<size_t>(blockIdx.x) * THREADS + tid; v < v1; v += 4 * stride) {
    uint4 t[4];  // four switch reductions in flight per thread: NVLink round trips are microseconds
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v + u * stride < v1) t[u] = multimem_ld_reduce_bf16x8(p.mc + (v + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (v + u * stride >= v1) continue;
      if (p.variant == 0) multimem_st_v4(p.mc + (v + u * stride) * 16, t[u]);  // the switch writes every replica
      else mine[v + u * stride] = t[u];
    }
  }

  // ---- phase 2: all blocks of all ranks have stored their reduced slices
  rank_barrier(p, SLOT_RS_DONE, SLOT_TICKET_A, true, true, 1);
  } else {
    // the one-shot variant has no reduce-scatter barrier, but it still counts itself in, so that calls of
    // different variants can be mixed on one set of flag pages (every counter advances on every call)
    rank_barrier(p, SLOT_RS_DONE, SLOT_TICKET_A, false, false, 1);
  }

  // ---- phase 3: epilogue over the complete sum; fused bias + dropout + residual.
  // Plain allreduce with out = NULL (multicast variant): the sum is already in every rank's copy, so
  // the symmetric buffer itself is the result and there is nothing to do here.
  const size_t total = p.out == nullptr ? 0 : static_cast<size_t>(p.rows) * hv;
  const size_t slice_v = static_cast<size_t>((p.rows + p.world - 1) / p.world) * hv;
  uint4* mine = reinterpret_cast<uint4*>(p.buf[p.rank]);
  for (size_t vb = static_cast<size_t>(blockIdx.x) * THREADS + tid; vb < total; vb += 4 * stride) {
    uint4 raws[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t v = vb + u * stride;
      if (v >= total) continue;
      if (p.variant == 0) {
        raws[u] = ld_weak_v4(mine + v);
      } else if (p.variant == 1) {
        raws[u] = ld_weak_v4(reinterpret_cast<const uint4*>(p.buf[v / slice_v]) + v);
      } else {  // two ranks, one shot: local + peer, added in fp32
        const uint4 a = ld_weak_v4(mine + v);
        const uint4 b = ld_weak_v4(reinterpret_cast<const uint4*>(p.buf[p.rank ^ 1]) + v);
        float fa[8], fb[8];
        unpack8(a, fa);
        unpack8(b, fb);
#pragma unroll
        for (int j = 0; j < 8; ++j) fa[j] += fb[j];
        raws[u] = pack8(fa);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t v = vb + u * stride;
      if (v >= total) continue;
      uint4 res = raws[u];
      if (p.residual != nullptr) {
        float y[8], r[8];
        unpack8(raws[u], y);
        unpack8(__ldg(reinterpret_cast<const uint4*>(p.residual) + v), r);
        if (p.bias != nullptr) {
          float b[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(p.bias) + (v % hv)), b);
#pragma unroll
          for (int j = 0; j < 8; ++j) y[j] += b[j];
        }
        float o[8];
        if (p.mask_in != nullptr) {
          const uint2 m = __ldg(reinterpret_cast<const uint2*>(p.mask_in) + v);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            o[j] = r[j] + (((m.x >> (8 * j)) & 0xFFu) ? y[j] * p.scale : 0.f);
            o[4 + j] = r[4 + j] + (((m.y >> (8 * j)) & 0xFFu) ? y[4 + j] * p.scale : 0.f);
          }
        } else if (p.p_drop > 0.f) {
          const Philox8 rnd = philox4x32_10(p.seed, p.offset, p.group0 + v);  // group index = vector index in [T, h]
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = r[j] + (philox_keep(rnd, j, p.thresh) ? y[j] * p.scale : 0.f);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = r[j] + y[j];
        }
        res = pack8(o);
      }
      reinterpret_cast<uint4*>(p.out)[v] = res;
    }
  }

  // ---- phase 4: every rank has finished reading every copy
  rank_barrier(p, SLOT_AG_DONE, SLOT_TICKET_B, true, false, 2);
}

// ---------------------------------------------------------------------------------------------------
// Sequence parallelism (sequence_parallel_utils.py:36-126, 150-326): between the attention and FFN blocks
// the activations are sharded along the sequence, rank r holding [s/n, b, h] (sequence-first, the
// reference's layout). The allreduce above then splits into its two halves, with the row permutation
// between the sharded sequence-first layout and the batch-first [b, s, h] layout of the GEMM / attention
// kernels folded into the transfer:
//   mode 0  all-gather (AllGatherOp forward, ReduceScatterOp backward): each rank multicasts its shard
//           rows into every rank's copy of the symmetric buffer (multimem.st), counter barrier;
//   mode 1  reduce-scatter (ReduceScatterOp forward, AllGatherOp backward) fused with bias + dropout +
//           residual: flag handshake, each rank sums ITS rows over all replicas in the switch
//           (multimem.ld_reduce) and applies the epilogue into an ordinary local shard tensor, counter
//           barrier before anyone overwrites the buffer.
// Dropout here is indexed by position in the LOCAL shard: the reference draws the mask of a [s/n, b, h]
// tensor under 'global_seed', which is the same on every mp rank (hybrid_model.py:546, :556).
struct SeqParParams {
  uint64_t sig[MAX_RANKS];
  uint64_t mc;               // multicast address of the [b*s, h] buffer
  int rank, world, mode;
  int b, s, h;               // full tensor [b, s, h]; shard = [s/world, b, h]
  uint32_t epoch;
  const __nv_bfloat16* src;  // mode 0: local shard to publish
  __nv_bfloat16* out;        // mode 1: local shard result
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  float p_drop, scale;
  uint32_t thresh;
  uint64_t seed, offset;
  const uint8_t* mask_in;
};

__global__ void __launch_bounds__(THREADS)
tp_seqpar_kernel(const SeqParParams p) {
  const int tid = threadIdx.x;
  // every call advances every flag and counter, whichever kernel or mode it is (epochs stay in lockstep).
  // Every rank has reached this call, i.e. (stream order) finished everything before it: mode 1 — all
  // partial sums are in place before the switch reads them; mode 0 — nobody still reads the previous
  // contents of its copy (a GEMM over an earlier gather) when the multicast stores arrive.
  ranks_ready(p);
  const int hv = p.h >> 3;
  const int sl = p.s / p.world;                       // sequence positions per rank
  const size_t total = static_cast<size_t>(sl) * p.b * hv;  // 16-byte vectors of the local shard
  const size_t stride = static_cast<size_t>(gridDim.x) * THREADS;
  for (size_t v = static_cast<size_t>(blockIdx.x) * THREADS + tid; v < total; v += stride) {
    const size_t row_l = v / hv;                      // local row = s_local * b + batch
    const int c = static_cast<int>(v - row_l * hv);
    const size_t s_local = row_l / p.b, batch = row_l - s_local * p.b;
    const size_t row_g = batch * p.s + static_cast<size_t>(p.rank) * sl + s_local;  // row in [b, s]
    const uint64_t addr = p.mc + (row_g * hv + c) * 16;
    if (p.mode == 0) {
      multimem_st_v4(addr, __ldg(reinterpret_cast<const uint4*>(p.src) + v));
    } else {
      uint4 res = multimem_ld_reduce_bf16x8(addr);
      if (p.residual != nullptr) {
        float y[8], r[8], o[8];
        unpack8(res, y);
        unpack8(__ldg(reinterpret_cast<const uint4*>(p.residual) + v), r);
        if (p.bias != nullptr) {
          float bb[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(p.bias) + c), bb);
#pragma unroll
          for (int j = 0; j < 8; ++j) y[j] += bb[j];
        }
        if (p.mask_in != nullptr) {
          const uint2 m = __ldg(reinterpret_cast<const uint2*>(p.mask_in) + v);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            o[j] = r[j] + (((m.x >> (8 * j)) & 0xFFu) ? y[j] * p.scale : 0.f);
            o[4 + j] = r[4 + j] + (((m.y >> (8 * j)) & 0xFFu) ? y[4 + j] * p.scale : 0.f);
          }
        } else if (p.p_drop > 0.f) {
          const Philox8 rnd = philox4x32_10(p.seed, p.offset, v);  // group index = vector index in the local shard
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = r[j] + (philox_keep(rnd, j, p.thresh) ? y[j] * p.scale : 0.f);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = r[j] + y[j];
        }
        res = pack8(o);
      }
      reinterpret_cast<uint4*>(p.out)[v] = res;
    }
  }
  // mode 0: the gathered tensor is complete once every block of every rank has stored its rows
  rank_barrier(p, SLOT_RS_DONE, SLOT_TICKET_A, p.mode == 0, true, 1);
  // mode 1: nobody may overwrite the buffer while a peer's switch reduction still reads it
  rank_barrier(p, SLOT_AG_DONE, SLOT_TICKET_B, p.mode == 1, false, 2);
}

static int launch_seqpar(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank, int world,
                         uint32_t epoch, int num_blocks, SeqParParams& q, int b, int s, int h, uint64_t buf_elem_offset,
                         const char* who) {
  FX_CHECK_ARG(world >= 2 && world <= MAX_RANKS && rank >= 0 && rank < world, "%s: bad rank/world (%d/%d)", who, rank, world);
  FX_CHECK_ARG(b > 0 && s > 0 && h > 0 && h % 8 == 0, "%s: b %d, s %d, h %d (h must be a multiple of 8)", who, b, s, h);
  FX_CHECK_ARG(s % world == 0, "%s: sequence length %d can't be divided exactly by sequence parallelism %d", who, s, world);
  FX_CHECK_ARG(peer_signal_pads && multicast_ptr, "%s: null pointer", who);
  FX_CHECK_ARG(epoch > 0, "%s: epoch must start at 1", who);
  FX_CHECK_ARG(buf_elem_offset % 8 == 0, "%s: element offsets must be multiples of 8", who);
  FX_CHECK_ARG(num_blocks > 0 && num_blocks <= num_sms(), "%s: num_blocks must be in [1, #SMs]", who);
  for (int i = 0; i < world; ++i) q.sig[i] = peer_signal_pads[i];
  q.mc = multicast_ptr + buf_elem_offset * 2;
  q.rank = rank; q.world = world;
  q.b = b; q.s = s; q.h = h;
  q.epoch = epoch;
  tp_seqpar_kernel<<<num_blocks, THREADS, 0, static_cast<cudaStream_t>(stream)>>>(q);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

}  // namespace comm
}  // namespace fx

extern "C" int fx_tp_allgather_rows(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank,
                                    int world, uint32_t epoch, int num_blocks, const void* shard, int b, int s, int h,
                                    uint64_t buf_elem_offset) {
  using namespace fx;
  FX_CHECK_ARG(shard != nullptr && (uintptr_t)shard % 16 == 0, "fx_tp_allgather_rows: null or misaligned shard");
  comm::SeqParParams q{};
  q.mode = 0;
  q.src = static_cast<const __nv_bfloat16*>(shard);
  return comm::launch_seqpar(stream, peer_signal_pads, multicast_ptr, rank, world, epoch, num_blocks, q, b, s, h,
                             buf_elem_offset, "fx_tp_allgather_rows");
}

extern "C" int fx_tp_reduce_scatter_bdr(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank,
                                        int world, uint32_t epoch, int num_blocks, void* out_shard, int b, int s, int h,
                                        const void* bias, const void* residual_shard, float p, uint64_t seed,
                                        uint64_t offset, const uint8_t* mask_in, uint64_t buf_elem_offset) {
  using namespace fx;
  FX_CHECK_ARG(out_shard != nullptr && ((uintptr_t)out_shard | (uintptr_t)residual_shard | (uintptr_t)bias) % 16 == 0,
               "fx_tp_reduce_scatter_bdr: null or misaligned pointer");
  FX_CHECK_ARG(p >= 0.f && p < 1.f, "fx_tp_reduce_scatter_bdr: dropout p must be in [0,1) (got %f)", p);
  FX_CHECK_ARG((bias == nullptr && mask_in == nullptr && p == 0.f) || residual_shard != nullptr,
               "fx_tp_reduce_scatter_bdr: bias / dropout need a residual");
  comm::SeqParParams q{};
  q.mode = 1;
  q.out = static_cast<__nv_bfloat16*>(out_shard);
  q.bias = static_cast<const __nv_bfloat16*>(bias);
  q.residual = static_cast<const __nv_bfloat16*>(residual_shard);
  q.p_drop = p; q.scale = 1.0f / (1.0f - p);
  q.thresh = dropout_threshold(p);
  q.seed = seed; q.offset = offset;
  q.mask_in = mask_in;
  return comm::launch_seqpar(stream, peer_signal_pads, multicast_ptr, rank, world, epoch, num_blocks, q, b, s, h,
                             buf_elem_offset, "fx_tp_reduce_scatter_bdr");
}

extern "C" int fx_tp_allreduce_bdr(void* stream, const uint64_t* peer_bufs, const uint64_t* peer_signal_pads,
                                   uint64_t multicast_ptr, int rank, int world, uint32_t epoch, int num_blocks,
                                   void* out, int rows, int h, const void* bias, const void* residual, float p,
                                   uint64_t seed, uint64_t offset, const uint8_t* mask_in, uint64_t buf_elem_offset,
                                   uint64_t dropout_elem_offset) {
  using namespace fx;
  FX_CHECK_ARG(world >= 2 && world <= comm::MAX_RANKS && rank >= 0 && rank < world,
               "fx_tp_allreduce_bdr: bad rank/world (%d/%d)", rank, world);
  FX_CHECK_ARG(rows > 0 && h > 0 && h % 8 == 0, "fx_tp_allreduce_bdr: rows %d, h %d (h must be a multiple of 8)", rows, h);
  FX_CHECK_ARG(peer_bufs && peer_signal_pads && multicast_ptr, "fx_tp_allreduce_bdr: null pointer");
  FX_CHECK_ARG(out != nullptr || residual == nullptr, "fx_tp_allreduce_bdr: out may only be null for a plain allreduce");
  FX_CHECK_ARG(p >= 0.f && p < 1.f, "fx_tp_allreduce_bdr: dropout p must be in [0,1) (got %f)", p);
  FX_CHECK_ARG(epoch > 0, "fx_tp_allreduce_bdr: epoch must start at 1");
  FX_CHECK_ARG(buf_elem_offset % 8 == 0 && dropout_elem_offset % 8 == 0,
               "fx_tp_allreduce_bdr: element offsets must be multiples of 8");
  FX_CHECK_ARG(num_blocks > 0 && num_blocks <= num_sms(), "fx_tp_allreduce_bdr: num_blocks must be in [1, #SMs] "
               "(all blocks of all ranks wait on each other, so they must be co-resident)");
  FX_CHECK_ARG((bias == nullptr && mask_in == nullptr && p == 0.f) || residual != nullptr,
               "fx_tp_allreduce_bdr: bias / dropout need a residual");
  comm::Params q{};
  for (int i = 0; i < world; ++i) {
    q.sig[i] = peer_signal_pads[i];
    q.buf[i] = peer_bufs[i] + buf_elem_offset * 2;
  }
  q.mc = multicast_ptr + buf_elem_offset * 2;
  q.group0 = dropout_elem_offset / 8;
  q.rank = rank; q.world = world;
  q.rows = rows; q.h = h;
  q.epoch = epoch;
  q.out = static_cast<__nv_bfloat16*>(out);
  q.bias = static_cast<const __nv_bfloat16*>(bias);
  q.residual = static_cast<const __nv_bfloat16*>(residual);
  q.p_drop = p; q.scale = 1.0f / (1.0f - p);
  q.thresh = dropout_threshold(p);
  q.seed = seed; q.offset = offset;
  q.mask_in = mask_in;
  // two ranks: one-shot peer loads beat a trip through the switch; otherwise reduce + multicast in the switch
  q.variant = (world == 2 && out != nullptr) ? 2 : 0;
  // tuning / test knob (read once, runtime.cu): force a variant wherever the call admits it
  const int kv = knobs().comm_variant;
  if (kv == 0 || (kv == 1 && out != nullptr) || (kv == 2 && out != nullptr && world == 2)) q.variant = kv;
  FX_CHECK_ARG(out != nullptr || q.variant == 0, "fx_tp_allreduce_bdr: in-place result needs the multicast variant");
  FX_CHECK_ARG(q.variant != 2 || world == 2, "fx_tp_allreduce_bdr: the one-shot variant is for 2 ranks");
  comm::tp_allreduce_kernel<<<num_blocks, comm::THREADS, 0, static_cast<cudaStream_t>(stream)>>>(q);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
