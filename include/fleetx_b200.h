/* fleetx_b200.h — C ABI of libfleetx_b200.so
 *
 * B200 (sm_100a) kernels for ONE path of PaddleFleetX (gongweibao/FleetX): the GPT hybrid-
 * parallel TransformerDecoderLayer, forward and backward
 * (ppfleetx/models/language_model/gpt/dygraph/hybrid_model.py:69-347, 439-571).
 *
 * The reference has no FFI for this path: its layer is Python over Paddle ops. This header is
 * therefore the boundary a Paddle custom op (PD_BUILD_OP), a ctypes stub or any other host would
 * bind; every entry point names the reference lines whose arithmetic it replaces.
 *
 * Conventions
 *  - plain pointers and sizes; all tensor pointers are DEVICE pointers unless a name says host;
 *  - `stream` is a cudaStream_t passed as void*; calls enqueue work and return without syncing;
 *  - activations / linear weights / linear biases are bf16, LayerNorm gamma/beta and all
 *    statistics are fp32; matrices are row-major; linear weights use Paddle's [in, out] layout;
 *  - the caller owns every buffer, including workspaces (sizes from the *_workspace_bytes
 *    functions); nothing is retained after return;
 *  - return value: FX_OK (0) or a negative FX_ERR_*; fx_last_error() gives thread-local text;
 *  - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef FLEETX_B200_H_
#define FLEETX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  FX_OK = 0,
  FX_ERR_INVALID_ARGUMENT = -1,
  FX_ERR_CUDA = -2,
  FX_ERR_UNSUPPORTED = -3,
  FX_ERR_COMM = -4
};

enum { FX_BF16 = 0, FX_F32 = 1 };

/* GEMM epilogues (fused into the TMEM drain of fx_gemm_bf16). */
enum {
  FX_EPI_NONE = 0,      /* C = acc                                                          */
  FX_EPI_BIAS = 1,      /* C = acc + bias[n]                                                */
  FX_EPI_BIAS_GELU = 2, /* Z = acc + bias[n] (stored to aux if non-null); C = gelu_tanh(Z)  */
  FX_EPI_DGELU = 3      /* C = acc * gelu_tanh'(aux[m,n])   (aux = saved pre-activation Z)  */
};

/* ---- runtime -------------------------------------------------------------------------------- */
const char* fx_last_error(void);
int fx_version(void);
/* Number of kernels this library has launched in this process (all threads). */
uint64_t fx_launch_count(void);

/* ---- LayerNorm: nn.LayerNorm(d_model, epsilon=1e-5), hybrid_model.py:513-514, used :534,:554 -- */
/* y = (x - mean) * rstd * gamma + beta over the last dim (biased variance). x,y bf16 [rows,h];
 * mean,rstd fp32 [rows] are saved for the backward pass. h % 8 == 0, h <= 16384. */
int fx_layernorm_fwd(void* stream, const void* x, const float* gamma, const float* beta, void* y,
                     float* mean, float* rstd, int rows, int h, float eps);
size_t fx_layernorm_bwd_workspace_bytes(int rows, int h);
/* dx = LN'(dy) (+ dresidual if non-null: the gradient arriving on the skip path, fused so the
 * residual fan-in add of hybrid_model.py:547/:561 costs no extra pass); dgamma,dbeta fp32 [h]. */
int fx_layernorm_bwd(void* stream, const void* dy, const void* x, const float* gamma,
                     const float* mean, const float* rstd, const void* dresidual, void* dx,
                     float* dgamma, float* dbeta, void* workspace, int rows, int h);

/* No counterpart in the reference. LayerNorm backward for callers that cannot allocate device
 * memory: a runtime that planned every buffer before the step and forbids cudaMalloc afterwards
 * (it synchronises the device), a static-graph custom op that receives only its declared inputs
 * and outputs, a latency-critical loop. Same arguments and results as fx_layernorm_bwd (up to fp32
 * summation order of dgamma/dbeta) minus the workspace: the partial sums live in a fixed array
 * inside the library's device image (148 x 2 x 4096 fp32 = 4,849,664 bytes, held by every GPU that
 * loads the library until the process exits). It never calls cudaMalloc/cudaFree or any allocating
 * routine.
 *  - h must be <= 4096 (a multiple of 8); larger h is rejected with FX_ERR_INVALID_ARGUMENT.
 *  - The first call on each device looks up the array's device address (with that device current);
 *    make that first call before starting any stream capture.
 *  - All calls on a device share the one array. Nothing in it needs clearing between calls, but
 *    two calls whose kernels overlap on the GPU (different streams, no dependency) race on it:
 *    keep calls on one stream or order them with events. */
int fx_layernorm_bwd_noalloc(void* stream, const void* dy, const void* x, const float* gamma,
                             const float* mean, const float* rstd, const void* dresidual, void* dx,
                             float* dgamma, float* dbeta, int rows, int h);

/* ---- bias + dropout + residual: hybrid_model.py:546-547 (dropout1) and :556-561 (dropout2) ---- */
/* out = residual + dropout(y + bias, p, upscale_in_train). bias may be null. The keep mask is a
 * pure function of (seed, offset, element index) (Philox4x32-10) so backward and recompute
 * (hybrid_model.py:332-335, 406-408, 537-539) regenerate it. mask_in (uint8 [rows*h], 1 = keep)
 * overrides the generator when non-null (parity tests inject masks); mask_out, when non-null,
 * receives the mask that was used. All of y, residual, out are bf16 [rows,h]; h % 8 == 0. */
int fx_bias_dropout_residual_fwd(void* stream, const void* y, const void* bias,
                                 const void* residual, void* out, int rows, int h, float p,
                                 uint64_t seed, uint64_t offset, const uint8_t* mask_in,
                                 uint8_t* mask_out);
/* dy = dout * mask / (1-p)  (bf16 [rows,h]); the residual branch gradient is dout itself. */
int fx_bias_dropout_residual_bwd(void* stream, const void* dout, void* dy, int rows, int h,
                                 float p, uint64_t seed, uint64_t offset, const uint8_t* mask_in);

/* Column sum of a bf16 [rows,n] matrix into fp32 or bf16 [n] (bias gradients of the four
 * linears: paddle autograd of `x·W + b`, hybrid_model.py:125-131,157-163,497-511). */
size_t fx_colsum_workspace_bytes(int rows, int n);
int fx_colsum_bf16(void* stream, const void* x, int ldx, void* out, int out_dtype, void* workspace,
                   int rows, int n);

/* ---- GEMM: the four linears and their gradients ------------------------------------------------
 * qkv_proj hybrid_model.py:125-131,166; out_proj :157-163,340; linear1 :497-503,559;
 * linear2 :505-511,558 (fleet.meta_parallel Column/RowParallelLinear local matmuls).
 * C[M,N] = op(A)[M,K] * op(B)[K,N], bf16 operands, fp32 accumulation in TMEM (tcgen05).
 *   trans_a = 0: A stored [M,K] row-major (lda >= K);  1: A stored [K,M] (lda >= M)
 *   trans_b = 0: B stored [K,N] row-major (ldb >= N);  1: B stored [N,K] (ldb >= K)
 * forward y = x·W: (0,0); dgrad dx = dy·W^T: (0,1); wgrad dW = x^T·dy: (1,0).
 * c_dtype FX_BF16 or FX_F32; accumulate != 0 adds into C. bias bf16 [N]; aux bf16 [M,ldaux].
 * All leading dimensions, N, K (and M when trans_a) must be multiples of 8; pointers 16-byte aligned. */
int fx_gemm_bf16(void* stream, int trans_a, int trans_b, int M, int N, int K, const void* A,
                 int lda, const void* B, int ldb, void* C, int ldc, int c_dtype, int accumulate,
                 int epilogue, const void* bias, void* aux, int ldaux);

/* The same product for outputs with few tiles and a long K (the weight gradients dW = x^T·dy at a high mp
 * degree: 256 x 2048 x 16384 is 16 tiles for 148 SMs): K is cut into `splits` ranges whose fp32 partial tiles
 * go to `workspace` (fx_gemm_splitk_workspace_bytes) and are summed — in a fixed order, so the result is
 * deterministic — into C (ldc = N, bf16 or fp32, overwritten). fx_gemm_splitk_splits proposes a split count for
 * a shape (1: use fx_gemm_bf16). No epilogue, no accumulate. */
int fx_gemm_splitk_splits(int M, int N, int K);
size_t fx_gemm_splitk_workspace_bytes(int M, int N, int splits);
int fx_gemm_bf16_splitk(void* stream, int trans_a, int trans_b, int M, int N, int K, const void* A, int lda,
                        const void* B, int ldb, void* C, int c_dtype, int splits, void* workspace);
/* The persistent GEMM kernels launched by the calling thread use at most `ctas` CTAs from now on (0: one per SM).
 * For a host that runs one of the cross-rank collectives on another stream at the same time: those kernels hold
 * their SMs until every rank is done, and a persistent GEMM CTA waiting for such an SM would hold its tiles back.
 * Returns the previous value. Host-side state only; no counterpart in the reference. */
int fx_set_sm_budget(int ctas);

/* ---- fused causal self-attention: hybrid_model.py:165-186 (_fuse_prepare_qkv) + :268-298 ------
 * qkv is the UNTOUCHED output of qkv_proj, bf16 [b, s, nh_local, 3, hd] (per head q|k|v, as the
 * reshape/transpose/split at :167-173 reads it); out is bf16 [b, s, nh_local*hd], i.e. already
 * in the merged-head layout of :293-296. softmax(scale·QK^T + causal) is computed online, the
 * s×s matrix never reaches HBM; lse fp32 [b, nh_local, s] is saved for backward. Dropout on the
 * probabilities (:279-285, 'local_seed' stream): keep bits are a pure function of (seed, offset,
 * batch, head, row, col) (Philox4x32-10); the forward call draws them once into dropout_mask
 * (fx_attn_dropout_mask_bytes bytes, 1 bit per (row, col); may be NULL when p_drop == 0) and the
 * backward call reads the same buffer. hd is 64 or 128. */
size_t fx_attn_dropout_mask_bytes(int b, int s, int nh);
int fx_attn_causal_fwd(void* stream, const void* qkv, void* out, float* lse, void* dropout_mask, int b,
                       int s, int nh, int hd, float scale, float p_drop, uint64_t seed, uint64_t offset);
/* The same call for sequences [batch0, batch0 + b) of a larger batch (the layer pipelines batch halves through
 * its collectives): qkv / out / lse / dropout_mask point at this part, and the keep bits are those the whole-batch
 * call draws for these sequences. */
int fx_attn_causal_fwd_at(void* stream, const void* qkv, void* out, float* lse, void* dropout_mask, int b, int s,
                          int nh, int hd, float scale, float p_drop, uint64_t seed, uint64_t offset, int batch0,
                          int mask_ready);
/* The keep bits alone — a pure function of (seed, offset, shape), so a host can draw them on another stream while
 * the QKV projection runs (integer work beside tensor-core work) and then call fx_attn_causal_fwd_at with
 * mask_ready = 1, which skips the draw. */
int fx_attn_dropout_mask(void* stream, void* dropout_mask, int b, int s, int nh, float p_drop, uint64_t seed,
                         uint64_t offset, int batch0);
size_t fx_attn_causal_bwd_workspace_bytes(int b, int s, int nh, int hd);
/* dqkv bf16, same packed layout as qkv, fully overwritten. dropout_mask: what the forward wrote.
 * hd == 64 runs ONE kernel that produces dK, dV and adds the dQ contribution of each KV tile into an
 * fp32 accumulator in the workspace (L2 reductions; their order is not fixed, so dQ may differ in the
 * last bf16 bit between runs — dK, dV are exact). With FX_ATTN_BWD=split in the environment, and always
 * for hd == 128, the deterministic pair of kernels (dK/dV, then dQ) runs instead. */
int fx_attn_causal_bwd(void* stream, const void* dout, const void* qkv, const void* out,
                       const float* lse, const void* dropout_mask, void* dqkv, void* workspace, int b,
                       int s, int nh, int hd, float scale, float p_drop);

/* ---- model-parallel allreduce over NVLink/NVSwitch peer memory, fused with what follows it ------
 * Replaces `_mp_allreduce` of fleet.meta_parallel.RowParallelLinear (call sites hybrid_model.py:
 * 157-163, 505-511) together with the bias add and the dropout + residual of :546-547 / :556-561,
 * and (with bias = residual = NULL, p = 0) the allreduce that is the backward of `_c_identity`
 * in front of ColumnParallelLinear.
 *
 * The partial result of this rank (bf16 [rows,h]) must sit in a SYMMETRIC buffer: allocated by the
 * host on every rank, mapped into every peer (peer_bufs[i] = rank i's copy as addressed from this
 * GPU) and into one NVSwitch multicast address (multicast_ptr). peer_signal_pads[i] is rank i's
 * zero-initialised flag page (>= 8 KiB), mapped likewise. All ranks call with the same rows, h,
 * num_blocks and epoch; epoch counts the calls on this set of buffers: 1, 2, 3, ...
 * out (ordinary local memory) = residual + dropout(sum + bias), or just the sum; with out = NULL
 * (plain allreduce only) the sum is left in every rank's copy of the symmetric buffer.
 * num_blocks <= #SMs: the blocks of all ranks wait on each other and must be co-resident. One
 * process per GPU. Row slices of a larger [T,h] tensor (the layer pipelines two batch halves so that
 * one half's transfer hides under the other half's GEMMs): buf_elem_offset is where this slice starts
 * inside the symmetric buffer, dropout_elem_offset where it starts inside [T,h] (the dropout mask is
 * indexed by position in the whole tensor); both in elements, multiples of 8; out / residual /
 * mask_in already point at the slice. */
int fx_tp_allreduce_bdr(void* stream, const uint64_t* peer_bufs, const uint64_t* peer_signal_pads,
                        uint64_t multicast_ptr, int rank, int world, uint32_t epoch, int num_blocks,
                        void* out, int rows, int h, const void* bias, const void* residual, float p,
                        uint64_t seed, uint64_t offset, const uint8_t* mask_in, uint64_t buf_elem_offset,
                        uint64_t dropout_elem_offset);

/* ---- sequence parallelism: sequence_parallel_utils.py:36-126 (scatter / all_gather / reduce_scatter and
 * the four PyLayers built on them), used by Column/RowSequenceParallelLinear (:150-326) ------------------
 * Between the attention and FFN blocks rank r holds the rows of ITS s/n sequence positions, sequence-first:
 * a shard is bf16 [s/n, b, h] (the reference's layout). The two halves of the allreduce above, with the
 * permutation to the batch-first [b, s, h] layout of the GEMM / attention kernels folded in; same symmetric
 * buffer, flag pages, epoch counting and num_blocks rule as fx_tp_allreduce_bdr (calls of all three may be
 * mixed on one set of buffers; every call counts as one epoch).
 *  - fx_tp_allgather_rows: AllGatherOp forward / ReduceScatterOp backward. Publishes this rank's shard;
 *    on return (stream order) every rank's copy of the symmetric buffer holds the full [b, s, h] tensor.
 *  - fx_tp_reduce_scatter_bdr: ReduceScatterOp forward / AllGatherOp backward, fused with what follows
 *    RowSequenceParallelLinear in the decoder layer: out_shard = residual_shard + dropout(sum + bias)
 *    over this rank's rows of the partial sums in the symmetric buffers (bias = residual = NULL, p = 0:
 *    the plain reduce-scatter). The dropout mask is indexed by position in the LOCAL shard — the reference
 *    draws the mask of a [s/n, b, h] tensor under 'global_seed', which is identical on all mp ranks. */
int fx_tp_allgather_rows(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank, int world,
                         uint32_t epoch, int num_blocks, const void* shard, int b, int s, int h,
                         uint64_t buf_elem_offset);
int fx_tp_reduce_scatter_bdr(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank,
                             int world, uint32_t epoch, int num_blocks, void* out_shard, int b, int s, int h,
                             const void* bias, const void* residual_shard, float p, uint64_t seed, uint64_t offset,
                             const uint8_t* mask_in, uint64_t buf_elem_offset);

/* ---- the allreduce with the row-wise work done by the OWNER of each row -----------------------------------
 * Same reference sites as fx_tp_allreduce_bdr (RowParallelLinear `_mp_allreduce` + bias, hybrid_model.py:157-163,
 * 505-511; dropout + residual :546-547, :556-561; the backward of `_c_identity`), plus what the layer does to the
 * summed tensor next: the LayerNorm of :553-554 in forward; LayerNorm backward, the skip-path add and dropout1's
 * backward (Paddle autograd of :531-554) in backward. Rank r owns rows [r*ceil(rows/world), ...). It pulls its
 * rows of the partial sums (symmetric buffer behind multicast_in) summed by the switch, applies the chain named
 * by `mode`, keeps the owner-only results in ordinary local memory and multicasts the tensor every rank needs
 * next into every rank's copy of a SECOND symmetric buffer (multicast_out); on return (stream order) that copy
 * is complete on this GPU. Flag pages / epoch / num_blocks rules as fx_tp_allreduce_bdr; h <= fx_tp_rows_max_h().
 * Tensors marked [rows, h] are indexed by global row, [own, h] by row - first own row.
 *   FX_ROWS_FWD_LN      x1 = residual[rows,h] + dropout(sum + bias) -> x_own[own,h]; mean, rstd [own];
 *                       LayerNorm(x1; gamma, beta, eps) -> multicast_out
 *   FX_ROWS_FWD_OUT     residual[own,h] + dropout(sum + bias) -> multicast_out
 *   FX_ROWS_BWD_LN_BDR  d_x1 = LayerNorm'(sum; x_own, mean, rstd [own], gamma) + dres[rows,h] -> dx_own[own,h];
 *                       dropout'(d_x1) -> multicast_out; acc[3,h] (fp32) += dgamma | dbeta | column sums of the
 *                       multicast tensor, over the own rows
 *   FX_ROWS_BWD_LN      LayerNorm'(sum; x_full[rows,h], mean, rstd [rows], gamma) + dres[own,h] -> multicast_out;
 *                       acc[2,h] += dgamma | dbeta over the own rows
 * Dropout masks are indexed by position in the whole [rows, h] tensor (the 'global_seed' stream: identical on all
 * ranks). No counterpart for the sharing of rows exists in the reference: there every
 * rank repeats the row-wise ops on the replicated tensor; the results are the same tensors. */
enum { FX_ROWS_FWD_LN = 0, FX_ROWS_FWD_OUT = 1, FX_ROWS_BWD_LN_BDR = 2, FX_ROWS_BWD_LN = 3 };
typedef struct fx_tp_rows_desc {
  const uint64_t* peer_signal_pads; /* host array [world]                                                  */
  uint64_t multicast_in;            /* NVSwitch multicast address of the partial sums, bf16 [rows, h]      */
  uint64_t multicast_out;           /* ... of the buffer that receives the gathered result, bf16 [rows, h] */
  int rank, world, num_blocks, mode;
  uint32_t epoch;
  int rows, h;
  const void* bias;                 /* bf16 [h] or NULL (FWD modes)                                        */
  const void* residual;             /* bf16, see mode                                                      */
  float p;                          /* dropout probability (FWD modes, BWD_LN_BDR)                         */
  uint64_t seed, offset;
  uint64_t dropout_elem_offset;     /* where this call's [rows, h] starts inside the tensor the mask is defined
                                       on (row slices of a larger [T, h]: the layer pipelines batch halves)  */
  const uint8_t* mask_in;           /* optional injected keep flags, uint8 [rows, h] (test hook)           */
  const float* gamma;               /* fp32 [h]                                                            */
  const float* beta;                /* fp32 [h] (FWD_LN)                                                   */
  float eps;
  void* x_own;                      /* bf16 [own, h]: FWD_LN out, BWD_LN_BDR in                            */
  const void* x_full;               /* bf16 [rows, h]: BWD_LN in                                           */
  float* mean;                      /* FWD_LN out / BWD_LN_BDR in: [own]; BWD_LN in: [rows]                */
  float* rstd;
  const void* dres;                 /* bf16, see mode                                                      */
  void* dx_own;                     /* bf16 [own, h]: BWD_LN_BDR out                                       */
  float* acc;                       /* fp32 [3, h] / [2, h] (BWD modes)                                    */
  int acc_accumulate;               /* 0: the call zeroes acc first; 1: it adds to what is there (a later
                                       batch chunk of the same backward pass)                               */
  uint64_t multicast_sum;           /* BWD_LN, optional (sum_out != NULL): once every rank's rows are done, the  */
  float* sum_out;                   /* fp32 vector behind multicast_sum (sum_n floats in a symmetric buffer: the */
  int sum_n;                        /* ranks' acc) is summed over the ranks into sum_out, as fx_tp_allreduce_f32  */
} fx_tp_rows_desc;
int fx_tp_rows_max_h(void);
int fx_tp_rows_fused(void* stream, const fx_tp_rows_desc* desc);
/* Sum over the ranks of a small fp32 vector that sits in a symmetric buffer (multicast_ptr; n a multiple of 4):
 * every rank gets the sums in `out` (ordinary local memory). Stands in for all_reduce_gradient_hook
 * (sequence_parallel_utils.py:136-140) and, with the row-owner kernels above, for the replicated computation of
 * the LayerNorm / row-parallel-bias gradients. Counts as one epoch on the flag pages. */
int fx_tp_allreduce_f32(void* stream, const uint64_t* peer_signal_pads, uint64_t multicast_ptr, int rank, int world,
                        uint32_t epoch, float* out, int n);

/* ---- the whole decoder layer: TransformerDecoderLayer.forward, hybrid_model.py:525-566 -------------
 * One call runs every kernel above in the order of the reference forward (LN -> qkv_proj -> core_attn
 * -> out_proj -> allreduce + bias + dropout1 + residual -> LN -> linear1 + GELU -> linear2 -> allreduce
 * + bias + dropout2 + residual; the post-LN order when normalize_before == 0), resp. its adjoint
 * (Paddle autograd in the reference), on ONE stream, with no host synchronisation and no allocation:
 * what a Paddle custom op or any non-Python host binds instead of sequencing the kernels itself.
 * All pointers are device pointers, per-rank shards in the reference's layouts (weights [in, out]). */
typedef struct fx_tp_comm {
  const uint64_t* peer_bufs;        /* host array [world]: every rank's symmetric buffer as mapped here  */
  const uint64_t* peer_signal_pads; /* host array [world]: every rank's flag page                        */
  uint64_t multicast_ptr;           /* NVSwitch multicast address of the symmetric buffer (0: none)      */
  void* local_buf;                  /* this rank's symmetric buffer, >= b*s*h bf16                       */
  int rank, world, num_blocks;
  uint32_t* epoch;                  /* host counter of allreduces issued on these buffers; updated       */
} fx_tp_comm;

typedef struct fx_layer_desc {
  int b, s, h;                /* micro-batch, sequence, d_model                                         */
  int nh_local, hd, f_local;  /* heads and FFN columns of THIS rank (nhead / n, dim_feedforward / n)    */
  int normalize_before;       /* hybrid_model.py:453; the reference configs use 1                       */
  float eps;                  /* LayerNorm epsilon (1e-5 at :513-514)                                   */
  float p_attn, p_hidden1, p_hidden2; /* attention dropout (:279-285), dropout1 (:546), dropout2 (:556); 0 in eval */
  uint64_t seed_attn, offset_attn;    /* 'local_seed' stream draw                                       */
  uint64_t seed_h1, offset_h1, seed_h2, offset_h2; /* 'global_seed' stream draws                        */
  const float *norm1_w, *norm1_b, *norm2_w, *norm2_b;                /* fp32 [h]                        */
  const void *qkv_w, *qkv_b;   /* bf16 [h, 3*nh_local*hd], [3*nh_local*hd]  (per head q|k|v)            */
  const void *out_w, *out_b;   /* bf16 [nh_local*hd, h], [h]                                             */
  const void *lin1_w, *lin1_b; /* bf16 [h, f_local], [f_local]                                           */
  const void *lin2_w, *lin2_b; /* bf16 [f_local, h], [h]                                                 */
  const fx_tp_comm* tp;        /* NULL: single rank (no collective)                                      */
} fx_layer_desc;

typedef struct fx_layer_grads { /* caller-allocated, fully overwritten; dtypes as the parameters         */
  float *norm1_w, *norm1_b, *norm2_w, *norm2_b;
  void *qkv_w, *qkv_b, *out_w, *out_b, *lin1_w, *lin1_b, *lin2_w, *lin2_b;
} fx_layer_grads;

/* Bytes of activations kept from forward to backward / of scratch either call may overwrite. */
size_t fx_decoder_layer_saved_bytes(const fx_layer_desc* d);
size_t fx_decoder_layer_workspace_bytes(const fx_layer_desc* d);
/* x, out: bf16 [b, s, h]. `saved` receives the activations backward needs (x itself is not copied). */
int fx_decoder_layer_fwd(void* stream, const fx_layer_desc* d, const void* x, void* out, void* saved,
                         void* workspace);
/* dout, dx: bf16 [b, s, h]; x and saved as given to / filled by the forward call with the same desc. */
int fx_decoder_layer_bwd(void* stream, const fx_layer_desc* d, const void* dout, const void* x,
                         const void* saved, void* dx, const fx_layer_grads* grads, void* workspace);

/* ---- around the decoder stack: GPTModelHybrid / GPTForPretrainingHybrid ----------------
 * GPTEmbeddings.forward, hybrid_model.py:603-622: out[t,:] = dropout(word[t,:] + wpe[pos[t],:]) with
 * word[t,:] = wte[ids[t] - vocab_start,:] on the rank that owns the id and zeros elsewhere
 * (fleet.meta_parallel.VocabParallelEmbedding, :589-593). ids / pos_ids int64 [T] (pos_ids NULL: t % s, the
 * default of :604-607); wte bf16 [vocab_rows, h] (this rank's slice), wpe bf16 [max_pos, h]. One rank: one
 * call. mp > 1: call once with wpe = NULL, p = 0 (the partial lookup), allreduce that tensor, then call
 * again with word_in = the sum (ids / wte ignored) for the position term and the dropout. */
int fx_embedding_fwd(void* stream, const int64_t* ids, const int64_t* pos_ids, const void* wte, int64_t vocab_start,
                     int64_t vocab_rows, const void* word_in, const void* wpe, void* out, int64_t T, int s, int h,
                     float p, uint64_t seed, uint64_t offset);
/* Adjoint: d = dropout'(dout); dwte[ids[t] - vocab_start,:] += d[t,:], dwpe[pos[t],:] += d[t,:], both fp32
 * accumulators (either may be NULL) that are ADDED to with L2 reductions, in no fixed order: zero them, or
 * let dwte already hold the logits GEMM's share of the tied weight's gradient (:779-781). */
int fx_embedding_bwd(void* stream, const void* dout, const int64_t* ids, const int64_t* pos_ids, float* dwte,
                     int64_t vocab_start, int64_t vocab_rows, float* dwpe, int64_t T, int s, int h, float p,
                     uint64_t seed, uint64_t offset);
int fx_f32_to_bf16(void* stream, const float* src, void* dst, int64_t n); /* n % 8 == 0 */

/* GPTPretrainingCriterionHybird.forward, hybrid_model.py:801-829: softmax cross entropy per token over the
 * vocabulary (CrossEntropyLoss(reduction="none"), or fleet.meta_parallel.ParallelCrossEntropy when the
 * logits are split over mp ranks: logits bf16 [T, vl] = columns [vocab_start, vocab_start + vl)), then
 * sum(loss * loss_mask) / sum(loss_mask).
 *   _stats:   lmax[t] = max_c l[t,c]; lsum[t] = sum_c exp(l[t,c] - lmax[t]); target[t] = l[t, label] or 0
 *   mp > 1:   gmax = allreduce_max(lmax); _rescale (lsum *= exp(lmax - gmax)); allreduce_sum(lsum, target)
 *   _loss:    out2[0] = sum_t (log gsum + gmax - target) * mask / sum_t mask, out2[1] = sum_t mask;
 *             loss_tok (may be NULL) receives the per-token losses; loss_mask NULL = all ones
 *   _bwd:     dlogits[t,c] = (exp(l - gmax) / gsum - [c == label]) * mask[t] / out2[1] * dloss[0]
 *             (dloss NULL = 1; dlogits may alias logits). */
int fx_cross_entropy_stats(void* stream, const void* logits, const int64_t* labels, int64_t vocab_start, int64_t T,
                           int vl, float* lmax, float* lsum, float* target);
int fx_cross_entropy_rescale(void* stream, const float* lmax, const float* gmax, float* lsum, int64_t T);
int fx_cross_entropy_loss(void* stream, const float* gmax, const float* gsum, const float* target,
                          const float* loss_mask, float* loss_tok, float* out2, int64_t T);
int fx_cross_entropy_bwd(void* stream, const void* logits, void* dlogits, const int64_t* labels, int64_t vocab_start,
                         int64_t T, int vl, const float* gmax, const float* gsum, const float* loss_mask,
                         const float* out2, const float* dloss);

/* ---- KV-cache self-attention in inference: MultiHeadAttention with use_cache=True,
 * hybrid_model.py:165-186 (k, v of the new tokens concatenated behind cache.k / cache.v) and :268-298 in
 * eval (no dropout). cache_k / cache_v: bf16 [b, nh, cap, hd], the first L rows per (batch, head) valid — the
 * reference's Cache(k, v) of shape [b, nh, L, hd] is the [:, :, :L] view. qkv: the packed qkv_proj output of
 * the t new tokens, bf16 [b, t, nh, 3, hd].
 *  - fx_kv_cache_append writes the new k, v into rows [L, L + t) (L + t <= cap).
 *  - fx_attn_decode (after the append): out[b, t, nh*hd]; new token i attends to rows [0, L + i]. Meant for
 *    small t (decoding); a long prompt runs fx_attn_causal_fwd and then fx_kv_cache_append. */
int fx_kv_cache_append(void* stream, const void* qkv, void* cache_k, void* cache_v, int b, int t, int nh, int hd,
                       int cap, int L);
int fx_attn_decode(void* stream, const void* qkv, const void* cache_k, const void* cache_v, void* out, int b, int t,
                   int nh, int hd, int cap, int L, float scale);

/* ---- FusedAdamW + ClipGradByGlobalNorm: ppfleetx/optims/optimizer.py:29-50,
 * pretrain_gpt_base.yaml:69-84. Every parameter of the model in two launches: the host keeps, in DEVICE
 * memory, a table of tensors and a table of chunks (pairs {tensor index, chunk index}, chunk_elems elements
 * each, covering every tensor). decay = weight_decay for parameters that decay (names without "bias" /
 * "norm", optimizer.py:40-44), else 0. master (fp32 copy of a bf16 parameter, multi_precision) may be NULL. */
typedef struct fx_adam_tensor {
  void* param;       /* bf16 or fp32 [n] */
  float* master;     /* fp32 [n] or NULL */
  const void* grad;  /* bf16 or fp32 [n] */
  float* m;          /* fp32 [n] first moment  */
  float* v;          /* fp32 [n] second moment */
  int64_t n;
  float decay;
  int32_t param_f32, grad_f32, is_distributed;
} fx_adam_tensor;
/* out_sq2[0] = sum of squared gradients of the mp-distributed parameters, out_sq2[1] = of the replicated
 * ones (the host allreduces [0] over the mp group: HybridParallelClipGrad). Zeroes out_sq2 first. */
int fx_multi_sq_norm(void* stream, const fx_adam_tensor* tensors_dev, const int32_t* chunks_dev, int n_chunks,
                     int chunk_elems, float* out_sq2);
/* g *= clip_norm / max(sqrt(sq_norm2[0] + sq_norm2[1]), clip_norm) (sq_norm2 NULL or clip_norm <= 0: no
 * clipping; read on the device, no host sync); p *= 1 - lr*decay; m, v updated; p -= lr/(1-b1^step) *
 * m / (sqrt(v)/sqrt(1-b2^step) + eps). step counts from 1. */
int fx_multi_adamw(void* stream, const fx_adam_tensor* tensors_dev, const int32_t* chunks_dev, int n_chunks,
                   int chunk_elems, float lr, float beta1, float beta2, float eps, int64_t step,
                   const float* sq_norm2, float clip_norm);

#ifdef __cplusplus
}
#endif
#endif /* FLEETX_B200_H_ */
