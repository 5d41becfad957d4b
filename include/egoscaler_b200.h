/* egoscaler_b200 -- C ABI of the B200-native (sm_100a) trajectory-VLM hot path.
 *
 * The reference (sk4161/EgoScaler, egoscaler/models/pointllm) has no FFI of its own: its hot path is a Python
 * nn.Module running on stock ATen/cuBLAS kernels.  This header is the lower boundary that a maintainer binds
 * (ctypes stub in INTEGRATION.md) to route each step of that module through hand-written CUDA.  Every entry
 * point names the reference code it replaces (paths relative to egoscaler/models/pointllm/ unless noted;
 * "HF" = transformers/models/llama/modeling_llama.py, the un-vendored dependency the reference calls into).
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name ends in _host; no torch types cross this boundary
 *   - bf16 tensors are raw uint16 storage (void*), row-major, leading dimensions in ELEMENTS
 *   - `stream` is a cudaStream_t passed as void*; every call is stream-ordered, never synchronises,
 *     never allocates device memory (workspaces are passed in)
 *   - return 0 on success, negative ESB_ERR_* otherwise; esb_last_error() gives the message (thread-local)
 */
#ifndef EGOSCALER_B200_H
#define EGOSCALER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ESB_OK 0
#define ESB_ERR_INVALID_ARG (-1)
#define ESB_ERR_UNSUPPORTED (-2)
#define ESB_ERR_CUDA (-3)
#define ESB_ERR_WORKSPACE (-4)
#define ESB_ERR_DRIVER (-5)

/* epilogue activation of esb_gemm_bf16 */
#define ESB_ACT_NONE 0
#define ESB_ACT_GELU 1 /* exact erf GELU: nn.GELU(), pointbert/point_encoder.py:16, pointllm.py:74 */
#define ESB_ACT_RELU 2 /* nn.ReLU in the mini-PointNet, pointbert/dvae.py:196,202 */

/* pose scaler applied after de-quantisation */
#define ESB_SCALE_NONE 0
#define ESB_SCALE_RT2 1    /* utils/utils.py:23-34 rt2_scaler */
#define ESB_SCALE_SIMPLE 2 /* utils/utils.py:36-45 simple_scaler */
#define ESB_SCALE_DENORM 3 /* dataset.py:139-146 denorm(do_norm) */

int esb_version(void);
const char* esb_last_error(void);
/* Launch mode of the single-token decode step (decode.py token_step; reference loop: HF generate driven by model_arch.py:94-108).
 * on != 0: the small kernels of the step (rmsnorm, rope, silu*mul, decode attention, embedding, arg-max, bookkeeping) are
 * launched with programmatic stream serialization, like the skinny GEMMs already are: each parks in griddepcontrol.wait
 * until its predecessor has completed, so results are unchanged and launch latency overlaps.  Process-global; default off. */
int esb_set_pdl_chain(int on);
int esb_device_sm_count(void);

/* ------------------------------------------------------------------------------------------------ GEMM
 * C[M,N] = act(alpha * A[M,K] . B[N,K]^T + bias[N]) + residual[M,N]        (tcgen05/TMEM, TMA-fed)
 * Replaces every nn.Linear / Conv1d(k=1) on the path: HF LlamaAttention q/k/v/o_proj and LlamaMLP
 * gate/up/down_proj (reached from pointllm/model/pointllm.py:173-178), lm_head (pointllm.py:228),
 * point_proj (pointllm.py:67-81,129), PointBERT qkv/proj/fc1/fc2/reduce_dim/pos_embed
 * (pointbert/point_encoder.py:11-55,123-131) and the mini-PointNet convs (pointbert/dvae.py:194-205).
 * B is the weight exactly as stored in the reference state_dict ([out_features, in_features]).
 * A, B: bf16, 16-byte aligned, lda/ldb multiples of 8.  C: bf16 (out_fp32=0) or fp32 (out_fp32=1;
 * accumulate=1 adds into C).  bias fp32 or NULL, residual bf16 or NULL.  force_bn: 0 = auto, or 64/128/256. */
int esb_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N, int K,
                  const float* bias, const void* residual, int64_t ldr, int act, int out_fp32, int accumulate,
                  float alpha, int force_bn, void* stream);

/* GEMMs with the HBM-bound neighbour kernel fused into the epilogue (same tcgen05 kernel, 128x256 tiles):
 * esb_gemm_swiglu_bf16 : HF LlamaMLP (modeling_llama.py:171-184) gate|up projection + act_fn(gate) * up in one launch.
 *     Wgu = [gate_proj ; up_proj] packed [2f, K]; act [M, f] = silu(A.Wg^T) * (A.Wu^T); gu [M, 2f] (gate | up, kept for the
 *     backward) is written when non-NULL.  Bit-identical to esb_gemm_bf16 + esb_silu_mul_fwd.
 * esb_gemm_dswiglu_bf16: backward of the same: dgu [M, 2f] = dSwiGLU(gu, dOut . down_proj); the weight is passed either as
 *     Wd = down_proj^T [f, K] (trans_b = 0) or as down_proj itself [K, f] (trans_b = 1); d(act) is never written.
 *     Bit-identical to esb_gemm_bf16 + esb_silu_mul_bwd.
 * esb_gemm_rope_bf16   : the packed q|k|v projection + apply_rotary_pos_emb (modeling_llama.py:138-168) on the heads (128
 *     columns each) that start below rope_cols, position pos0 + (row % L).  Bit-identical to esb_gemm_bf16 + esb_rope. */
int esb_gemm_swiglu_bf16(const void* A, int64_t lda, const void* Wgu, int64_t ldw, void* act, int64_t ld_act, void* gu,
                         int64_t ld_gu, int M, int f, int K, void* stream);
int esb_gemm_dswiglu_bf16(const void* dOut, int64_t lda, const void* Wd, int64_t ldb, int trans_b, const void* gu, int64_t ld_gu,
                          void* dgu, int64_t ld_dgu, int M, int f, int K, void* stream);
int esb_gemm_rope_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                       const void* cos_tab, const void* sin_tab, int L, int pos0, int rope_cols, int head_dim, void* stream);

/* C[M,N] (+)= alpha * op(A) . op(B) for the two contractions of a Linear's backward (autograd of the nn.Linear call sites
 * above; train.py:183), read from the tensors exactly as they sit in HBM - MN-major tcgen05 operands, no transposed copies:
 *   trans_a = 1, trans_b = 1: A [Kc, M], B [Kc, N] row-major  -> weight gradient dW[out,in] = dY[T,out]^T . X[T,in]
 *   trans_a = 0, trans_b = 1: A [M, Kc], B [Kc, N] row-major  -> data gradient   dX[T,in]  = dY[T,out] . W[out,in]
 * (trans_a = trans_b = 0 is esb_gemm_bf16 without epilogue.)  C bf16 (out_fp32=0) or fp32 (out_fp32=1; accumulate=1 adds). */
int esb_gemm_t_bf16(const void* A, int64_t lda, int trans_a, const void* B, int64_t ldb, int trans_b, void* C, int64_t ldc, int M,
                    int N, int Kc, int out_fp32, int accumulate, float alpha, void* stream);

/* Route the large NT GEMMs (M >= 1024, N >= 512; plain and fused epilogues) through the CTA-pair kernel: clusters of two CTAs,
 * tcgen05.mma.cta_group::2 on 256 x 256 tiles, each SM stages half of the weight tile.  Returns the previous setting.  Default:
 * on (environment variable ESB_GEMM_2CTA=0 turns it off). */
int esb_set_gemm_2cta(int on);
/* Data-parallel training (train.py:92-125, DeepSpeed's gradient all-reduce overlapped with the backward): leave k SMs out of the
 * persistent GEMM grids while a collective is in flight so that NCCL's CTAs get SMs of their own; returns the previous k. */
int esb_set_gemm_sm_reserve(int k);

/* Weight-streaming variant for the cached single-token decode step (M = sequences per GPU <= 64):
 * C[M,N] = X[M,K] . W[N,K]^T + residual[M,N], bf16 in/out, fp32 accumulation, split-K over the SMs with a
 * deterministic in-kernel reduction.  Same Linear call sites as esb_gemm_bf16, reached through generate()
 * (model_arch.py:94-108 -> pointllm.py:255-275).  workspace: device buffer of at least
 * esb_gemm_skinny_workspace_bytes(M,N,K) bytes whose first 16 KiB are ZERO on first use (the kernel leaves them
 * zero); it may be shared by all launches on one stream.  force_splits: 0 = auto.
 * w_tiled = 0: W is the state_dict layout [N,K] (ldw).  w_tiled = 1: W was re-laid out once by
 * esb_retile_weight_bf16 into [ceil(N/128)][ceil(K/64)][128][64] (zero padded) so that every pipeline stage is one
 * contiguous 16 KiB read - the HBM-friendly layout for streaming each weight exactly once (ldw ignored).
 * w_tiled bit 1 (value 2): the caller promises that W is NOT written by the kernel launched immediately before on this
 * stream; the launch then uses programmatic dependent launch and prefetches its first weight stages while that kernel
 * drains (X, residual and workspace are only touched after griddepcontrol.wait). */
int64_t esb_gemm_skinny_workspace_bytes(int M, int N, int K);
int esb_gemm_skinny_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                         const void* residual, int64_t ldr, void* workspace, int64_t workspace_bytes, int force_splits,
                         int w_tiled, void* stream);
/* No counterpart in the reference: added for callers that cannot allocate device memory
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  esb_gemm_skinny_bf16 with its split-K
 * workspace in a file-scope device array of the library (16 KiB of tile counters + 8 MiB of fp32 partials per GPU, held until
 * exit) instead of a caller buffer; inputs and outputs are still the caller's.  Limits: M <= 64, N <= 16384, K <= 16384 -
 * larger sizes are rejected with ESB_ERR_UNSUPPORTED (use esb_gemm_skinny_bf16).  Calls no cudaMalloc / cudaFree / allocating
 * library routine; the counters are cleared by a cudaMemsetAsync on `stream` in every call, and w_tiled bit 1 (programmatic
 * launch) is ignored.  ALL calls on one GPU share the array: calls that overlap on the GPU (different streams) race - serialise
 * them.  The first call on a device resolves the array's address (cudaGetSymbolAddress); make it outside graph capture. */
int64_t esb_gemm_skinny_static_scratch_bytes(void);
int esb_gemm_skinny_static_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                                const void* residual, int64_t ldr, int force_splits, int w_tiled, void* stream);
int64_t esb_retiled_weight_elems(int N, int K);
int esb_retile_weight_bf16(const void* W, int64_t ldw, int N, int K, void* out, void* stream);
/* The decode step with LlamaRMSNorm (HF modeling_llama.py LlamaRMSNorm, called at the three norm sites of
 * pointllm.py:173-178 / :227) folded around the Linear that consumes it:  Linear(norm(x)) = rstd(x) * (x . (W*g)^T).
 *   esb_retile_weight_scaled_bf16: like esb_retile_weight_bf16 with column k multiplied by the norm weight g[k] (bf16).
 *   esb_gemm_skinny_normed_bf16:   like esb_gemm_skinny_bf16, plus
 *     scale_src != NULL: accumulator row m is multiplied by rstd[m] before the residual is added;  scale_parts == 0:
 *       scale_src[m] is rstd itself (esb_rmsnorm_fwd's rstd output); scale_parts > 0: scale_src is [scale_parts][64] partial
 *       sums of squares of x (as left by ssq_out of the GEMM that produced x) and rstd = rsqrt(sum / K + eps);
 *     ssq_out != NULL: [ceil(N/128)][64] fp32, entry [t][m] = sum over the 128 output columns of tile t of bf16(C[m,:])^2,
 *       summed in a fixed order (deterministic) - the next norm's statistics without a separate pass over C. */
int esb_retile_weight_scaled_bf16(const void* W, int64_t ldw, int N, int K, const void* col_scale, void* out, void* stream);
int esb_gemm_skinny_normed_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                                const void* residual, int64_t ldr, void* workspace, int64_t workspace_bytes, int force_splits,
                                int w_tiled, const float* scale_src, int scale_parts, float eps, float* ssq_out, void* stream);

/* The Linear layers BETWEEN two attention calls of the cached token step as ONE persistent launch (csrc/gemm_chain.cu):
 *   o-proj (+residual) -> gate|up (+SwiGLU) -> down-proj (+residual) -> next layer's q|k|v projection (or lm_head),
 * HF LlamaDecoderLayer.forward / LlamaMLP.forward / lm_head as reached from generate() (model_arch.py:94-108 ->
 * pointllm.py:255-275, :173-178, :228).  Up to 4 phases; phase i+1 may read what phase i wrote (grid-wide hand-off inside the
 * kernel), the weight stream runs ahead across the hand-off.  Each phase is an esb_gemm_skinny_normed_bf16 (same meaning of
 * residual / scale_src / scale_parts / eps / ssq_out, K = the length the mean of squares is taken over) on a TILED weight image;
 * swiglu != 0: W_tiles is the esb_retile_weight_gateup_bf16 image of the packed [2f, K] gate|up weight (N = 2f) and
 * C[M, f] = bf16(bf16(silu(gate)) * up) of the bf16-rounded projections (HF LlamaMLP act_fn(gate_proj(x)) * up_proj(x)).
 * workspace: esb_gemm_chain_workspace_bytes(M, largest N) bytes, first 16 KiB ZERO on first use (left zero), private to this
 * entry point.  flags bit 1 (value 2): programmatic dependent launch as for esb_gemm_skinny_bf16.  All CTAs of the launch must
 * be co-resident: do not run it concurrently with other work on the same GPU. */
typedef struct esb_chain_phase {
    const void* X;          /* [M, K] bf16 activations */
    int64_t ldx;
    const void* W_tiles;
    void* C;                /* [M, N] bf16 (swiglu: [M, N/2]) */
    int64_t ldc;
    const void* residual;   /* NULL or [M, N] bf16 */
    int64_t ldr;
    const float* scale_src;
    float* ssq_out;
    int32_t N, K, scale_parts, swiglu;
    float eps;
    int32_t reserved;
} esb_chain_phase;
int64_t esb_gemm_chain_workspace_bytes(int M, int max_N);
/* diagnostics: buf = NULL (off) or a device buffer of (2 * SM count) * 4 * 8 int64 that later launches fill with %globaltimer
 * stamps per CTA and phase (tools/chain_trace.py) */
int esb_gemm_chain_debug_buffer(void* buf);
int64_t esb_gemm_chain_ctas_per_sm(int M);      /* resident CTAs per SM the grid is sized from; 0 = the kernel cannot run here */
int esb_gemm_chain_bf16(const void* phases, int n_phases, int M, void* workspace, int64_t workspace_bytes, int flags, void* stream);
/* [2f, K] packed gate|up weight -> [ceil(f/64)][ceil(K/64)][128][64] tiles whose rows 0..63 are gate rows t*64.. and rows 64..127
 * the up rows f + t*64.. (zero padded); col_scale (NULL or bf16 [K]) folds a norm weight into the columns. */
int64_t esb_retiled_gateup_elems(int f, int K);
int esb_retile_weight_gateup_bf16(const void* W, int64_t ldw, int f, int K, const void* col_scale, void* out, void* stream);

/* ------------------------------------------------------------------------------------------------ trajectory tokens
 * utils/utils.py:13-16 discretize_action: out[i] = np.digitize(x[i], np.linspace(-1,1,num_bins)) - 1 + id_offset
 * (bit-exact incl. edge values, NaN -> num_bins-1, x < -1 -> -1).  x: float32 (is_f64=0) or float64. */
int esb_traj_quantize(const void* x, int is_f64, int64_t n, int num_bins, int64_t id_offset, int64_t* out, void* stream);
/* utils/utils.py:18-21 token_to_action + :96 float32 cast + scaler (:23-45 / dataset.py:139-146).
 * ids [n_poses, 6] int64 (token id = id_offset + bin) -> out [n_poses, 6] float32.  Out-of-range ids give NaN. */
int esb_traj_dequantize(const int64_t* ids, int64_t n_poses, int num_bins, int64_t id_offset, int scale_mode,
                        double d_max, double d_min, float* out, void* stream);
/* In-place scaler on float32 [n_poses, 6] (utils/utils.py:23-45, dataset.py:139-146). */
int esb_traj_scale(float* poses, int64_t n_poses, int scale_mode, double d_max, double d_min, void* stream);
/* Generated ids -> poses: the <tsep>-segment / first-run-of-6-bins / repeat-previous logic of
 * utils/utils.py:47-104 str_to_float(rt2=True) on token ids, EOS cut (train.py:240-244), optional
 * repeat-last padding to max_steps (train.py:251-255).  ids [B,T] int64 -> out [B,max_steps,6] f32, count [B]. */
int esb_traj_ids_to_poses(const int64_t* ids, int B, int T, int num_bins, int64_t base_id, int64_t tsep_id,
                          int64_t eos_id, int scale_mode, double d_max, double d_min, int max_steps, int pad_repeat,
                          float* out, int32_t* count, void* stream);

/* ------------------------------------------------------------------------------------------------ norms
 * HF LlamaRMSNorm.forward (modeling_llama.py:53-67; call sites input_layernorm / post_attention_layernorm / model.norm
 * reached from pointllm/model/pointllm.py:173-178): y = w * bf16(x * rsqrt(mean(x^2) + eps)).  rstd [rows] fp32 is kept
 * for the backward.  x, w, y bf16; d, ldx, ldy multiples of 8. */
int esb_rmsnorm_fwd(const void* x, int64_t ldx, const void* w, void* y, int64_t ldy, float* rstd, int64_t rows, int d,
                    float eps, void* stream);
/* Backward of the above (autograd of modeling_llama.py:53-67): dx = d(rmsnorm)/dx . dy (+ dres if not NULL);
 * dw (fp32 [d], accumulated with atomics) only when not NULL (model.norm is trainable, model_arch.py:33-51). */
int esb_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres, void* dx, float* dw,
                    int64_t rows, int d, void* stream);
/* nn.LayerNorm(eps) of PointBERT (pointbert/point_encoder.py:62,66,142) with the per-block `x + pos`
 * (point_encoder.py:96-97) fused: s = x + add[row % add_period] (written to sum_out when not NULL), y = LN(s). */
int esb_layernorm_fwd(const void* x, const void* add, int64_t add_period, const float* gamma, const float* beta,
                      void* sum_out, void* y, int64_t rows, int d, float eps, void* stream);

/* ------------------------------------------------------------------------------------------------ element-wise
 * HF apply_rotary_pos_emb (modeling_llama.py:138-168) in place on `n_heads` heads of width hd starting at column 0 of
 * buf [rows, ld] (q and k heads of the packed qkv projection); position of row r is pos0 + r % L; cos/sin tables
 * [max_pos, hd/2] bf16 (LlamaRotaryEmbedding, :117-135).  inverse=1 applies the transposed rotation (backward).
 * When kcache != NULL the rotated K heads [kv_head0, kv_head0+kv_heads) and the V heads at column v_col0 are also
 * written to kcache/vcache [B, cache_len, kv_heads*hd] at position pos0 + r % L (DynamicCache.update, :269-270).
 * pos_ptr (device, optional): *pos_ptr is added to pos0 at run time, so a captured CUDA graph of the decode step can
 * be replayed for every token. */
int esb_rope(void* buf, int64_t ld, int64_t rows, int L, int n_heads, int hd, const void* cos_tab, const void* sin_tab,
             int pos0, int inverse, void* kcache, void* vcache, int kv_head0, int kv_heads, int64_t v_col0, int cache_len,
             const int32_t* pos_ptr, void* stream);
/* HF LlamaMLP act_fn(gate_proj(x)) * up_proj(x) (modeling_llama.py:171-184) on gu [rows, 2f] = (gate | up). */
int esb_silu_mul_fwd(const void* gu, void* act, int64_t rows, int f, void* stream);
int esb_silu_mul_bwd(const void* gu, const void* dact, void* dgu, int64_t rows, int f, void* stream);
/* nn.GELU() (exact erf) of point_proj (pointllm/model/pointllm.py:72-79): backward=0: y = gelu(z);
 * backward=1: y = dy * gelu'(z).  n multiple of 8. */
int esb_gelu(const void* z, const void* dy, void* y, int64_t n, int backward, void* stream);

/* ------------------------------------------------------------------------------------------------ embedding + splice
 * Device-side version of the validation scans at pointllm/model/pointllm.py:138-164.
 * info[b] = {n_start, n_end, n_patch, first_start_pos, end_follows_start, patches_consecutive, first_patch_pos, 0}. */
int esb_point_layout(const int64_t* ids, int B, int L, int64_t patch_id, int64_t start_id, int64_t end_id, int n_point,
                     int32_t* info, void* stream);
/* embed_tokens(input_ids) (pointllm.py:107) + splice of the projected point tokens (pointllm.py:134-171):
 * x[b,l] = feats[b, l - pos[b]] inside [pos[b], pos[b]+n_point), embed[ids[b,l]] elsewhere (pos[b] < 0: no splice).
 * ids outside [0, vocab_size) never touch memory: zero row forward, skipped backward. */
int esb_embed_splice_fwd(const int64_t* ids, const void* embed, const void* feats, const int32_t* splice_pos, int n_point,
                         void* x, int B, int L, int d, int64_t vocab_size, void* stream);
/* Backward: point rows -> dfeats (bf16 copy), other rows scatter-added into dembed (fp32, row padding_idx skipped as
 * nn.Embedding(padding_idx) does). */
int esb_embed_splice_bwd(const int64_t* ids, const void* dx, const int32_t* splice_pos, int n_point, float* dembed,
                         void* dfeats, int64_t padding_idx, int B, int L, int d, int64_t vocab_size, void* stream);
/* Sparse exchange of the embed_tokens gradient between data-parallel ranks (replaces the dense [V,d] all-reduce DeepSpeed
 * performs at train.py:183-184): esb_embed_rows_pack compacts the non-point rows of dx [B,L,d] into rows_out [B*rmax, d] bf16 with
 * their token ids (ids_out [B*rmax]; unused slots = pad_id; rmax = L - n_point when every sample carries a point block, else L);
 * after an all-gather, esb_embed_scatter_rows adds the rows of every id into dembed [V,d] fp32 in increasing row order - a
 * deterministic reduction, so every rank ends with bit-identical gradients.  R <= 65536 rows per call. */
int esb_embed_rows_pack(const int64_t* ids, const void* dx, const int32_t* splice_pos, int n_point, int rmax, void* rows_out,
                        int64_t* ids_out, int64_t pad_id, int B, int L, int d, void* stream);
int esb_embed_scatter_rows(const int64_t* ids, const void* rows, float* dembed, int64_t pad_id, int64_t vocab_size, int R, int d,
                           void* stream);
/* esb_embed_scatter_rows for ids the caller has sorted (stable, ascending): order[p] = index into rows of the p-th sorted id.  Same
 * sums in the same order (rows of one id in increasing original index), no limit on R, no O(R^2) scan. */
int esb_embed_scatter_sorted(const int64_t* sorted_ids, const int64_t* order, const void* rows, float* dembed, int64_t pad_id,
                             int64_t vocab_size, int R, int d, void* stream);
/* Batched strided row copy (logits slice train.py:174-175, hidden-row gathers); all byte strides multiples of 16. */
int esb_copy_rows(const void* src, int64_t src_bstride_bytes, int64_t src_ld_bytes, void* dst, int64_t dst_bstride_bytes,
                  int64_t dst_ld_bytes, int batches, int rows, int row_bytes, void* stream);
/* out[c, r] = in[r, c] (bf16): operand re-layout for the weight-gradient GEMMs. */
int esb_transpose_bf16(const void* in, int64_t ld_in, void* out, int64_t ld_out, int R, int C, void* stream);
/* out[n] += sum_m in[m, n]: bias gradients of point_proj. */
int esb_colsum_bf16(const void* in, int64_t ld, int M, int N, float* out, void* stream);
/* mini-PointNet pooling (pointbert/dvae.py:211-215,219): f [groups*n, C] -> concat=0: max over n [groups, C];
 * concat=1: [groups*n, 2C] = (max over n broadcast | f). */
int esb_group_maxpool(const void* f, void* out, int64_t groups, int n, int C, int concat, void* stream);

/* ------------------------------------------------------------------------------------------------ loss / sampling / optimizer
 * train.py:174-181: mean cross-entropy over rows whose target != ignore_index; logits [rows, ld] bf16 (V valid columns).
 * write_grad=1 overwrites logits with dloss/dlogits (bf16).  row_loss [rows], scratch [1], loss [1] fp32 device. */
int esb_cross_entropy(void* logits, int64_t ld, int64_t rows, int V, const int64_t* targets, int64_t ignore_index,
                      float* row_loss, float* scratch, float* loss, int write_grad, void* stream);
/* Greedy step (HF generation/utils.py:2762-2806 via model_arch.py:94-108): first-max argmax per row of bf16 logits,
 * optional fp32 `scores` copy, optional suppressed column (-1: none).  With step_ptr (device) the scores block written
 * is scores + (*step_ptr + step_off) * scores_step_stride (one [rows, V] block per generated token). */
int esb_argmax_rows(const void* logits, int64_t ld, int64_t rows, int V, int64_t* out_idx, float* scores,
                    int64_t scores_ld, int suppress_col, const int32_t* step_ptr, int step_off, int64_t scores_step_stride,
                    void* stream);
/* Sampling step, the reference's default decode mode (model_arch.py:77-108: do_sample=True, top_k=50, top_p=0.95,
 * temperature=1, repetition_penalty=1; HF generation/logits_process.py RepetitionPenalty :302 -> Temperature :236 ->
 * TopK :536 -> TopP :469, then softmax + draw, generation/utils.py _sample).  logits [rows, ld] bf16.  hist (or NULL):
 * ids already in each row, [rows, hist_ld], first hist_len (+ *pos_ptr) entries valid.  u: uniform [0,1) draws,
 * u[step * u_step_stride + row] with step = (*step_ptr or 0) + step_off.  scores (or NULL): processed fp32 scores,
 * -inf where filtered, block `step` of stride scores_step_stride.  out_idx [rows]; n_kept (or NULL) [rows] = survivors.
 * top_k in 1..1024.  The survivor set / scores are exactly HF's on identical logits; the draw is an inverse-CDF walk
 * over survivors in descending-score order (distribution-equal to torch.multinomial, not draw-equal). */
int esb_sample_rows(const void* logits, int64_t ld, int64_t rows, int V, float temperature, int top_k, float top_p,
                    float penalty, const int64_t* hist, int64_t hist_ld, int hist_len, const int32_t* pos_ptr,
                    const float* u, int64_t u_step_stride, const int32_t* step_ptr, int step_off, float* scores,
                    int64_t scores_ld, int64_t scores_step_stride, int suppress_col, int64_t* out_idx, int32_t* n_kept,
                    void* stream);
/* Finished-row bookkeeping of the same loop: finished rows emit pad, EOS finishes a row, token appended to
 * sequences[:, cur_len], attention mask grown by one.  With pos_ptr (device) the column is cur_len + *pos_ptr and
 * *pos_ptr is incremented afterwards (the cache-length counter of the graph-replayed decode step). */
int esb_greedy_update(const int64_t* next, int64_t* tokens_out, int32_t* unfinished, int64_t* sequences, int64_t seq_ld,
                      int cur_len, uint8_t* mask, int64_t mask_ld, int64_t eos, int64_t pad, int B, int32_t* pos_ptr,
                      void* stream);
/* torch.optim.AdamW step (train.py:107-113,184) on fp32 master weights; refreshes the bf16 model copy. */
int esb_adamw(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float lr, float beta1, float beta2,
              float eps, float weight_decay, int step, float grad_scale, void* stream);
int esb_cast_f32_bf16(const float* in, void* out, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------------ attention
 * softmax(Q K^T * scale + mask) V, fp32 softmax.  LLaMA: HF sdpa path (integrations/sdpa_attention.py:40-98) with
 * causal AND key-padding mask; PointBERT: pointbert/point_encoder.py:43-55 (causal=0, kmask=NULL).
 * Element (b,l,h,e) at ptr + b*bs + l*rs + h*hd + e.  lse [B,H,Lq] fp32 (NULL to skip).  hd in {64,128}. */
int esb_attn_fwd(const void* q, int64_t q_bs, int64_t q_rs, const void* k, int64_t k_bs, int64_t k_rs, const void* v,
                 int64_t v_bs, int64_t v_rs, void* o, int64_t o_bs, int64_t o_rs, float* lse, const uint8_t* kmask,
                 int64_t kmask_ld, int B, int H, int Lq, int Lk, int hd, int causal, int q_pos0, float scale, void* stream);
/* Backward of esb_attn_fwd (autograd of the sdpa attention above; needed through all frozen layers because
 * embed_tokens / point_proj are trainable, model_arch.py:33-51).  dvec is fp32 scratch of 2*B*H*round_up(Lq,64) floats
 * (per-row (lse, rowsum(dO o O)) pairs).
 * dq/dk/dv use the same (bs, rs) addressing as q/k/v, so they can be the three slices of one packed [B,L,3,H,hd] buffer. */
int esb_attn_bwd(const void* q, int64_t q_bs, int64_t q_rs, const void* k, int64_t k_bs, int64_t k_rs, const void* v,
                 int64_t v_bs, int64_t v_rs, const void* o, int64_t o_bs, int64_t o_rs, const void* dout, int64_t do_bs,
                 int64_t do_rs, const float* lse, float* dvec, void* dq, int64_t dq_bs, int64_t dq_rs, void* dk, int64_t dk_bs,
                 int64_t dk_rs, void* dv, int64_t dv_bs, int64_t dv_rs, const uint8_t* kmask, int64_t kmask_ld, int B, int H,
                 int Lq, int Lk, int hd, int causal, int q_pos0, float scale, void* stream);
/* esb_attn_bwd with the backward of apply_rotary_pos_emb (modeling_llama.py:138-168) fused into the dq / dk epilogues: the
 * gradients leave the kernels w.r.t. the PRE-rotation q and k (position q_pos0 + query index / key index); rope_cos / rope_sin
 * [max_pos, 64] bf16; head_dim 128 only.  Bit-identical to esb_attn_bwd followed by esb_rope(inverse). */
int esb_attn_bwd_rope(const void* q, int64_t q_bs, int64_t q_rs, const void* k, int64_t k_bs, int64_t k_rs, const void* v,
                      int64_t v_bs, int64_t v_rs, const void* o, int64_t o_bs, int64_t o_rs, const void* dout, int64_t do_bs,
                      int64_t do_rs, const float* lse, float* dvec, void* dq, int64_t dq_bs, int64_t dq_rs, void* dk, int64_t dk_bs,
                      int64_t dk_rs, void* dv, int64_t dv_bs, int64_t dv_rs, const uint8_t* kmask, int64_t kmask_ld, int B, int H,
                      int Lq, int Lk, int hd, int causal, int q_pos0, float scale, const void* rope_cos, const void* rope_sin,
                      void* stream);
/* One query token per sequence against the KV cache (the per-token step of generate, pointllm.py:255-275).
 * Context length = len_static + (*len_ptr when not NULL), so a captured graph follows a device-side counter. */
int esb_attn_decode(const void* q, int64_t q_bs, const void* k, const void* v, int64_t kv_bs, int64_t kv_rs, void* o,
                    int64_t o_bs, const uint8_t* kmask, int64_t kmask_ld, const int32_t* len_ptr, int len_static, int B,
                    int H, int hd, float scale, void* stream);

/* ------------------------------------------------------------------------------------------------ point cloud
 * pointbert/misc.py:40-60 fps: pts [B,N,C] fp32 (xyz first), start [B] = the torch.randint draw of misc.py:52;
 * centers [B,G,3] fp32, center_idx [B,G].  Index-exact with the reference's fp32 arithmetic.
 * centers_bf16_pad (optional) [B*G, pad_ld] bf16 = (x, y, z, 0...) rows, the A operand of pos_embed.0
 * (pointbert/point_encoder.py:126-130). */
int esb_fps(const float* pts, int B, int N, int C, int G, const int64_t* start, float* centers, int32_t* center_idx,
            void* centers_bf16_pad, int pad_ld, void* stream);
/* pointbert/dvae.py:107-140,150-187: k nearest points of every centre, xyz re-centred, emitted as bf16 rows
 * [B*G*K, out_ld] (columns >= C zero) in ascending point-index order; idx_out [B,G,K] optional. */
int esb_knn_group(const float* pts, const float* centers, int B, int N, int C, int G, int K, int out_ld, void* nb_out,
                  int32_t* idx_out, void* stream);

/* ------------------------------------------------------------------------------------------------ trainable point encoder
 * Only used with --unfreeze_pc_encoder (model_arch.py:36-38,121-122): PointBERT then runs in train mode.
 * nn.LayerNorm backward (point_encoder.py:58,64,142): x = LayerNorm input [rows,d] bf16; dx = dLN + dres (dres or NULL);
 * dgamma/dbeta fp32 [d] are ACCUMULATED (or NULL). */
int esb_layernorm_bwd(const void* dy, const void* x, const float* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                      int64_t rows, int d, float eps, void* stream);
/* out0[c] += sum_r a[r,c]; out1[c] += sum_r a[r,c]*b[r,c] (b NULL: a*a).  BatchNorm1d batch statistics, dvae.py:196,202.
 * Deterministic (fixed summation order): the row slices meet in a 16 MiB scratch array of the library, shared by all calls on a
 * GPU - calls on different streams must not overlap.  N <= 4096. */
int esb_colsum2_bf16(const void* a, int64_t lda, const void* b, int64_t ldb, int64_t M, int N, float* out0, float* out1, void* stream);
/* Per-channel BatchNorm coefficients from those sums.  mode 0 (forward): y = A*z + C, saves mean/rstd, updates the running
 * statistics (momentum, unbiased variance).  mode 1 (backward): dz = A*dy + B*z + C, dgamma/dbeta accumulated. */
int esb_bn_coeffs(int mode, const float* sum0, const float* sum1, int64_t R, int C, const float* gamma, const float* beta, float eps,
                  float momentum, float* mean_io, float* rstd_io, float* running_mean, float* running_var, float* A, float* B,
                  float* Cc, float* dgamma, float* dbeta, void* stream);
/* out[r,c] = act(A[c]*x[r,c] + B[c]*y[r,c] + C[c]); y/B may be NULL; relu=1 applies nn.ReLU.  bf16 [rows,C] contiguous. */
int esb_affine_cols(const void* x, const void* y, const float* A, const float* B, const float* Cc, void* out, int64_t rows, int C,
                    int relu, void* stream);
/* nn.ReLU backward: out = y > 0 ? dy : 0 (y = ReLU output). */
int esb_relu_mask(const void* dy, const void* y, void* out, int64_t n, void* stream);
/* torch.max(dim) backward of the mini-PointNet pools (dvae.py:217-220), first arg-max receives the gradient.
 * f [groups*n, C]; dpool [groups, C] (plain) XOR dcat [groups*n, 2C] (gradient of cat(global.expand, f)). */
int esb_group_maxpool_bwd(const void* f, const void* dpool, const void* dcat, void* df, int64_t groups, int n, int C, void* stream);
/* DropPath (timm, point_encoder.py:62,74-75): out = x + scale[row / rows_per_batch] * y; x NULL: out = scale * y (backward). */
int esb_scale_rows_add(const void* x, const void* y, const float* scale, void* out, int64_t rows, int d, int64_t rows_per_batch,
                       void* stream);

/* Data-side tokenisation (inverse of esb_traj_ids_to_poses): poses [B,T,6] f32, lens [B] (NULL: T) ->
 * ids [B,out_ld] int64 = <ts> (6 bins <tsep>) x len <te> eos pad..., out_len [B] (or NULL); bins by utils/utils.py:13-16
 * discretize_action, token layout as parsed by utils/utils.py:47-104 (constant.py templates).  norm=1 first maps physical
 * poses to [-1,1] with the inverse of dataset.py:139-146 (range6 = device {min_x,max_x,min_y,max_y,min_z,max_z}, rot/pi).
 * out_ld >= 3 + 7*T. */
int esb_traj_poses_to_ids(const float* poses, const int32_t* lens, int B, int T, int num_bins, int64_t base_id, int64_t ts_id,
                          int64_t tsep_id, int64_t te_id, int64_t eos_id, int64_t pad_id, int norm, const double* range6, int out_ld,
                          int64_t* out, int32_t* out_len, void* stream);

/* egoscaler/models/utils/traj_utils.py: smoothing_traj :42-96 (smooth != 0; position columns 0..2, the reference's edge rules,
 * fp32) then preprocess_traj :3-40 to num_steps (linspace-index gather when len >= num_steps, else pad with the last point).
 * traj [B,T,D] fp32, lens [B] (NULL: T) -> out [B,num_steps,D], mask [B,num_steps] int32 (1 = real step, 0 = padding; or NULL). */
int esb_traj_preprocess(const float* traj, const int32_t* lens, int B, int T, int D, int num_steps, int smooth, float* out,
                        int32_t* mask, void* stream);

/* ------------------------------------------------------------------------------------------------ RGB-D front-end
 * Producer of `pcrgbs`.  esb_rgbd_to_points replaces egoscaler/data/tools/pcm_tools.py:68-96 get_points_colors
 * (call sites data/train/7_get_object_trajectory.py:191-200,244-253; unprojection also depth.py:55-60): rgbd [H,W,4] fp32
 * (r,g,b in 0..255, metric depth), boxes [n_boxes,4] int32 {xmin,ymin,xmax,ymax} (non-negative; pixels inside are dropped),
 * pinhole (principal_point, focal_x, focal_y; configs/camera.py:6-9), optional z < d_thres.  out [cap,6] fp32 =
 * (x,y,z,r/255,g/255,b/255) of the surviving pixels in row-major order, *out_count = number of survivors (may exceed cap;
 * rows >= cap are not written).  workspace >= esb_rgbd_workspace_bytes(H,W) device bytes. */
int64_t esb_rgbd_workspace_bytes(int H, int W);
int esb_rgbd_to_points(const float* rgbd, int H, int W, const int32_t* boxes, int n_boxes, double principal_point, double focal_x,
                       double focal_y, int has_d_thres, double d_thres, float* out, int64_t cap, int64_t* out_count,
                       void* workspace, int64_t workspace_bytes, void* stream);
/* pointllm/data/utils.py:146-157 pc_norm over a caller-chosen subset: points [B][n,6] fp32 (cloud_stride elements apart),
 * idx [B,S] int64 row indices (NULL: the first S rows) -> out [B,S,6]: xyz - mean, / max L2 norm (fp64 arithmetic), rgb copied. */
int esb_pc_sample_norm(const float* points, int64_t cloud_stride, int64_t n, const int64_t* idx, int B, int S, float* out,
                       void* stream);

/* ------------------------------------------------------------------------------------------------ metrics
 * egoscaler/models/utils/metrics.py: average_displacement_error :37-54, final_displacement_error :7-26,
 * initial_displacement_error :28-35, anglar_distance :60-88, per sample (train.py:258-260, evaluate.py:145-147).
 * gen [B,Tg,D], gt [B,Tt,D] fp32; gen_len / gt_len [B] valid steps (NULL = all); a shorter gen repeats its last pose, a
 * longer one is truncated.  out [B,4] fp64 = (ADE, FDE, IDE, angular distance); displacement over columns [c0,c1),
 * rotation vector in columns [r0,r0+3) (r0 < 0: no angular distance).  gen_len <= 0 -> NaN row (sample skipped). */
int esb_traj_metrics(const float* gen, const float* gt, const int32_t* gen_len, const int32_t* gt_len, int B, int Tg, int Tt, int D,
                     int c0, int c1, int r0, double* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EGOSCALER_B200_H */
