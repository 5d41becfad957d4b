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

#ifdef __cplusplus
}
#endif
#endif /* EGOSCALER_B200_H */
