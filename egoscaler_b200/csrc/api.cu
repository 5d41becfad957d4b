// extern "C" surface of libegoscaler_b200.so (declared in include/egoscaler_b200.h) + error plumbing.
#include "esb_internal.h"
#include "esb_ops.h"
#include "../../include/egoscaler_b200.h"

#include <cstdarg>
#include <cstdio>

namespace esb {

static thread_local char g_err[512] = "";
int g_pdl_chain = 0;

void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* last_error() { return g_err; }

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_last_error("%s: %s", what, cudaGetErrorString(e));
        return -3;
    }
    return 0;
}

int device_sm_count() {
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev] = v;
    }
    return sms[dev];
}

}  // namespace esb

using namespace esb;
#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int esb_version(void) { return 100; }
const char* esb_last_error(void) { return last_error(); }
int esb_device_sm_count(void) { return device_sm_count(); }

int esb_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N, int K,
                  const float* bias, const void* residual, int64_t ldr, int act, int out_fp32, int accumulate,
                  float alpha, int force_bn, void* stream) {
    return gemm_bf16(A, lda, B, ldb, C, ldc, M, N, K, bias, residual, ldr, act, out_fp32, accumulate, alpha, force_bn,
                     ST(stream));
}

int esb_gemm_swiglu_bf16(const void* A, int64_t lda, const void* Wgu, int64_t ldw, void* act, int64_t ld_act, void* gu,
                         int64_t ld_gu, int M, int f, int K, void* stream) {
    return gemm_swiglu_bf16(A, lda, Wgu, ldw, act, ld_act, gu, ld_gu, M, f, K, ST(stream));
}
int esb_gemm_dswiglu_bf16(const void* dOut, int64_t lda, const void* Wd, int64_t ldb, int trans_b, const void* gu, int64_t ld_gu,
                          void* dgu, int64_t ld_dgu, int M, int f, int K, void* stream) {
    return gemm_dswiglu_bf16(dOut, lda, Wd, ldb, trans_b, gu, ld_gu, dgu, ld_dgu, M, f, K, ST(stream));
}
int esb_gemm_rope_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                       const void* cos_tab, const void* sin_tab, int L, int pos0, int rope_cols, int head_dim, void* stream) {
    return gemm_rope_bf16(A, lda, W, ldw, C, ldc, M, N, K, cos_tab, sin_tab, L, pos0, rope_cols, head_dim, ST(stream));
}

int esb_gemm_t_bf16(const void* A, int64_t lda, int trans_a, const void* B, int64_t ldb, int trans_b, void* C, int64_t ldc, int M,
                    int N, int Kc, int out_fp32, int accumulate, float alpha, void* stream) {
    return gemm_t_bf16(A, lda, trans_a, B, ldb, trans_b, C, ldc, M, N, Kc, out_fp32, accumulate, alpha, ST(stream));
}
int esb_set_gemm_2cta(int on) { return set_gemm_2cta(on); }
int esb_set_gemm_sm_reserve(int k) { return set_gemm_sm_reserve(k); }
int64_t esb_gemm_skinny_workspace_bytes(int M, int N, int K) { return gemm_skinny_workspace_bytes(M, N, K); }
int esb_gemm_skinny_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                         const void* residual, int64_t ldr, void* workspace, int64_t workspace_bytes, int force_splits,
                         int w_tiled, void* stream) {
    return gemm_skinny_bf16(X, ldx, W, ldw, C, ldc, M, N, K, residual, ldr, workspace, workspace_bytes, force_splits,
                            w_tiled, ST(stream));
}
int64_t esb_gemm_skinny_static_scratch_bytes(void) { return gemm_skinny_static_scratch_bytes(); }
int esb_gemm_skinny_static_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                                const void* residual, int64_t ldr, int force_splits, int w_tiled, void* stream) {
    return gemm_skinny_static_bf16(X, ldx, W, ldw, C, ldc, M, N, K, residual, ldr, force_splits, w_tiled, ST(stream));
}
int64_t esb_retiled_weight_elems(int N, int K) { return retiled_weight_elems(N, K); }
int esb_retile_weight_bf16(const void* W, int64_t ldw, int N, int K, void* out, void* stream) {
    return retile_weight_bf16(W, ldw, N, K, out, ST(stream));
}

int esb_retile_weight_scaled_bf16(const void* W, int64_t ldw, int N, int K, const void* col_scale, void* out, void* stream) {
    return retile_weight_scaled_bf16(W, ldw, N, K, col_scale, out, ST(stream));
}
int esb_gemm_skinny_normed_bf16(const void* X, int64_t ldx, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, int K,
                                const void* residual, int64_t ldr, void* workspace, int64_t workspace_bytes, int force_splits,
                                int w_tiled, const float* scale_src, int scale_parts, float eps, float* ssq_out, void* stream) {
    return gemm_skinny_normed_bf16(X, ldx, W, ldw, C, ldc, M, N, K, residual, ldr, workspace, workspace_bytes, force_splits, w_tiled,
                                   scale_src, scale_parts, eps, ssq_out, ST(stream));
}

int64_t esb_gemm_chain_workspace_bytes(int M, int max_N) { return gemm_chain_workspace_bytes(M, max_N); }
int64_t esb_gemm_chain_ctas_per_sm(int M) { return gemm_chain_ctas_per_sm(M); }
int esb_gemm_chain_debug_buffer(void* buf) { return gemm_chain_debug_buffer(buf); }
int esb_gemm_chain_bf16(const void* phases, int n_phases, int M, void* workspace, int64_t workspace_bytes, int flags, void* stream) {
    return gemm_chain_bf16(static_cast<const esb_chain_phase*>(phases), n_phases, M, workspace, workspace_bytes, flags, ST(stream));
}
int64_t esb_retiled_gateup_elems(int f, int K) { return retiled_gateup_elems(f, K); }
int esb_retile_weight_gateup_bf16(const void* W, int64_t ldw, int f, int K, const void* col_scale, void* out, void* stream) {
    return retile_weight_gateup_bf16(W, ldw, f, K, col_scale, out, ST(stream));
}

int esb_traj_quantize(const void* x, int is_f64, int64_t n, int num_bins, int64_t id_offset, int64_t* out, void* stream) {
    return traj_quantize(x, is_f64, n, num_bins, id_offset, reinterpret_cast<long long*>(out), ST(stream));
}
int esb_traj_dequantize(const int64_t* ids, int64_t n_poses, int num_bins, int64_t id_offset, int scale_mode,
                        double d_max, double d_min, float* out, void* stream) {
    return traj_dequantize(reinterpret_cast<const long long*>(ids), n_poses, num_bins, id_offset, scale_mode, d_max,
                           d_min, out, ST(stream));
}
int esb_traj_scale(float* poses, int64_t n_poses, int scale_mode, double d_max, double d_min, void* stream) {
    return traj_scale(poses, n_poses, scale_mode, d_max, d_min, ST(stream));
}
int esb_traj_ids_to_poses(const int64_t* ids, int B, int T, int num_bins, int64_t base_id, int64_t tsep_id,
                          int64_t eos_id, int scale_mode, double d_max, double d_min, int max_steps, int pad_repeat,
                          float* out, int32_t* count, void* stream) {
    return traj_ids_to_poses(reinterpret_cast<const long long*>(ids), B, T, num_bins, base_id, tsep_id, eos_id,
                             scale_mode, d_max, d_min, max_steps, pad_repeat, out, count, ST(stream));
}

int esb_traj_poses_to_ids(const float* poses, const int32_t* lens, int B, int T, int num_bins, int64_t base_id, int64_t ts_id,
                          int64_t tsep_id, int64_t te_id, int64_t eos_id, int64_t pad_id, int norm, const double* range6, int out_ld,
                          int64_t* out, int32_t* out_len, void* stream) {
    return traj_poses_to_ids(poses, lens, B, T, num_bins, base_id, ts_id, tsep_id, te_id, eos_id, pad_id, norm, range6, out_ld,
                             reinterpret_cast<long long*>(out), out_len, ST(stream));
}

int esb_traj_preprocess(const float* traj, const int32_t* lens, int B, int T, int D, int num_steps, int smooth, float* out,
                        int32_t* mask, void* stream) {
    return traj_preprocess(traj, lens, B, T, D, num_steps, smooth, out, mask, ST(stream));
}

int esb_set_pdl_chain(int on) {
    g_pdl_chain = on ? 1 : 0;
    return ESB_OK;
}

int64_t esb_rgbd_workspace_bytes(int H, int W) { return rgbd_workspace_bytes(H, W); }

int esb_rgbd_to_points(const float* rgbd, int H, int W, const int32_t* boxes, int n_boxes, double principal_point, double focal_x,
                       double focal_y, int has_d_thres, double d_thres, float* out, int64_t cap, int64_t* out_count,
                       void* workspace, int64_t workspace_bytes, void* stream) {
    return rgbd_to_points(rgbd, H, W, boxes, n_boxes, principal_point, focal_x, focal_y, has_d_thres, d_thres, out, cap,
                          reinterpret_cast<long long*>(out_count), workspace, workspace_bytes, ST(stream));
}

int esb_pc_sample_norm(const float* points, int64_t cloud_stride, int64_t n, const int64_t* idx, int B, int S, float* out,
                       void* stream) {
    return pc_sample_norm(points, cloud_stride, n, reinterpret_cast<const long long*>(idx), B, S, out, ST(stream));
}

}  // extern "C"
