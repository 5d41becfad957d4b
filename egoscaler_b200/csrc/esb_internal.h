// Internal (C++) declarations shared by the translation units of libegoscaler_b200.so.
// The public C ABI is include/egoscaler_b200.h; everything here is namespace esb.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>

namespace esb {

void set_last_error(const char* fmt, ...);
const char* last_error();
int check_launch(const char* what);
int device_sm_count();

// Decode-step launch chain (esb_set_pdl_chain): when on, launch_chain() adds cudaLaunchAttributeProgrammaticStreamSerialization,
// so the kernel's CTAs are scheduled while its predecessor drains and park in griddepcontrol.wait (launch latency and, for
// the skinny GEMM, the weight-ring prefetch overlap the predecessor).  Only kernels that begin with griddep_wait() use it.
extern int g_pdl_chain;

template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_pdl_chain ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows);

// 3-D bf16 tensor {cols (contiguous), rows, batches} with row / batch strides in elements; box = {box_cols, box_rows, 1},
// 128-byte swizzle (box_cols * 2 must be 128).
int make_tma_3d_bf16(CUtensorMap* map, const void* ptr, long long cols, long long rows, long long batches, long long row_stride,
                     long long batch_stride, int box_cols, int box_rows);

int attn_fwd_tc128(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
                   long long v_bs, long long v_rs, void* o, long long o_bs, long long o_rs, float* lse, const uint8_t* kmask,
                   long long kmask_ld, int B, int H, int Lq, int Lk, int causal, int q_pos0, float scale, cudaStream_t s);

int attn_bwd_tc128(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
                   long long v_bs, long long v_rs, const void* o, long long o_bs, long long o_rs, const void* dout, long long do_bs,
                   long long do_rs, const float* lse, float* stats_ws, void* dq, long long dq_bs, long long dq_rs, void* dk,
                   long long dk_bs, long long dk_rs, void* dv, long long dv_bs, long long dv_rs, const uint8_t* kmask,
                   long long kmask_ld, int B, int H, int Lq, int Lk, int causal, int q_pos0, float scale, cudaStream_t s,
                   const void* rope_cos = nullptr, const void* rope_sin = nullptr);

int gemm_bf16(const void* A, long long lda, const void* B, long long ldb, void* C, long long ldc, int M, int N, int K,
              const float* bias, const void* residual, long long ldr, int act, int out_fp32, int accumulate, float alpha,
              int force_bn, cudaStream_t stream);

int gemm_swiglu_bf16(const void* A, long long lda, const void* Wgu, long long ldw, void* act, long long ld_act, void* gu,
                     long long ld_gu, int M, int f, int K, cudaStream_t stream);
int gemm_dswiglu_bf16(const void* dOut, long long lda, const void* WdownT, long long ldb, int trans_b, const void* gu, long long ld_gu,
                      void* dgu, long long ld_dgu, int M, int f, int K, cudaStream_t stream);
int gemm_rope_bf16(const void* A, long long lda, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                   const void* cos_tab, const void* sin_tab, int L, int pos0, int rope_cols, int head_dim, cudaStream_t stream);

int gemm_t_bf16(const void* A, long long lda, int trans_a, const void* B, long long ldb, int trans_b, void* C, long long ldc, int M,
                int N, int Kc, int out_fp32, int accumulate, float alpha, cudaStream_t stream);
int set_gemm_2cta(int on);
int set_gemm_sm_reserve(int k);
long long gemm_skinny_workspace_bytes(int M, int N, int K);
int gemm_skinny_bf16(const void* X, long long ldx, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                     const void* residual, long long ldr, void* workspace, long long workspace_bytes, int force_splits,
                     int w_tiled, cudaStream_t stream);
int gemm_skinny_normed_bf16(const void* X, long long ldx, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                            const void* residual, long long ldr, void* workspace, long long workspace_bytes, int force_splits,
                            int w_tiled, const float* scale_src, int scale_parts, float eps, float* ssq_out, cudaStream_t stream);
long long gemm_skinny_static_scratch_bytes();
int gemm_skinny_static_bf16(const void* X, long long ldx, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                            const void* residual, long long ldr, int force_splits, int w_tiled, cudaStream_t stream);
int retile_weight_scaled_bf16(const void* W, long long ldw, int N, int K, const void* col_scale, void* out, cudaStream_t stream);
long long retiled_weight_elems(int N, int K);
int retile_weight_bf16(const void* W, long long ldw, int N, int K, void* out, cudaStream_t stream);
long long retiled_gateup_elems(int f, int K);
int retile_weight_gateup_bf16(const void* W, long long ldw, int f, int K, const void* col_scale, void* out, cudaStream_t stream);
long long gemm_chain_workspace_bytes(int M, int max_N);
int gemm_chain_ctas_per_sm(int M);
int gemm_chain_debug_buffer(void* buf);
}  // namespace esb
struct esb_chain_phase;      // include/egoscaler_b200.h
namespace esb {
int gemm_chain_bf16(const esb_chain_phase* phases, int n_phases, int M, void* workspace, long long workspace_bytes, int flags,
                    cudaStream_t stream);

}  // namespace esb
