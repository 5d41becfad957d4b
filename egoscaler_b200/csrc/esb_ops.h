// Internal declarations of the non-GEMM operators (namespace esb); see include/egoscaler_b200.h for semantics.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace esb {

// traj_tokens.cu
int traj_quantize(const void* x, int is_f64, long long n, int num_bins, long long id_offset, long long* out, cudaStream_t s);
int traj_dequantize(const long long* ids, long long n_poses, int num_bins, long long id_offset, int mode, double d_max,
                    double d_min, float* out, cudaStream_t s);
int traj_scale(float* poses, long long n_poses, int mode, double d_max, double d_min, cudaStream_t s);
int traj_ids_to_poses(const long long* ids, int B, int T, int num_bins, long long base, long long tsep, long long eos,
                      int mode, double d_max, double d_min, int max_steps, int pad_repeat, float* out, int* count,
                      cudaStream_t s);


// metrics.cu
int traj_metrics(const float* gen, const float* gt, const int* gen_len, const int* gt_len, int B, int Tg, int Tt, int D, int c0,
                 int c1, int r0, double* out, cudaStream_t s);
int traj_poses_to_ids(const float* poses, const int* lens, int B, int T, int num_bins, long long base, long long ts, long long tsep,
                      long long te, long long eos, long long pad, int norm, const double* range6, int out_ld, long long* out,
                      int* out_len, cudaStream_t s);

int traj_preprocess(const float* traj, const int* lens, int B, int T, int D, int num_steps, int smooth, float* out, int* mask,
                    cudaStream_t s);

// rgbd_frontend.cu
long long rgbd_workspace_bytes(int H, int W);
int rgbd_to_points(const float* rgbd, int H, int W, const int* boxes, int n_boxes, double principal, double fx, double fy,
                   int has_thres, double d_thres, float* out, long long cap, long long* out_count, void* workspace,
                   long long workspace_bytes, cudaStream_t s);
int pc_sample_norm(const float* pts, long long cloud_stride, long long n, const long long* idx, int B, int S, float* out,
                   cudaStream_t s);

// norms.cu
int rmsnorm_fwd(const void* x, long long ldx, const void* w, void* y, long long ldy, float* rstd, long long rows, int d,
                float eps, cudaStream_t s);
int rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres, void* dx, float* dw,
                long long rows, int d, cudaStream_t s);
int layernorm_fwd(const void* x, const void* add, long long add_period, const float* gamma, const float* beta, void* sum_out,
                  void* y, long long rows, int d, float eps, cudaStream_t s);

// elementwise.cu
int rope(void* buf, long long ld, long long rows, int L, int n_heads, int hd, const void* cos_tab, const void* sin_tab,
         int pos0, int inverse, void* kcache, void* vcache, int kv_head0, int kv_heads, long long v_col0, int cache_len,
         const int* pos_ptr, cudaStream_t s);
int silu_mul_fwd(const void* gu, void* act, long long rows, int f, cudaStream_t s);
int silu_mul_bwd(const void* gu, const void* dact, void* dgu, long long rows, int f, cudaStream_t s);
int gelu(const void* z, const void* dy, void* y, long long n, int backward, cudaStream_t s);

// gather.cu
int point_layout(const long long* ids, int B, int L, long long patch_id, long long start_id, long long end_id, int n_point,
                 int* info, cudaStream_t s);
int embed_splice_fwd(const long long* ids, const void* embed, const void* feats, const int* splice_pos, int n_point, void* x,
                     int B, int L, int d, long long vocab, cudaStream_t s);
int embed_splice_bwd(const long long* ids, const void* dx, const int* splice_pos, int n_point, float* dembed, void* dfeats,
                     long long padding_idx, int B, int L, int d, long long vocab, cudaStream_t s);
int embed_rows_pack(const long long* ids, const void* dx, const int* splice_pos, int n_point, int rmax, void* rows_out,
                    long long* ids_out, long long pad, int B, int L, int d, cudaStream_t s);
int embed_scatter_rows(const long long* ids, const void* rows, float* dembed, long long pad, long long vocab, int R, int d, cudaStream_t s);
int embed_scatter_sorted(const long long* sorted_ids, const long long* order, const void* rows, float* dembed, long long pad,
                         long long vocab, int R, int d, cudaStream_t s);
int copy_rows(const void* src, long long src_bstride_bytes, long long src_ld_bytes, void* dst, long long dst_bstride_bytes,
              long long dst_ld_bytes, int batches, int rows, int row_bytes, cudaStream_t s);
int transpose_bf16(const void* in, long long ld_in, void* out, long long ld_out, int R, int C, cudaStream_t s);
int colsum_bf16(const void* in, long long ld, int M, int N, float* out, cudaStream_t s);
int group_maxpool(const void* f, void* out, long long groups, int n, int C, int concat, cudaStream_t s);

// loss_optim.cu
int cross_entropy(void* logits, long long ld, long long rows, int V, const long long* targets, long long ignore_index,
                  float* row_loss, float* scratch_inv_count, float* loss, int write_grad, cudaStream_t s);
int argmax_rows(const void* logits, long long ld, long long rows, int V, long long* out_idx, float* scores, long long scores_ld,
                int suppress_col, const int* step_ptr, int step_off, long long scores_step_stride, cudaStream_t s);
// sampling.cu
int sample_rows(const void* logits, long long ld, long long rows, int V, float temperature, int top_k, float top_p, float penalty,
                const long long* hist, long long hist_ld, int hist_len, const int* pos_ptr, const float* u,
                long long u_step_stride, const int* step_ptr, int step_off, float* scores, long long scores_ld,
                long long scores_step_stride, int suppress_col, long long* out_idx, int* n_kept, cudaStream_t s);
int greedy_update(const long long* next, long long* tokens_out, int* unfinished, long long* sequences, long long seq_ld,
                  int cur_len, uint8_t* mask, long long mask_ld, long long eos, long long pad, int B, int* pos_ptr, cudaStream_t s);
int adamw(float* p, const float* g, float* m, float* v, void* p_bf16, long long n, float lr, float beta1, float beta2, float eps,
          float wd, int step, float grad_scale, cudaStream_t s);
int cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t s);

// attention_fwd.cu / attention_bwd.cu / attention_decode.cu
int attn_fwd(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
             long long v_bs, long long v_rs, void* o, long long o_bs, long long o_rs, float* lse, const uint8_t* kmask,
             long long kmask_ld, int B, int H, int Lq, int Lk, int hd, int causal, int q_pos0, float scale, cudaStream_t s);
int attn_bwd(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
             long long v_bs, long long v_rs, const void* o, long long o_bs, long long o_rs, const void* dout, long long do_bs,
             long long do_rs, const float* lse, float* dvec, void* dq, long long dq_bs, long long dq_rs, void* dk, long long dk_bs,
             long long dk_rs, void* dv, long long dv_bs, long long dv_rs, const uint8_t* kmask, long long kmask_ld, int B, int H,
             int Lq, int Lk, int hd, int causal, int q_pos0, float scale, cudaStream_t s, const void* rope_cos = nullptr,
             const void* rope_sin = nullptr);
int attn_decode(const void* q, long long q_bs, const void* k, const void* v, long long kv_bs, long long kv_rs, void* o,
                long long o_bs, const uint8_t* kmask, long long kmask_ld, const int* len_ptr, int len_static, int B, int H,
                int hd, float scale, cudaStream_t s);

// encoder_train.cu (--unfreeze_pc_encoder)
int layernorm_bwd(const void* dy, const void* x, const float* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                  long long rows, int d, float eps, cudaStream_t s);
int colsum2_bf16(const void* a, long long lda, const void* b, long long ldb, long long M, int N, float* out0, float* out1,
                 cudaStream_t s);
int bn_coeffs(int mode, const float* sum0, const float* sum1, long long R, int C, const float* gamma, const float* beta, float eps,
              float momentum, float* mean_io, float* rstd_io, float* running_mean, float* running_var, float* A, float* B, float* Cc,
              float* dgamma, float* dbeta, cudaStream_t s);
int affine_cols(const void* x, const void* y, const float* A, const float* B, const float* Cc, void* out, long long rows, int C,
                int relu, cudaStream_t s);
int relu_mask(const void* dy, const void* y, void* out, long long n, cudaStream_t s);
int group_maxpool_bwd(const void* f, const void* dpool, const void* dcat, void* df, long long groups, int n, int C, cudaStream_t s);
int scale_rows_add(const void* x, const void* y, const float* scale, void* out, long long rows, int d, long long rows_per_batch,
                   cudaStream_t s);

// pointcloud.cu
int fps(const float* pts, int B, int N, int C, int G, const long long* start, float* centers, int* center_idx, void* centers_pad,
        int pad_ld, cudaStream_t s);
int knn_group(const float* pts, const float* centers, int B, int N, int C, int G, int K, int out_ld, void* nb_out, int* idx_out,
              cudaStream_t s);

}  // namespace esb
