// extern "C" wrappers (part 2): norms, element-wise, gather, loss/optimizer, attention, point cloud.
#include "esb_internal.h"
#include "esb_ops.h"
#include "../../include/egoscaler_b200.h"

using namespace esb;
#define ST(s) reinterpret_cast<cudaStream_t>(s)
typedef long long ll;
#define LLP(p) reinterpret_cast<ll*>(p)
#define CLLP(p) reinterpret_cast<const ll*>(p)

extern "C" {

int esb_rmsnorm_fwd(const void* x, int64_t ldx, const void* w, void* y, int64_t ldy, float* rstd, int64_t rows, int d,
                    float eps, void* stream) { return rmsnorm_fwd(x, ldx, w, y, ldy, rstd, rows, d, eps, ST(stream)); }
int esb_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres, void* dx, float* dw,
                    int64_t rows, int d, void* stream) { return rmsnorm_bwd(dy, x, w, rstd, dres, dx, dw, rows, d, ST(stream)); }
int esb_layernorm_fwd(const void* x, const void* add, int64_t add_period, const float* gamma, const float* beta,
                      void* sum_out, void* y, int64_t rows, int d, float eps, void* stream) {
    return layernorm_fwd(x, add, add_period, gamma, beta, sum_out, y, rows, d, eps, ST(stream));
}
int esb_rope(void* buf, int64_t ld, int64_t rows, int L, int n_heads, int hd, const void* cos_tab, const void* sin_tab,
             int pos0, int inverse, void* kcache, void* vcache, int kv_head0, int kv_heads, int64_t v_col0, int cache_len,
             const int32_t* pos_ptr, void* stream) {
    return rope(buf, ld, rows, L, n_heads, hd, cos_tab, sin_tab, pos0, inverse, kcache, vcache, kv_head0, kv_heads, v_col0,
                cache_len, pos_ptr, ST(stream));
}
int esb_silu_mul_fwd(const void* gu, void* act, int64_t rows, int f, void* stream) { return silu_mul_fwd(gu, act, rows, f, ST(stream)); }
int esb_silu_mul_bwd(const void* gu, const void* dact, void* dgu, int64_t rows, int f, void* stream) {
    return silu_mul_bwd(gu, dact, dgu, rows, f, ST(stream));
}
int esb_gelu(const void* z, const void* dy, void* y, int64_t n, int backward, void* stream) { return gelu(z, dy, y, n, backward, ST(stream)); }

int esb_point_layout(const int64_t* ids, int B, int L, int64_t patch_id, int64_t start_id, int64_t end_id, int n_point,
                     int32_t* info, void* stream) {
    return point_layout(CLLP(ids), B, L, patch_id, start_id, end_id, n_point, info, ST(stream));
}
int esb_embed_splice_fwd(const int64_t* ids, const void* embed, const void* feats, const int32_t* splice_pos, int n_point,
                         void* x, int B, int L, int d, int64_t vocab_size, void* stream) {
    return embed_splice_fwd(CLLP(ids), embed, feats, splice_pos, n_point, x, B, L, d, vocab_size, ST(stream));
}
int esb_embed_splice_bwd(const int64_t* ids, const void* dx, const int32_t* splice_pos, int n_point, float* dembed,
                         void* dfeats, int64_t padding_idx, int B, int L, int d, int64_t vocab_size, void* stream) {
    return embed_splice_bwd(CLLP(ids), dx, splice_pos, n_point, dembed, dfeats, padding_idx, B, L, d, vocab_size, ST(stream));
}
int esb_embed_rows_pack(const int64_t* ids, const void* dx, const int32_t* splice_pos, int n_point, int rmax, void* rows_out,
                        int64_t* ids_out, int64_t pad_id, int B, int L, int d, void* stream) {
    return embed_rows_pack(CLLP(ids), dx, splice_pos, n_point, rmax, rows_out, reinterpret_cast<long long*>(ids_out), pad_id, B, L, d, ST(stream));
}
int esb_embed_scatter_rows(const int64_t* ids, const void* rows, float* dembed, int64_t pad_id, int64_t vocab_size, int R, int d,
                           void* stream) {
    return embed_scatter_rows(CLLP(ids), rows, dembed, pad_id, vocab_size, R, d, ST(stream));
}
int esb_embed_scatter_sorted(const int64_t* sorted_ids, const int64_t* order, const void* rows, float* dembed, int64_t pad_id,
                             int64_t vocab_size, int R, int d, void* stream) {
    return embed_scatter_sorted(CLLP(sorted_ids), CLLP(order), rows, dembed, pad_id, vocab_size, R, d, ST(stream));
}
int esb_copy_rows(const void* src, int64_t src_bstride_bytes, int64_t src_ld_bytes, void* dst, int64_t dst_bstride_bytes,
                  int64_t dst_ld_bytes, int batches, int rows, int row_bytes, void* stream) {
    return copy_rows(src, src_bstride_bytes, src_ld_bytes, dst, dst_bstride_bytes, dst_ld_bytes, batches, rows, row_bytes, ST(stream));
}
int esb_transpose_bf16(const void* in, int64_t ld_in, void* out, int64_t ld_out, int R, int C, void* stream) {
    return transpose_bf16(in, ld_in, out, ld_out, R, C, ST(stream));
}
int esb_colsum_bf16(const void* in, int64_t ld, int M, int N, float* out, void* stream) { return colsum_bf16(in, ld, M, N, out, ST(stream)); }
int esb_group_maxpool(const void* f, void* out, int64_t groups, int n, int C, int concat, void* stream) {
    return group_maxpool(f, out, groups, n, C, concat, ST(stream));
}

int esb_cross_entropy(void* logits, int64_t ld, int64_t rows, int V, const int64_t* targets, int64_t ignore_index,
                      float* row_loss, float* scratch, float* loss, int write_grad, void* stream) {
    return cross_entropy(logits, ld, rows, V, CLLP(targets), ignore_index, row_loss, scratch, loss, write_grad, ST(stream));
}
int esb_argmax_rows(const void* logits, int64_t ld, int64_t rows, int V, int64_t* out_idx, float* scores,
                    int64_t scores_ld, int suppress_col, const int32_t* step_ptr, int step_off, int64_t scores_step_stride,
                    void* stream) {
    return argmax_rows(logits, ld, rows, V, LLP(out_idx), scores, scores_ld, suppress_col, step_ptr, step_off, scores_step_stride,
                       ST(stream));
}
int esb_sample_rows(const void* logits, int64_t ld, int64_t rows, int V, float temperature, int top_k, float top_p,
                    float penalty, const int64_t* hist, int64_t hist_ld, int hist_len, const int32_t* pos_ptr,
                    const float* u, int64_t u_step_stride, const int32_t* step_ptr, int step_off, float* scores,
                    int64_t scores_ld, int64_t scores_step_stride, int suppress_col, int64_t* out_idx, int32_t* n_kept,
                    void* stream) {
    return sample_rows(logits, ld, rows, V, temperature, top_k, top_p, penalty, CLLP(hist), hist_ld, hist_len, pos_ptr, u,
                       u_step_stride, step_ptr, step_off, scores, scores_ld, scores_step_stride, suppress_col, LLP(out_idx), n_kept,
                       ST(stream));
}
int esb_greedy_update(const int64_t* next, int64_t* tokens_out, int32_t* unfinished, int64_t* sequences, int64_t seq_ld,
                      int cur_len, uint8_t* mask, int64_t mask_ld, int64_t eos, int64_t pad, int B, int32_t* pos_ptr, void* stream) {
    return greedy_update(CLLP(next), LLP(tokens_out), unfinished, LLP(sequences), seq_ld, cur_len, mask, mask_ld, eos, pad, B, pos_ptr,
                         ST(stream));
}
int esb_adamw(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float lr, float beta1, float beta2,
              float eps, float weight_decay, int step, float grad_scale, void* stream) {
    return adamw(p, g, m, v, p_bf16, n, lr, beta1, beta2, eps, weight_decay, step, grad_scale, ST(stream));
}
int esb_cast_f32_bf16(const float* in, void* out, int64_t n, void* stream) { return cast_f32_bf16(in, out, n, ST(stream)); }

int esb_attn_fwd(const void* q, int64_t q_bs, int64_t q_rs, const void* k, int64_t k_bs, int64_t k_rs, const void* v,
                 int64_t v_bs, int64_t v_rs, void* o, int64_t o_bs, int64_t o_rs, float* lse, const uint8_t* kmask,
                 int64_t kmask_ld, int B, int H, int Lq, int Lk, int hd, int causal, int q_pos0, float scale, void* stream) {
    return attn_fwd(q, q_bs, q_rs, k, k_bs, k_rs, v, v_bs, v_rs, o, o_bs, o_rs, lse, kmask, kmask_ld, B, H, Lq, Lk, hd, causal,
                    q_pos0, scale, ST(stream));
}
int esb_attn_bwd(const void* q, int64_t q_bs, int64_t q_rs, const void* k, int64_t k_bs, int64_t k_rs, const void* v,
                 int64_t v_bs, int64_t v_rs, const void* o, int64_t o_bs, int64_t o_rs, const void* dout, int64_t do_bs,
                 int64_t do_rs, const float* lse, float* dvec, void* dq, int64_t dq_bs, int64_t dq_rs, void* dk, int64_t dk_bs,
                 int64_t dk_rs, void* dv, int64_t dv_bs, int64_t dv_rs, const uint8_t* kmask, int64_t kmask_ld, int B, int H,
                 int Lq, int Lk, int hd, int causal, int q_pos0, float scale, void* stream) {
    return attn_bwd(q, q_bs, q_rs, k, k_bs, k_rs, v, v_bs, v_rs, o, o_bs, o_rs, dout, do_bs, do_rs, lse, dvec, dq, dq_bs, dq_rs, dk,
                    dk_bs, dk_rs, dv, dv_bs, dv_rs, kmask, kmask_ld, B, H, Lq, Lk, hd, causal, q_pos0, scale, ST(stream));
}
int esb_attn_bwd_rope(const void* q, int64_t q_bs, int64_t q_rs, const void* k, int64_t k_bs, int64_t k_rs, const void* v,
                      int64_t v_bs, int64_t v_rs, const void* o, int64_t o_bs, int64_t o_rs, const void* dout, int64_t do_bs,
                      int64_t do_rs, const float* lse, float* dvec, void* dq, int64_t dq_bs, int64_t dq_rs, void* dk, int64_t dk_bs,
                      int64_t dk_rs, void* dv, int64_t dv_bs, int64_t dv_rs, const uint8_t* kmask, int64_t kmask_ld, int B, int H,
                      int Lq, int Lk, int hd, int causal, int q_pos0, float scale, const void* rope_cos, const void* rope_sin,
                      void* stream) {
    return attn_bwd(q, q_bs, q_rs, k, k_bs, k_rs, v, v_bs, v_rs, o, o_bs, o_rs, dout, do_bs, do_rs, lse, dvec, dq, dq_bs, dq_rs, dk,
                    dk_bs, dk_rs, dv, dv_bs, dv_rs, kmask, kmask_ld, B, H, Lq, Lk, hd, causal, q_pos0, scale, ST(stream), rope_cos, rope_sin);
}
int esb_attn_decode(const void* q, int64_t q_bs, const void* k, const void* v, int64_t kv_bs, int64_t kv_rs, void* o,
                    int64_t o_bs, const uint8_t* kmask, int64_t kmask_ld, const int32_t* len_ptr, int len_static, int B,
                    int H, int hd, float scale, void* stream) {
    return attn_decode(q, q_bs, k, v, kv_bs, kv_rs, o, o_bs, kmask, kmask_ld, len_ptr, len_static, B, H, hd, scale, ST(stream));
}

int esb_fps(const float* pts, int B, int N, int C, int G, const int64_t* start, float* centers, int32_t* center_idx,
            void* centers_bf16_pad, int pad_ld, void* stream) {
    return fps(pts, B, N, C, G, CLLP(start), centers, center_idx, centers_bf16_pad, pad_ld, ST(stream));
}
int esb_knn_group(const float* pts, const float* centers, int B, int N, int C, int G, int K, int out_ld, void* nb_out,
                  int32_t* idx_out, void* stream) {
    return knn_group(pts, centers, B, N, C, G, K, out_ld, nb_out, idx_out, ST(stream));
}

/* ---- trainable point encoder (--unfreeze_pc_encoder) */
int esb_layernorm_bwd(const void* dy, const void* x, const float* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                      int64_t rows, int d, float eps, void* stream) {
    return layernorm_bwd(dy, x, gamma, dres, dx, dgamma, dbeta, rows, d, eps, ST(stream));
}
int esb_colsum2_bf16(const void* a, int64_t lda, const void* b, int64_t ldb, int64_t M, int N, float* out0, float* out1, void* stream) {
    return colsum2_bf16(a, lda, b, ldb, M, N, out0, out1, ST(stream));
}
int esb_bn_coeffs(int mode, const float* sum0, const float* sum1, int64_t R, int C, const float* gamma, const float* beta, float eps,
                  float momentum, float* mean_io, float* rstd_io, float* running_mean, float* running_var, float* A, float* B,
                  float* Cc, float* dgamma, float* dbeta, void* stream) {
    return bn_coeffs(mode, sum0, sum1, R, C, gamma, beta, eps, momentum, mean_io, rstd_io, running_mean, running_var, A, B, Cc, dgamma,
                     dbeta, ST(stream));
}
int esb_affine_cols(const void* x, const void* y, const float* A, const float* B, const float* Cc, void* out, int64_t rows, int C,
                    int relu, void* stream) {
    return affine_cols(x, y, A, B, Cc, out, rows, C, relu, ST(stream));
}
int esb_relu_mask(const void* dy, const void* y, void* out, int64_t n, void* stream) { return relu_mask(dy, y, out, n, ST(stream)); }
int esb_group_maxpool_bwd(const void* f, const void* dpool, const void* dcat, void* df, int64_t groups, int n, int C, void* stream) {
    return group_maxpool_bwd(f, dpool, dcat, df, groups, n, C, ST(stream));
}
int esb_scale_rows_add(const void* x, const void* y, const float* scale, void* out, int64_t rows, int d, int64_t rows_per_batch,
                       void* stream) {
    return scale_rows_add(x, y, scale, out, rows, d, rows_per_batch, ST(stream));
}

int esb_traj_metrics(const float* gen, const float* gt, const int32_t* gen_len, const int32_t* gt_len, int B, int Tg, int Tt, int D,
                     int c0, int c1, int r0, double* out, void* stream) {
    return traj_metrics(gen, gt, gen_len, gt_len, B, Tg, Tt, D, c0, c1, r0, out, ST(stream));
}

}  // extern "C"
