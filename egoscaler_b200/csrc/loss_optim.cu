// Cross-entropy over the trajectory positions, greedy argmax, fused AdamW.
//
// cross_entropy : train.py:174-181  F.cross_entropy(logits[:, P-1:-1], tokens[:, P:], ignore_index=pad), mean over
//                 the non-ignored targets; forward and backward in one pass over the logits (dlogits overwrite them).
// argmax_rows   : HF greedy step torch.argmax (generation/utils.py:2762-2806), first maximum wins.
// adamw         : torch.optim.AdamW as constructed at train.py:107-113 (fp32 master weights, bf16 model copy).
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

struct MaxSum { float m, s; };
ESB_DEVICE MaxSum ms_combine(MaxSum a, MaxSum b) {
    const float m = fmaxf(a.m, b.m);
    if (m == -INFINITY) return {m, 0.0f};
    return {m, a.s * __expf(a.m - m) + b.s * __expf(b.m - m)};
}

// one CTA per row. logits [rows, ld] bf16 (V valid columns). row_loss[row] = lse - logit[target] (0 if ignored).
// if dlogits_scale_ptr != null: logits <- (softmax - onehot) * (*scale) (ignored rows -> 0), in place, bf16.
__global__ void __launch_bounds__(256) ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, long long ld, int V,
                                                         const long long* __restrict__ targets, long long ignore_index,
                                                         float* __restrict__ row_loss, const float* __restrict__ inv_count,
                                                         int write_grad) {
    __shared__ MaxSum red[32];
    const long long row = blockIdx.x;
    __nv_bfloat16* lp = logits + row * ld;
    const long long tgt = targets[row];
    // a target outside [0, V) that is not ignore_index (torch raises a device assert there) is treated as ignored: the raw-pointer
    // ABI must never read or write out of bounds (e.g. the reference's IGNORE_INDEX = -100 while ignore_index = pad = 0)
    const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= V;
    MaxSum acc = {-INFINITY, 0.0f};
    if (!ignored) {
        for (int i = threadIdx.x; i < V; i += blockDim.x) acc = ms_combine(acc, MaxSum{__bfloat162float(lp[i]), 1.0f});
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            MaxSum other;
            other.m = __shfl_xor_sync(0xffffffffu, acc.m, o);
            other.s = __shfl_xor_sync(0xffffffffu, acc.s, o);
            acc = ms_combine(acc, other);
        }
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        acc = red[0];
        for (int w = 1; w < (blockDim.x >> 5); ++w) acc = ms_combine(acc, red[w]);
    }
    const float lse = ignored ? 0.0f : acc.m + __logf(acc.s);
    if (threadIdx.x == 0) row_loss[row] = ignored ? 0.0f : lse - __bfloat162float(lp[tgt]);
    if (write_grad) {
        const float sc = *inv_count;
        __syncthreads();                                   // everyone has read lp[...] needed above
        for (int i = threadIdx.x; i < V; i += blockDim.x) {
            float g = 0.0f;
            if (!ignored) {
                g = __expf(__bfloat162float(lp[i]) - lse);
                if (i == tgt) g -= 1.0f;
                g *= sc;
            }
            lp[i] = __float2bfloat16_rn(g);
        }
    }
}

// count non-ignored targets -> inv_count = 1/max(count,1)... (count==0 -> NaN mean like torch: 0/0) ; single block
__global__ void ce_count_kernel(const long long* __restrict__ targets, long long n, long long ignore_index, int V, float* inv_count) {
    __shared__ int cnt;
    if (threadIdx.x == 0) cnt = 0;
    __syncthreads();
    int c = 0;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) c += (targets[i] != ignore_index && targets[i] >= 0 && targets[i] < V);
    atomicAdd(&cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) *inv_count = 1.0f / static_cast<float>(cnt);
}
// loss = sum(row_loss) * inv_count ; deterministic single-block tree
__global__ void ce_reduce_kernel(const float* __restrict__ row_loss, long long n, const float* __restrict__ inv_count, float* loss) {
    __shared__ float red[32];
    float s = 0.0f;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) s += row_loss[i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0f;
        t = warp_sum(t);
        if (threadIdx.x == 0) *loss = t * (*inv_count);
    }
}

int cross_entropy(void* logits, long long ld, long long rows, int V, const long long* targets, long long ignore_index,
                  float* row_loss, float* scratch_inv_count, float* loss, int write_grad, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    ce_count_kernel<<<1, 1024, 0, s>>>(targets, rows, ignore_index, V, scratch_inv_count);
    ce_fwd_bwd_kernel<<<static_cast<unsigned>(rows), 256, 0, s>>>(static_cast<__nv_bfloat16*>(logits), ld, V, targets, ignore_index,
                                                                  row_loss, scratch_inv_count, write_grad);
    ce_reduce_kernel<<<1, 1024, 0, s>>>(row_loss, rows, scratch_inv_count, loss);
    return check_launch("cross_entropy");
}

// out_idx[row] = first index of the maximum of logits[row, :V]; optional fp32 copy of the row (HF `scores`),
// optional suppressed column (set to -inf before the argmax, e.g. EOS in the throughput benchmark).
__global__ void __launch_bounds__(1024) argmax_rows_kernel(const __nv_bfloat16* __restrict__ logits, long long ld, int V,
                                                          long long* __restrict__ out_idx, float* __restrict__ scores,
                                                          long long scores_ld, int suppress_col, const int* __restrict__ step_ptr,
                                                          int step_off, long long scores_step_stride) {
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    const long long row = blockIdx.x;
    const __nv_bfloat16* lp = logits + row * ld;
    if (scores && step_ptr) scores += (static_cast<long long>(*step_ptr) + step_off) * scores_step_stride;
    float* sp = scores ? scores + row * scores_ld : nullptr;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    auto take = [&](float v, int i) {
        if (i == suppress_col) v = -INFINITY;
        if (sp) sp[i] = v;
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    };
    // 16-byte loads over the aligned body of the row (one CTA streams 64 KB; scalar 2-byte loads made this 64 us per token)
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(lp) & 15) == 0);
    const int nvec = vec_ok ? V >> 3 : 0;
    for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(lp) + vi);
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
        const int i0 = vi << 3;
        take(a.x, i0); take(a.y, i0 + 1); take(b.x, i0 + 2); take(b.y, i0 + 3);
        take(c.x, i0 + 4); take(c.y, i0 + 5); take(d.x, i0 + 6); take(d.y, i0 + 7);
    }
    for (int i = (nvec << 3) + threadIdx.x; i < V; i += blockDim.x) take(__bfloat162float(lp[i]), i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { s_val[threadIdx.x >> 5] = best; s_idx[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (blockDim.x >> 5); ++w)
            if (s_val[w] > best || (s_val[w] == best && s_idx[w] < bi)) { best = s_val[w]; bi = s_idx[w]; }
        out_idx[row] = (bi == 0x7fffffff) ? 0 : bi;       // all -inf / NaN row -> 0 like torch.argmax on equal values
    }
}
int argmax_rows(const void* logits, long long ld, long long rows, int V, long long* out_idx, float* scores, long long scores_ld,
                int suppress_col, const int* step_ptr, int step_off, long long scores_step_stride, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    argmax_rows_kernel<<<static_cast<unsigned>(rows), 1024, 0, s>>>(static_cast<const __nv_bfloat16*>(logits), ld, V, out_idx, scores, scores_ld, suppress_col,
                                                                   step_ptr, step_off, scores_step_stride);
    return check_launch("argmax_rows");
}

// HF greedy bookkeeping on device (generation/utils.py:2790-2806): finished rows emit pad, rows that emit EOS finish;
// appends the token to `sequences[:, cur_len]` and sets attention mask 1.
__global__ void greedy_update_kernel(const long long* __restrict__ next, long long* __restrict__ tokens_out, int* __restrict__ unfinished,
                                     long long* __restrict__ sequences, long long seq_ld, int cur_len,
                                     uint8_t* __restrict__ mask, long long mask_ld, long long eos, long long pad, int B,
                                     int* __restrict__ pos_ptr) {
    // single CTA: with pos_ptr the write position is cur_len + *pos_ptr and the counter is advanced afterwards
    if (pos_ptr) cur_len += *pos_ptr;
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        const int uf = unfinished[b];
        const long long t = uf ? next[b] : pad;
        tokens_out[b] = t;
        sequences[b * seq_ld + cur_len] = t;
        if (mask) mask[b * mask_ld + cur_len] = 1;
        unfinished[b] = uf && (t != eos);
    }
    __syncthreads();
    if (pos_ptr && threadIdx.x == 0) *pos_ptr += 1;
}
int greedy_update(const long long* next, long long* tokens_out, int* unfinished, long long* sequences, long long seq_ld,
                  int cur_len, uint8_t* mask, long long mask_ld, long long eos, long long pad, int B, int* pos_ptr, cudaStream_t s) {
    if (B <= 0) return ESB_OK;
    greedy_update_kernel<<<1, 256, 0, s>>>(next, tokens_out, unfinished, sequences, seq_ld, cur_len, mask, mask_ld, eos, pad, B, pos_ptr);
    return check_launch("greedy_update");
}

// AdamW (decoupled weight decay), fp32 master p/m/v, grad fp32 * grad_scale, writes bf16 copy (optional).
struct AdamWConst { float lr, beta1, beta2, eps, wd, bc1, bc2_sqrt, grad_scale; };

ESB_DEVICE void adamw_one(float& p, float g, float& m, float& v, const AdamWConst& c) {
    // explicit roundings: the 16-byte and the scalar loop must agree bit for bit (no compiler-chosen FMA contraction)
    const float gi = __fmul_rn(g, c.grad_scale);
    const float mi = __fmaf_rn(c.beta1, m, __fmul_rn(1.0f - c.beta1, gi));
    const float vi = __fmaf_rn(c.beta2, v, __fmul_rn(__fmul_rn(1.0f - c.beta2, gi), gi));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), c.bc2_sqrt), c.eps);
    p = __fmaf_rn(-(c.lr / c.bc1), __fdiv_rn(mi, denom), __fmul_rn(p, 1.0f - c.lr * c.wd));
    m = mi; v = vi;
}

// 30 B per parameter (p, m, v read + written, g read, bf16 copy written).  vec != 0: all arrays 16-byte aligned (bf16 copy 8-byte),
// four parameters per thread with 16-byte accesses; the n % 4 tail and unaligned tensors take the scalar loop.
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                    float* __restrict__ v, __nv_bfloat16* __restrict__ p_bf16, long long n,
                                                    const AdamWConst c, int vec) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    const long long n4 = vec ? (n >> 2) : 0;
    for (long long i = tid; i < n4; i += stride) {
        float4 pv = reinterpret_cast<float4*>(p)[i];
        const float4 gv = __ldcs(reinterpret_cast<const float4*>(g) + i);
        float4 mv = reinterpret_cast<float4*>(m)[i];
        float4 vv = reinterpret_cast<float4*>(v)[i];
        adamw_one(pv.x, gv.x, mv.x, vv.x, c);
        adamw_one(pv.y, gv.y, mv.y, vv.y, c);
        adamw_one(pv.z, gv.z, mv.z, vv.z, c);
        adamw_one(pv.w, gv.w, mv.w, vv.w, c);
        reinterpret_cast<float4*>(p)[i] = pv;
        reinterpret_cast<float4*>(m)[i] = mv;
        reinterpret_cast<float4*>(v)[i] = vv;
        if (p_bf16) reinterpret_cast<uint2*>(p_bf16)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
    }
    for (long long i = (n4 << 2) + tid; i < n; i += stride) {
        float pi = p[i], mi = m[i], vi = v[i];
        adamw_one(pi, g[i], mi, vi, c);
        p[i] = pi; m[i] = mi; v[i] = vi;
        if (p_bf16) p_bf16[i] = __float2bfloat16_rn(pi);
    }
}
int adamw(float* p, const float* g, float* m, float* v, void* p_bf16, long long n, float lr, float beta1, float beta2, float eps,
          float wd, int step, float grad_scale, cudaStream_t s) {
    if (n <= 0) return ESB_OK;
    if (step < 1) { set_last_error("adamw: step must be >= 1"); return ESB_ERR_INVALID_ARG; }
    const float bc1 = 1.0f - powf(beta1, static_cast<float>(step));
    const float bc2s = sqrtf(1.0f - powf(beta2, static_cast<float>(step)));
    const auto al = [](const void* q, uintptr_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; };
    const int vec = al(p, 16) && al(g, 16) && al(m, 16) && al(v, 16) && (!p_bf16 || al(p_bf16, 8)) ? 1 : 0;
    const long long items = vec ? (n + 3) / 4 : n;
    const long long want = (items + 255) / 256, cap = static_cast<long long>(device_sm_count()) * 32;
    const AdamWConst c{lr, beta1, beta2, eps, wd, bc1, bc2s, grad_scale};
    adamw_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, s>>>(p, g, m, v, static_cast<__nv_bfloat16*>(p_bf16), n, c, vec);
    return check_launch("adamw");
}

// fp32 -> bf16 cast (weights upload / gradient down-cast), n % 4 == 0 fast path
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = __float2bfloat16_rn(in[i]);
}
int cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t s) {
    if (n <= 0) return ESB_OK;
    const long long want = (n + 255) / 256, cap = static_cast<long long>(device_sm_count()) * 16;
    cast_f32_bf16_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, s>>>(in, static_cast<__nv_bfloat16*>(out), n);
    return check_launch("cast_f32_bf16");
}

}  // namespace esb
