// Fused multi-head attention forward (flash-style online softmax, fp32 statistics, bf16 tensor-core MMAs).
//
// Reference semantics:
//   LLaMA trunk : HF LlamaAttention -> sdpa_attention_forward (transformers integrations/sdpa_attention.py:40-98,
//                 modeling_llama.py:253-290): softmax(Q K^T * hd^-0.5 + mask) V with mask = causal AND key-padding
//                 (masking_utils.py), reached from pointllm/model/pointllm.py:173-178.
//   PointBERT   : pointbert/point_encoder.py:43-55, non-causal, no mask, 513 tokens, head_dim 64.
// Layout: element (b, l, h, e) of Q/K/V/O lives at ptr + b*bstride + l*rstride + h*HD + e (elements), so the packed
// qkv GEMM output, the KV cache and plain [B,L,d] tensors are all addressed without a transpose.
#include "attn_common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

struct AttnParams {
    const __nv_bfloat16 *q, *k, *v;
    __nv_bfloat16* o;
    float* lse;                 // [B, H, Lq] or null
    const uint8_t* kmask;       // [B, kmask_ld] 1 = attend, or null
    long long q_bs, q_rs, k_bs, k_rs, v_bs, v_rs, o_bs, o_rs, kmask_ld;
    int B, H, Lq, Lk;
    int causal;                 // key j visible to query i iff j <= i + q_pos0
    int q_pos0;
    float scale;
};

constexpr int ATT_BM = 64, ATT_BN = 64, ATT_THREADS = 128;

template <int HD>
__global__ void __launch_bounds__(ATT_THREADS) attn_fwd_kernel(const AttnParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int TILE = ATT_BM * tile_row_bytes<HD>();
    uint8_t* sQ = smem;
    uint8_t* sK = smem + TILE;            // 2 stages
    uint8_t* sV = smem + 3 * TILE;        // 2 stages
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int mblk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int m0 = mblk * ATT_BM;

    const __nv_bfloat16* qb = p.q + b * p.q_bs + static_cast<long long>(h) * HD;
    const __nv_bfloat16* kb = p.k + b * p.k_bs + static_cast<long long>(h) * HD;
    const __nv_bfloat16* vb = p.v + b * p.v_bs + static_cast<long long>(h) * HD;
    const uint8_t* km = p.kmask ? p.kmask + b * p.kmask_ld : nullptr;

    int kv_end = p.Lk;
    if (p.causal) kv_end = min(p.Lk, m0 + ATT_BM + p.q_pos0);
    const int nblk = (kv_end + ATT_BN - 1) / ATT_BN;

    load_tile_async<HD, ATT_BM, ATT_THREADS>(sQ, qb, p.q_rs, m0, p.Lq);
    if (nblk > 0) {
        load_tile_async<HD, ATT_BN, ATT_THREADS>(sK, kb, p.k_rs, 0, p.Lk);
        load_tile_async<HD, ATT_BN, ATT_THREADS>(sV, vb, p.v_rs, 0, p.Lk);
    }
    cp_async_commit();

    float o_acc[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.0f;
    float row_m[2] = {-INFINITY, -INFINITY}, row_l[2] = {0.0f, 0.0f};
    const float sl2 = p.scale * 1.4426950408889634f;
    const int qrow0 = m0 + warp * 16 + g;      // this thread's rows: qrow0, qrow0 + 8

    for (int nb = 0; nb < nblk; ++nb) {
        const int st = nb & 1;
        if (nb + 1 < nblk) {
            load_tile_async<HD, ATT_BN, ATT_THREADS>(sK + (st ^ 1) * TILE, kb, p.k_rs, (nb + 1) * ATT_BN, p.Lk);
            load_tile_async<HD, ATT_BN, ATT_THREADS>(sV + (st ^ 1) * TILE, vb, p.v_rs, (nb + 1) * ATT_BN, p.Lk);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const uint32_t aQ = smem_u32(sQ), aK = smem_u32(sK + st * TILE), aV = smem_u32(sV + st * TILE);

        // ---- S = Q K^T  (16 x 64 per warp)
        float s[ATT_BN / 8][4];
#pragma unroll
        for (int i = 0; i < ATT_BN / 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.0f;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
            uint32_t a[4];
            ldsm_x4(aQ + tile_off<HD>(warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, kk * 2 + (lane >> 4)), a[0], a[1], a[2], a[3]);
#pragma unroll
            for (int np = 0; np < ATT_BN / 16; ++np) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4(aK + tile_off<HD>(np * 16 + (lane & 7) + (lane >> 4) * 8, kk * 2 + ((lane >> 3) & 1)), b0, b1, b2, b3);
                mma16816(s[np * 2], a, b0, b1);
                mma16816(s[np * 2 + 1], a, b2, b3);
            }
        }
        // ---- mask + online softmax
        const int kbase = nb * ATT_BN;
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < ATT_BN / 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int key = kbase + nt * 8 + 2 * t + (e & 1);
                const int qr = qrow0 + (e >> 1) * 8;
                bool ok = key < p.Lk;
                if (p.causal) ok = ok && (key <= qr + p.q_pos0);
                if (km && ok) ok = km[key] != 0;
                if (!ok) s[nt][e] = -INFINITY;
                mx[e >> 1] = fmaxf(mx[e >> 1], s[nt][e]);
            }
        }
        float corr[2], msub[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
            const float mnew = fmaxf(row_m[r], mx[r]);
            corr[r] = (mnew == -INFINITY) ? 1.0f : exp2f((row_m[r] - mnew) * sl2);
            msub[r] = (mnew == -INFINITY) ? 0.0f : mnew * sl2;
            row_m[r] = mnew;
        }
        float ls[2] = {0.0f, 0.0f};
        uint32_t pfrag[ATT_BN / 16][4];
#pragma unroll
        for (int nt = 0; nt < ATT_BN / 8; ++nt) {
            const float p0 = exp2f(s[nt][0] * sl2 - msub[0]);
            const float p1 = exp2f(s[nt][1] * sl2 - msub[0]);
            const float p2 = exp2f(s[nt][2] * sl2 - msub[1]);
            const float p3 = exp2f(s[nt][3] * sl2 - msub[1]);
            ls[0] += p0 + p1;
            ls[1] += p2 + p3;
            pfrag[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16x2(p0, p1);
            pfrag[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16x2(p2, p3);
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) row_l[r] = row_l[r] * corr[r] + ls[r];
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) {
            o_acc[i][0] *= corr[0]; o_acc[i][1] *= corr[0];
            o_acc[i][2] *= corr[1]; o_acc[i][3] *= corr[1];
        }
        // ---- O += P V
#pragma unroll
        for (int kk = 0; kk < ATT_BN / 16; ++kk) {
#pragma unroll
            for (int np = 0; np < HD / 16; ++np) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4_t(aV + tile_off<HD>(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, np * 2 + (lane >> 4)), b0, b1, b2, b3);
                mma16816(o_acc[np * 2], pfrag[kk], b0, b1);
                mma16816(o_acc[np * 2 + 1], pfrag[kk], b2, b3);
            }
        }
        __syncthreads();       // stage `st` is overwritten by the prefetch issued in the next iteration
    }
    if (nblk == 0) cp_async_wait<0>();

    // ---- finalize
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        row_l[r] += __shfl_xor_sync(0xffffffffu, row_l[r], 1);
        row_l[r] += __shfl_xor_sync(0xffffffffu, row_l[r], 2);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int qr = qrow0 + r * 8;
        if (qr >= p.Lq) continue;
        const float inv = row_l[r] > 0.0f ? 1.0f / row_l[r] : 0.0f;
        __nv_bfloat16* op = p.o + b * p.o_bs + static_cast<long long>(qr) * p.o_rs + static_cast<long long>(h) * HD;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i)
            *reinterpret_cast<uint32_t*>(op + i * 8 + 2 * t) = pack_bf16x2(o_acc[i][r * 2] * inv, o_acc[i][r * 2 + 1] * inv);
        if (p.lse && t == 0)
            p.lse[(static_cast<long long>(b) * p.H + h) * p.Lq + qr] =
                row_l[r] > 0.0f ? row_m[r] * p.scale + __logf(row_l[r]) : -INFINITY;
    }
}

template <int HD>
static int launch_attn_fwd(const AttnParams& p, cudaStream_t s) {
    constexpr int SMEM = 5 * ATT_BM * tile_row_bytes<HD>();
    static bool set = false;
    if (!set) {
        cudaError_t e = cudaFuncSetAttribute(attn_fwd_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) { set_last_error("attn_fwd attr: %s", cudaGetErrorString(e)); return ESB_ERR_CUDA; }
        set = true;
    }
    dim3 grid((p.Lq + ATT_BM - 1) / ATT_BM, p.H, p.B);
    attn_fwd_kernel<HD><<<grid, ATT_THREADS, SMEM, s>>>(p);
    return check_launch("attn_fwd");
}

int attn_fwd(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
             long long v_bs, long long v_rs, void* o, long long o_bs, long long o_rs, float* lse, const uint8_t* kmask,
             long long kmask_ld, int B, int H, int Lq, int Lk, int hd, int causal, int q_pos0, float scale, cudaStream_t s) {
    if (B <= 0 || H <= 0 || Lq <= 0) return ESB_OK;
    if ((q_rs & 7) || (k_rs & 7) || (v_rs & 7) || (o_rs & 1) || (q_bs & 7) || (k_bs & 7) || (v_bs & 7)) {
        set_last_error("attn_fwd: strides must keep 16-byte alignment");
        return ESB_ERR_INVALID_ARG;
    }
    AttnParams p;
    p.q = static_cast<const __nv_bfloat16*>(q); p.k = static_cast<const __nv_bfloat16*>(k); p.v = static_cast<const __nv_bfloat16*>(v);
    p.o = static_cast<__nv_bfloat16*>(o); p.lse = lse; p.kmask = kmask;
    p.q_bs = q_bs; p.q_rs = q_rs; p.k_bs = k_bs; p.k_rs = k_rs; p.v_bs = v_bs; p.v_rs = v_rs; p.o_bs = o_bs; p.o_rs = o_rs;
    p.kmask_ld = kmask_ld; p.B = B; p.H = H; p.Lq = Lq; p.Lk = Lk; p.causal = causal; p.q_pos0 = q_pos0; p.scale = scale;
    if (hd == 64) return launch_attn_fwd<64>(p, s);
    if (hd == 96) return launch_attn_fwd<96>(p, s);     // PointBERT base encoder (trans_dim 1152, 12 heads)
    if (hd == 128)      // tcgen05 / TMEM / TMA kernel (attention_fwd_tc.cu); the mma.sync kernel below serves head_dim 64 only
        return attn_fwd_tc128(q, q_bs, q_rs, k, k_bs, k_rs, v, v_bs, v_rs, o, o_bs, o_rs, lse, kmask, kmask_ld, B, H, Lq, Lk, causal,
                              q_pos0, scale, s);
    set_last_error("attn_fwd: head_dim %d unsupported (64, 96, 128)", hd);
    return ESB_ERR_UNSUPPORTED;
}

}  // namespace esb
