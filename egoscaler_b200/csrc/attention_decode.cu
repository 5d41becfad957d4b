// Single-query (KV-cache) attention for autoregressive decode: one new token per sequence attends to the cached
// keys/values of its own sequence.  HBM-bound: every cached K/V byte is read once per step.
//
// Reference: the per-token step of HF generation reached through TrajPointLLMForCausalLM.generate
// (model_arch.py:77-108, pointllm.py:255-275): LlamaAttention with a DynamicCache (keys stored post-RoPE,
// modeling_llama.py:269-270), mask = key-padding over the grown sequence, fp32 softmax.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

constexpr int DEC_THREADS = 128;
constexpr int DEC_GROUPS = DEC_THREADS / 8;      // 8 lanes cooperate on one key

// q: (b,h,e) at q + b*q_bs + h*HD + e ; k/v cache: (b,j,h,e) at k + b*k_bs + j*k_rs + h*HD + e ; o like q.
template <int HD>
__global__ void __launch_bounds__(DEC_THREADS) attn_decode_kernel(const __nv_bfloat16* __restrict__ q, long long q_bs,
                                                                  const __nv_bfloat16* __restrict__ k, const __nv_bfloat16* __restrict__ v,
                                                                  long long kv_bs, long long kv_rs, __nv_bfloat16* __restrict__ o,
                                                                  long long o_bs, const uint8_t* __restrict__ kmask, long long kmask_ld,
                                                                  const int* __restrict__ len_ptr, int len_static, float scale) {
    constexpr int EPL = HD / 8;                  // elements per lane
    constexpr int VPL = EPL / 8;                 // 16-byte vectors per lane
    __shared__ float s_acc[DEC_GROUPS][HD];
    __shared__ float s_m[DEC_GROUPS], s_l[DEC_GROUPS];
    griddep_launch();
    griddep_wait();
    const int h = blockIdx.x, b = blockIdx.y;
    const int grp = threadIdx.x >> 3, sub = threadIdx.x & 7;
    const int Lk = len_static + (len_ptr ? *len_ptr : 0);

    float qv[EPL];
    {
        const __nv_bfloat16* qp = q + b * q_bs + static_cast<long long>(h) * HD + sub * EPL;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const uint4 u = *reinterpret_cast<const uint4*>(qp + i * 8);
            const float2 a = unpack_bf16x2(u.x), bb = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
            qv[i * 8 + 0] = a.x; qv[i * 8 + 1] = a.y; qv[i * 8 + 2] = bb.x; qv[i * 8 + 3] = bb.y;
            qv[i * 8 + 4] = c.x; qv[i * 8 + 5] = c.y; qv[i * 8 + 6] = d.x; qv[i * 8 + 7] = d.y;
        }
    }
    float acc[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) acc[i] = 0.0f;
    float m = -INFINITY, l = 0.0f;
    const __nv_bfloat16* kb = k + b * kv_bs + static_cast<long long>(h) * HD + sub * EPL;
    const __nv_bfloat16* vb = v + b * kv_bs + static_cast<long long>(h) * HD + sub * EPL;
    const uint8_t* km = kmask ? kmask + b * kmask_ld : nullptr;

    // DEC_UNROLL keys per group per trip: all K/V vectors of the trip are requested before the first one is consumed, which
    // doubles the bytes in flight per thread (the kernel is a pure HBM stream: 2 * Lk * d * 2 B per sequence per layer).
    constexpr int DEC_UNROLL = 2;
    for (int j0 = 0; j0 < Lk; j0 += DEC_GROUPS * DEC_UNROLL) {   // uniform trip count: the shuffles below need every lane
        uint4 ku[DEC_UNROLL][VPL], vu[DEC_UNROLL][VPL];
        bool okv[DEC_UNROLL];
#pragma unroll
        for (int u = 0; u < DEC_UNROLL; ++u) {
            const int jj = j0 + u * DEC_GROUPS + grp;
            const bool in_range = jj < Lk;
            const int j = in_range ? jj : (Lk - 1);
            okv[u] = in_range && (km ? (km[j] != 0) : true);       // uniform over the 8 lanes of a group
#pragma unroll
            for (int i = 0; i < VPL; ++i) {
                ku[u][i] = __ldg(reinterpret_cast<const uint4*>(kb + static_cast<long long>(j) * kv_rs + i * 8));
                vu[u][i] = __ldg(reinterpret_cast<const uint4*>(vb + static_cast<long long>(j) * kv_rs + i * 8));
            }
        }
#pragma unroll
        for (int u = 0; u < DEC_UNROLL; ++u) {
            float dot = 0.0f;
#pragma unroll
            for (int i = 0; i < VPL; ++i) {
                const uint4 w = ku[u][i];
                const float2 a = unpack_bf16x2(w.x), bb = unpack_bf16x2(w.y), c = unpack_bf16x2(w.z), d = unpack_bf16x2(w.w);
                dot += qv[i * 8 + 0] * a.x + qv[i * 8 + 1] * a.y + qv[i * 8 + 2] * bb.x + qv[i * 8 + 3] * bb.y +
                       qv[i * 8 + 4] * c.x + qv[i * 8 + 5] * c.y + qv[i * 8 + 6] * d.x + qv[i * 8 + 7] * d.y;
            }
            dot += __shfl_xor_sync(0xffffffffu, dot, 1);
            dot += __shfl_xor_sync(0xffffffffu, dot, 2);
            dot += __shfl_xor_sync(0xffffffffu, dot, 4);
            if (!okv[u]) continue;
            const float sc = dot * scale;
            const float mn = fmaxf(m, sc);
            const float corr = __expf(m - mn);            // m = -inf -> 0
            const float pr = __expf(sc - mn);
            // the reference rounds P to bf16 before P@V (softmax(...).to(q.dtype) @ V); normalisation happens after here
            l = l * corr + pr;
            m = mn;
#pragma unroll
            for (int i = 0; i < VPL; ++i) {
                const uint4 w = vu[u][i];
                const float2 a = unpack_bf16x2(w.x), bb = unpack_bf16x2(w.y), c = unpack_bf16x2(w.z), d = unpack_bf16x2(w.w);
                acc[i * 8 + 0] = acc[i * 8 + 0] * corr + pr * a.x; acc[i * 8 + 1] = acc[i * 8 + 1] * corr + pr * a.y;
                acc[i * 8 + 2] = acc[i * 8 + 2] * corr + pr * bb.x; acc[i * 8 + 3] = acc[i * 8 + 3] * corr + pr * bb.y;
                acc[i * 8 + 4] = acc[i * 8 + 4] * corr + pr * c.x; acc[i * 8 + 5] = acc[i * 8 + 5] * corr + pr * c.y;
                acc[i * 8 + 6] = acc[i * 8 + 6] * corr + pr * d.x; acc[i * 8 + 7] = acc[i * 8 + 7] * corr + pr * d.y;
            }
        }
    }
    if (sub == 0) { s_m[grp] = m; s_l[grp] = l; }
#pragma unroll
    for (int i = 0; i < EPL; ++i) s_acc[grp][sub * EPL + i] = acc[i];
    __syncthreads();
    for (int e = threadIdx.x; e < HD; e += DEC_THREADS) {
        float mm = -INFINITY;
#pragma unroll
        for (int gI = 0; gI < DEC_GROUPS; ++gI) mm = fmaxf(mm, s_m[gI]);
        float num = 0.0f, den = 0.0f;
#pragma unroll
        for (int gI = 0; gI < DEC_GROUPS; ++gI) {
            const float w = (s_m[gI] == -INFINITY) ? 0.0f : __expf(s_m[gI] - mm);
            num += w * s_acc[gI][e];
            den += w * s_l[gI];
        }
        o[b * o_bs + static_cast<long long>(h) * HD + e] = __float2bfloat16_rn(den > 0.0f ? num / den : 0.0f);
    }
}

int attn_decode(const void* q, long long q_bs, const void* k, const void* v, long long kv_bs, long long kv_rs, void* o,
                long long o_bs, const uint8_t* kmask, long long kmask_ld, const int* len_ptr, int len_static, int B, int H,
                int hd, float scale, cudaStream_t s) {
    if (B <= 0 || H <= 0) return ESB_OK;
    if ((q_bs & 7) || (kv_bs & 7) || (kv_rs & 7)) { set_last_error("attn_decode: strides must keep 16-byte alignment"); return ESB_ERR_INVALID_ARG; }
    dim3 grid(H, B);
    const __nv_bfloat16 *qq = static_cast<const __nv_bfloat16*>(q), *kk = static_cast<const __nv_bfloat16*>(k), *vv = static_cast<const __nv_bfloat16*>(v);
    if (hd == 128) launch_chain(attn_decode_kernel<128>, grid, dim3(DEC_THREADS), 0, s, qq, q_bs, kk, vv, kv_bs, kv_rs, static_cast<__nv_bfloat16*>(o), o_bs, kmask, kmask_ld, len_ptr, len_static, scale);
    else if (hd == 64) launch_chain(attn_decode_kernel<64>, grid, dim3(DEC_THREADS), 0, s, qq, q_bs, kk, vv, kv_bs, kv_rs, static_cast<__nv_bfloat16*>(o), o_bs, kmask, kmask_ld, len_ptr, len_static, scale);
    else { set_last_error("attn_decode: head_dim %d unsupported", hd); return ESB_ERR_UNSUPPORTED; }
    return check_launch("attn_decode");
}

}  // namespace esb
