// Attention backward (flash-style recomputation; no [L,L] matrix is ever stored).
//
// Autograd of the forward in attention_fwd.cu, i.e. of HF sdpa attention (integrations/sdpa_attention.py:40-98)
// as used by LlamaAttention (modeling_llama.py:253-290).  The data gradient must cross all 32 frozen layers to reach
// embed_tokens / point_proj (model_arch.py:33-51), so this runs in every training step.
//   P  = exp(S*scale - lse),  D_i = sum_e dO_ie O_ie,  dS = P o (dO V^T - D) * scale
//   dV = P^T dO,  dK = dS^T Q,  dQ = dS K
// Three kernels: row-dot preprocess (D), dK/dV (one CTA per 64 keys, loops over query blocks), dQ (one CTA per 64
// queries, loops over key blocks).  S is recomputed in both; nothing is accumulated with atomics.
#include "attn_common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

struct AttnBwdParams {
    const __nv_bfloat16 *q, *k, *v, *o, *dout;
    __nv_bfloat16 *dq, *dk, *dv;
    const float* lse;      // [B,H,Lq]
    float* dvec;           // [B,H,Lq]  D = rowsum(dO o O)
    const uint8_t* kmask;
    long long q_bs, q_rs, k_bs, k_rs, v_bs, v_rs, o_bs, o_rs, do_bs, do_rs, dq_bs, dq_rs, dk_bs, dk_rs, dv_bs, dv_rs, kmask_ld;
    int B, H, Lq, Lk, causal, q_pos0;
    float scale;
};

constexpr int BWD_BM = 64, BWD_BN = 64, BWD_THREADS = 128;

// D[b,h,i] = sum_e dO[b,i,h,e] * O[b,i,h,e]; one warp per (b,i,h)
template <int HD>
__global__ void __launch_bounds__(128) attn_bwd_dot_kernel(const AttnBwdParams p) {
    const long long gw = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long total = static_cast<long long>(p.B) * p.Lq * p.H;
    if (gw >= total) return;
    const int h = static_cast<int>(gw % p.H);
    const long long t = gw / p.H;
    const int i = static_cast<int>(t % p.Lq);
    const int b = static_cast<int>(t / p.Lq);
    const __nv_bfloat16* op = p.o + b * p.o_bs + static_cast<long long>(i) * p.o_rs + static_cast<long long>(h) * HD;
    const __nv_bfloat16* dp = p.dout + b * p.do_bs + static_cast<long long>(i) * p.do_rs + static_cast<long long>(h) * HD;
    float acc = 0.0f;
#pragma unroll
    for (int e = lane * 2; e < HD; e += 64) {
        const float2 a = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(op + e));
        const float2 d = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(dp + e));
        acc += a.x * d.x + a.y * d.y;
    }
    acc = warp_sum(acc);
    if (lane == 0) p.dvec[(static_cast<long long>(b) * p.H + h) * p.Lq + i] = acc;
}

// ------------------------------------------------------------------------------------------------ dK / dV
template <int HD>
__global__ void __launch_bounds__(BWD_THREADS) attn_bwd_dkdv_kernel(const AttnBwdParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int TILE = 64 * tile_row_bytes<HD>();
    uint8_t* sK = smem;
    uint8_t* sV = smem + TILE;
    uint8_t* sQ = smem + 2 * TILE;          // 2 stages
    uint8_t* sDO = smem + 4 * TILE;         // 2 stages
    float* sLse = reinterpret_cast<float*>(smem + 6 * TILE);   // 2 x 64
    float* sD = sLse + 128;                                     // 2 x 64
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int nblk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int k0 = nblk * BWD_BN;
    const __nv_bfloat16* qb = p.q + b * p.q_bs + static_cast<long long>(h) * HD;
    const __nv_bfloat16* kb = p.k + b * p.k_bs + static_cast<long long>(h) * HD;
    const __nv_bfloat16* vb = p.v + b * p.v_bs + static_cast<long long>(h) * HD;
    const __nv_bfloat16* dob = p.dout + b * p.do_bs + static_cast<long long>(h) * HD;
    const float* lseb = p.lse + (static_cast<long long>(b) * p.H + h) * p.Lq;
    const float* db = p.dvec + (static_cast<long long>(b) * p.H + h) * p.Lq;
    const uint8_t* km = p.kmask ? p.kmask + b * p.kmask_ld : nullptr;

    int mb0 = 0;
    if (p.causal) { const int first = k0 - p.q_pos0; mb0 = first > 0 ? first / BWD_BM : 0; }
    const int mb_end = (p.Lq + BWD_BM - 1) / BWD_BM;

    load_tile_async<HD, 64, BWD_THREADS>(sK, kb, p.k_rs, k0, p.Lk);
    load_tile_async<HD, 64, BWD_THREADS>(sV, vb, p.v_rs, k0, p.Lk);
    auto load_q = [&](int mb, int st) {
        load_tile_async<HD, 64, BWD_THREADS>(sQ + st * TILE, qb, p.q_rs, mb * BWD_BM, p.Lq);
        load_tile_async<HD, 64, BWD_THREADS>(sDO + st * TILE, dob, p.do_rs, mb * BWD_BM, p.Lq);
        if (threadIdx.x < 64) {
            const int i = mb * BWD_BM + threadIdx.x;
            sLse[st * 64 + threadIdx.x] = i < p.Lq ? lseb[i] : INFINITY;
            sD[st * 64 + threadIdx.x] = i < p.Lq ? db[i] : 0.0f;
        }
    };
    if (mb0 < mb_end) load_q(mb0, 0);
    cp_async_commit();

    bool key_ok[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int key = k0 + warp * 16 + g + r * 8;
        key_ok[r] = key < p.Lk && (km ? km[key] != 0 : true);
    }
    float dk_acc[HD / 8][4], dv_acc[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) dk_acc[i][e] = dv_acc[i][e] = 0.0f;
    const float sl2 = p.scale * 1.4426950408889634f;
    const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);

    for (int mb = mb0; mb < mb_end; ++mb) {
        const int st = (mb - mb0) & 1;
        if (mb + 1 < mb_end) { load_q(mb + 1, st ^ 1); cp_async_commit(); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
        const uint32_t aQ = smem_u32(sQ + st * TILE), aDO = smem_u32(sDO + st * TILE);
        const float* lse_s = sLse + st * 64;
        const float* d_s = sD + st * 64;
        const int q0 = mb * BWD_BM;
#pragma unroll
        for (int hq = 0; hq < 2; ++hq) {
            float stt[4][4], dpt[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int e = 0; e < 4; ++e) stt[i][e] = dpt[i][e] = 0.0f;
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk) {
                uint32_t ak[4], av[4];
                const int arow = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ach = kk * 2 + (lane >> 4);
                ldsm_x4(aK + tile_off<HD>(arow, ach), ak[0], ak[1], ak[2], ak[3]);
                ldsm_x4(aV + tile_off<HD>(arow, ach), av[0], av[1], av[2], av[3]);
#pragma unroll
                for (int np = 0; np < 2; ++np) {
                    const int brow = hq * 32 + np * 16 + (lane & 7) + (lane >> 4) * 8, bch = kk * 2 + ((lane >> 3) & 1);
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4(aQ + tile_off<HD>(brow, bch), b0, b1, b2, b3);
                    mma16816(stt[np * 2], ak, b0, b1);
                    mma16816(stt[np * 2 + 1], ak, b2, b3);
                    ldsm_x4(aDO + tile_off<HD>(brow, bch), b0, b1, b2, b3);
                    mma16816(dpt[np * 2], av, b0, b1);
                    mma16816(dpt[np * 2 + 1], av, b2, b3);
                }
            }
            uint32_t pf[2][4], dsf[2][4];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                float pv[4], dsv[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int ql = hq * 32 + nt * 8 + 2 * t + (e & 1);
                    const int qi = q0 + ql;
                    const int key = k0 + warp * 16 + g + (e >> 1) * 8;
                    bool ok = key_ok[e >> 1] && qi < p.Lq;
                    if (p.causal) ok = ok && (key <= qi + p.q_pos0);
                    const float pr = ok ? exp2f(stt[nt][e] * sl2 - lse_s[ql] * 1.4426950408889634f) : 0.0f;
                    pv[e] = pr;
                    dsv[e] = pr * (dpt[nt][e] - d_s[ql]) * p.scale;
                }
                pf[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16x2(pv[0], pv[1]);
                pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16x2(pv[2], pv[3]);
                dsf[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16x2(dsv[0], dsv[1]);
                dsf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16x2(dsv[2], dsv[3]);
            }
#pragma unroll
            for (int kk2 = 0; kk2 < 2; ++kk2) {
#pragma unroll
                for (int np = 0; np < HD / 16; ++np) {
                    const int brow = hq * 32 + kk2 * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, bch = np * 2 + (lane >> 4);
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4_t(aDO + tile_off<HD>(brow, bch), b0, b1, b2, b3);
                    mma16816(dv_acc[np * 2], pf[kk2], b0, b1);
                    mma16816(dv_acc[np * 2 + 1], pf[kk2], b2, b3);
                    ldsm_x4_t(aQ + tile_off<HD>(brow, bch), b0, b1, b2, b3);
                    mma16816(dk_acc[np * 2], dsf[kk2], b0, b1);
                    mma16816(dk_acc[np * 2 + 1], dsf[kk2], b2, b3);
                }
            }
        }
        __syncthreads();
    }
    if (mb0 >= mb_end) cp_async_wait<0>();
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int key = k0 + warp * 16 + g + r * 8;
        if (key >= p.Lk) continue;
        __nv_bfloat16* dkp = p.dk + b * p.dk_bs + static_cast<long long>(key) * p.dk_rs + static_cast<long long>(h) * HD;
        __nv_bfloat16* dvp = p.dv + b * p.dv_bs + static_cast<long long>(key) * p.dv_rs + static_cast<long long>(h) * HD;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) {
            *reinterpret_cast<uint32_t*>(dkp + i * 8 + 2 * t) = pack_bf16x2(dk_acc[i][r * 2], dk_acc[i][r * 2 + 1]);
            *reinterpret_cast<uint32_t*>(dvp + i * 8 + 2 * t) = pack_bf16x2(dv_acc[i][r * 2], dv_acc[i][r * 2 + 1]);
        }
    }
}

#include "attention_bwd_dq.inc"

}  // namespace esb
