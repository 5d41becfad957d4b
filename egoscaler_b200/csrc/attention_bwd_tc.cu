// Attention backward on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), head_dim 128.
//
// Autograd of attention_fwd_tc.cu, i.e. of HF sdpa attention (integrations/sdpa_attention.py:40-98) as used by
// LlamaAttention (modeling_llama.py:253-290).  The data gradient crosses all 32 frozen layers to reach embed_tokens /
// point_proj (model_arch.py:33-51), so this runs in every training step.
//     P  = exp(S*scale - lse),   D_i = sum_e dO_ie O_ie,   dS = P o (dO V^T - D) * scale
//     dV = P^T dO,   dK = dS^T Q,   dQ = dS K
// Three kernels:
//   stats : (lse * log2 e, D) per query row, padded to 64-row blocks (so the dK/dV kernel can bulk-copy a block)
//   dq    : one CTA per 128 queries, loops over 64-key blocks:  S = Q K^T, dP = dO V^T (TMEM) -> dS (smem) -> dQ += dS K
//   dkdv  : one CTA per 128 keys, loops over 64-query blocks:   S^T = K Q^T, dP^T = V dO^T (TMEM) -> P^T, dS^T (smem)
//                                                               -> dV += P^T dO, dK += dS^T Q   (accumulators in TMEM)
// Both compute kernels: 8 element-wise warps (TMEM lane = tile row, two warps share a lane quarter and split the 64
// block columns) + 1 warp whose lane 0 issues every TMA load and tcgen05.mma; hand-offs are mbarriers only.  The
// [key, hd] / [query, hd] tiles serve as K-major operands for the S / dP products and, unchanged, as MN-major B
// operands for the dV / dK / dQ products, so nothing is transposed in shared memory.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

namespace {

constexpr int BT_HD = 128;
constexpr int BT_TILE = 128;                 // rows owned by a CTA (queries in dq, keys in dkdv)
constexpr int BT_BLK = 64;                   // rows of the streamed operand per step
constexpr int BT_COMPUTE = 256;
constexpr int BT_THREADS = BT_COMPUTE + 32;
constexpr int TILE_BYTES = BT_TILE * BT_HD * 2;      // 32 KB
constexpr int BLK_BYTES = BT_BLK * BT_HD * 2;        // 16 KB
constexpr int OP_BYTES = BT_TILE * BT_BLK * 2;       // 16 KB  ([128 x 64] bf16 operand written by the compute warps)
constexpr float LOG2E = 1.4426950408889634f;

struct BwdTcParams {
    __nv_bfloat16 *dq, *dk, *dv;
    const float2* stats;                      // [B, H, Lqp] (lse * log2e  (+inf: dead row), D)
    const uint8_t* kmask;
    long long dq_bs, dq_rs, dk_bs, dk_rs, dv_bs, dv_rs, kmask_ld;
    int H, Lq, Lqp, Lk, causal, q_pos0;
    float scale, scale_log2;
    // not NULL: dq / dk leave the kernels already multiplied by the TRANSPOSED rotary embedding (the backward of
    // apply_rotary_pos_emb, modeling_llama.py:138-168) - tables [max_pos, 64] bf16, same rounding order as rope_kernel(inverse)
    const __nv_bfloat16 *rope_cos, *rope_sin;
};

// Backward of the rotary embedding on one row: x1 = 32 columns of the first half of the head, x2 = the paired 32 columns of the
// second half (fp32 accumulators, scaled and rounded to bf16 first like the materialised gradient was), cs/sn = the 32 table
// entries.  out1 = x1 cos + x2 sin, out2 = x2 cos - x1 sin with every product rounded to bf16 (rope_kernel, sign = -1).
ESB_DEVICE void rope_bwd_store(__nv_bfloat16* g1, __nv_bfloat16* g2, const uint32_t* a1, const uint32_t* a2, float mul,
                               const __nv_bfloat16* cs, const __nv_bfloat16* sn) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint4 cu = __ldg(reinterpret_cast<const uint4*>(cs) + g), su = __ldg(reinterpret_cast<const uint4*>(sn) + g);
        const uint32_t cw[4] = {cu.x, cu.y, cu.z, cu.w}, sw[4] = {su.x, su.y, su.z, su.w};
        uint32_t o1[4], o2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 x1 = unpack_bf16x2(pack_bf16x2(__uint_as_float(a1[g * 8 + 2 * j]) * mul, __uint_as_float(a1[g * 8 + 2 * j + 1]) * mul));
            const float2 x2 = unpack_bf16x2(pack_bf16x2(__uint_as_float(a2[g * 8 + 2 * j]) * mul, __uint_as_float(a2[g * 8 + 2 * j + 1]) * mul));
            const float2 cc = unpack_bf16x2(cw[j]), ss = unpack_bf16x2(sw[j]);
            o1[j] = pack_bf16x2(bf16_round(x1.x * cc.x) + bf16_round(x2.x * ss.x), bf16_round(x1.y * cc.y) + bf16_round(x2.y * ss.y));
            o2[j] = pack_bf16x2(bf16_round(x2.x * cc.x) - bf16_round(x1.x * ss.x), bf16_round(x2.y * cc.y) - bf16_round(x1.y * ss.y));
        }
        *reinterpret_cast<uint4*>(g1 + g * 8) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
        *reinterpret_cast<uint4*>(g2 + g * 8) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
    }
}

ESB_DEVICE void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// D = rowsum(dO o O), packed with lse * log2(e).  One warp per (b, h, 8 consecutive rows of the padded length Lqp): a 16-lane half
// covers one 256-byte row with 16-byte loads, and all 8 loads of a thread are issued before the first use (HBM-bound kernel).
__global__ void __launch_bounds__(128) attn_bwd_stats_kernel(const __nv_bfloat16* __restrict__ o, long long o_bs, long long o_rs,
                                                             const __nv_bfloat16* __restrict__ dout, long long do_bs, long long do_rs,
                                                             const float* __restrict__ lse, float2* __restrict__ stats, int B, int H,
                                                             int Lq, int Lqp) {
    const long long gw = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int groups = Lqp >> 3;
    const long long total = static_cast<long long>(B) * H * groups;
    if (gw >= total) return;
    const int g8 = static_cast<int>(gw % groups);
    const long long t = gw / groups;
    const int h = static_cast<int>(t % H);
    const int b = static_cast<int>(t / H);
    const int half = lane >> 4, l16 = lane & 15;
    uint4 av[4], dv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int i = g8 * 8 + 2 * k + half;
        av[k] = dv[k] = make_uint4(0u, 0u, 0u, 0u);
        if (i < Lq) {
            av[k] = *reinterpret_cast<const uint4*>(o + b * o_bs + static_cast<long long>(i) * o_rs + static_cast<long long>(h) * BT_HD + l16 * 8);
            dv[k] = *reinterpret_cast<const uint4*>(dout + b * do_bs + static_cast<long long>(i) * do_rs + static_cast<long long>(h) * BT_HD + l16 * 8);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int i = g8 * 8 + 2 * k + half;
        const float2 a0 = unpack_bf16x2(av[k].x), a1 = unpack_bf16x2(av[k].y), a2 = unpack_bf16x2(av[k].z), a3 = unpack_bf16x2(av[k].w);
        const float2 d0 = unpack_bf16x2(dv[k].x), d1 = unpack_bf16x2(dv[k].y), d2 = unpack_bf16x2(dv[k].z), d3 = unpack_bf16x2(dv[k].w);
        float acc = a0.x * d0.x + a0.y * d0.y + a1.x * d1.x + a1.y * d1.y + a2.x * d2.x + a2.y * d2.y + a3.x * d3.x + a3.y * d3.y;
#pragma unroll
        for (int sft = 8; sft > 0; sft >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, sft);
        if (l16 == 0) {
            float2 out = make_float2(INFINITY, 0.0f);
            if (i < Lq) {
                const float l = lse[(static_cast<long long>(b) * H + h) * Lq + i];
                out = make_float2(l == -INFINITY ? INFINITY : l * LOG2E, acc);
            }
            stats[(static_cast<long long>(b) * H + h) * Lqp + i] = out;
        }
    }
}

// Write 32 consecutive bf16 (16 packed words) of row r, columns [32*half, 32*half+32), into a [128 x 64] K-major
// SW128 operand tile (row pitch 128 B, 16-byte unit u stored at unit u ^ (r & 7)).
// Explicit st.shared / ld.shared: through a generic pointer these compiled to ST.E / LD.E plus a MEMBAR.ALL.CTA in front of the proxy
// fence (7 % of the dK/dV kernel's stall samples, ncu source view).
ESB_DEVICE void store_operand_half_row(uint8_t* tile, int r, int half, const uint32_t* w) {
    const uint32_t row = smem_u32(tile) + static_cast<uint32_t>(r * 128);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int u = half * 4 + i;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + static_cast<uint32_t>((u ^ (r & 7)) << 4)), "r"(w[i * 4 + 0]),
                     "r"(w[i * 4 + 1]), "r"(w[i * 4 + 2]), "r"(w[i * 4 + 3])
                     : "memory");
    }
}
ESB_DEVICE float4 lds_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// ================================================================================================ dQ
// K / V rings one stage deeper than the MMA look-ahead (4 / 3): a ring slot is only freed by an MMA that retires one block time
// before the slot's next use, and with 3 / 2 stages that left the TMA round trip on the critical path (ncu: 36 % of the compute
// warps' samples waiting for S / dP, tensor pipe 16 % - profiles/r02_attn_bwd_ncu.txt)
constexpr int DQ_K_STAGES = 4;
constexpr int DQ_V_STAGES = 3;
constexpr int DQ_SMEM = 2 * TILE_BYTES + DQ_K_STAGES * BLK_BYTES + DQ_V_STAGES * BLK_BYTES + 2 * OP_BYTES + 256 + 1024;
// TMEM columns: S[2] at 0/64, dP[2] at 128/192, dQ at 256..383

__global__ void __launch_bounds__(BT_THREADS, 1)
attn_bwd_dq_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tdo,
                      const __grid_constant__ CUtensorMap tk, const __grid_constant__ CUtensorMap tv, const BwdTcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sDO = sQ + TILE_BYTES;
    uint8_t* sK = sDO + TILE_BYTES;                      // DQ_K_STAGES stages
    uint8_t* sV = sK + DQ_K_STAGES * BLK_BYTES;          // DQ_V_STAGES stages
    uint8_t* sDS = sV + DQ_V_STAGES * BLK_BYTES;         // 2 buffers
    uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + 2 * OP_BYTES);
    uint64_t* bar_qdo = bars;
    uint64_t* bar_k = bars + 1;        // [4]
    uint64_t* bar_v = bars + 5;        // [3]
    uint64_t* bar_s = bars + 8;        // [2] S_j and dP_j computed
    uint64_t* bar_sfree = bars + 10;   // [2] all compute threads have read them
    uint64_t* bar_ds = bars + 12;      // [2] dS_j written
    uint64_t* bar_dq = bars + 14;      // [2] dQ += dS_j K_j retired
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int qt = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);
    const int h = blockIdx.y, b = blockIdx.z;
    const int m0 = qt * BT_TILE;
    int kv_end = p.Lk;
    if (p.causal) kv_end = min(p.Lk, m0 + BT_TILE + p.q_pos0);
    const int nblk = max(0, (kv_end + BT_BLK - 1) / BT_BLK);

    const int hcol = h * BT_HD;
    if (warp == 8) {
        // the thread that initialises the barriers also requests the first operand tiles right away: the TMA round trip then runs
        // under the TMEM allocation and the CTA-wide barrier instead of after them (every CTA pays this prologue, 10.4 waves)
        if (lane == 0) {
            mbar_init(bar_qdo, 1);
            for (int i = 0; i < DQ_K_STAGES; ++i) mbar_init(&bar_k[i], 1);
            for (int i = 0; i < DQ_V_STAGES; ++i) mbar_init(&bar_v[i], 1);
            for (int i = 0; i < 2; ++i) {
                mbar_init(&bar_s[i], 1); mbar_init(&bar_sfree[i], BT_COMPUTE / 32);
                mbar_init(&bar_ds[i], BT_COMPUTE / 32); mbar_init(&bar_dq[i], 1);
            }
            fence_barrier_init();
            if (nblk > 0) {
                tma_prefetch_desc(&tq); tma_prefetch_desc(&tdo); tma_prefetch_desc(&tk); tma_prefetch_desc(&tv);
                mbar_arrive_expect_tx(bar_qdo, 2 * TILE_BYTES);
                tma_load_3d(sQ, &tq, bar_qdo, hcol, m0, b);
                tma_load_3d(sQ + TILE_BYTES / 2, &tq, bar_qdo, hcol + 64, m0, b);
                tma_load_3d(sDO, &tdo, bar_qdo, hcol, m0, b);
                tma_load_3d(sDO + TILE_BYTES / 2, &tdo, bar_qdo, hcol + 64, m0, b);
                for (int j = 0; j < min(nblk, DQ_K_STAGES); ++j) {
                    mbar_arrive_expect_tx(&bar_k[j], BLK_BYTES);
                    tma_load_3d(sK + j * BLK_BYTES, &tk, &bar_k[j], hcol, j * BT_BLK, b);
                    tma_load_3d(sK + j * BLK_BYTES + BLK_BYTES / 2, &tk, &bar_k[j], hcol + 64, j * BT_BLK, b);
                }
                for (int j = 0; j < min(nblk, DQ_V_STAGES); ++j) {
                    mbar_arrive_expect_tx(&bar_v[j], BLK_BYTES);
                    tma_load_3d(sV + j * BLK_BYTES, &tv, &bar_v[j], hcol, j * BT_BLK, b);
                    tma_load_3d(sV + j * BLK_BYTES + BLK_BYTES / 2, &tv, &bar_v[j], hcol + 64, j * BT_BLK, b);
                }
            }
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        if (lane == 0 && nblk > 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(BT_TILE, BT_BLK);          // 128 x 64, K-major x K-major
            constexpr uint32_t idesc_q = umma_idesc_bf16_bmn(BT_TILE, BT_HD);       // 128 x 128, B MN-major
            auto load_k = [&](int j) {
                const int st = j % DQ_K_STAGES;
                mbar_arrive_expect_tx(&bar_k[st], BLK_BYTES);
                tma_load_3d(sK + st * BLK_BYTES, &tk, &bar_k[st], hcol, j * BT_BLK, b);
                tma_load_3d(sK + st * BLK_BYTES + BLK_BYTES / 2, &tk, &bar_k[st], hcol + 64, j * BT_BLK, b);
            };
            auto load_v = [&](int j) {
                const int st = j % DQ_V_STAGES;
                mbar_arrive_expect_tx(&bar_v[st], BLK_BYTES);
                tma_load_3d(sV + st * BLK_BYTES, &tv, &bar_v[st], hcol, j * BT_BLK, b);
                tma_load_3d(sV + st * BLK_BYTES + BLK_BYTES / 2, &tv, &bar_v[st], hcol + 64, j * BT_BLK, b);
            };
            auto issue_sdp = [&](int j) {             // S_j = Q K_j^T, dP_j = dO V_j^T
                const int st = j & 1, ks3 = j % DQ_K_STAGES, vs3 = j % DQ_V_STAGES;
                mbar_wait(&bar_k[ks3], (j / DQ_K_STAGES) & 1);
                mbar_wait(&bar_v[vs3], (j / DQ_V_STAGES) & 1);
                tc_fence_after();
                const uint32_t aq = smem_u32(sQ), ado = smem_u32(sDO), ak = smem_u32(sK + ks3 * BLK_BYTES), av = smem_u32(sV + vs3 * BLK_BYTES);
#pragma unroll
                for (int ks = 0; ks < BT_HD / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(aq + (ks >> 2) * (TILE_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    const uint64_t db = umma_desc_kmajor_sw128(ak + (ks >> 2) * (BLK_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    umma_bf16_ss(tmem_base + static_cast<uint32_t>(st * 64), da, db, idesc_s, ks != 0 ? 1u : 0u);
                }
#pragma unroll
                for (int ks = 0; ks < BT_HD / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(ado + (ks >> 2) * (TILE_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    const uint64_t db = umma_desc_kmajor_sw128(av + (ks >> 2) * (BLK_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    umma_bf16_ss(tmem_base + static_cast<uint32_t>(128 + st * 64), da, db, idesc_s, ks != 0 ? 1u : 0u);
                }
                umma_commit(&bar_s[st]);
            };
            mbar_wait(bar_qdo, 0);                       // Q, dO and the first K / V blocks were requested in the prologue
            issue_sdp(0);
            if (nblk > 1) issue_sdp(1);
            for (int j = 0; j < nblk; ++j) {
                const int st = j & 1;
                if (j + DQ_V_STAGES < nblk) {            // V stage j % 3 is free once dP_j has been computed
                    mbar_wait(&bar_s[st], (j >> 1) & 1);
                    load_v(j + DQ_V_STAGES);
                }
                if (j >= 1 && j - 1 + DQ_K_STAGES < nblk) {   // K stage (j-1) % 4 is free once dQ += dS_{j-1} K_{j-1} retired
                    mbar_wait(&bar_dq[(j - 1) & 1], ((j - 1) >> 1) & 1);
                    load_k(j - 1 + DQ_K_STAGES);
                }
                // S / dP of block j+2 go to the tensor pipe as soon as the compute warps have READ buffer st (early in their work
                // on block j) - not after dS_j is finished: the MMAs then run under the exp / pack / store of block j instead of
                // after them (the loop is bound by these hand-off latencies, not by MMA throughput)
                if (j + 2 < nblk) {
                    mbar_wait(&bar_sfree[st], (j >> 1) & 1);
                    issue_sdp(j + 2);
                }
                mbar_wait(&bar_ds[st], (j >> 1) & 1);    // dS_j is in smem
                tc_fence_after();
                const uint32_t ads = smem_u32(sDS + st * OP_BYTES), ak = smem_u32(sK + (j % DQ_K_STAGES) * BLK_BYTES);
#pragma unroll
                for (int ks = 0; ks < BT_BLK / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(ads) + static_cast<uint64_t>(2 * ks);
                    const uint64_t db = umma_desc_mnmajor_sw128(ak + ks * 2048, BLK_BYTES / 2);
                    umma_bf16_ss(tmem_base + 256, da, db, idesc_q, (j | ks) != 0 ? 1u : 0u);
                }
                umma_commit(&bar_dq[st]);
            }
        }
    } else {
        // ===================================================== compute warps
        const int quarter = warp & 3, half = warp >> 2;
        const int r = quarter * 32 + lane;                  // tile row = TMEM lane
        const uint32_t tmem_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
        const int qrow = m0 + r;
        const int qpos = qrow + p.q_pos0;
        const int qpos_warp_min = m0 + quarter * 32 + p.q_pos0;
        const uint8_t* km = p.kmask ? p.kmask + static_cast<long long>(b) * p.kmask_ld : nullptr;
        const float2 stt = p.stats[(static_cast<long long>(b) * p.H + h) * p.Lqp + min(qrow, p.Lqp - 1)];
        const float neg_lse2 = -stt.x, drow = stt.y;        // dead / out-of-range rows carry lse2 = +inf -> P = 0
        const float sc = p.scale_log2;
        // the mask byte of the NEXT block is requested one iteration ahead and only compared at the top of that iteration: a
        // compare right behind the load stalled the warp for a global-memory round trip per block (10 % of this kernel's samples)
        auto key_raw = [&](int key) -> uint32_t { return key < p.Lk ? (km ? static_cast<uint32_t>(km[key]) : 1u) : 0u; };
        uint32_t mk_n = nblk > 0 ? key_raw(half * 32 + lane) : 0u;

        for (int j = 0; j < nblk; ++j) {
            const int st = j & 1;
            const int cbase = j * BT_BLK + half * 32;       // first key of this warp's 32 columns
            uint32_t vb = __ballot_sync(0xffffffffu, mk_n != 0u);
            if (j + 1 < nblk) mk_n = key_raw(cbase + BT_BLK + lane);
            const bool need_mask = vb != 0xffffffffu || (p.causal && cbase + 31 > qpos_warp_min);
            if (p.causal) {
                const int nvis = qpos - cbase + 1;
                vb &= nvis >= 32 ? 0xffffffffu : (nvis <= 0 ? 0u : ((1u << nvis) - 1u));
            }
            mbar_wait(&bar_s[st], (j >> 1) & 1);
            tc_fence_after();
            uint32_t sr[32], dpr[32];
            tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(st * 64 + half * 32), sr);
            tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(128 + st * 64 + half * 32), dpr);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_sfree[st]);      // one arrive per warp: 8 instead of 256 serialised arrivals
            uint32_t w[16];
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
                float p0 = ex2_approx(fmaf(__uint_as_float(sr[c]), sc, neg_lse2));
                float p1 = ex2_approx(fmaf(__uint_as_float(sr[c + 1]), sc, neg_lse2));
                if (need_mask) {
                    if (!((vb >> c) & 1u)) p0 = 0.0f;
                    if (!((vb >> (c + 1)) & 1u)) p1 = 0.0f;
                }
                w[c >> 1] = pack_bf16x2(p0 * (__uint_as_float(dpr[c]) - drow), p1 * (__uint_as_float(dpr[c + 1]) - drow));
            }
            if (j >= 2) mbar_wait(&bar_dq[st], ((j - 2) >> 1) & 1);     // dS buffer st was last read by block j-2
            store_operand_half_row(sDS + st * OP_BYTES, r, half, w);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_ds[st]);
        }
        // ---- epilogue: dQ * scale -> bf16
        if (nblk > 0) {
            mbar_wait(&bar_dq[(nblk - 1) & 1], ((nblk - 1) >> 1) & 1);
            tc_fence_after();
        }
        __nv_bfloat16* gp = p.dq + static_cast<long long>(b) * p.dq_bs + static_cast<long long>(qrow) * p.dq_rs + static_cast<long long>(h) * BT_HD + half * 64;
        if (p.rope_cos) {
            // fused inverse RoPE: this thread takes columns [32 half, +32) and their rotation partners [64 + 32 half, +32)
            uint32_t a1[32], a2[32];
            if (nblk > 0) {
                tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(256 + half * 32), a1);
                tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(256 + 64 + half * 32), a2);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) a1[i] = a2[i] = 0u;
            }
            if (qrow < p.Lq) {
                __nv_bfloat16* g0 = gp - half * 64 + half * 32;
                rope_bwd_store(g0, g0 + 64, a1, a2, p.scale, p.rope_cos + static_cast<long long>(qpos) * 64 + half * 32,
                               p.rope_sin + static_cast<long long>(qpos) * 64 + half * 32);
            }
        } else {
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            uint32_t acc[32];
            if (nblk > 0) {
                tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(256 + half * 64 + c * 32), acc);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = 0u;
            }
            if (qrow < p.Lq) {
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    uint4 u;
                    u.x = pack_bf16x2(__uint_as_float(acc[g * 8 + 0]) * p.scale, __uint_as_float(acc[g * 8 + 1]) * p.scale);
                    u.y = pack_bf16x2(__uint_as_float(acc[g * 8 + 2]) * p.scale, __uint_as_float(acc[g * 8 + 3]) * p.scale);
                    u.z = pack_bf16x2(__uint_as_float(acc[g * 8 + 4]) * p.scale, __uint_as_float(acc[g * 8 + 5]) * p.scale);
                    u.w = pack_bf16x2(__uint_as_float(acc[g * 8 + 6]) * p.scale, __uint_as_float(acc[g * 8 + 7]) * p.scale);
                    *reinterpret_cast<uint4*>(gp + c * 32 + g * 8) = u;
                }
            }
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ================================================================================================ dK / dV
constexpr int KV_Q_STAGES = 3;
constexpr int KV_STAT_BYTES = BT_BLK * 8;            // 64 x float2
// P^T / dS^T double-buffered: with one buffer the compute warps of block j+1 stalled until the dV / dK MMAs of block j had retired
// (ncu: 18 % of their samples on that wait, profiles/r02_attn_bwd_ncu.txt)
constexpr int KV_SMEM = 2 * TILE_BYTES + 2 * KV_Q_STAGES * BLK_BYTES + 4 * OP_BYTES + KV_Q_STAGES * KV_STAT_BYTES + 256 + 1024;
static_assert(KV_SMEM <= 232448, "dK/dV kernel shared memory");
// TMEM columns: S^T[2] at 0/64, dP^T[2] at 128/192, dV at 256..383, dK at 384..511

__global__ void __launch_bounds__(BT_THREADS, 1)
attn_bwd_dkdv_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tdo,
                        const __grid_constant__ CUtensorMap tk, const __grid_constant__ CUtensorMap tv, const BwdTcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem;
    uint8_t* sV = sK + TILE_BYTES;
    uint8_t* sQ = sV + TILE_BYTES;                       // 3 stages
    uint8_t* sDO = sQ + KV_Q_STAGES * BLK_BYTES;         // 3 stages
    uint8_t* sPT = sDO + KV_Q_STAGES * BLK_BYTES;        // 2 buffers
    uint8_t* sDST = sPT + 2 * OP_BYTES;                  // 2 buffers
    float2* sStat = reinterpret_cast<float2*>(sDST + 2 * OP_BYTES);    // 3 x 64
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sStat) + KV_Q_STAGES * KV_STAT_BYTES);
    uint64_t* bar_kv = bars;
    uint64_t* bar_q = bars + 1;        // [3] Q_i, dO_i, stats_i landed
    uint64_t* bar_s = bars + 4;        // [2] S^T, dP^T computed
    uint64_t* bar_sfree = bars + 6;    // [2]
    uint64_t* bar_p = bars + 8;        // [2] P^T, dS^T written (buffer jj & 1)
    uint64_t* bar_acc = bars + 10;     // [2] dV / dK MMAs of the block retired (buffer jj & 1 reusable)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = blockIdx.y, b = blockIdx.z;
    const int n0 = blockIdx.x * BT_TILE;                 // first key of this tile (early tiles see the most queries)
    const int nq_total = (p.Lq + BT_BLK - 1) / BT_BLK;
    int i_begin = 0;
    if (p.causal) i_begin = max(0, n0 - p.q_pos0) / BT_BLK;
    const int nq = max(0, nq_total - i_begin);

    const int hcol = h * BT_HD;
    const float2* stat_row = p.stats + (static_cast<long long>(b) * p.H + h) * p.Lqp;
    if (warp == 8) {
        if (lane == 0) {                                     // barriers, then the first tiles at once (see the dQ kernel)
            mbar_init(bar_kv, 1);
            for (int i = 0; i < KV_Q_STAGES; ++i) mbar_init(&bar_q[i], 1);
            for (int i = 0; i < 2; ++i) { mbar_init(&bar_s[i], 1); mbar_init(&bar_sfree[i], BT_COMPUTE / 32); }
            for (int i = 0; i < 2; ++i) { mbar_init(&bar_p[i], BT_COMPUTE / 32); mbar_init(&bar_acc[i], 1); }
            fence_barrier_init();
            if (nq > 0) {
                tma_prefetch_desc(&tq); tma_prefetch_desc(&tdo); tma_prefetch_desc(&tk); tma_prefetch_desc(&tv);
                mbar_arrive_expect_tx(bar_kv, 2 * TILE_BYTES);
                tma_load_3d(sK, &tk, bar_kv, hcol, n0, b);
                tma_load_3d(sK + TILE_BYTES / 2, &tk, bar_kv, hcol + 64, n0, b);
                tma_load_3d(sV, &tv, bar_kv, hcol, n0, b);
                tma_load_3d(sV + TILE_BYTES / 2, &tv, bar_kv, hcol + 64, n0, b);
                for (int jj = 0; jj < min(nq, KV_Q_STAGES); ++jj) {
                    const int row0 = (i_begin + jj) * BT_BLK;
                    mbar_arrive_expect_tx(&bar_q[jj], 2 * BLK_BYTES + KV_STAT_BYTES);
                    tma_load_3d(sQ + jj * BLK_BYTES, &tq, &bar_q[jj], hcol, row0, b);
                    tma_load_3d(sQ + jj * BLK_BYTES + BLK_BYTES / 2, &tq, &bar_q[jj], hcol + 64, row0, b);
                    tma_load_3d(sDO + jj * BLK_BYTES, &tdo, &bar_q[jj], hcol, row0, b);
                    tma_load_3d(sDO + jj * BLK_BYTES + BLK_BYTES / 2, &tdo, &bar_q[jj], hcol + 64, row0, b);
                    bulk_load_1d(sStat + jj * BT_BLK, stat_row + row0, KV_STAT_BYTES, &bar_q[jj]);
                }
            }
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        if (lane == 0 && nq > 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(BT_TILE, BT_BLK);
            constexpr uint32_t idesc_a = umma_idesc_bf16_bmn(BT_TILE, BT_HD);
            auto load_q = [&](int jj) {
                const int st = jj % KV_Q_STAGES, row0 = (i_begin + jj) * BT_BLK;
                mbar_arrive_expect_tx(&bar_q[st], 2 * BLK_BYTES + KV_STAT_BYTES);
                tma_load_3d(sQ + st * BLK_BYTES, &tq, &bar_q[st], hcol, row0, b);
                tma_load_3d(sQ + st * BLK_BYTES + BLK_BYTES / 2, &tq, &bar_q[st], hcol + 64, row0, b);
                tma_load_3d(sDO + st * BLK_BYTES, &tdo, &bar_q[st], hcol, row0, b);
                tma_load_3d(sDO + st * BLK_BYTES + BLK_BYTES / 2, &tdo, &bar_q[st], hcol + 64, row0, b);
                bulk_load_1d(sStat + st * BT_BLK, stat_row + row0, KV_STAT_BYTES, &bar_q[st]);
            };
            auto issue_s = [&](int jj) {              // S^T = K Q_i^T, dP^T = V dO_i^T
                const int st = jj & 1, q3 = jj % KV_Q_STAGES;
                mbar_wait(&bar_q[q3], (jj / KV_Q_STAGES) & 1);
                tc_fence_after();
                const uint32_t ak = smem_u32(sK), av = smem_u32(sV), aq = smem_u32(sQ + q3 * BLK_BYTES), ado = smem_u32(sDO + q3 * BLK_BYTES);
#pragma unroll
                for (int ks = 0; ks < BT_HD / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(ak + (ks >> 2) * (TILE_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    const uint64_t db = umma_desc_kmajor_sw128(aq + (ks >> 2) * (BLK_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    umma_bf16_ss(tmem_base + static_cast<uint32_t>(st * 64), da, db, idesc_s, ks != 0 ? 1u : 0u);
                }
#pragma unroll
                for (int ks = 0; ks < BT_HD / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(av + (ks >> 2) * (TILE_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    const uint64_t db = umma_desc_kmajor_sw128(ado + (ks >> 2) * (BLK_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    umma_bf16_ss(tmem_base + static_cast<uint32_t>(128 + st * 64), da, db, idesc_s, ks != 0 ? 1u : 0u);
                }
                umma_commit(&bar_s[st]);
            };
            mbar_wait(bar_kv, 0);
            issue_s(0);
            if (nq > 1) issue_s(1);
            for (int jj = 0; jj < nq; ++jj) {
                const int st = jj & 1, q3 = jj % KV_Q_STAGES;
                if (jj >= 1 && jj + 2 < nq) {            // stage (jj-1)%3 is free once the dV/dK MMAs of block jj-1 retired
                    mbar_wait(&bar_acc[(jj - 1) & 1], ((jj - 1) >> 1) & 1);
                    load_q(jj + 2);
                }
                mbar_wait(&bar_p[st], (jj >> 1) & 1);    // P^T_jj, dS^T_jj are in smem (buffer st)
                tc_fence_after();
                const uint32_t apt = smem_u32(sPT + st * OP_BYTES), adst = smem_u32(sDST + st * OP_BYTES), aq = smem_u32(sQ + q3 * BLK_BYTES), ado = smem_u32(sDO + q3 * BLK_BYTES);
#pragma unroll
                for (int ks = 0; ks < BT_BLK / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(apt) + static_cast<uint64_t>(2 * ks);
                    const uint64_t db = umma_desc_mnmajor_sw128(ado + ks * 2048, BLK_BYTES / 2);
                    umma_bf16_ss(tmem_base + 256, da, db, idesc_a, (jj | ks) != 0 ? 1u : 0u);
                }
#pragma unroll
                for (int ks = 0; ks < BT_BLK / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(adst) + static_cast<uint64_t>(2 * ks);
                    const uint64_t db = umma_desc_mnmajor_sw128(aq + ks * 2048, BLK_BYTES / 2);
                    umma_bf16_ss(tmem_base + 384, da, db, idesc_a, (jj | ks) != 0 ? 1u : 0u);
                }
                umma_commit(&bar_acc[st]);
                // S^T / dP^T of block jj+2 go out BEHIND the accumulate MMAs of block jj: their Q / dO stage was requested at the
                // top of this iteration, and with only 3 stages its TMA round trip (~1 us) used to stall this thread in front of
                // the accumulate MMAs every iteration (ncu: the compute warps of this kernel waited 17 % + 9 % on S and Q)
                if (jj + 2 < nq) {
                    mbar_wait(&bar_sfree[st], (jj >> 1) & 1);
                    issue_s(jj + 2);
                }
            }
        }
    } else {
        // ===================================================== compute warps
        const int quarter = warp & 3, half = warp >> 2;
        const int r = quarter * 32 + lane;                  // tile row = TMEM lane = key
        const uint32_t tmem_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
        const int key = n0 + r;
        const bool kok = key < p.Lk && (p.kmask == nullptr || p.kmask[static_cast<long long>(b) * p.kmask_ld + key] != 0);
        const bool any_dead = __any_sync(0xffffffffu, !kok);
        const int key_warp_max = n0 + quarter * 32 + 31;
        const float sc = p.scale_log2;

        for (int jj = 0; jj < nq; ++jj) {
            const int st = jj & 1, q3 = jj % KV_Q_STAGES;
            const int cbase = (i_begin + jj) * BT_BLK + half * 32;      // first query of this warp's 32 columns
            uint32_t vb = kok ? 0xffffffffu : 0u;
            bool need_mask = any_dead;
            if (p.causal) {
                need_mask = need_mask || (key_warp_max - p.q_pos0 > cbase);
                const int cmin = key - p.q_pos0 - cbase;                // columns c >= cmin are visible
                vb &= cmin <= 0 ? 0xffffffffu : (cmin >= 32 ? 0u : ~((1u << cmin) - 1u));
            }
            mbar_wait(&bar_q[q3], (jj / KV_Q_STAGES) & 1);              // stats_i visible
            mbar_wait(&bar_s[st], (jj >> 1) & 1);
            tc_fence_after();
            uint32_t sr[32], dpr[32];
            tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(st * 64 + half * 32), sr);
            tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(128 + st * 64 + half * 32), dpr);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_sfree[st]);      // one arrive per warp: 8 instead of 256 serialised arrivals
            const uint32_t stq = smem_u32(sStat + q3 * BT_BLK + half * 32);
            uint32_t wp[16], wd[16];
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
                const float4 s01 = lds_f4(stq + static_cast<uint32_t>(c * 8));          // (lse2, D) of queries c and c + 1
                const float2 s0 = make_float2(s01.x, s01.y), s1 = make_float2(s01.z, s01.w);
                float p0 = ex2_approx(fmaf(__uint_as_float(sr[c]), sc, -s0.x));
                float p1 = ex2_approx(fmaf(__uint_as_float(sr[c + 1]), sc, -s1.x));
                if (need_mask) {
                    if (!((vb >> c) & 1u)) p0 = 0.0f;
                    if (!((vb >> (c + 1)) & 1u)) p1 = 0.0f;
                }
                wp[c >> 1] = pack_bf16x2(p0, p1);
                wd[c >> 1] = pack_bf16x2(p0 * (__uint_as_float(dpr[c]) - s0.y), p1 * (__uint_as_float(dpr[c + 1]) - s1.y));
            }
            if (jj >= 2) mbar_wait(&bar_acc[st], ((jj - 2) >> 1) & 1);  // the dV/dK MMAs of block jj-2 no longer read buffer st
            store_operand_half_row(sPT + st * OP_BYTES, r, half, wp);
            store_operand_half_row(sDST + st * OP_BYTES, r, half, wd);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_p[st]);
        }
        // ---- epilogue: dV, dK * scale -> bf16
        if (nq > 0) {
            mbar_wait(&bar_acc[(nq - 1) & 1], ((nq - 1) >> 1) & 1);
            tc_fence_after();
        }
        __nv_bfloat16* gv = p.dv + static_cast<long long>(b) * p.dv_bs + static_cast<long long>(key) * p.dv_rs + static_cast<long long>(h) * BT_HD + half * 64;
        __nv_bfloat16* gk = p.dk + static_cast<long long>(b) * p.dk_bs + static_cast<long long>(key) * p.dk_rs + static_cast<long long>(h) * BT_HD + half * 64;
        if (p.rope_cos) {
            uint32_t a1[32], a2[32];
            if (nq > 0) {
                tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(384 + half * 32), a1);
                tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(384 + 64 + half * 32), a2);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) a1[i] = a2[i] = 0u;
            }
            if (key < p.Lk) {
                __nv_bfloat16* g0 = gk - half * 64 + half * 32;
                rope_bwd_store(g0, g0 + 64, a1, a2, p.scale, p.rope_cos + static_cast<long long>(key) * 64 + half * 32,
                               p.rope_sin + static_cast<long long>(key) * 64 + half * 32);
            }
        }
#pragma unroll 1
        for (int c = 0; c < (p.rope_cos ? 2 : 4); ++c) {
            const bool is_k = c >= 2;
            const float mul = is_k ? p.scale : 1.0f;
            uint32_t acc[32];
            if (nq > 0) {
                tmem_ld_32x32(tmem_lane + static_cast<uint32_t>((is_k ? 384 : 256) + half * 64 + (c & 1) * 32), acc);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = 0u;
            }
            if (key < p.Lk) {
                __nv_bfloat16* gp = (is_k ? gk : gv) + (c & 1) * 32;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    uint4 u;
                    u.x = pack_bf16x2(__uint_as_float(acc[g * 8 + 0]) * mul, __uint_as_float(acc[g * 8 + 1]) * mul);
                    u.y = pack_bf16x2(__uint_as_float(acc[g * 8 + 2]) * mul, __uint_as_float(acc[g * 8 + 3]) * mul);
                    u.z = pack_bf16x2(__uint_as_float(acc[g * 8 + 4]) * mul, __uint_as_float(acc[g * 8 + 5]) * mul);
                    u.w = pack_bf16x2(__uint_as_float(acc[g * 8 + 6]) * mul, __uint_as_float(acc[g * 8 + 7]) * mul);
                    *reinterpret_cast<uint4*>(gp + g * 8) = u;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}


}  // namespace

// Backward of attn_fwd_tc128.  stats_ws: scratch of 2 * B * H * round_up(Lq, 64) floats.
int attn_bwd_tc128(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
                   long long v_bs, long long v_rs, const void* o, long long o_bs, long long o_rs, const void* dout, long long do_bs,
                   long long do_rs, const float* lse, float* stats_ws, void* dq, long long dq_bs, long long dq_rs, void* dk,
                   long long dk_bs, long long dk_rs, void* dv, long long dv_bs, long long dv_rs, const uint8_t* kmask,
                   long long kmask_ld, int B, int H, int Lq, int Lk, int causal, int q_pos0, float scale, cudaStream_t s,
                   const void* rope_cos, const void* rope_sin) {
    if ((dq_rs & 7) || (dk_rs & 7) || (dv_rs & 7) || (dq_bs & 7) || (dk_bs & 7) || (dv_bs & 7) || (o_rs & 7) || (o_bs & 7) ||
        (do_rs & 7) || (do_bs & 7) || (reinterpret_cast<uintptr_t>(o) & 15) || (reinterpret_cast<uintptr_t>(dout) & 15) ||
        (reinterpret_cast<uintptr_t>(dq) & 15) || (reinterpret_cast<uintptr_t>(dk) & 15) || (reinterpret_cast<uintptr_t>(dv) & 15) ||
        (reinterpret_cast<uintptr_t>(stats_ws) & 15)) {
        set_last_error("attn_bwd: o / dout / gradients / workspace must keep 16-byte alignment");
        return ESB_ERR_INVALID_ARG;
    }
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e1 = cudaFuncSetAttribute(attn_bwd_dq_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DQ_SMEM);
        cudaError_t e2 = cudaFuncSetAttribute(attn_bwd_dkdv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, KV_SMEM);
        if (e1 != cudaSuccess || e2 != cudaSuccess) {
            set_last_error("attn_bwd_tc attr: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
            return ESB_ERR_CUDA;
        }
        attr_set = true;
    }
    const int Lqp = (Lq + BT_BLK - 1) / BT_BLK * BT_BLK;
    const long long cols = static_cast<long long>(H) * BT_HD;
    CUtensorMap tq128, tdo128, tk64, tv64, tq64, tdo64, tk128, tv128;
    int rc = 0;
    if ((rc = make_tma_3d_bf16(&tq128, q, cols, Lq, B, q_rs, q_bs, 64, BT_TILE))) return rc;
    if ((rc = make_tma_3d_bf16(&tdo128, dout, cols, Lq, B, do_rs, do_bs, 64, BT_TILE))) return rc;
    if ((rc = make_tma_3d_bf16(&tk64, k, cols, Lk, B, k_rs, k_bs, 64, BT_BLK))) return rc;
    if ((rc = make_tma_3d_bf16(&tv64, v, cols, Lk, B, v_rs, v_bs, 64, BT_BLK))) return rc;
    if ((rc = make_tma_3d_bf16(&tq64, q, cols, Lq, B, q_rs, q_bs, 64, BT_BLK))) return rc;
    if ((rc = make_tma_3d_bf16(&tdo64, dout, cols, Lq, B, do_rs, do_bs, 64, BT_BLK))) return rc;
    if ((rc = make_tma_3d_bf16(&tk128, k, cols, Lk, B, k_rs, k_bs, 64, BT_TILE))) return rc;
    if ((rc = make_tma_3d_bf16(&tv128, v, cols, Lk, B, v_rs, v_bs, 64, BT_TILE))) return rc;

    float2* stats = reinterpret_cast<float2*>(stats_ws);
    const long long warps = static_cast<long long>(B) * H * (Lqp / 8);          // Lqp is a multiple of 128
    attn_bwd_stats_kernel<<<static_cast<unsigned>((warps + 3) / 4), 128, 0, s>>>(
        static_cast<const __nv_bfloat16*>(o), o_bs, o_rs, static_cast<const __nv_bfloat16*>(dout), do_bs, do_rs, lse, stats, B, H, Lq, Lqp);

    BwdTcParams p;
    p.dq = static_cast<__nv_bfloat16*>(dq); p.dk = static_cast<__nv_bfloat16*>(dk); p.dv = static_cast<__nv_bfloat16*>(dv);
    p.stats = stats; p.kmask = kmask;
    p.dq_bs = dq_bs; p.dq_rs = dq_rs; p.dk_bs = dk_bs; p.dk_rs = dk_rs; p.dv_bs = dv_bs; p.dv_rs = dv_rs; p.kmask_ld = kmask_ld;
    p.rope_cos = static_cast<const __nv_bfloat16*>(rope_cos); p.rope_sin = static_cast<const __nv_bfloat16*>(rope_sin);
    p.H = H; p.Lq = Lq; p.Lqp = Lqp; p.Lk = Lk; p.causal = causal; p.q_pos0 = q_pos0; p.scale = scale; p.scale_log2 = scale * LOG2E;
    dim3 gkv((Lk + BT_TILE - 1) / BT_TILE, H, B);
    attn_bwd_dkdv_tc_kernel<<<gkv, BT_THREADS, KV_SMEM, s>>>(tq64, tdo64, tk128, tv128, p);
    dim3 gq((Lq + BT_TILE - 1) / BT_TILE, H, B);
    attn_bwd_dq_tc_kernel<<<gq, BT_THREADS, DQ_SMEM, s>>>(tq128, tdo128, tk64, tv64, p);
    return check_launch("attn_bwd_tc");
}


}  // namespace esb
