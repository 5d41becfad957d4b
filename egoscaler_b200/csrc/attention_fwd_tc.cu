// Fused attention forward on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), head_dim 128.
//
// Reference semantics: HF LlamaAttention -> sdpa_attention_forward (transformers integrations/sdpa_attention.py:40-98,
// modeling_llama.py:253-290) reached from pointllm/model/pointllm.py:173-178: softmax(Q K^T * hd^-0.5 + mask) V with
// mask = causal AND key-padding, fp32 softmax statistics, probabilities rounded to bf16 before P@V.
//
// One CTA = one (batch, head, 128-query tile); 128 threads, thread t owns query row t (= TMEM lane t), so the softmax
// needs no shuffles.  Per 64-key block j:
//     S_j = Q K_j^T        tcgen05.mma 128x64x16 (x8), both operands K-major from TMA-written 128B-swizzled smem,
//                          fp32 result in TMEM (two S buffers, so S_{j+1} is computed while block j is in softmax)
//     P_j = exp2(S_j*c-m)  tcgen05.ld -> registers -> bf16 -> smem in the K-major SW128 operand layout
//     O  += P_j V_j        tcgen05.mma 128x128x16 (x4), A = P (K-major), B = V_j as stored ([key, hd] = MN-major)
// O stays in TMEM for the whole tile; it is only rescaled when a row maximum grows by more than 2^8 (lazy rescale),
// which keeps exp2 arguments bounded and makes the correction pass rare.  Two CTAs are co-resident per SM (112 KB smem,
// 256 TMEM columns each), so one CTA's softmax overlaps the other's MMAs.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

namespace {

constexpr int FA_BM = 128;          // queries per CTA
constexpr int FA_BN = 64;           // keys per block
constexpr int FA_HD = 128;
constexpr int FA_SOFTMAX_THREADS = 128;
constexpr int FA_THREADS = 160;        // 4 softmax warps + 1 producer / MMA-issuer warp
constexpr int Q_BYTES = FA_BM * FA_HD * 2;        // 32 KB : 2 chunks [128 rows x 64 hd]
constexpr int K_BYTES = FA_BN * FA_HD * 2;        // 16 KB : 2 chunks [64 keys x 64 hd]
constexpr int V_BYTES = FA_BN * FA_HD * 2;        // 16 KB : 2 chunks [64 keys x 64 hd] (chunk = hd half)
constexpr int P_BYTES = FA_BM * FA_BN * 2;        // 16 KB : [128 rows x 64 keys]
// 2 CTAs/SM: 2 * (FA_SMEM + 1 KB driver reserve) must stay <= 228 KB -> exactly 1 KB for barriers (128 B) + alignment slack
// (the dynamic smem base is at least 128-byte aligned, so 896 B of slack always reaches the next 1 KB boundary).
constexpr int FA_SMEM = Q_BYTES + 2 * K_BYTES + 2 * V_BYTES + P_BYTES + 128 + 896;
constexpr uint32_t TMEM_COLS = 256;               // S0 [0,64) S1 [64,128) O [128,256)
constexpr float RESCALE_THRESHOLD = 8.0f;         // log2 units

struct FaParams {
    __nv_bfloat16* o;
    float* lse;
    const uint8_t* kmask;
    long long o_bs, o_rs, kmask_ld;
    int H, Lq, Lk, causal, q_pos0;
    float scale_log2;        // scale * log2(e)
};

// Warp roles: warps 0-3 = softmax (thread t owns query row t = TMEM lane t), warp 4 lane 0 = TMA producer + MMA issuer.
// Hand-offs are mbarriers only (no CTA-wide barrier inside the key loop):
//   bar_s[2]     issuer -> softmax : S buffer computed (tcgen05.commit)
//   bar_sfree[2] softmax -> issuer : all 128 threads have read the S buffer (it may be overwritten by S_{j+2})
//   bar_p        softmax -> issuer : P_j is in smem (and any O rescale is done) -> issue O += P_j V_j
//   bar_o        issuer -> softmax : P V of block j retired (P / O / the other V stage may be touched)
__global__ void __launch_bounds__(FA_THREADS, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                   const __grid_constant__ CUtensorMap tv, const FaParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + Q_BYTES;                 // 2 stages
    uint8_t* sV = sK + 2 * K_BYTES;             // 2 stages
    uint8_t* sP = sV + 2 * V_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + P_BYTES);
    uint64_t* bar_q = bars;
    uint64_t* bar_k = bars + 1;                 // [2]
    uint64_t* bar_v = bars + 3;                 // [2]
    uint64_t* bar_s = bars + 5;                 // [2]
    uint64_t* bar_sfree = bars + 7;             // [2]
    uint64_t* bar_o = bars + 9;
    uint64_t* bar_p = bars + 10;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int qt = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);   // heavy (late) tiles first
    const int h = blockIdx.y, b = blockIdx.z;
    const int m0 = qt * FA_BM;

    int kv_end = p.Lk;
    if (p.causal) kv_end = min(p.Lk, m0 + FA_BM + p.q_pos0);
    const int nblk = (kv_end + FA_BN - 1) / FA_BN;

    const int hcol = h * FA_HD;
    if (warp == 4) {
        // barriers, then the first tiles at once: the TMA round trip runs under the TMEM allocation and the CTA barrier
        if (lane == 0) {
            mbar_init(bar_q, 1);
            for (int i = 0; i < 2; ++i) {
                mbar_init(&bar_k[i], 1); mbar_init(&bar_v[i], 1); mbar_init(&bar_s[i], 1); mbar_init(&bar_sfree[i], FA_SOFTMAX_THREADS);
            }
            mbar_init(bar_o, 1);
            mbar_init(bar_p, FA_SOFTMAX_THREADS);
            fence_barrier_init();
            if (nblk > 0) {
                tma_prefetch_desc(&tq); tma_prefetch_desc(&tk); tma_prefetch_desc(&tv);
                mbar_arrive_expect_tx(bar_q, Q_BYTES);
                tma_load_3d(sQ, &tq, bar_q, hcol, m0, b);
                tma_load_3d(sQ + Q_BYTES / 2, &tq, bar_q, hcol + 64, m0, b);
                for (int j = 0; j < min(nblk, 2); ++j) {
                    mbar_arrive_expect_tx(&bar_k[j], K_BYTES);
                    tma_load_3d(sK + j * K_BYTES, &tk, &bar_k[j], hcol, j * FA_BN, b);
                    tma_load_3d(sK + j * K_BYTES + K_BYTES / 2, &tk, &bar_k[j], hcol + 64, j * FA_BN, b);
                }
                mbar_arrive_expect_tx(&bar_v[0], V_BYTES);
                tma_load_3d(sV, &tv, &bar_v[0], hcol, 0, b);
                tma_load_3d(sV + V_BYTES / 2, &tv, &bar_v[0], hcol + 64, 0, b);
            }
        }
        __syncwarp();
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    constexpr uint32_t tmem_o = 128;

    if (warp == 4) {
        if (lane == 0 && nblk > 0) {
            // ================================================= TMA producer + MMA issuer
            constexpr uint32_t idesc_s = umma_idesc_bf16(FA_BM, FA_BN);
            constexpr uint32_t idesc_o = umma_idesc_bf16_bmn(FA_BM, FA_HD);
            auto load_k = [&](int j) {
                const int st = j & 1;
                mbar_arrive_expect_tx(&bar_k[st], K_BYTES);
                tma_load_3d(sK + st * K_BYTES, &tk, &bar_k[st], hcol, j * FA_BN, b);
                tma_load_3d(sK + st * K_BYTES + K_BYTES / 2, &tk, &bar_k[st], hcol + 64, j * FA_BN, b);
            };
            auto load_v = [&](int j) {
                const int st = j & 1;
                mbar_arrive_expect_tx(&bar_v[st], V_BYTES);
                tma_load_3d(sV + st * V_BYTES, &tv, &bar_v[st], hcol, j * FA_BN, b);
                tma_load_3d(sV + st * V_BYTES + V_BYTES / 2, &tv, &bar_v[st], hcol + 64, j * FA_BN, b);
            };
            auto issue_s = [&](int j) {              // S_j = Q K_j^T -> TMEM S buffer j&1
                const int st = j & 1;
                mbar_wait(&bar_k[st], (j >> 1) & 1);
                tc_fence_after();
                const uint32_t aq = smem_u32(sQ), ak = smem_u32(sK + st * K_BYTES);
#pragma unroll
                for (int ks = 0; ks < FA_HD / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(aq + (ks >> 2) * (Q_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    const uint64_t db = umma_desc_kmajor_sw128(ak + (ks >> 2) * (K_BYTES / 2)) + static_cast<uint64_t>(2 * (ks & 3));
                    umma_bf16_ss(tmem_base + static_cast<uint32_t>(st * FA_BN), da, db, idesc_s, ks != 0 ? 1u : 0u);
                }
                umma_commit(&bar_s[st]);
            };
            mbar_wait(bar_q, 0);                         // Q, K_0, K_1 and V_0 were requested in the prologue
            issue_s(0);
            if (nblk > 1) issue_s(1);
            for (int j = 0; j < nblk; ++j) {
                const int st = j & 1;
                if (j + 2 < nblk) {                      // K stage st is free once S_j has been computed
                    mbar_wait(&bar_s[st], (j >> 1) & 1);
                    load_k(j + 2);
                }
                if (j + 1 < nblk) {                      // V stage (j+1)&1 is free once P V of block j-1 retired
                    if (j > 0) mbar_wait(bar_o, (j - 1) & 1);
                    load_v(j + 1);
                }
                mbar_wait(bar_p, j & 1);                 // P_j written by all softmax threads
                mbar_wait(&bar_v[st], (j >> 1) & 1);
                tc_fence_after();
                const uint32_t ap = smem_u32(sP), av = smem_u32(sV + st * V_BYTES);
#pragma unroll
                for (int ks = 0; ks < FA_BN / 16; ++ks) {
                    const uint64_t da = umma_desc_kmajor_sw128(ap) + static_cast<uint64_t>(2 * ks);
                    const uint64_t db = umma_desc_mnmajor_sw128(av + ks * 2048, V_BYTES / 2);
                    umma_bf16_ss(tmem_base + tmem_o, da, db, idesc_o, (j | ks) != 0 ? 1u : 0u);
                }
                umma_commit(bar_o);
                if (j + 2 < nblk) {
                    mbar_wait(&bar_sfree[st], (j >> 1) & 1);   // every softmax thread has read S_j
                    issue_s(j + 2);
                }
            }
        }
    } else {
        // ===================================================== softmax warps
        const uint32_t tmem_lane = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
        const int qrow = m0 + tid;                          // this thread's query row
        const int qpos = qrow + p.q_pos0;                   // last visible key index (causal)
        const int qpos_warp_min = m0 + warp * 32 + p.q_pos0;
        const uint8_t* km = p.kmask ? p.kmask + static_cast<long long>(b) * p.kmask_ld : nullptr;
        float m_used = -INFINITY, l_run = 0.0f;
        const float sc = p.scale_log2;
        // key-validity prefetch (length + padding mask): lane l looks at keys kbase + l and kbase + 32 + l
        auto key_ok = [&](int key) -> bool { return key < p.Lk && (km == nullptr || km[key] != 0); };
        bool ok_a = nblk > 0 ? key_ok(lane) : false, ok_b = nblk > 0 ? key_ok(32 + lane) : false;

        for (int j = 0; j < nblk; ++j) {
            const int st = j & 1;
            const int kbase = j * FA_BN;
            uint32_t vb0 = __ballot_sync(0xffffffffu, ok_a), vb1 = __ballot_sync(0xffffffffu, ok_b);
            if (j + 1 < nblk) { ok_a = key_ok(kbase + FA_BN + lane); ok_b = key_ok(kbase + FA_BN + 32 + lane); }
            const bool need_mask = (vb0 & vb1) != 0xffffffffu || (p.causal && kbase + FA_BN - 1 > qpos_warp_min);   // warp-uniform
            if (p.causal) {
                const int nvis = qpos - kbase + 1;
                const uint32_t c0 = nvis >= 32 ? 0xffffffffu : (nvis <= 0 ? 0u : ((1u << nvis) - 1u));
                const uint32_t c1 = nvis >= 64 ? 0xffffffffu : (nvis <= 32 ? 0u : ((1u << (nvis - 32)) - 1u));
                vb0 &= c0; vb1 &= c1;
            }
            mbar_wait(&bar_s[st], (j >> 1) & 1);
            tc_fence_after();
            uint32_t sr[FA_BN];
            tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(st * FA_BN), sr);
            tmem_ld_32x32(tmem_lane + static_cast<uint32_t>(st * FA_BN + 32), sr + 32);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&bar_sfree[st]);
            if (need_mask) {
#pragma unroll
                for (int c = 0; c < FA_BN; ++c)
                    if (!(((c < 32 ? vb0 : vb1) >> (c & 31)) & 1u)) sr[c] = 0xff800000u;     // -inf
            }
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
            for (int c = 0; c < FA_BN; c += 4) {
                mx0 = fmaxf(mx0, __uint_as_float(sr[c])); mx1 = fmaxf(mx1, __uint_as_float(sr[c + 1]));
                mx2 = fmaxf(mx2, __uint_as_float(sr[c + 2])); mx3 = fmaxf(mx3, __uint_as_float(sr[c + 3]));
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * sc;       // scale > 0: max commutes with it
            // ---- lazy rescale decision
            float corr = 1.0f;
            if ((mx > m_used + RESCALE_THRESHOLD || m_used == -INFINITY) && mx != -INFINITY) {
                corr = (m_used == -INFINITY) ? 0.0f : ex2_approx(m_used - mx);
                m_used = mx;
            }
            const float neg_m = (m_used == -INFINITY) ? 0.0f : -m_used;          // all-masked so far: exp2(-inf) = 0 anyway
            float ls0 = 0.0f, ls1 = 0.0f;
            uint32_t pk[FA_BN / 2];
#pragma unroll
            for (int c = 0; c < FA_BN; c += 2) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(sr[c]), sc, neg_m));
                const float p1 = ex2_approx(fmaf(__uint_as_float(sr[c + 1]), sc, neg_m));
                ls0 += p0; ls1 += p1;
                pk[c >> 1] = pack_bf16x2(p0, p1);
            }
            l_run = l_run * corr + (ls0 + ls1);
            // ---- previous P V must have retired before P / O are touched
            if (j > 0) {
                mbar_wait(bar_o, (j - 1) & 1);
                tc_fence_after();
                if (__any_sync(0xffffffffu, corr != 1.0f)) {
#pragma unroll 1
                    for (int c = 0; c < FA_HD / 32; ++c) {
                        uint32_t orr[32];
                        tmem_ld_32x32(tmem_lane + tmem_o + static_cast<uint32_t>(c * 32), orr);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) orr[i] = __float_as_uint(__uint_as_float(orr[i]) * corr);
                        tmem_st_32x32(tmem_lane + tmem_o + static_cast<uint32_t>(c * 32), orr);
                    }
                    tmem_st_wait();
                }
            }
            // ---- P -> smem (K-major, 128B swizzle: 16-byte unit u of row r lives at unit u ^ (r & 7))
            {
                const uint32_t prow = smem_u32(sP) + static_cast<uint32_t>(tid * 128);     // explicit st.shared (not a generic ST.E)
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + static_cast<uint32_t>((u ^ (tid & 7)) << 4)),
                                 "r"(pk[u * 4 + 0]), "r"(pk[u * 4 + 1]), "r"(pk[u * 4 + 2]), "r"(pk[u * 4 + 3])
                                 : "memory");
            }
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(bar_p);
        }

        // ---- epilogue: O / l -> bf16 global, lse
        if (nblk > 0) {
            mbar_wait(bar_o, (nblk - 1) & 1);
            tc_fence_after();
        }
        const float inv = l_run > 0.0f ? 1.0f / l_run : 0.0f;
        __nv_bfloat16* op = p.o + static_cast<long long>(b) * p.o_bs + static_cast<long long>(qrow) * p.o_rs + static_cast<long long>(h) * FA_HD;
#pragma unroll 1
        for (int c = 0; c < FA_HD / 32; ++c) {
            uint32_t orr[32];
            if (nblk > 0) {
                tmem_ld_32x32(tmem_lane + tmem_o + static_cast<uint32_t>(c * 32), orr);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) orr[i] = 0u;
            }
            if (qrow < p.Lq) {
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    uint4 u;
                    u.x = pack_bf16x2(__uint_as_float(orr[g * 8 + 0]) * inv, __uint_as_float(orr[g * 8 + 1]) * inv);
                    u.y = pack_bf16x2(__uint_as_float(orr[g * 8 + 2]) * inv, __uint_as_float(orr[g * 8 + 3]) * inv);
                    u.z = pack_bf16x2(__uint_as_float(orr[g * 8 + 4]) * inv, __uint_as_float(orr[g * 8 + 5]) * inv);
                    u.w = pack_bf16x2(__uint_as_float(orr[g * 8 + 6]) * inv, __uint_as_float(orr[g * 8 + 7]) * inv);
                    *reinterpret_cast<uint4*>(op + c * 32 + g * 8) = u;
                }
            }
        }
        if (p.lse && qrow < p.Lq)
            p.lse[(static_cast<long long>(b) * p.H + h) * p.Lq + qrow] =
                l_run > 0.0f ? (m_used + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace

// q/k/v element (b, l, h, e) at ptr + b*bs + l*rs + h*128 + e.
int attn_fwd_tc128(const void* q, long long q_bs, long long q_rs, const void* k, long long k_bs, long long k_rs, const void* v,
                   long long v_bs, long long v_rs, void* o, long long o_bs, long long o_rs, float* lse, const uint8_t* kmask,
                   long long kmask_ld, int B, int H, int Lq, int Lk, int causal, int q_pos0, float scale, cudaStream_t s) {
    if ((o_rs & 7) || (o_bs & 7) || (reinterpret_cast<uintptr_t>(o) & 15)) {
        set_last_error("attn_fwd: output must keep 16-byte alignment");
        return ESB_ERR_INVALID_ARG;
    }
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM);
        if (e != cudaSuccess) { set_last_error("attn_fwd_tc attr: %s", cudaGetErrorString(e)); return ESB_ERR_CUDA; }
        attr_set = true;
    }
    CUtensorMap tq, tk, tv;
    int rc = make_tma_3d_bf16(&tq, q, static_cast<long long>(H) * FA_HD, Lq, B, q_rs, q_bs, 64, FA_BM);
    if (rc) return rc;
    rc = make_tma_3d_bf16(&tk, k, static_cast<long long>(H) * FA_HD, Lk, B, k_rs, k_bs, 64, FA_BN);
    if (rc) return rc;
    rc = make_tma_3d_bf16(&tv, v, static_cast<long long>(H) * FA_HD, Lk, B, v_rs, v_bs, 64, FA_BN);
    if (rc) return rc;
    FaParams p;
    p.o = static_cast<__nv_bfloat16*>(o); p.lse = lse; p.kmask = kmask; p.o_bs = o_bs; p.o_rs = o_rs; p.kmask_ld = kmask_ld;
    p.H = H; p.Lq = Lq; p.Lk = Lk; p.causal = causal; p.q_pos0 = q_pos0; p.scale_log2 = scale * 1.4426950408889634f;
    dim3 grid((Lq + FA_BM - 1) / FA_BM, H, B);
    attn_fwd_tc_kernel<<<grid, FA_THREADS, FA_SMEM, s>>>(tq, tk, tv, p);
    return check_launch("attn_fwd_tc");
}

}  // namespace esb
