// A CHAIN of skinny bf16 GEMMs in ONE persistent launch - the Linear layers of the single-token decode step between two
// attention calls:      o-proj (+residual)  ->  gate|up (+SwiGLU)  ->  down-proj (+residual)  ->  next layer's q|k|v  (or lm_head)
//
// Reference call sites: HF LlamaDecoderLayer / LlamaMLP / lm_head when generate() feeds ONE new token per sequence against
// the KV cache (pointllm/model/pointllm.py:255-275 -> :173-178, :228; callers model_arch.py:94-108).
//
// Why one launch: with M = 32 tokens every Linear is a weight stream (HBM bound), and as separate launches each of the 129
// GEMMs of a token step pays pipeline fill, split-K tail and drain (~8 us against 8-28 us of streaming,
// profiles/r01_decode_launches_v3.txt).  Here the weight stream never stops:
//   * one persistent CTA per SM with a 10-stage ring (200 KiB of weights in flight per SM); its TMA producer warp runs ahead
//     ACROSS phases - weights do not depend on the previous phase, so while a phase drains (reduction tail, grid-wide hand-off) the ring already holds the next phase's tiles;
//   * stream-K: each phase's (128-row weight tile x 64-column k-block) units are cut into G equal contiguous ranges, so all CTAs
//     stream the same number of bytes whatever N and K are; a tile split over several CTAs is summed in fixed order by the last
//     contributor (fp32 partials through L2, per-tile counters) - deterministic for a given SM count;
//   * the hand-off between phases is a monotonically increasing global counter (release / acquire); only the loads of the
//     ACTIVATION operand wait for it;
//   * RMSNorm is folded as in gemm_skinny.cu (norm weight folded into W at retile time, 1/rms applied to the accumulator from
//     the sums of squares the producing phase left behind); SwiGLU runs in the gate|up epilogue (tiles hold 64 gate rows + the
//     64 matching up rows, esb_retile_weight_gateup_bf16).
#include "common.cuh"
#include "esb_internal.h"
#include "../../include/egoscaler_b200.h"
#include <cstring>
#include <cstdlib>

namespace esb {

namespace {

constexpr int CH_ROWS = 128;
constexpr int CH_BK = 64;
constexpr int CH_THREADS = 192;          // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue
constexpr int CH_MAX_PHASES = 4;
constexpr int CH_W_BYTES = CH_ROWS * CH_BK * 2;
constexpr int CH_COUNTER_BYTES = 16384;  // [0, 12 KiB): per-tile counters; int at byte 12288: the phase hand-off counter

struct ChainPhase {
    const uint8_t* w_tiles;      // [n_tiles][num_kb][128][64] bf16, 128B-swizzled images
    __nv_bfloat16* C;
    long long ldc;
    const __nv_bfloat16* res;
    long long ldr;
    const float* scale_src;      // NULL | rstd[M] (scale_parts == 0) | [scale_parts][64] sums of squares
    float* ssq_out;              // NULL | [n_tiles][64]
    int scale_parts;
    float scale_inv_d, scale_eps;
    int N;                       // output columns (SwiGLU: f)
    int n_tiles, num_kb;
    int swiglu;
    int vec4;                    // C / residual rows allow 8-byte accesses at columns that are multiples of 4
    int total;                   // n_tiles * num_kb
    int g_eff;                   // CTAs that take part in this phase = min(grid, total)
};

struct ChainParams {
    int M, n_phases, grid;
    long long* dbg;              // NULL, or [grid][CH_MAX_PHASES][8] globaltimer stamps (esb_gemm_chain_debug_buffer)
    int pf_dist;                 // weight blocks (16 KiB) requested into L2 ahead of the ring, per CTA
    ChainPhase ph[CH_MAX_PHASES];
    float* partial;              // [n_tiles_max + grid][BT][128] fp32
    int* counters;
    int* grid_ctr;
};

struct alignas(64) ChainMaps {
    CUtensorMap x[CH_MAX_PHASES];
};

template <int BT, int STAGES>
struct ChainSmem {
    static constexpr int X_BYTES = BT * CH_BK * 2;
    static constexpr int STAGE_BYTES = CH_W_BYTES + X_BYTES;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;             // barriers + flags + s_scale[64]
    static constexpr int RED_OFF = BAR_OFF + 1024;                   // s_red [4][64] fp32
    static constexpr int X_OFF = RED_OFF + 1024;                     // s_x [16][64] fp32 (SwiGLU exchange)
    static constexpr int TOTAL = X_OFF + 4096 + 1024;                // + alignment slack
};

ESB_DEVICE void ch_epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
ESB_DEVICE int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
ESB_DEVICE void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
ESB_DEVICE void wait_counter(const int* ctr, int target) {
    while (ld_acquire_gpu(ctr) < target) __nanosleep(32);
}

// CTA c of g_eff owns units [c * total / g_eff, (c + 1) * total / g_eff); the owner of unit b follows from the same formula
ESB_DEVICE int range_start(int c, int total, int g) { return static_cast<int>(static_cast<long long>(c) * total / g); }
ESB_DEVICE int unit_owner(int b, int total, int g) { return static_cast<int>((static_cast<long long>(b + 1) * g - 1) / total); }

ESB_DEVICE long long gtime() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define CH_STAMP(pi, ev) do { if (p.dbg) p.dbg[(static_cast<size_t>(cta) * CH_MAX_PHASES + (pi)) * 8 + (ev)] = gtime(); } while (0)

ESB_DEVICE void l2_prefetch(const void* gsrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}

// Finish 16 tokens (m0 .. m0+15) of weight row `nl` of `tile`: v[] holds the complete fp32 sums.  Called by all 128 epilogue
// threads with the same (tile, m0) - it contains named barriers.
ESB_DEVICE void emit16(const ChainPhase& ph, int M, float (&v)[16], int m0, int tile, int nl, int q, int lane, bool scaled,
                       const float* s_scale, float (*s_red)[64], float (*s_x)[64]) {
    if (!ph.swiglu) {
        const int n = tile * CH_ROWS + nl;
        const bool n_ok = n < ph.N;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int m = m0 + j;
            float sq = 0.0f;
            if (n_ok && m < M) {
                float o = v[j];
                if (scaled) o *= s_scale[m];
                if (ph.res)          // written by an earlier phase / kernel on another SM: read through L2
                    o += __bfloat162float(__ushort_as_bfloat16(__ldcg(reinterpret_cast<const unsigned short*>(ph.res) + static_cast<long long>(m) * ph.ldr + n)));
                const __nv_bfloat16 ob = __float2bfloat16_rn(o);
                ph.C[static_cast<long long>(m) * ph.ldc + n] = ob;
                const float of = __bfloat162float(ob);
                sq = of * of;
            }
            if (ph.ssq_out) {
                sq = warp_sum(sq);
                if (lane == 0) s_red[q][m] = sq;
            }
        }
    } else {
        // rows 0..63 of the tile: gate columns tile*64 + r, rows 64..127: the matching up columns.  act = bf16(bf16(silu(g)) * u)
        // on the bf16-rounded projections, exactly like silu_mul_fwd_kernel (elementwise.cu) after a bf16 gate|up GEMM.
        const int col = tile * 64 + (nl & 63);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            float o = v[j];
            if (scaled) o *= s_scale[min(m0 + j, 63)];
            v[j] = bf16_round(o);
        }
        if (nl >= 64) {
#pragma unroll
            for (int j = 0; j < 16; ++j) s_x[j][nl - 64] = v[j];
        }
        ch_epi_sync();
        if (nl < 64 && col < ph.N) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int m = m0 + j;
                if (m < M) ph.C[static_cast<long long>(m) * ph.ldc + col] = __float2bfloat16_rn(bf16_round(silu(v[j])) * s_x[j][nl]);
            }
        }
        ch_epi_sync();
    }
}

// Finish up to 4 tokens (mm[0..3]; >= M: none) of a tile whose sums were just rebuilt from the partials: this thread holds the four weight
// rows 4*lane .. 4*lane+3 of the tile for each token (one warp covers all 128 rows of a token, so a token's sum of squares is a
// warp reduction).  Shuffles only - warps may call it independently.
ESB_DEVICE void emit_rows(const ChainPhase& ph, int M, const float4 (&acc)[4], const int (&mm)[4], int tile, int lane, bool scaled,
                          const float* s_scale) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int m = mm[r];
        const bool m_ok = m < M;                                   // warp-uniform
        const float sc = (scaled && m_ok) ? s_scale[m] : 1.0f;
        float v[4] = {acc[r].x * sc, acc[r].y * sc, acc[r].z * sc, acc[r].w * sc};
        if (!ph.swiglu) {
            const int n0 = tile * CH_ROWS + 4 * lane;
            float sq = 0.0f;
            if (m_ok && n0 < ph.N) {
                __nv_bfloat16* cp = ph.C + static_cast<long long>(m) * ph.ldc + n0;
                if (ph.vec4 && n0 + 3 < ph.N) {
                    if (ph.res) {
                        const uint2 ru = __ldcg(reinterpret_cast<const uint2*>(ph.res + static_cast<long long>(m) * ph.ldr + n0));
                        const float2 r01 = unpack_bf16x2(ru.x), r23 = unpack_bf16x2(ru.y);
                        v[0] += r01.x; v[1] += r01.y; v[2] += r23.x; v[3] += r23.y;
                    }
                    uint2 ou;
                    ou.x = pack_bf16x2(v[0], v[1]);
                    ou.y = pack_bf16x2(v[2], v[3]);
                    *reinterpret_cast<uint2*>(cp) = ou;
                    const float2 o01 = unpack_bf16x2(ou.x), o23 = unpack_bf16x2(ou.y);
                    sq = o01.x * o01.x + o01.y * o01.y + o23.x * o23.x + o23.y * o23.y;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (n0 + e >= ph.N) break;
                        float o = v[e];
                        if (ph.res)
                            o += __bfloat162float(__ushort_as_bfloat16(__ldcg(reinterpret_cast<const unsigned short*>(ph.res) + static_cast<long long>(m) * ph.ldr + n0 + e)));
                        const __nv_bfloat16 ob = __float2bfloat16_rn(o);
                        cp[e] = ob;
                        const float of = __bfloat162float(ob);
                        sq += of * of;
                    }
                }
            }
            if (ph.ssq_out) {
                sq = warp_sum(sq);
                if (lane == 0 && m_ok) ph.ssq_out[tile * 64 + m] = sq;
            }
        } else {
            // lanes 0..15 hold gate columns tile*64 + 4*lane .., lane + 16 the matching up columns
            float u[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                v[e] = bf16_round(v[e]);
                u[e] = __shfl_xor_sync(0xffffffffu, v[e], 16);
            }
            const int col = tile * 64 + 4 * lane;
            if (m_ok && lane < 16 && col < ph.N) {
                __nv_bfloat16* cp = ph.C + static_cast<long long>(m) * ph.ldc + col;
                float a4[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) a4[e] = bf16_round(silu(v[e])) * u[e];
                if (ph.vec4 && col + 3 < ph.N) {
                    uint2 ou;
                    ou.x = pack_bf16x2(a4[0], a4[1]);
                    ou.y = pack_bf16x2(a4[2], a4[3]);
                    *reinterpret_cast<uint2*>(cp) = ou;
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (col + e < ph.N) cp[e] = __float2bfloat16_rn(a4[e]);
                }
            }
        }
    }
}

template <int BT, int STAGES>
__global__ void __launch_bounds__(CH_THREADS, 1)
gemm_chain_tcgen05_kernel(const __grid_constant__ ChainMaps maps, const __grid_constant__ ChainParams p) {
    using S = ChainSmem<BT, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + STAGES;
    uint64_t* tmem_full = bars + 2 * STAGES;
    uint64_t* tmem_empty = bars + 2 * STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    volatile int* last_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);
    float* s_scale = reinterpret_cast<float*>(smem + S::BAR_OFF + 512);           // [64]
    float (*s_red)[64] = reinterpret_cast<float (*)[64]>(smem + S::RED_OFF);        // [4][64]
    float (*s_x)[64] = reinterpret_cast<float (*)[64]>(smem + S::X_OFF);            // [16][64]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int cta = blockIdx.x;
    griddep_launch();

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < p.n_phases; ++i) tma_prefetch_desc(&maps.x[i]);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 128);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 2 * BT);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ===================== TMA producer.  Weight tiles are requested as soon as a ring stage is free, also across a
            // phase boundary; the activation tile of a stage follows once the phase's input is complete (phase 0: the previous
            // kernel of the stream, griddepcontrol.wait; later phases: every CTA has finished the previous phase).
            // While this thread waits for a phase's input (~8 us of dependent latencies: reduction tail, counter, fences), the
            // next pf_dist weight blocks beyond the ring are requested into L2, so that HBM keeps streaming through the hand-off.
            int stage = 0;
            uint32_t phase = 0;
            for (int pi = 0; pi < p.n_phases; ++pi) {
                const ChainPhase& ph = p.ph[pi];
                int b = 0, end = 0;
                if (cta < ph.g_eff) { b = range_start(cta, ph.total, ph.g_eff); end = range_start(cta + 1, ph.total, ph.g_eff); }
                if (b >= end) continue;
                int tile = b / ph.num_kb, kb = b - tile * ph.num_kb;
                CH_STAMP(pi, 0);
                bool ready = false;
                int n_pend = 0;
                int pend_stage[STAGES], pend_kb[STAGES];
                auto flush = [&]() {
                    for (int i = 0; i < p.pf_dist && b + i < end; ++i)
                        l2_prefetch(ph.w_tiles + static_cast<size_t>(b + i) * CH_W_BYTES, CH_W_BYTES);
                    if (pi == 0) griddep_wait();
                    else { wait_counter(p.grid_ctr, pi * p.grid); fence_proxy_async_all(); }
                    CH_STAMP(pi, 1);
                    for (int i = 0; i < n_pend; ++i)
                        tma_load_2d_hint(smem + pend_stage[i] * S::STAGE_BYTES + CH_W_BYTES, &maps.x[pi], &full_bar[pend_stage[i]],
                                         pend_kb[i] * CH_BK, 0, kEvictLast);
                    n_pend = 0;
                    ready = true;
                };
                for (; b < end; ++b) {
                    if (!ready && n_pend == STAGES) flush();
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    uint8_t* sw = smem + stage * S::STAGE_BYTES;
                    mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                    bulk_load_1d_hint(sw, ph.w_tiles + (static_cast<size_t>(tile) * ph.num_kb + kb) * CH_W_BYTES, CH_W_BYTES,
                                      &full_bar[stage], kEvictFirst);
                    if (ready) tma_load_2d_hint(sw + CH_W_BYTES, &maps.x[pi], &full_bar[stage], kb * CH_BK, 0, kEvictLast);
                    else { pend_stage[n_pend] = stage; pend_kb[n_pend] = kb; ++n_pend; }

                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                    if (++kb == ph.num_kb) { kb = 0; ++tile; }
                }
                if (!ready) flush();
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===================== MMA issuer: D[128 weight rows, BT tokens] += W_block * X_block^T, one accumulator per segment
            constexpr uint32_t idesc = umma_idesc_bf16(CH_ROWS, BT);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int pi = 0; pi < p.n_phases; ++pi) {
                const ChainPhase& ph = p.ph[pi];
                int b = 0, end = 0;
                if (cta < ph.g_eff) { b = range_start(cta, ph.total, ph.g_eff); end = range_start(cta + 1, ph.total, ph.g_eff); }
                bool first = true;
                while (b < end) {
                    const int tile = b / ph.num_kb, kb0 = b - tile * ph.num_kb;
                    const int nkb = min(ph.num_kb - kb0, end - b);
                    mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BT);
                    for (int i = 0; i < nkb; ++i) {
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        if (first) { CH_STAMP(pi, 2); first = false; }
                        const uint32_t sw = smem_u32(smem + stage * S::STAGE_BYTES);
                        const uint64_t da = umma_desc_kmajor_sw128(sw);
                        const uint64_t db = umma_desc_kmajor_sw128(sw + CH_W_BYTES);
#pragma unroll
                        for (int k = 0; k < CH_BK / 16; ++k)
                            umma_bf16_ss(d_tmem, da + static_cast<uint64_t>(2 * k), db + static_cast<uint64_t>(2 * k), idesc,
                                         (i != 0 || k != 0) ? 1u : 0u);
                        umma_commit(&empty_bar[stage]);
                        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                    }
                    umma_commit(&tmem_full[acc]);
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1u;
                    b += nkb;
                }
                if (!first) CH_STAMP(pi, 3);
            }
        }
    } else {
        // ===================== epilogue (4 warps): thread <-> one weight row of the tile, registers <-> tokens
        const int q = warp & 3;
        const int nl = q * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int pi = 0; pi < p.n_phases; ++pi) {
            const ChainPhase& ph = p.ph[pi];
            // inputs of this phase written by other CTAs (residual, sums of squares): acquire the hand-off counter once
            if (pi == 0) griddep_wait();
            else {
                if (nl == 0) wait_counter(p.grid_ctr, pi * p.grid);
                ch_epi_sync();
            }
            if (nl == 0) CH_STAMP(pi, 4);
            int b = 0, end = 0;
            if (cta < ph.g_eff) { b = range_start(cta, ph.total, ph.g_eff); end = range_start(cta + 1, ph.total, ph.g_eff); }
            const bool scaled = ph.scale_src != nullptr;
            if (scaled && b < end) {
                if (nl < 64) {
                    float sc = 1.0f;
                    if (nl < p.M) {
                        if (ph.scale_parts == 0) sc = __ldcg(ph.scale_src + nl);
                        else {
                            float t = 0.0f;
                            for (int j = 0; j < ph.scale_parts; ++j) t += __ldcg(ph.scale_src + j * 64 + nl);   // fixed order
                            sc = rsqrtf(t * ph.scale_inv_d + ph.scale_eps);
                        }
                    }
                    s_scale[nl] = sc;
                }
                ch_epi_sync();
            }
            // split tiles this CTA contributed to: at most its first and its last tile
            int n_def = 0;
            int def_tile[2], def_c0[2], def_n[2];
            while (b < end) {
                const int tile = b / ph.num_kb, kb0 = b - tile * ph.num_kb;
                const int nkb = min(ph.num_kb - kb0, end - b);
                b += nkb;
                const int c0 = unit_owner(tile * ph.num_kb, ph.total, ph.g_eff);
                const int c1 = unit_owner((tile + 1) * ph.num_kb - 1, ph.total, ph.g_eff);
                const int n_contrib = c1 - c0 + 1;
                mbar_wait(&tmem_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BT);
                if (n_contrib == 1) {
#pragma unroll 1
                    for (int c = 0; c < BT / 32; ++c) {
                        if (c * 32 >= p.M) break;
                        uint32_t r[32];
                        tmem_ld_32x32(t_addr + static_cast<uint32_t>(c * 32), r);
                        tmem_ld_wait();
                        if (c == BT / 32 - 1 || (c + 1) * 32 >= p.M) {      // accumulator fully read: hand it back to the MMA warp
                            tc_fence_before();
                            mbar_arrive(&tmem_empty[acc]);
                        }
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            if (c * 32 + hh * 16 >= p.M) break;
                            float v[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[hh * 16 + j]);
                            emit16(ph, p.M, v, c * 32 + hh * 16, tile, nl, q, lane, scaled, s_scale, s_red, s_x);
                        }
                    }
                } else {
                    // slot (tile + cta) is unique: contributors of a tile are consecutive CTAs and c0 never decreases with the tile
                    float* part = p.partial + static_cast<size_t>(tile + cta) * (BT * CH_ROWS) + nl;
#pragma unroll 1
                    for (int c = 0; c < BT / 32; ++c) {
                        if (c * 32 >= p.M) break;
                        uint32_t r[32];
                        tmem_ld_32x32(t_addr + static_cast<uint32_t>(c * 32), r);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (c * 32 + j < p.M) part[(c * 32 + j) * CH_ROWS] = __uint_as_float(r[j]);
                    }
                    tc_fence_before();
                    mbar_arrive(&tmem_empty[acc]);
                    __threadfence();
                    ch_epi_sync();
                    if (nl == 0) atomicAdd(&p.counters[tile], 1);      // arrival; the sums are rebuilt after the last segment (below)
                    def_tile[n_def] = tile; def_c0[n_def] = c0; def_n[n_def] = n_contrib;
                    ++n_def;
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1u;
                    continue;
                }
                if (ph.ssq_out) {                                 // direct path: the tile was finished from TMEM by emit16
                    ch_epi_sync();
                    if (nl < p.M) ph.ssq_out[tile * 64 + nl] = s_red[0][nl] + s_red[1][nl] + s_red[2][nl] + s_red[3][nl];
                    ch_epi_sync();
                }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1u;
            }
            // Every split tile is rebuilt by ALL its contributors, contributor k taking the tokens m = k (mod n): the phase ends
            // with 148 CTAs each summing a few token rows (16-byte loads, fixed order s = 0..n-1) instead of 32..172 CTAs each
            // summing a whole tile (8-28 us of dependent L2 round trips at the tail of every phase, tools/chain_trace.py).
            // counters[tile]: n arrivals, then n completions; the last completion leaves it at zero.
            for (int di = 0; di < n_def; ++di) {
                const int tile = def_tile[di], c0 = def_c0[di], n_contrib = def_n[di];
                if (nl == 0) wait_counter(&p.counters[tile], n_contrib);
                ch_epi_sync();
                const float* base = p.partial + static_cast<size_t>(tile + c0) * (BT * CH_ROWS) + 4 * lane;
                const int my = cta - c0;
#pragma unroll 1
                for (int i0 = 4 * q; ; i0 += 16) {                      // warp q: the share's tokens 4q .. 4q+3, 4q+16 .., ...
                    int mm[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int m = my + n_contrib * (i0 + r);
                        mm[r] = m < p.M ? m : p.M;
                    }
                    if (mm[0] >= p.M) break;
                    float4 a4[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) a4[r] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll 1
                    for (int s2 = 0; s2 < n_contrib; s2 += 4) {
                        float4 t[4][4];
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int r = 0; r < 4; ++r)
                                t[i][r] = (s2 + i < n_contrib && mm[r] < p.M)
                                              ? __ldcg(reinterpret_cast<const float4*>(base + (static_cast<size_t>(s2 + i) * BT + mm[r]) * CH_ROWS))
                                              : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int r = 0; r < 4; ++r) {
                                a4[r].x += t[i][r].x; a4[r].y += t[i][r].y; a4[r].z += t[i][r].z; a4[r].w += t[i][r].w;
                            }
                    }
                    emit_rows(ph, p.M, a4, mm, tile, lane, scaled, s_scale);
                }
                ch_epi_sync();                                        // all loads of the share are done
                if (nl == 0) {
                    const int prev = atomicAdd(&p.counters[tile], 1);
                    if (prev == 2 * n_contrib - 1) p.counters[tile] = 0;      // every contributor has arrived and finished
                }
            }
            // hand-off: everything this CTA wrote in the phase is visible before the counter moves
            if (nl == 0) CH_STAMP(pi, 5);
            __threadfence();
            fence_proxy_async_all();                              // the next phase reads these rows through TMA (async proxy)
            ch_epi_sync();
            if (nl == 0) {
                const int prev = atomicAdd(p.grid_ctr, 1);
                if (prev + 1 == p.n_phases * p.grid) *p.grid_ctr = 0;      // last arrival of the launch: nobody polls any more
                CH_STAMP(pi, 6);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * BT);
    }
}

// CTAs of this instantiation the runtime keeps resident per SM (0: the kernel cannot launch).  The phase hand-off needs every CTA
// of the grid resident, so the grid is sized from this number.  Measured on B200: the occupancy query answers 1 for this kernel
// whatever its shared-memory / register footprint (80 KB and 124 registers included; the same query on gemm_skinny's <32, 4>
// kernel also answers 1), hence ONE CTA per SM with a ring deep enough to hold the bytes two CTAs would have had in flight.
template <int BT, int STAGES>
int chain_ctas_per_sm() {
    using S = ChainSmem<BT, STAGES>;
    static int per_sm = -1;
    if (per_sm < 0) {
        auto kernel = gemm_chain_tcgen05_kernel<BT, STAGES>;
        int n = 0;
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, CH_THREADS, S::TOTAL);
        if (e != cudaSuccess) {
            set_last_error("gemm_chain<%d,%d>: %s", BT, STAGES, cudaGetErrorString(e));
            (void)cudaGetLastError();
            n = 0;
        }
        per_sm = n;
    }
    return per_sm;
}

template <int BT, int STAGES>
int launch_chain_kernel(const ChainMaps& maps, const ChainParams& p, bool pdl, cudaStream_t stream) {
    using S = ChainSmem<BT, STAGES>;
    auto kernel = gemm_chain_tcgen05_kernel<BT, STAGES>;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.grid);
    cfg.blockDim = dim3(CH_THREADS);
    cfg.dynamicSmemBytes = S::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, maps, p);
    if (e != cudaSuccess) {
        set_last_error("gemm_chain launch: %s", cudaGetErrorString(e));
        return ESB_ERR_CUDA;
    }
    return check_launch("gemm_chain_tcgen05");
}

constexpr int CH_STAGES_32 = 10;      // 10 x 20 KiB
constexpr int CH_STAGES_64 = 8;       //  8 x 24 KiB

}  // namespace

static long long* g_chain_dbg = nullptr;
int gemm_chain_debug_buffer(void* buf) { g_chain_dbg = static_cast<long long*>(buf); return ESB_OK; }

int gemm_chain_ctas_per_sm(int M) { return M <= 32 ? chain_ctas_per_sm<32, CH_STAGES_32>() : chain_ctas_per_sm<64, CH_STAGES_64>(); }

long long gemm_chain_workspace_bytes(int M, int max_N) {
    const int bt = M <= 32 ? 32 : 64;
    const long long slots = (max_N + 63) / 64 + 2 * device_sm_count();      // grid <= 2 CTAs per SM
    return CH_COUNTER_BYTES + slots * bt * CH_ROWS * 4;
}

int gemm_chain_bf16(const esb_chain_phase* phases, int n_phases, int M, void* workspace, long long workspace_bytes, int flags,
                    cudaStream_t stream) {
    if (n_phases <= 0 || M <= 0) return ESB_OK;
    if (n_phases > CH_MAX_PHASES || M > 64 || !phases || !workspace) {
        set_last_error("gemm_chain_bf16: bad arguments (n_phases=%d (max %d) M=%d (max 64))", n_phases, CH_MAX_PHASES, M);
        return ESB_ERR_INVALID_ARG;
    }
    const int bt = M <= 32 ? 32 : 64;
    const int per_sm = gemm_chain_ctas_per_sm(M);
    if (per_sm < 1) return ESB_ERR_CUDA;
    const int grid = device_sm_count() * (per_sm < 2 ? per_sm : 2);
    ChainMaps maps;
    ChainParams p;
    memset(&maps, 0, sizeof(maps));
    memset(&p, 0, sizeof(p));
    p.M = M; p.n_phases = n_phases; p.grid = grid;
    static int pf_dist = -1;
    if (pf_dist < 0) {
        const char* e = getenv("ESB_CHAIN_PREFETCH");      // tuning aid; default measured with tools/chain_bench.py
        pf_dist = e ? atoi(e) : 0;           // measured: 109 / 111 / 112 / 116 us per layer at 0 / 8 / 16 / 32 blocks - off
        if (pf_dist < 0) pf_dist = 0;
        if (pf_dist > 256) pf_dist = 256;
    }
    p.pf_dist = pf_dist;
    p.dbg = g_chain_dbg;
    long long max_tiles = 0;
    for (int i = 0; i < n_phases; ++i) {
        const esb_chain_phase& s = phases[i];
        ChainPhase& d = p.ph[i];
        if (!s.X || !s.W_tiles || !s.C || s.N <= 0 || s.K <= 0 || (s.ldx & 7) || (reinterpret_cast<uintptr_t>(s.X) & 15) ||
            (reinterpret_cast<uintptr_t>(s.W_tiles) & 15) || (s.swiglu && ((s.N & 1) || s.residual || s.ssq_out)) ||
            s.scale_parts < 0 || s.scale_parts > 1024) {
            set_last_error("gemm_chain_bf16: phase %d has bad arguments (N=%d K=%d ldx=%lld)", i, s.N, s.K, static_cast<long long>(s.ldx));
            return ESB_ERR_INVALID_ARG;
        }
        d.w_tiles = reinterpret_cast<const uint8_t*>(s.W_tiles);
        d.C = reinterpret_cast<__nv_bfloat16*>(s.C); d.ldc = s.ldc;
        d.res = reinterpret_cast<const __nv_bfloat16*>(s.residual); d.ldr = s.ldr;
        d.scale_src = s.scale_src; d.ssq_out = s.ssq_out; d.scale_parts = s.scale_parts;
        d.scale_inv_d = 1.0f / static_cast<float>(s.K); d.scale_eps = s.eps;
        d.swiglu = s.swiglu ? 1 : 0;
        d.vec4 = ((s.ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(s.C) & 7) == 0 &&
                  (!s.residual || ((s.ldr & 3) == 0 && (reinterpret_cast<uintptr_t>(s.residual) & 7) == 0))) ? 1 : 0;
        d.N = s.swiglu ? s.N / 2 : s.N;
        d.n_tiles = s.swiglu ? (d.N + 63) / 64 : (d.N + CH_ROWS - 1) / CH_ROWS;
        d.num_kb = (s.K + CH_BK - 1) / CH_BK;
        const long long total = static_cast<long long>(d.n_tiles) * d.num_kb;
        if (total > (1ll << 30) || d.n_tiles * 4 > 12288) {
            set_last_error("gemm_chain_bf16: phase %d too large (N=%d K=%d)", i, s.N, s.K);
            return ESB_ERR_UNSUPPORTED;
        }
        d.total = static_cast<int>(total);
        d.g_eff = d.total < grid ? d.total : grid;
        if (d.n_tiles > max_tiles) max_tiles = d.n_tiles;
        const int rc = make_tma_2d_bf16(&maps.x[i], s.X, M, s.K, s.ldx, bt);
        if (rc) return rc;
    }
    const long long need = CH_COUNTER_BYTES + (max_tiles + grid) * bt * CH_ROWS * 4;
    if (workspace_bytes < need) {
        set_last_error("gemm_chain_bf16: workspace of %lld bytes, %lld needed", workspace_bytes, need);
        return ESB_ERR_WORKSPACE;
    }
    p.counters = reinterpret_cast<int*>(workspace);
    p.grid_ctr = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(workspace) + 12288);
    p.partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + CH_COUNTER_BYTES);
    const bool pdl = (flags & 2) != 0;            // caller promises no weight image is written by the kernel launched just before
    return bt == 32 ? launch_chain_kernel<32, CH_STAGES_32>(maps, p, pdl, stream) : launch_chain_kernel<64, CH_STAGES_64>(maps, p, pdl, stream);
}

}  // namespace esb
