// bf16 GEMM on the 5th-generation tensor cores:  C[M,N] = epilogue(A[M,K] * B[N,K]^T)
//
// This is the contraction every Linear on the trajectory-VLM hot path reduces to (reference call sites:
// HF LlamaAttention/LlamaMLP reached from pointllm/model/pointllm.py:173-178, lm_head pointllm.py:228,
// point_proj pointllm.py:67-81,129, PointBERT Linear/Conv1d(k=1) layers point_encoder.py:11-55, dvae.py:194-205).
// Weights are stored [out_features, in_features] exactly like the reference state_dict, i.e. B is K-major.
//
// Structure (persistent, warp-specialised, one CTA per SM):
//   warp 0 lane 0 : TMA producer   (cp.async.bulk.tensor 2D, 128B swizzle, mbarrier complete_tx)
//   warp 1 lane 0 : MMA issuer     (tcgen05.mma cta_group::1 kind::f16, 128 x BN x 16, fp32 accum in TMEM)
//   warp 2        : TMEM allocator (2 accumulator stages = 2*BN columns)
//   warps 4..7    : epilogue       (tcgen05.ld 32x32b -> bias / activation / residual -> bf16|fp32 global)
// Pipelines: smem ring full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue), static tile scheduler.
// Ragged M/N/K edges rely on TMA zero fill for loads and on masked stores in the epilogue.
#include <cstdlib>
#include "common.cuh"
#include "esb_internal.h"

#include <cudaTypedefs.h>
#include <mutex>

namespace esb {

constexpr int BM = 128;
constexpr int BK = 64;   // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 256;

struct GemmParams {
    int M, N, K;
    void* C;
    long long ldc;
    const float* bias;        // [N] or null
    const __nv_bfloat16* res; // [M, ldr] or null
    long long ldr;
    int act;                  // 0 none, 1 gelu(erf), 2 relu
    int out_fp32;             // 0 -> bf16 out, 1 -> fp32 out
    int accumulate;           // fp32 out only: C += result
    float alpha;              // scales the product before bias
    int tiles_m, tiles_n;
    // Tile walk + L2 policy, chosen per launch on the host (launch_gemm) to keep DRAM traffic near A + B + C:
    //   raster_mode 0: M-groups of `group` row tiles; inside a group tiles are walked M-fastest across all of N, so the group's
    //                  A panels (group * BM * K * 2 bytes, sized to fit L2) are re-read from L2 by every wave while each
    //                  B panel is used by one wave only;  A is loaded evict_last, B evict_normal.
    //   raster_mode 1: the mirror image - N-groups of `group` column panels walked N-fastest across all of M; the group's B
    //                  panels persist in L2 (evict_last) and A streams through.
    // C larger than L2 is written with streaming stores so that the output does not push the persisting operand out.
    int raster_mode, group;
    uint64_t a_policy, b_policy;
    int c_stream;
    // Fused epilogues (template parameter EPI; VERDICT r1 next #5 - the HBM-bound kernels that ran behind the GEMMs):
    //   EPI_SWIGLU  : B = packed gate|up weight [2f, K]; a tile multiplies 128 gate rows AND the matching 128 up rows (two TMA
    //                 boxes per stage), the epilogue writes act = silu(gate) * up [M, f] and, when C2 != null, gate|up [M, 2f]
    //                 (kept for the backward) - replaces silu_mul_fwd (HF LlamaMLP, modeling_llama.py:171-184).
    //   EPI_DSWIGLU : the d(act) = dOut . Wdown GEMM; the epilogue reads gate|up (aux) and writes d(gate|up) to C2 [M, 2f]
    //                 without materialising d(act) - replaces silu_mul_bwd.
    //   EPI_ROPE    : the q|k|v projection; heads of 128 columns below rope_cols get apply_rotary_pos_emb
    //                 (modeling_llama.py:138-168) with the same bf16 rounding order as rope_kernel - replaces rope.
    void* C2;
    long long ldc2;
    const __nv_bfloat16* aux;
    long long ld_aux;
    int f;
    const __nv_bfloat16* cos_tab;
    const __nv_bfloat16* sin_tab;
    int rope_L, rope_pos0, rope_cols;
    int dbg;                  // CTA-pair kernel experiments (ESB_GEMM2_DBG): bit 0 peer also arrives on full[s], bit 1 spin on test_wait, bit 2 per-thread tempty arrives
};
enum : int { EPI_PLAIN = 0, EPI_SWIGLU = 1, EPI_DSWIGLU = 2, EPI_ROPE = 3 };
static_assert(EPI_DSWIGLU == 2, "GemmSmem keys its staging buffers on this value");

ESB_DEVICE void tile_coords(const GemmParams& p, int t, int& tm, int& tn) {
    if (p.raster_mode == 0) {
        const int per_group = p.group * p.tiles_n;
        const int g = t / per_group, r = t - g * per_group;
        const int base = g * p.group;
        const int gm = min(p.group, p.tiles_m - base);
        tn = r / gm;
        tm = base + (r - tn * gm);
    } else {
        const int per_group = p.group * p.tiles_m;
        const int g = t / per_group, r = t - g * per_group;
        const int base = g * p.group;
        const int gn = min(p.group, p.tiles_n - base);
        tm = r / gn;
        tn = base + (r - tm * gn);
    }
}

constexpr int EPI_DSWIGLU_ID = 2;
template <int BN, int EPI = 0>
struct GemmSmem {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    // the dSwiGLU epilogue stages the saved gate|up tile through shared memory (two buffers of [128 rows x 64 cols] gate + up =
    // 32 KB each, TMA-loaded by warp 3): one operand stage fewer pays for them
    static constexpr int STG_BUF_BYTES = 2 * BM * 64 * 2;
    static constexpr int STG_BYTES = (EPI == EPI_DSWIGLU_ID) ? 2 * STG_BUF_BYTES : 0;
    static constexpr int STAGES = (EPI == EPI_DSWIGLU_ID) ? 3 : ((BN == 256) ? 4 : (BN == 128 ? 6 : 8));
    static constexpr int BAR_BYTES = 1024;
    static constexpr int TOTAL = STAGES * STAGE_BYTES + STG_BYTES + BAR_BYTES + 1024;   // +1024 alignment slack
};

// bf16 stores of 32 consecutive columns of one row (vector path when aligned and complete)
ESB_DEVICE void store_row32_bf16(__nv_bfloat16* cp, const uint32_t* packed, int ncols, bool vec_ok, int stream) {
    if (vec_ok && ncols == 32) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint4 u = make_uint4(packed[g * 4 + 0], packed[g * 4 + 1], packed[g * 4 + 2], packed[g * 4 + 3]);
            if (stream) __stcs(reinterpret_cast<uint4*>(cp) + g, u);
            else reinterpret_cast<uint4*>(cp)[g] = u;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) {
                const uint32_t w = packed[j >> 1];
                const unsigned short h = (j & 1) ? static_cast<unsigned short>(w >> 16) : static_cast<unsigned short>(w & 0xffffu);
                cp[j] = __ushort_as_bfloat16(h);
            }
    }
}
// 32 consecutive bf16 columns of one row -> 16 packed words
ESB_DEVICE void load_row32_bf16(const __nv_bfloat16* rp, uint32_t* packed, int ncols, bool vec_ok) {
    if (vec_ok && ncols == 32) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint4 u = __ldg(reinterpret_cast<const uint4*>(rp) + g);
            packed[g * 4 + 0] = u.x; packed[g * 4 + 1] = u.y; packed[g * 4 + 2] = u.z; packed[g * 4 + 3] = u.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const unsigned short lo = (2 * j < ncols) ? __bfloat16_as_ushort(rp[2 * j]) : 0;
            const unsigned short hi = (2 * j + 1 < ncols) ? __bfloat16_as_ushort(rp[2 * j + 1]) : 0;
            packed[j] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
        }
    }
}

// One accumulator tile: TMEM -> registers -> (epilogue arithmetic) -> global.  Called by the four epilogue warps with
// t_addr = TMEM address of this warp's lane quarter of the accumulator stage, row = the global row this thread owns.
template <int BN, int EPI>
ESB_DEVICE void epilogue_tile(const GemmParams& p, const uint32_t t_addr, const int row, const bool row_ok, const int n0) {
    const bool vec_ok_bf16 = (!p.out_fp32) && ((p.ldc & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
    const bool vec_ok_f32 = (p.out_fp32) && ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
    const bool res_vec_ok = p.res && ((p.ldr & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.res) & 15) == 0);
    if constexpr (EPI == EPI_SWIGLU) {
        // act = bf16(bf16(silu(g)) * u) exactly like silu_mul_fwd_kernel on the bf16-rounded projections
        __nv_bfloat16* act = reinterpret_cast<__nv_bfloat16*>(p.C);
        __nv_bfloat16* gu = reinterpret_cast<__nv_bfloat16*>(p.C2);
        const bool v_act = ((p.ldc & 7) == 0) && ((reinterpret_cast<uintptr_t>(act) & 15) == 0);
        const bool v_gu = gu && ((p.ldc2 & 7) == 0) && ((p.f & 7) == 0) && ((reinterpret_cast<uintptr_t>(gu) & 15) == 0);
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            const int col0 = n0 + c * 32;
            if (col0 >= p.f) break;
            uint32_t rg[32], ru[32];
            tmem_ld_32x32(t_addr + static_cast<uint32_t>(c * 32), rg);
            tmem_ld_32x32(t_addr + static_cast<uint32_t>(128 + c * 32), ru);
            tmem_ld_wait();
            const int ncols = min(32, p.f - col0);
            uint32_t pg[16], pu[16], pa[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                pg[j] = pack_bf16x2(__uint_as_float(rg[2 * j]), __uint_as_float(rg[2 * j + 1]));
                pu[j] = pack_bf16x2(__uint_as_float(ru[2 * j]), __uint_as_float(ru[2 * j + 1]));
                const float2 gv = unpack_bf16x2(pg[j]), uv = unpack_bf16x2(pu[j]);
                pa[j] = pack_bf16x2(bf16_round(silu(gv.x)) * uv.x, bf16_round(silu(gv.y)) * uv.y);
            }
            if (row_ok) {
                store_row32_bf16(act + static_cast<long long>(row) * p.ldc + col0, pa, ncols, v_act, 0);
                if (gu) {
                    store_row32_bf16(gu + static_cast<long long>(row) * p.ldc2 + col0, pg, ncols, v_gu, p.c_stream);
                    store_row32_bf16(gu + static_cast<long long>(row) * p.ldc2 + p.f + col0, pu, ncols, v_gu, p.c_stream);
                }
            }
        }
    } else if constexpr (EPI == EPI_DSWIGLU) {
        // d(gate|up) from d(act) (this tile, rounded to bf16 like the materialised tensor was) and the saved gate|up:
        // the arithmetic of silu_mul_bwd_kernel
        const __nv_bfloat16* gu = p.aux;
        __nv_bfloat16* dgu = reinterpret_cast<__nv_bfloat16*>(p.C2);
        const bool v_in = ((p.ld_aux & 7) == 0) && ((p.f & 7) == 0) && ((reinterpret_cast<uintptr_t>(gu) & 15) == 0);
        const bool v_out = ((p.ldc2 & 7) == 0) && ((p.f & 7) == 0) && ((reinterpret_cast<uintptr_t>(dgu) & 15) == 0);
        // The gate|up rows of a warp lie 2f elements apart (32 different lines per load instruction) and the L1 next to 194 KB
        // of shared memory is tiny, so every thread reads WHOLE 128-byte lines (64 columns of gate and of up per group) exactly
        // once and the next group's loads are in flight while this group's arithmetic runs (two register buffers).  History:
        // per-32-column loads without prefetch ran the GEMM at 810 TF/s; prefetching half lines two chunks ahead thrashed the L1
        // (404 TF/s).
        constexpr int NGRP = BN / 64;
        uint32_t bg[2][32], bu[2][32];
        const __nv_bfloat16* grow = gu + static_cast<long long>(row_ok ? row : 0) * p.ld_aux;
        auto issue = [&](int g, uint32_t* pg_, uint32_t* pu_) {
            const int col0 = n0 + g * 64;
            if (row_ok && col0 < p.N) {
                const int n64 = min(64, p.N - col0);
                load_row32_bf16(grow + col0, pg_, min(32, n64), v_in);
                load_row32_bf16(grow + p.f + col0, pu_, min(32, n64), v_in);
                if (n64 > 32) {
                    load_row32_bf16(grow + col0 + 32, pg_ + 16, n64 - 32, v_in);
                    load_row32_bf16(grow + p.f + col0 + 32, pu_ + 16, n64 - 32, v_in);
                }
            }
        };
        issue(0, bg[0], bu[0]);
#pragma unroll
        for (int g = 0; g < NGRP; ++g) {
            if (g + 1 < NGRP) issue(g + 1, bg[(g + 1) & 1], bu[(g + 1) & 1]);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int col0 = n0 + g * 64 + h * 32;
                if (col0 < p.N) {                                        // warp-uniform
                    uint32_t r[32];
                    tmem_ld_32x32(t_addr + static_cast<uint32_t>(g * 64 + h * 32), r);
                    tmem_ld_wait();
                    const int ncols = min(32, p.N - col0);
                    if (row_ok) {
                        const uint32_t* qg = bg[g & 1] + h * 16;
                        const uint32_t* qu = bu[g & 1] + h * 16;
                        uint32_t og[16], ou[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float2 gv = unpack_bf16x2(qg[j]), uv = unpack_bf16x2(qu[j]);
                            const float2 dv = unpack_bf16x2(pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1])));
                            const float s0 = 1.0f / (1.0f + __expf(-gv.x)), s1 = 1.0f / (1.0f + __expf(-gv.y));
                            og[j] = pack_bf16x2(dv.x * uv.x * s0 * (1.0f + gv.x * (1.0f - s0)), dv.y * uv.y * s1 * (1.0f + gv.y * (1.0f - s1)));
                            ou[j] = pack_bf16x2(dv.x * gv.x * s0, dv.y * gv.y * s1);
                        }
                        store_row32_bf16(dgu + static_cast<long long>(row) * p.ldc2 + col0, og, ncols, v_out, p.c_stream);
                        store_row32_bf16(dgu + static_cast<long long>(row) * p.ldc2 + p.f + col0, ou, ncols, v_out, p.c_stream);
                    }
                }
            }
        }
    } else if constexpr (EPI == EPI_ROPE) {
        // two heads of 128 columns per tile; column j pairs with j + 64 (rotate_half)
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.C);
        const bool v_out = ((p.ldc & 7) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
        const int pos = p.rope_pos0 + (row_ok ? row % p.rope_L : 0);
        const __nv_bfloat16* cp_ = p.cos_tab + static_cast<long long>(pos) * 64;
        const __nv_bfloat16* sp_ = p.sin_tab + static_cast<long long>(pos) * 64;
#pragma unroll 1
        for (int hc = 0; hc < 4; ++hc) {                     // (head in tile, 32-column chunk of the first half)
            const int h = hc >> 1, c = hc & 1;
            const int head0 = n0 + h * 128;
            if (head0 >= p.N) break;
            uint32_t lo[32], hi[32];
            tmem_ld_32x32(t_addr + static_cast<uint32_t>(h * 128 + c * 32), lo);
            tmem_ld_32x32(t_addr + static_cast<uint32_t>(h * 128 + 64 + c * 32), hi);
            tmem_ld_wait();
            uint32_t o1[16], o2[16];
            const bool rot = head0 < p.rope_cols;
            uint32_t cw[16], sw[16];
            if (rot && row_ok) {
                load_row32_bf16(cp_ + c * 32, cw, 32, true);
                load_row32_bf16(sp_ + c * 32, sw, 32, true);
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                o1[j] = pack_bf16x2(__uint_as_float(lo[2 * j]), __uint_as_float(lo[2 * j + 1]));
                o2[j] = pack_bf16x2(__uint_as_float(hi[2 * j]), __uint_as_float(hi[2 * j + 1]));
                if (rot && row_ok) {
                    const float2 x1 = unpack_bf16x2(o1[j]), x2 = unpack_bf16x2(o2[j]);
                    const float2 cc = unpack_bf16x2(cw[j]), ss = unpack_bf16x2(sw[j]);
                    o1[j] = pack_bf16x2(bf16_round(x1.x * cc.x) - bf16_round(x2.x * ss.x), bf16_round(x1.y * cc.y) - bf16_round(x2.y * ss.y));
                    o2[j] = pack_bf16x2(bf16_round(x2.x * cc.x) + bf16_round(x1.x * ss.x), bf16_round(x2.y * cc.y) + bf16_round(x1.y * ss.y));
                }
            }
            if (row_ok) {
                __nv_bfloat16* rp = out + static_cast<long long>(row) * p.ldc + head0 + c * 32;
                store_row32_bf16(rp, o1, 32, v_out, 0);
                store_row32_bf16(rp + 64, o2, 32, v_out, 0);
            }
        }
    } else {
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;                    // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(t_addr + static_cast<uint32_t>(c * 32), r);
        tmem_ld_wait();
        const int ncols = min(32, p.N - col0);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
        if (p.bias) {
            if (ncols == 32) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + col0 + j);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < ncols) v[j] += __ldg(p.bias + col0 + j);
            }
        }
        if (p.act == 1) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
        } else if (p.act == 2) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
        }
        if (row_ok) {
            if (p.res) {
                const __nv_bfloat16* rp = p.res + static_cast<long long>(row) * p.ldr + col0;
                if (res_vec_ok && ncols == 32) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const uint4 u = __ldg(reinterpret_cast<const uint4*>(rp) + g);
                        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z),
                                     d = unpack_bf16x2(u.w);
                        v[g * 8 + 0] += a.x; v[g * 8 + 1] += a.y; v[g * 8 + 2] += b.x; v[g * 8 + 3] += b.y;
                        v[g * 8 + 4] += cc.x; v[g * 8 + 5] += cc.y; v[g * 8 + 6] += d.x; v[g * 8 + 7] += d.y;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < ncols) v[j] += __bfloat162float(rp[j]);
                }
            }
            if (!p.out_fp32) {
                __nv_bfloat16* cp = reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
                if (vec_ok_bf16 && ncols == 32) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        uint4 u;
                        u.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
                        u.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
                        u.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
                        u.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
                        if (p.c_stream) __stcs(reinterpret_cast<uint4*>(cp) + g, u);
                        else reinterpret_cast<uint4*>(cp)[g] = u;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < ncols) cp[j] = __float2bfloat16_rn(v[j]);
                }
            } else {
                float* cp = reinterpret_cast<float*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
                if (vec_ok_f32 && ncols == 32) {
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        float4 u = make_float4(v[g * 4 + 0], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
                        if (p.accumulate) {
                            const float4 o = reinterpret_cast<float4*>(cp)[g];
                            u.x += o.x; u.y += o.y; u.z += o.z; u.w += o.w;
                        }
                        reinterpret_cast<float4*>(cp)[g] = u;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < ncols) cp[j] = p.accumulate ? cp[j] + v[j] : v[j];
                }
            }
        }
    }
    }   // EPI_PLAIN
}

// dSwiGLU epilogue with the gate|up tile staged through shared memory: warp 3 TMA-loads [128 rows x 64 cols] of gate and of up
// (128B swizzle) two groups ahead; each epilogue thread reads ITS row (128 bytes per stream, 16-byte units XOR-swizzled by row & 7)
// instead of issuing 32-line global loads.  Arithmetic identical to silu_mul_bwd_kernel / epilogue_tile<EPI_DSWIGLU>.
template <int BN>
ESB_DEVICE void epilogue_dswiglu_staged(const GemmParams& p, const uint32_t t_addr, const int row, const bool row_ok, const int n0,
                                        const int r_in_tile, uint8_t* stg, uint64_t* gu_full, uint64_t* gu_empty, uint32_t& gcount,
                                        const int lane) {
    constexpr int STG_BUF = 2 * BM * 64 * 2;
    __nv_bfloat16* dgu = reinterpret_cast<__nv_bfloat16*>(p.C2);
    const bool v_out = ((p.ldc2 & 7) == 0) && ((p.f & 7) == 0) && ((reinterpret_cast<uintptr_t>(dgu) & 15) == 0);
#pragma unroll 1
    for (int g = 0; g < BN / 64; ++g, ++gcount) {
        const int buf = static_cast<int>(gcount & 1u);
        mbar_wait(&gu_full[buf], (gcount >> 1) & 1u);
        const uint8_t* sg = stg + buf * STG_BUF + r_in_tile * 128;
        const uint8_t* su = sg + BM * 128;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int col0 = n0 + g * 64 + h * 32;
            if (col0 < p.N) {                                            // warp-uniform
                uint32_t r[32];
                tmem_ld_32x32(t_addr + static_cast<uint32_t>(g * 64 + h * 32), r);
                tmem_ld_wait();
                const int ncols = min(32, p.N - col0);
                uint32_t qg[16], qu[16], og[16], ou[16];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int unit = ((h * 4 + u) ^ (r_in_tile & 7)) << 4;
                    const uint4 a = *reinterpret_cast<const uint4*>(sg + unit);
                    const uint4 b = *reinterpret_cast<const uint4*>(su + unit);
                    qg[u * 4 + 0] = a.x; qg[u * 4 + 1] = a.y; qg[u * 4 + 2] = a.z; qg[u * 4 + 3] = a.w;
                    qu[u * 4 + 0] = b.x; qu[u * 4 + 1] = b.y; qu[u * 4 + 2] = b.z; qu[u * 4 + 3] = b.w;
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 gv = unpack_bf16x2(qg[j]), uv = unpack_bf16x2(qu[j]);
                    const float2 dv = unpack_bf16x2(pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1])));
                    const float s0 = 1.0f / (1.0f + __expf(-gv.x)), s1 = 1.0f / (1.0f + __expf(-gv.y));
                    og[j] = pack_bf16x2(dv.x * uv.x * s0 * (1.0f + gv.x * (1.0f - s0)), dv.y * uv.y * s1 * (1.0f + gv.y * (1.0f - s1)));
                    ou[j] = pack_bf16x2(dv.x * gv.x * s0, dv.y * gv.y * s1);
                }
                if (row_ok) {
                    store_row32_bf16(dgu + static_cast<long long>(row) * p.ldc2 + col0, og, ncols, v_out, p.c_stream);
                    store_row32_bf16(dgu + static_cast<long long>(row) * p.ldc2 + p.f + col0, ou, ncols, v_out, p.c_stream);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&gu_empty[buf]);                      // this warp is done with the staging buffer
    }
}

// TA / TB = true: that operand is stored with the CONTRACTION index outermost (A [Kc, M] / B [Kc, N] row-major) instead of
// innermost.  TA & TB: C = A^T . B, the weight-gradient contraction dW = dY^T . X on the activations exactly as they sit in HBM;
// TB only: C = A . B with B = a weight [in = Kc, out = N]... i.e. the data-gradient dX = dY . W on the state_dict layout - no
// transposed copies either way.  Such an operand is staged as [64 contraction rows x 64 columns] TMA boxes (128-byte rows, 128B
// swizzle) = an MN-major UMMA operand: the 64-wide column chunks of a tile lie 8 KB apart (LBO), 16 contraction rows (one MMA)
// are 2 KB apart.
template <int BN, int EPI, bool TA = false, bool TB = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c, const GemmParams p) {
    static_assert(EPI == EPI_PLAIN || BN == 256, "fused epilogues are built on the 128x256 tile");
    static_assert(!TA || EPI == EPI_PLAIN, "a transposed A operand comes with the plain epilogue only");
    static_assert(!TB || EPI == EPI_PLAIN || EPI == EPI_DSWIGLU, "a transposed B operand: plain or dSwiGLU epilogue");
    constexpr int BN_OUT = (EPI == EPI_SWIGLU) ? 128 : BN;      // output columns per tile
    using S = GemmSmem<BN, EPI>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* tiles = smem;
    uint8_t* stg = smem + S::STAGES * S::STAGE_BYTES;                                  // dSwiGLU gate|up staging (else empty)
    uint64_t* bars = reinterpret_cast<uint64_t*>(stg + S::STG_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + S::STAGES;
    uint64_t* tmem_full = bars + 2 * S::STAGES;
    uint64_t* tmem_empty = bars + 2 * S::STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S::STAGES + 4);
    uint64_t* gu_full = bars + 2 * S::STAGES + 6;
    uint64_t* gu_empty = bars + 2 * S::STAGES + 8;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int num_kb = (p.K + BK - 1) / BK;
    const int num_tiles = p.tiles_m * p.tiles_n;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        if constexpr (EPI == EPI_DSWIGLU) tma_prefetch_desc(&tma_c);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < S::STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 128);
            mbar_init(&gu_full[i], 1);
            mbar_init(&gu_empty[i], 4);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_slot, 2 * BN);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ===================== TMA producer
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                int tm, tn;
                tile_coords(p, t, tm, tn);
                const int m0 = tm * BM, n0 = tn * BN_OUT;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    uint8_t* sa = tiles + stage * S::STAGE_BYTES;
                    uint8_t* sb = sa + S::A_BYTES;
                    mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                    if constexpr (TA) {
#pragma unroll
                        for (int i = 0; i < BM / 64; ++i)
                            tma_load_2d_hint(sa + i * 8192, &tma_a, &full_bar[stage], m0 + i * 64, kb * BK, p.a_policy);
                    } else {
                        tma_load_2d_hint(sa, &tma_a, &full_bar[stage], kb * BK, m0, p.a_policy);
                    }
                    if constexpr (TB) {
#pragma unroll
                        for (int i = 0; i < BN / 64; ++i)
                            tma_load_2d_hint(sb + i * 8192, &tma_b, &full_bar[stage], n0 + i * 64, kb * BK, p.b_policy);
                    } else if constexpr (EPI == EPI_SWIGLU) {
                        // rows [n0, n0+128) of the gate half and the SAME rows of the up half (box = 128 rows): accumulator
                        // columns [0,128) = gate, [128,256) = up of output columns n0..n0+127.  A gate box that runs past f
                        // reads up rows (never stored: columns >= f are masked); an up box past 2f is zero-filled.
                        tma_load_2d_hint(sb, &tma_b, &full_bar[stage], kb * BK, n0, p.b_policy);
                        tma_load_2d_hint(sb + 128 * BK * 2, &tma_b, &full_bar[stage], kb * BK, p.f + n0, p.b_policy);
                    } else {
                        tma_load_2d_hint(sb, &tma_b, &full_bar[stage], kb * BK, n0, p.b_policy);
                    }
                    if (++stage == S::STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===================== MMA issuer
            constexpr uint32_t idesc = umma_idesc_bf16(BM, BN) | (TA ? (1u << 15) : 0u) | (TB ? (1u << 16) : 0u);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(tiles + stage * S::STAGE_BYTES);
                    const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        // K-major operand: advance 16 bf16 = 32 B inside the 128 B swizzle row (+2 in the addr >> 4 field);
                        // MN-major operand: 16 contraction rows = 2 KB further down the box
                        const uint64_t da = TA ? umma_desc_mnmajor_sw128(sa + k * 2048, 8192)
                                               : umma_desc_kmajor_sw128(sa) + static_cast<uint64_t>(2 * k);
                        const uint64_t db = TB ? umma_desc_mnmajor_sw128(sb + k * 2048, 8192)
                                               : umma_desc_kmajor_sw128(sb) + static_cast<uint64_t>(2 * k);
                        umma_bf16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[stage]);       // smem slot reusable once these MMAs retire
                    if (++stage == S::STAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit(&tmem_full[acc]);             // accumulator complete -> epilogue
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1u;
            }
        }
    } else if (warp == 3) {
        if constexpr (EPI == EPI_DSWIGLU) {
            if (lane == 0) {
                // ===================== gate|up loader: [128 rows x 64 cols] of gate and of up per group, two groups ahead
                uint32_t gc = 0;
                for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                    int tm, tn;
                    tile_coords(p, t, tm, tn);
                    const int m0 = tm * BM, n0 = tn * BN;
                    for (int g = 0; g < BN / 64; ++g, ++gc) {
                        const int buf = static_cast<int>(gc & 1u);
                        mbar_wait(&gu_empty[buf], ((gc >> 1) & 1u) ^ 1u);
                        uint8_t* dst = stg + buf * S::STG_BUF_BYTES;
                        mbar_arrive_expect_tx(&gu_full[buf], S::STG_BUF_BYTES);
                        // read once: evict-first, so the 245 MB of gate|up do not push the GEMM operands out of L2 (first ncu
                        // capture: L2 hit rate 68 % against 86 % for the plain launch, 834 MB of DRAM reads)
                        tma_load_2d_hint(dst, &tma_c, &gu_full[buf], n0 + g * 64, m0, kEvictFirst);
                        tma_load_2d_hint(dst + BM * 128, &tma_c, &gu_full[buf], p.f + n0 + g * 64, m0, kEvictFirst);
                    }
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: TMEM -> registers -> global
        const int q = warp & 3;                            // TMEM lane quarter this warp may access
        int acc = 0;
        uint32_t acc_phase = 0;
        [[maybe_unused]] uint32_t gcount = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            int tm, tn;
            tile_coords(p, t, tm, tn);
            const int m0 = tm * BM, n0 = tn * BN_OUT;
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < p.M;
            const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BN);
            if constexpr (EPI == EPI_DSWIGLU)
                epilogue_dswiglu_staged<BN>(p, t_addr, row, row_ok, n0, q * 32 + lane, stg, gu_full, gu_empty, gcount, lane);
            else
                epilogue_tile<BN, EPI>(p, t_addr, row, row_ok, n0);
            tc_fence_before();
            mbar_arrive(&tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1u;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 2 * BN);
    }
}

// ============================================================================================ CTA-pair GEMM (cta_group::2)
// Two CTAs of a cluster (the two SMs of a TPC) compute one 256 x 256 tile: each CTA stages ITS 128 rows of A and ITS 128 of the
// 256 weight rows (32 KB per stage instead of 48 KB -> 6 stages), the leader CTA issues tcgen05.mma.cta_group::2 (M = 256), which
// reads A from both CTAs and gives every SM the B half it does not hold through the pair's shared-memory path: half the B
// traffic from L2 and half the shared-memory operand bytes per flop of the single-CTA kernel (the step is
// power-capped, so bytes per flop are the lever).  Accumulators: 128 lanes x 256 columns in EACH CTA's TMEM, two stages.
//   full[s]   (leader)  : count 1 = the leader's arrive.expect_tx(64 KB); both CTAs' TMA credit their bytes to it.  (A second,
//                         remote arrive by the peer per stage - DeepGEMM's form - measured 640 TF/s instead of 1250: the phase
//                         cannot complete before the leader's own arrive anyway, and the peer never runs a stage ahead because
//                         it waits for the multicast empty[s] first.)
//   empty[s]  (both)    : tcgen05.commit multicast to both CTAs once the MMAs reading stage s retired
//   tfull[a]  (both)    : tcgen05.commit multicast - accumulator stage a complete
//   tempty[a] (leader)  : count 8 = one elected lane of each epilogue warp of both CTAs (the peer's four arrive remotely)
ESB_DEVICE void mbar_wait_dbg(int spin, uint64_t* bar, uint32_t parity) {
    if (!spin) { mbar_wait(bar, parity); return; }
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    }
}
constexpr int P2_STAGES = 6;
constexpr int P2_STAGE_BYTES = 2 * 128 * BK * 2;
constexpr int P2_SMEM = P2_STAGES * P2_STAGE_BYTES + 1024 + 1024;

template <int EPI, bool TA = false, bool TB = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm2_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
    static_assert(!(TA || TB) || EPI == EPI_PLAIN, "transposed operands: plain epilogue");
    constexpr int BN = 256;
    constexpr int BN_OUT = (EPI == EPI_SWIGLU) ? 128 : BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* tiles = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + P2_STAGES * P2_STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + P2_STAGES;
    uint64_t* tmem_full = bars + 2 * P2_STAGES;
    uint64_t* tmem_empty = bars + 2 * P2_STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * P2_STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = static_cast<int>(blockIdx.x >> 1), npairs = static_cast<int>(gridDim.x >> 1);
    const int num_kb = (p.K + BK - 1) / BK;
    const int num_tiles = p.tiles_m * p.tiles_n;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < P2_STAGES; ++i) {
            mbar_init(&full_bar[i], (p.dbg & 1) ? 2 : 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], (p.dbg & 4) ? 256 : 8);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc_2cta(tmem_slot, 2 * BN);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    cluster_sync_all();                                    // barrier inits + TMEM allocation visible to both CTAs
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            // ===================== TMA producer (both CTAs): own 128 rows of A, own 128 weight rows
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t leader_full0 = mapa_shared(smem_u32(&full_bar[0]), 0);
            for (int t = pair; t < num_tiles; t += npairs) {
                int tm, tn;
                tile_coords(p, t, tm, tn);
                const int m0 = tm * 256 + static_cast<int>(rank) * 128;
                const int nb = (EPI == EPI_SWIGLU) ? (tn * 128 + (rank ? p.f : 0)) : (tn * 256 + static_cast<int>(rank) * 128);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait_dbg(p.dbg & 2, &empty_bar[stage], phase ^ 1u);
                    uint8_t* sa = tiles + stage * P2_STAGE_BYTES;
                    uint8_t* sb = sa + 128 * BK * 2;
                    if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * P2_STAGE_BYTES);
                    if constexpr (TA) {                    // [64 contraction rows x 64 columns] boxes, MN-major operand (see the single-CTA kernel)
                        tma_load_2d_2sm_hint(sa, &tma_a, &full_bar[stage], m0, kb * BK, p.a_policy);
                        tma_load_2d_2sm_hint(sa + 8192, &tma_a, &full_bar[stage], m0 + 64, kb * BK, p.a_policy);
                    } else {
                        tma_load_2d_2sm_hint(sa, &tma_a, &full_bar[stage], kb * BK, m0, p.a_policy);
                    }
                    if constexpr (TB) {
                        tma_load_2d_2sm_hint(sb, &tma_b, &full_bar[stage], nb, kb * BK, p.b_policy);
                        tma_load_2d_2sm_hint(sb + 8192, &tma_b, &full_bar[stage], nb + 64, kb * BK, p.b_policy);
                    } else {
                        tma_load_2d_2sm_hint(sb, &tma_b, &full_bar[stage], kb * BK, nb, p.b_policy);
                    }
                    if (rank != 0 && (p.dbg & 1)) mbar_arrive_cluster(leader_full0 + static_cast<uint32_t>(stage) * 8u);
                    if (++stage == P2_STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // ===================== MMA issuer (leader CTA only)
            constexpr uint32_t idesc = umma_idesc_bf16(256, BN) | (TA ? (1u << 15) : 0u) | (TB ? (1u << 16) : 0u);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = pair; t < num_tiles; t += npairs) {
                mbar_wait_dbg(p.dbg & 2, &tmem_empty[acc], acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait_dbg(p.dbg & 2, &full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(tiles + stage * P2_STAGE_BYTES);
                    const uint32_t sb = sa + 128 * BK * 2;
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t da = TA ? umma_desc_mnmajor_sw128(sa + k * 2048, 8192)
                                               : umma_desc_kmajor_sw128(sa) + static_cast<uint64_t>(2 * k);
                        const uint64_t db = TB ? umma_desc_mnmajor_sw128(sb + k * 2048, 8192)
                                               : umma_desc_kmajor_sw128(sb) + static_cast<uint64_t>(2 * k);
                        umma_bf16_ss_2cta(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit_2cta(&empty_bar[stage], 3);          // both CTAs may refill stage `stage`
                    if (++stage == P2_STAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit_2cta(&tmem_full[acc], 3);                // accumulator complete in both CTAs -> epilogues
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1u;
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue (both CTAs): own 128 rows of the tile
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        const uint32_t leader_tempty0 = mapa_shared(smem_u32(&tmem_empty[0]), 0);
        for (int t = pair; t < num_tiles; t += npairs) {
            int tm, tn;
            tile_coords(p, t, tm, tn);
            const int m0 = tm * 256 + static_cast<int>(rank) * 128, n0 = tn * BN_OUT;
            mbar_wait_dbg(p.dbg & 2, &tmem_full[acc], acc_phase);
            tc_fence_after();
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < p.M;
            const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BN);
            epilogue_tile<BN, EPI>(p, t_addr, row, row_ok, n0);
            tc_fence_before();
            if (p.dbg & 4) {
                mbar_arrive_cluster(leader_tempty0 + static_cast<uint32_t>(acc) * 8u);
            } else {                                   // one arrive per warp: a remote mbarrier arrive costs ~100x a local one
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(leader_tempty0 + static_cast<uint32_t>(acc) * 8u);
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1u;
        }
    }

    tc_fence_before();
    cluster_sync_all();                                    // nobody frees TMEM / exits while the partner still works
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, 2 * BN);
    }
}


// ----------------------------------------------------------------------------- host side
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static std::once_flag g_encode_once;

static void load_encode() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
        g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
}

// 2-D bf16 row-major [rows, cols] tensor with leading dimension ld (elements); box = {64 cols, box_rows}.
int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows) {
    std::call_once(g_encode_once, load_encode);
    if (!g_encode) {
        set_last_error("cuTensorMapEncodeTiled entry point unavailable");
        return ESB_ERR_DRIVER;
    }
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(BK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    // L2 promotion 256 B vs 128 B vs none: no difference in DRAM bytes or time (ncu, profiles/r01_gemm_traffic_raster.txt)
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld ptr=%p", (int)r, rows, cols, ld, ptr);
        return ESB_ERR_DRIVER;
    }
    return ESB_OK;
}

int make_tma_3d_bf16(CUtensorMap* map, const void* ptr, long long cols, long long rows, long long batches, long long row_stride,
                     long long batch_stride, int box_cols, int box_rows) {
    std::call_once(g_encode_once, load_encode);
    if (!g_encode) {
        set_last_error("cuTensorMapEncodeTiled entry point unavailable");
        return ESB_ERR_DRIVER;
    }
    if (batches <= 1) batch_stride = rows * row_stride;        // any legal stride: the coordinate is always 0
    cuuint64_t gdim[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(batches < 1 ? 1 : batches)};
    cuuint64_t gstride[2] = {static_cast<cuuint64_t>(row_stride) * 2, static_cast<cuuint64_t>(batch_stride) * 2};
    cuuint32_t box[3] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled(3d) failed (%d) cols=%lld rows=%lld batches=%lld rs=%lld bs=%lld ptr=%p", (int)r, cols,
                       rows, batches, row_stride, batch_stride, ptr);
        return ESB_ERR_DRIVER;
    }
    return ESB_OK;
}

// Tile grid + walk order + L2 policies of one launch (tile = bm rows x bn_out output columns, B panel = bn weight rows).
// Modelled DRAM bytes = persisting operand once + streamed operand once per group + C.
static void choose_walk(GemmParams& p, int bm, int bn, int bn_out, int c_mult) {
    p.tiles_m = (p.M + bm - 1) / bm;
    p.tiles_n = (p.N + bn_out - 1) / bn_out;
    static const int hints = [] { const char* e = getenv("ESB_GEMM_L2_HINTS"); return e ? atoi(e) : 7; }();
    static const double budget_mb = [] { const char* e = getenv("ESB_GEMM_L2_BUDGET_MB"); return e ? atof(e) : 48.0; }();
    const double budget = budget_mb * (1 << 20);              // bytes of one operand group kept L2-resident (126 MB L2, two halves)
    const double a_panel = 2.0 * bm * p.K, b_panel = 2.0 * bn * p.K;
    const double a_bytes = a_panel * p.tiles_m, b_bytes = b_panel * p.tiles_n;
    const double c_bytes = static_cast<double>(p.M) * p.N * (p.out_fp32 ? 4 : 2) * c_mult;
    int gm = static_cast<int>(budget / a_panel);
    gm = gm < 1 ? 1 : (gm > p.tiles_m ? p.tiles_m : gm);
    const int groups_m = (p.tiles_m + gm - 1) / gm;
    gm = (p.tiles_m + groups_m - 1) / groups_m;               // even groups
    int gn = static_cast<int>(budget / b_panel);
    gn = gn < 1 ? 1 : (gn > p.tiles_n ? p.tiles_n : gn);
    const int groups_n = (p.tiles_n + gn - 1) / gn;
    gn = (p.tiles_n + groups_n - 1) / groups_n;
    const double cost_m = a_bytes + b_bytes * groups_m, cost_n = b_bytes + a_bytes * groups_n;
    p.raster_mode = 0; p.group = p.tiles_m;                   // ESB_GEMM_L2_HINTS bit 2 off: the plain M-fastest walk
    if (hints & 4) {
        if (cost_m <= cost_n) { p.raster_mode = 0; p.group = gm; }
        else { p.raster_mode = 1; p.group = gn; }
    }
    const bool persist_fits = (p.raster_mode == 0 ? a_panel * p.group : b_panel * p.group) <= 64.0 * (1 << 20);
    p.a_policy = ((hints & 1) && p.raster_mode == 0 && persist_fits) ? kEvictLast : kEvictNormal;
    p.b_policy = ((hints & 1) && p.raster_mode == 1 && persist_fits) ? kEvictLast : kEvictNormal;
    p.c_stream = ((hints & 2) && !p.out_fp32 && c_bytes >= 96.0 * (1 << 20)) ? 1 : 0;
}

template <int BN, int EPI = EPI_PLAIN, bool TA = false, bool TB = false>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, GemmParams p, cudaStream_t stream, int sms,
                       const CUtensorMap* tc = nullptr) {
    using S = GemmSmem<BN, EPI>;
    constexpr int BN_OUT = (EPI == EPI_SWIGLU) ? 128 : BN;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, EPI, TA, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
        if (e != cudaSuccess) {
            set_last_error("cudaFuncSetAttribute(gemm<%d>): %s", BN, cudaGetErrorString(e));
            return ESB_ERR_CUDA;
        }
        attr_set = true;
    }
    choose_walk(p, BM, BN, BN_OUT, (EPI == EPI_SWIGLU || EPI == EPI_DSWIGLU) ? 2 : 1);
    const int tiles = p.tiles_m * p.tiles_n;
    const int grid = tiles < sms ? tiles : sms;
    gemm_bf16_tcgen05_kernel<BN, EPI, TA, TB><<<grid, GEMM_THREADS, S::TOTAL, stream>>>(ta, tb, tc ? *tc : ta, p);
    return check_launch("gemm_bf16_tcgen05");
}

// CTA-pair variant (gemm2_bf16_tcgen05_kernel): ON by default since it measured 1450-1520 TF/s against 1385-1430 for the
// single-CTA kernel on the four trunk shapes and 143.2-143.8 ms against 147.7-148.8 ms per training step on the same box
// (profiles/r02_v6_*); esb_set_gemm_2cta(0) / ESB_GEMM_2CTA=0 switch back
// SMs left out of the persistent grids (esb_set_gemm_sm_reserve): while a gradient all-reduce is in flight NCCL's CTAs need SMs of
// their own - next to a 1-CTA-per-SM kernel with ~200 KB of shared memory they only get scheduled in the gaps between launches and
// the collective is still running when the backward ends (bench.py dp_timeline: 3.2 ms of exchange not hidden at N=2).
static int g_gemm_sm_reserve = 0;
int set_gemm_sm_reserve(int k) {
    const int old = g_gemm_sm_reserve;
    g_gemm_sm_reserve = k < 0 ? 0 : (k > 64 ? 64 : k);
    return old;
}
static int gemm_sm_count() {
    const int n = device_sm_count() - g_gemm_sm_reserve;
    return n < 2 ? 2 : n;
}
static int g_gemm_2cta = [] { const char* e = getenv("ESB_GEMM_2CTA"); return e ? atoi(e) : 1; }();
int set_gemm_2cta(int on) { const int old = g_gemm_2cta; g_gemm_2cta = on ? 1 : 0; return old; }
static bool use_2cta(int M, int N, int K) { return g_gemm_2cta && M >= 1024 && N >= 512 && K >= 256; }

template <int EPI, bool TA = false, bool TB = false>
static int launch_gemm2(const CUtensorMap& ta, const CUtensorMap& tb, GemmParams p, cudaStream_t stream, int sms) {
    constexpr int BN_OUT = (EPI == EPI_SWIGLU) ? 128 : 256;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm2_bf16_tcgen05_kernel<EPI, TA, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, P2_SMEM);
        if (e != cudaSuccess) {
            set_last_error("cudaFuncSetAttribute(gemm2): %s", cudaGetErrorString(e));
            return ESB_ERR_CUDA;
        }
        attr_set = true;
    }
    choose_walk(p, 256, 256, BN_OUT, (EPI == EPI_SWIGLU || EPI == EPI_DSWIGLU) ? 2 : 1);
    { const char* e = getenv("ESB_GEMM2_DBG"); p.dbg = e ? atoi(e) : 0; }
    const int tiles = p.tiles_m * p.tiles_n;
    int pairs = sms / 2;
    if (tiles < pairs) pairs = tiles;
    gemm2_bf16_tcgen05_kernel<EPI, TA, TB><<<2 * pairs, GEMM_THREADS, P2_SMEM, stream>>>(ta, tb, p);
    return check_launch("gemm2_bf16_tcgen05");
}

int gemm_bf16(const void* A, long long lda, const void* B, long long ldb, void* C, long long ldc, int M, int N, int K,
              const float* bias, const void* residual, long long ldr, int act, int out_fp32, int accumulate, float alpha,
              int force_bn, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return ESB_OK;
    if (K <= 0 || !A || !B || !C) {
        set_last_error("gemm_bf16: bad arguments (M=%d N=%d K=%d)", M, N, K);
        return ESB_ERR_INVALID_ARG;
    }
    if ((lda & 7) || (ldb & 7) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) {
        set_last_error("gemm_bf16: A/B must be 16-byte aligned with leading dimensions that are multiples of 8 (lda=%lld ldb=%lld)",
                       lda, ldb);
        return ESB_ERR_INVALID_ARG;
    }
    if (accumulate && !out_fp32) {
        set_last_error("gemm_bf16: accumulate requires fp32 output");
        return ESB_ERR_INVALID_ARG;
    }
    const int sms = gemm_sm_count();
    int bn = force_bn;
    if (bn == 0) {
        if (N <= 64) bn = 64;
        else if (N <= 128) bn = 128;
        else {
            const long long t256 = static_cast<long long>((M + BM - 1) / BM) * ((N + 255) / 256);
            bn = (t256 >= sms) ? 256 : 128;
            if (bn == 128) {
                const long long t128 = static_cast<long long>((M + BM - 1) / BM) * ((N + 127) / 128);
                if (t128 < sms / 2) bn = 64;
            }
        }
    }
    const bool pair = force_bn == 0 && use_2cta(M, N, K);
    CUtensorMap ta, tb;
    int rc = make_tma_2d_bf16(&ta, A, M, K, lda, BM);
    if (rc) return rc;
    rc = make_tma_2d_bf16(&tb, B, N, K, ldb, pair ? 128 : bn);
    if (rc) return rc;
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.C = C; p.ldc = ldc; p.bias = bias;
    p.res = reinterpret_cast<const __nv_bfloat16*>(residual); p.ldr = ldr;
    p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate; p.alpha = alpha;
    p.tiles_m = p.tiles_n = 0;
    p.raster_mode = 0; p.group = 1; p.a_policy = p.b_policy = kEvictNormal; p.c_stream = 0;      // set by launch_gemm
    p.C2 = nullptr; p.ldc2 = 0; p.aux = nullptr; p.ld_aux = 0; p.f = 0; p.cos_tab = p.sin_tab = nullptr;
    p.rope_L = 1; p.rope_pos0 = 0; p.rope_cols = 0; p.dbg = 0;
    if (pair) return launch_gemm2<EPI_PLAIN>(ta, tb, p, stream, sms);
    switch (bn) {
        case 256: return launch_gemm<256>(ta, tb, p, stream, sms);
        case 128: return launch_gemm<128>(ta, tb, p, stream, sms);
        case 64: return launch_gemm<64>(ta, tb, p, stream, sms);
        default:
            set_last_error("gemm_bf16: unsupported BN %d", bn);
            return ESB_ERR_INVALID_ARG;
    }
}

// ----------------------------------------------------------------------------- fused-epilogue entry points
static int fused_common(GemmParams& p, const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, const char* who) {
    if (K <= 0 || !A || !B) { set_last_error("%s: bad arguments (M=%d N=%d K=%d)", who, M, N, K); return ESB_ERR_INVALID_ARG; }
    if ((lda & 7) || (ldb & 7) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) {
        set_last_error("%s: A/B must be 16-byte aligned with leading dimensions that are multiples of 8", who);
        return ESB_ERR_INVALID_ARG;
    }
    p.M = M; p.N = N; p.K = K; p.C = nullptr; p.ldc = 0; p.bias = nullptr; p.res = nullptr; p.ldr = 0;
    p.act = 0; p.out_fp32 = 0; p.accumulate = 0; p.alpha = 1.0f; p.tiles_m = p.tiles_n = 0;
    p.raster_mode = 0; p.group = 1; p.a_policy = p.b_policy = kEvictNormal; p.c_stream = 0;
    p.C2 = nullptr; p.ldc2 = 0; p.aux = nullptr; p.ld_aux = 0; p.f = 0; p.cos_tab = p.sin_tab = nullptr;
    p.rope_L = 1; p.rope_pos0 = 0; p.rope_cols = 0; p.dbg = 0;
    return ESB_OK;
}

// act [M, f] = silu(A . Wg^T) * (A . Wu^T), Wgu = [Wg ; Wu] packed [2f, K]; gu [M, 2f] optional (saved for the backward)
int gemm_swiglu_bf16(const void* A, long long lda, const void* Wgu, long long ldw, void* act, long long ld_act, void* gu,
                     long long ld_gu, int M, int f, int K, cudaStream_t stream) {
    if (M <= 0 || f <= 0) return ESB_OK;
    GemmParams p;
    int rc = fused_common(p, A, lda, Wgu, ldw, M, f, K, "gemm_swiglu_bf16");
    if (rc) return rc;
    if (!act) { set_last_error("gemm_swiglu_bf16: act is null"); return ESB_ERR_INVALID_ARG; }
    CUtensorMap ta, tb;
    rc = make_tma_2d_bf16(&ta, A, M, K, lda, BM);
    if (rc) return rc;
    rc = make_tma_2d_bf16(&tb, Wgu, 2LL * f, K, ldw, 128);
    if (rc) return rc;
    p.C = act; p.ldc = ld_act; p.C2 = gu; p.ldc2 = ld_gu; p.f = f;
    if (use_2cta(M, f, K)) return launch_gemm2<EPI_SWIGLU>(ta, tb, p, stream, gemm_sm_count());      // same 128-row boxes
    return launch_gemm<256, EPI_SWIGLU>(ta, tb, p, stream, gemm_sm_count());
}

// dgu [M, 2f] = dSwiGLU(gu, dOut . Wdown) with the weight given either as WdownT [f, K=d] (contraction-major copy, trans_b = 0)
// or as down_proj itself [K=d, f] (state_dict layout, trans_b = 1: MN-major B operand)
int gemm_dswiglu_bf16(const void* dOut, long long lda, const void* WdownT, long long ldb, int trans_b, const void* gu, long long ld_gu,
                      void* dgu, long long ld_dgu, int M, int f, int K, cudaStream_t stream) {
    if (M <= 0 || f <= 0) return ESB_OK;
    GemmParams p;
    int rc = fused_common(p, dOut, lda, WdownT, ldb, M, f, K, "gemm_dswiglu_bf16");
    if (rc) return rc;
    if (!gu || !dgu) { set_last_error("gemm_dswiglu_bf16: gu / dgu is null"); return ESB_ERR_INVALID_ARG; }
    CUtensorMap ta, tb;
    rc = make_tma_2d_bf16(&ta, dOut, M, K, lda, BM);
    if (rc) return rc;
    const bool pair = !trans_b && use_2cta(M, f, K);
    rc = trans_b ? make_tma_2d_bf16(&tb, WdownT, K, f, ldb, 64) : make_tma_2d_bf16(&tb, WdownT, f, K, ldb, pair ? 128 : 256);
    if (rc) return rc;
    p.aux = static_cast<const __nv_bfloat16*>(gu); p.ld_aux = ld_gu; p.C2 = dgu; p.ldc2 = ld_dgu; p.f = f;
    if (pair) return launch_gemm2<EPI_DSWIGLU>(ta, tb, p, stream, gemm_sm_count());
    if ((ld_gu & 7) || (reinterpret_cast<uintptr_t>(gu) & 15)) {
        set_last_error("gemm_dswiglu_bf16: gu must be 16-byte aligned with a leading dimension that is a multiple of 8");
        return ESB_ERR_INVALID_ARG;
    }
    CUtensorMap tc;                                            // saved gate|up [M, 2f]: box = 64 columns x 128 rows for the staged epilogue
    rc = make_tma_2d_bf16(&tc, gu, M, 2LL * f, ld_gu, BM);
    if (rc) return rc;
    if (trans_b) return launch_gemm<256, EPI_DSWIGLU, false, true>(ta, tb, p, stream, gemm_sm_count(), &tc);
    return launch_gemm<256, EPI_DSWIGLU>(ta, tb, p, stream, gemm_sm_count(), &tc);
}

// C [M, N] = rope(A . W^T): heads of 128 columns starting below rope_cols are rotated at position pos0 + (row % L)
int gemm_rope_bf16(const void* A, long long lda, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                   const void* cos_tab, const void* sin_tab, int L, int pos0, int rope_cols, int head_dim, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return ESB_OK;
    GemmParams p;
    int rc = fused_common(p, A, lda, W, ldw, M, N, K, "gemm_rope_bf16");
    if (rc) return rc;
    if (head_dim != 128 || (N & 127) || (rope_cols & 127) || !C || !cos_tab || !sin_tab || L <= 0) {
        set_last_error("gemm_rope_bf16: needs head_dim 128, N and rope_cols multiples of 128 (N=%d rope_cols=%d hd=%d)", N, rope_cols, head_dim);
        return ESB_ERR_INVALID_ARG;
    }
    CUtensorMap ta, tb;
    rc = make_tma_2d_bf16(&ta, A, M, K, lda, BM);
    if (rc) return rc;
    const bool pair = use_2cta(M, N, K);
    rc = make_tma_2d_bf16(&tb, W, N, K, ldw, pair ? 128 : 256);
    if (rc) return rc;
    p.C = C; p.ldc = ldc; p.cos_tab = static_cast<const __nv_bfloat16*>(cos_tab); p.sin_tab = static_cast<const __nv_bfloat16*>(sin_tab);
    p.rope_L = L; p.rope_pos0 = pos0; p.rope_cols = rope_cols;
    if (pair) return launch_gemm2<EPI_ROPE>(ta, tb, p, stream, gemm_sm_count());
    return launch_gemm<256, EPI_ROPE>(ta, tb, p, stream, gemm_sm_count());
}

// C[M,N] (+)= alpha * op(A) . op(B): trans_a = 1: A is [Kc, M] row-major (else [M, Kc]); trans_b = 1: B is [Kc, N] row-major
// (else [N, Kc], the state_dict layout the plain entry point takes).  trans_a = trans_b = 0 is gemm_bf16 itself.
template <bool TA, bool TB>
static int launch_t(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int M, int N, cudaStream_t stream) {
    const int sms = gemm_sm_count();
    if (use_2cta(M, N, p.K)) return launch_gemm2<EPI_PLAIN, TA, TB>(ta, tb, p, stream, sms);   // same 64 x 64 boxes / 128-row A box
    const long long t256 = static_cast<long long>((M + BM - 1) / BM) * ((N + 255) / 256);
    if (N > 128 && t256 >= sms / 2) return launch_gemm<256, EPI_PLAIN, TA, TB>(ta, tb, p, stream, sms);
    if (N > 64) return launch_gemm<128, EPI_PLAIN, TA, TB>(ta, tb, p, stream, sms);
    return launch_gemm<64, EPI_PLAIN, TA, TB>(ta, tb, p, stream, sms);
}
int gemm_t_bf16(const void* A, long long lda, int trans_a, const void* B, long long ldb, int trans_b, void* C, long long ldc, int M,
                int N, int Kc, int out_fp32, int accumulate, float alpha, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return ESB_OK;
    if (!trans_a && !trans_b)
        return gemm_bf16(A, lda, B, ldb, C, ldc, M, N, Kc, nullptr, nullptr, 0, 0, out_fp32, accumulate, alpha, 0, stream);
    GemmParams p;
    int rc = fused_common(p, A, lda, B, ldb, M, N, Kc, "gemm_t_bf16");
    if (rc) return rc;
    if (!C || (accumulate && !out_fp32)) { set_last_error("gemm_t_bf16: C is null or accumulate without fp32 output"); return ESB_ERR_INVALID_ARG; }
    CUtensorMap ta, tb;
    rc = trans_a ? make_tma_2d_bf16(&ta, A, Kc, M, lda, 64)     // rows = contraction index, box = 64 columns x 64 rows
                 : make_tma_2d_bf16(&ta, A, M, Kc, lda, BM);
    if (rc) return rc;
    p.C = C; p.ldc = ldc; p.out_fp32 = out_fp32; p.accumulate = accumulate; p.alpha = alpha;
    if (trans_b) {
        rc = make_tma_2d_bf16(&tb, B, Kc, N, ldb, 64);
        if (rc) return rc;
        return trans_a ? launch_t<true, true>(ta, tb, p, M, N, stream) : launch_t<false, true>(ta, tb, p, M, N, stream);
    }
    set_last_error("gemm_t_bf16: trans_a without trans_b is not built (no call site)");
    return ESB_ERR_UNSUPPORTED;
}

}  // namespace esb
