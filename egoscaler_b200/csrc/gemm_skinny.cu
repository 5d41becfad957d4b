// Skinny bf16 GEMM for the single-token decode step:  C[M,N] = X[M,K] * W[N,K]^T (+ residual),  M <= 64.
//
// Reference call sites: the seven Linear layers of HF LlamaDecoderLayer and lm_head when generate() feeds ONE new
// token per sequence against the KV cache (pointllm/model/pointllm.py:255-275 -> :173-178, :228; callers
// model_arch.py:94-108).  With M = batch (32 per GPU in BASELINE configs[3]) the contraction is a weight-streaming
// problem: every byte of W is read once, so the bound is HBM, not the tensor pipe.
//
// Design (differs from gemm_tcgen05.cu on purpose):
//   * operands swapped: the 128-row WEIGHT tile is the UMMA "A" operand (M=128) and the <=64 tokens are the "B"
//     operand (N = BT), so no tensor-core rows are wasted on padding and a smem stage is 16 KB of weights + 4 KB of X;
//   * split-K: work unit = (128 weight rows, K/S slice); the host picks S so that units ~ a multiple of the SM count
//     (N = 4096 has only 32 row tiles - without split-K 116 of 148 SMs would idle and HBM could not be saturated);
//   * two persistent CTAs per SM walk units round-robin, each with a 4-stage TMA ring that runs ahead across unit
//     boundaries; two TMEM accumulator stages decouple the MMA from the epilogue;
//   * deterministic reduction without a second launch: every unit stores its fp32 partial to a workspace, bumps a
//     per-tile counter, and the LAST arriving unit of a tile sums the S partials in fixed order s = 0..S-1, adds the
//     residual and writes bf16.  Counters are left at zero, so the kernel is CUDA-graph replay safe.
//   * W is streamed with an evict-first L2 hint, X (re-read by every CTA) with evict-last.
#include "common.cuh"
#include "esb_internal.h"


namespace esb {

namespace {

constexpr int SK_ROWS = 128;      // weight rows per tile = UMMA M
constexpr int SK_BK = 64;         // 64 bf16 = one 128-byte swizzle row
constexpr int SK_THREADS = 192;   // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue
constexpr int SK_COUNTER_BYTES = 16384;

struct SkinnyParams {
    int M, N, K;
    __nv_bfloat16* C;
    long long ldc;
    const __nv_bfloat16* res;
    long long ldr;
    int n_tiles, splits, kb_per, num_kb;
    const uint8_t* w_tiles;
    int w_tiled;      // 1: W is pre-tiled [n_tiles][num_kb][128][64] (esb_retile_weight_bf16), every stage one contiguous 16 KB
    float* partial;   // [n_tiles][splits][BT][128] fp32
    int* counters;    // [n_tiles], zero on entry and on exit
    // RMSNorm folded around the GEMM (decode step): the norm weight is folded into W at retile time, the per-token 1/rms
    // is applied to the accumulator here, and the GEMM that PRODUCES the next norm's input leaves its sum of squares behind.
    const float* scale_src;   // NULL: none; scale_parts == 0: rstd[m] itself; else per-tile sums of squares [scale_parts][64]
    int scale_parts;
    float scale_inv_d, scale_eps;
    float* ssq_out;           // NULL or [n_tiles][64]: sum over this tile's 128 columns of bf16(out[m, :])^2
};

template <int BT, int SK_STAGES>
struct SkinnySmem {
    static constexpr int W_BYTES = SK_ROWS * SK_BK * 2;
    static constexpr int X_BYTES = BT * SK_BK * 2;
    static constexpr int STAGE_BYTES = W_BYTES + X_BYTES;
    static constexpr int TOTAL = SK_STAGES * STAGE_BYTES + 1024 + 1024 + 1024;     // + s_red [4][64] fp32
};

ESB_DEVICE void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// NORMED = false compiles the scale / sum-of-squares epilogue out (the plain token step keeps its 114-register kernel).
template <int BT, int SK_STAGES, bool NORMED>
__global__ void __launch_bounds__(SK_THREADS, SK_STAGES <= 4 ? 2 : 1)
gemm_skinny_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_w, const __grid_constant__ CUtensorMap tma_x,
                           const SkinnyParams p) {
    using S = SkinnySmem<BT, SK_STAGES>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SK_STAGES * S::STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + SK_STAGES;
    uint64_t* tmem_full = bars + 2 * SK_STAGES;
    uint64_t* tmem_empty = bars + 2 * SK_STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * SK_STAGES + 4);
    volatile int* last_flag = reinterpret_cast<volatile int*>(tmem_slot + 1);
    float* s_scale = reinterpret_cast<float*>(tmem_slot + 4);                       // [64] per-token scale of the accumulator
    float (*s_red)[64] = reinterpret_cast<float (*)[64]>(smem + SK_STAGES * S::STAGE_BYTES + 1024);   // [4][64]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int units = p.n_tiles * p.splits;
    griddep_launch();                                     // dependents park in their own griddepcontrol.wait

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_w);
        tma_prefetch_desc(&tma_x);
        for (int i = 0; i < SK_STAGES; ++i) {
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
            // ===================== TMA producer: runs ahead across unit boundaries.
            // Programmatic dependent launch: the weights do not depend on the previous kernel of the stream, so the first
            // ring of WEIGHT stages is requested before griddepcontrol.wait; the X tiles of those stages (written by the
            // previous kernel) follow after it.  Launch latency, prologue and the first DRAM round trip of the weight
            // stream thereby overlap the tail of the preceding small kernel (rmsnorm / rope+attention / silu).
            int stage = 0;
            uint32_t phase = 0;
            int n_pend = 0;
            int pend_kb[SK_STAGES];
            bool dep_ready = false;
            for (int u = blockIdx.x; u < units; u += gridDim.x) {
                const int tile = u / p.splits, s = u - tile * p.splits;
                const int kb0 = s * p.kb_per;
                const int kb1 = min(p.num_kb, kb0 + p.kb_per);
                for (int kb = kb0; kb < kb1; ++kb) {
                    if (!dep_ready && n_pend == SK_STAGES) {
                        griddep_wait();
                        for (int i = 0; i < n_pend; ++i)
                            tma_load_2d_hint(smem + i * S::STAGE_BYTES + S::W_BYTES, &tma_x, &full_bar[i], pend_kb[i] * SK_BK, 0, kEvictLast);
                        n_pend = 0;
                        dep_ready = true;
                    }
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    uint8_t* sw = smem + stage * S::STAGE_BYTES;
                    mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                    if (p.w_tiled)
                        bulk_load_1d_hint(sw, p.w_tiles + (static_cast<size_t>(tile) * p.num_kb + kb) * S::W_BYTES, S::W_BYTES,
                                          &full_bar[stage], kEvictFirst);
                    else
                        tma_load_2d_hint(sw, &tma_w, &full_bar[stage], kb * SK_BK, tile * SK_ROWS, kEvictFirst);
                    if (dep_ready) tma_load_2d_hint(sw + S::W_BYTES, &tma_x, &full_bar[stage], kb * SK_BK, 0, kEvictLast);
                    else pend_kb[n_pend++] = kb;             // first ring pass: stage index == n_pend
                    if (++stage == SK_STAGES) { stage = 0; phase ^= 1u; }
                }
            }
            if (!dep_ready) {
                griddep_wait();
                for (int i = 0; i < n_pend; ++i)
                    tma_load_2d_hint(smem + i * S::STAGE_BYTES + S::W_BYTES, &tma_x, &full_bar[i], pend_kb[i] * SK_BK, 0, kEvictLast);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            griddep_wait();
            // ===================== MMA issuer: D[128 weight rows, BT tokens] += W_tile * X_tile^T
            constexpr uint32_t idesc = umma_idesc_bf16(SK_ROWS, BT);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int u = blockIdx.x; u < units; u += gridDim.x) {
                const int tile = u / p.splits, s = u - tile * p.splits;
                const int kb0 = s * p.kb_per;
                const int kb1 = min(p.num_kb, kb0 + p.kb_per);
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BT);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sw = smem_u32(smem + stage * S::STAGE_BYTES);
                    const uint64_t da = umma_desc_kmajor_sw128(sw);
                    const uint64_t db = umma_desc_kmajor_sw128(sw + S::W_BYTES);
#pragma unroll
                    for (int k = 0; k < SK_BK / 16; ++k)
                        umma_bf16_ss(d_tmem, da + static_cast<uint64_t>(2 * k), db + static_cast<uint64_t>(2 * k), idesc,
                                     (kb != kb0 || k != 0) ? 1u : 0u);
                    umma_commit(&empty_bar[stage]);
                    if (++stage == SK_STAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit(&tmem_full[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1u;
            }
        }
    } else {
        // ===================== epilogue (4 warps): thread <-> one weight row n, registers <-> tokens
        griddep_wait();                                   // residual / workspace / counters belong to earlier kernels
        const int q = warp & 3;
        const int nl = q * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        const bool scaled = NORMED && p.scale_src != nullptr;
        float* const ssq_out = NORMED ? p.ssq_out : nullptr;
        if (scaled) {
            if (nl < 64) {
                float sc = 1.0f;
                if (nl < p.M) {
                    if (p.scale_parts == 0) sc = __ldcg(p.scale_src + nl);
                    else {
                        float t = 0.0f;
                        for (int j = 0; j < p.scale_parts; ++j) t += __ldcg(p.scale_src + j * 64 + nl);   // fixed order
                        sc = rsqrtf(t * p.scale_inv_d + p.scale_eps);
                    }
                }
                s_scale[nl] = sc;
            }
            epi_bar_sync();
        }
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int tile = u / p.splits, s = u - tile * p.splits;
            const int n = tile * SK_ROWS + nl;
            const bool n_ok = n < p.N;
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BT);
            float* part = p.partial + (static_cast<size_t>(tile) * p.splits + s) * (BT * SK_ROWS) + nl;
#pragma unroll 1
            for (int c = 0; c < BT / 32; ++c) {
                if (c * 32 >= p.M) break;
                uint32_t r[32];
                tmem_ld_32x32(t_addr + static_cast<uint32_t>(c * 32), r);
                tmem_ld_wait();
                if (p.splits == 1) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int m = c * 32 + j;
                        float sq = 0.0f;
                        if (n_ok && m < p.M) {
                            float v = __uint_as_float(r[j]);
                            if (scaled) v *= s_scale[m];
                            if (p.res) v += __bfloat162float(p.res[static_cast<long long>(m) * p.ldr + n]);
                            const __nv_bfloat16 ob = __float2bfloat16_rn(v);
                            p.C[static_cast<long long>(m) * p.ldc + n] = ob;
                            const float of = __bfloat162float(ob);
                            sq = of * of;
                        }
                        if (ssq_out) {
                            sq = warp_sum(sq);
                            if (lane == 0) s_red[q][m] = sq;
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c * 32 + j < p.M) part[(c * 32 + j) * SK_ROWS] = __uint_as_float(r[j]);
                }
            }
            tc_fence_before();
            mbar_arrive(&tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1u;
            if (p.splits == 1 && ssq_out) {
                epi_bar_sync();
                if (nl < p.M) ssq_out[tile * 64 + nl] = s_red[0][nl] + s_red[1][nl] + s_red[2][nl] + s_red[3][nl];
                epi_bar_sync();
            }
            if (p.splits > 1) {
                __threadfence();
                epi_bar_sync();
                if (nl == 0) {
                    const int prev = atomicAdd(&p.counters[tile], 1);
                    const int last = (prev == p.splits - 1) ? 1 : 0;
                    if (last) p.counters[tile] = 0;           // nobody else touches this counter again in this launch
                    *last_flag = last;
                }
                epi_bar_sync();
                if (*last_flag) {
                    __threadfence();
                    if (NORMED || n_ok) {
                        // fixed summation order s = 0..S-1 per element; 16 tokens x 4 slices of loads in flight
                        const float* base = p.partial + static_cast<size_t>(tile) * p.splits * (BT * SK_ROWS) + nl;
#pragma unroll 1
                        for (int m0 = 0; m0 < p.M; m0 += 16) {
                            float v[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = 0.0f;
                            int s2 = 0;
#pragma unroll 1
                            for (; s2 + 4 <= p.splits; s2 += 4) {
                                float t[4][16];
#pragma unroll
                                for (int i = 0; i < 4; ++i)
#pragma unroll
                                    for (int j = 0; j < 16; ++j)
                                        t[i][j] = n_ok ? __ldcg(base + (static_cast<size_t>(s2 + i) * BT + m0 + j) * SK_ROWS) : 0.0f;
#pragma unroll
                                for (int i = 0; i < 4; ++i)
#pragma unroll
                                    for (int j = 0; j < 16; ++j) v[j] += t[i][j];
                            }
#pragma unroll 1
                            for (; s2 < p.splits; ++s2) {
                                float t[16];
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    t[j] = n_ok ? __ldcg(base + (static_cast<size_t>(s2) * BT + m0 + j) * SK_ROWS) : 0.0f;
#pragma unroll
                                for (int j = 0; j < 16; ++j) v[j] += t[j];
                            }
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int m = m0 + j;
                                float sq = 0.0f;
                                if (n_ok && m < p.M) {
                                    float o = v[j];
                                    if (scaled) o *= s_scale[m];
                                    if (p.res) o += __bfloat162float(p.res[static_cast<long long>(m) * p.ldr + n]);
                                    const __nv_bfloat16 ob = __float2bfloat16_rn(o);
                                    p.C[static_cast<long long>(m) * p.ldc + n] = ob;
                                    const float of = __bfloat162float(ob);
                                    sq = of * of;
                                }
                                if (ssq_out) {
                                    sq = warp_sum(sq);
                                    if (lane == 0 && m < 64) s_red[q][m] = sq;
                                }
                            }
                        }
                    }
                    if (ssq_out) {
                        epi_bar_sync();
                        if (nl < p.M) ssq_out[tile * 64 + nl] = s_red[0][nl] + s_red[1][nl] + s_red[2][nl] + s_red[3][nl];
                        epi_bar_sync();
                    }
                }
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

template <int BT, int SK_STAGES>
int launch_skinny(const CUtensorMap& tw, const CUtensorMap& tx, const SkinnyParams& p, int grid, bool pdl, cudaStream_t stream) {
    using S = SkinnySmem<BT, SK_STAGES>;
    const bool normed = p.scale_src != nullptr || p.ssq_out != nullptr;
    auto kernel = normed ? gemm_skinny_tcgen05_kernel<BT, SK_STAGES, true> : gemm_skinny_tcgen05_kernel<BT, SK_STAGES, false>;
    static bool attr_set[2] = {false, false};
    if (!attr_set[normed]) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
        if (e != cudaSuccess) {
            set_last_error("cudaFuncSetAttribute(gemm_skinny<%d>): %s", BT, cudaGetErrorString(e));
            return ESB_ERR_CUDA;
        }
        attr_set[normed] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(SK_THREADS);
    cfg.dynamicSmemBytes = S::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, tw, tx, p);
    if (e != cudaSuccess) {
        set_last_error("gemm_skinny launch: %s", cudaGetErrorString(e));
        return ESB_ERR_CUDA;
    }
    return check_launch("gemm_skinny_tcgen05");
}

}  // namespace

namespace {
// tile (tile, kb) = W[tile*128 .. +128][kb*64 .. +64] (zero beyond N / K), 16 KiB contiguous, stored as the 128-byte-swizzled
// shared-memory image tcgen05 expects (16-byte chunk c of row r at chunk c ^ (r & 7)), so a 1-D bulk copy stages it.
// gu_f > 0: W is a packed [2 gu_f, K] gate|up weight and tile t holds gate rows t*64.. (tile rows 0..63) and up rows gu_f + t*64..
// (tile rows 64..127), so that one accumulator tile carries both factors of 64 SwiGLU outputs (gemm_chain.cu).
__global__ void retile_weight_kernel(const __nv_bfloat16* __restrict__ W, long long ldw, int N, int K, int num_kb,
                                     long long total_vec, uint4* __restrict__ out, const __nv_bfloat16* __restrict__ col_scale,
                                     int gu_f) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total_vec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int c8 = static_cast<int>(i & 7);
        const int r = static_cast<int>((i >> 3) & 127);
        const long long blk = i >> 10;
        const int kb = static_cast<int>(blk % num_kb);
        const long long tile = blk / num_kb;
        long long n = tile * SK_ROWS + r;
        if (gu_f > 0) {
            const long long col = tile * 64 + (r & 63);
            n = col < gu_f ? col + (r >= 64 ? gu_f : 0) : N;          // N: out of range -> zero row
        }
        const int k = kb * SK_BK + c8 * 8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (n < N && k + 8 <= K) v = __ldg(reinterpret_cast<const uint4*>(W + n * ldw + k));
        else if (n < N && k < K) {
            __nv_bfloat16 t[8];
            for (int j = 0; j < 8; ++j) t[j] = (k + j < K) ? W[n * ldw + k + j] : __float2bfloat16(0.0f);
            v = *reinterpret_cast<uint4*>(t);
        }
        if (col_scale && n < N && k < K) {                               // fold a norm weight g[k] into the column: bf16(W * g)
            __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(&v);
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (k + j < K) e[j] = __float2bfloat16_rn(__bfloat162float(e[j]) * __bfloat162float(col_scale[k + j]));
        }
        out[(i & ~7ll) | static_cast<long long>(c8 ^ (r & 7))] = v;     // SWIZZLE_128B image of the smem tile
    }
}
}  // namespace

long long retiled_weight_elems(int N, int K) {
    return static_cast<long long>((N + SK_ROWS - 1) / SK_ROWS) * ((K + SK_BK - 1) / SK_BK) * SK_ROWS * SK_BK;
}

int retile_weight_bf16(const void* W, long long ldw, int N, int K, void* out, cudaStream_t stream) {
    return retile_weight_scaled_bf16(W, ldw, N, K, nullptr, out, stream);
}

int retile_weight_scaled_bf16(const void* W, long long ldw, int N, int K, const void* col_scale, void* out, cudaStream_t stream) {
    if (N <= 0 || K <= 0 || !W || !out || (ldw & 7) || (reinterpret_cast<uintptr_t>(W) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) {
        set_last_error("retile_weight_bf16: bad arguments (N=%d K=%d ldw=%lld)", N, K, ldw);
        return ESB_ERR_INVALID_ARG;
    }
    const int num_kb = (K + SK_BK - 1) / SK_BK;
    const long long total_vec = retiled_weight_elems(N, K) / 8;
    const int grid = device_sm_count() * 8;
    retile_weight_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(W), ldw, N, K, num_kb, total_vec,
                                                   reinterpret_cast<uint4*>(out), reinterpret_cast<const __nv_bfloat16*>(col_scale), 0);
    return check_launch("retile_weight");
}

long long retiled_gateup_elems(int f, int K) {
    return static_cast<long long>((f + 63) / 64) * ((K + SK_BK - 1) / SK_BK) * SK_ROWS * SK_BK;
}

int retile_weight_gateup_bf16(const void* W, long long ldw, int f, int K, const void* col_scale, void* out, cudaStream_t stream) {
    if (f <= 0 || K <= 0 || !W || !out || (ldw & 7) || (reinterpret_cast<uintptr_t>(W) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) {
        set_last_error("retile_weight_gateup_bf16: bad arguments (f=%d K=%d ldw=%lld)", f, K, ldw);
        return ESB_ERR_INVALID_ARG;
    }
    const int num_kb = (K + SK_BK - 1) / SK_BK;
    const long long total_vec = retiled_gateup_elems(f, K) / 8;
    const int grid = device_sm_count() * 8;
    retile_weight_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(W), ldw, 2 * f, K, num_kb, total_vec,
                                                   reinterpret_cast<uint4*>(out), reinterpret_cast<const __nv_bfloat16*>(col_scale), f);
    return check_launch("retile_weight_gateup");
}

long long gemm_skinny_workspace_bytes(int M, int N, int K) {
    const int bt = M <= 32 ? 32 : 64;
    const long long n_tiles = (N + SK_ROWS - 1) / SK_ROWS;
    return SK_COUNTER_BYTES + n_tiles * 16 * bt * SK_ROWS * 4;     // 16 = largest split factor ever chosen
}

int gemm_skinny_bf16(const void* X, long long ldx, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                     const void* residual, long long ldr, void* workspace, long long workspace_bytes, int force_splits,
                     int w_tiled, cudaStream_t stream) {
    return gemm_skinny_normed_bf16(X, ldx, W, ldw, C, ldc, M, N, K, residual, ldr, workspace, workspace_bytes, force_splits, w_tiled,
                                   nullptr, 0, 0.0f, nullptr, stream);
}

int gemm_skinny_normed_bf16(const void* X, long long ldx, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                            const void* residual, long long ldr, void* workspace, long long workspace_bytes, int force_splits,
                            int w_tiled, const float* scale_src, int scale_parts, float eps, float* ssq_out, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return ESB_OK;
    if (scale_parts < 0 || scale_parts > 1024) {
        set_last_error("gemm_skinny_normed_bf16: scale_parts=%d out of range", scale_parts);
        return ESB_ERR_INVALID_ARG;
    }
    if (M > 64 || K <= 0 || !X || !W || !C) {
        set_last_error("gemm_skinny_bf16: bad arguments (M=%d (max 64) N=%d K=%d)", M, N, K);
        return ESB_ERR_INVALID_ARG;
    }
    if ((ldx & 7) || (ldw & 7) || (reinterpret_cast<uintptr_t>(X) & 15) || (reinterpret_cast<uintptr_t>(W) & 15)) {
        set_last_error("gemm_skinny_bf16: X/W must be 16-byte aligned with leading dimensions that are multiples of 8");
        return ESB_ERR_INVALID_ARG;
    }
    const int bt = M <= 32 ? 32 : 64;
    const int sms = device_sm_count();
    const int n_tiles = (N + SK_ROWS - 1) / SK_ROWS;
    const int num_kb = (K + SK_BK - 1) / SK_BK;
    if (n_tiles * 4 > SK_COUNTER_BYTES) {
        set_last_error("gemm_skinny_bf16: N=%d too large for the counter block", N);
        return ESB_ERR_UNSUPPORTED;
    }
    const long long part_bytes_per_split = static_cast<long long>(n_tiles) * bt * SK_ROWS * 4;
    long long max_splits_ws = 1;
    if (workspace && workspace_bytes > SK_COUNTER_BYTES)
        max_splits_ws = (workspace_bytes - SK_COUNTER_BYTES) / part_bytes_per_split;
    // Split factor from a measured model of this kernel on B200 (tools/skinny_bench.py, profiles/r01_skinny_bench.txt):
    // two CTAs are resident per SM; a CTA streams one 16 KiB k-block in ~0.36 us when few CTAs run and in
    // ~0.78 us * (resident CTAs / slots) once HBM is the limit; a split costs ~1.5 us + 0.3 us per slice of reduction tail.
    const int slots = 2 * sms;
    int best_s = 1;
    if (force_splits > 0) {
        best_s = force_splits > 16 ? 16 : force_splits;
        if (best_s > num_kb) best_s = num_kb;
        if (best_s > max_splits_ws) best_s = static_cast<int>(max_splits_ws);
        while (best_s > 1 && static_cast<long long>(best_s - 1) * ((num_kb + best_s - 1) / best_s) >= num_kb) --best_s;
    } else {
        double best_cost = -1.0;
        for (int s = 1; s <= 16 && s <= num_kb; ++s) {
            const int kb_per = (num_kb + s - 1) / s;
            if (static_cast<long long>(s - 1) * kb_per >= num_kb) continue;      // an empty slice
            if (s > 1 && (kb_per < 4 || s > max_splits_ws)) continue;
            const long long units = static_cast<long long>(n_tiles) * s;
            const long long rounds = (units + slots - 1) / slots;
            const double conc = static_cast<double>(units < slots ? units : slots) / slots;
            const double t_kb = 0.78 * conc > 0.36 ? 0.78 * conc : 0.36;
            const double cost = rounds * kb_per * t_kb + (s > 1 ? 1.5 + 0.3 * s : 0.0);
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_s = s; }
        }
    }
    const int kb_per = (num_kb + best_s - 1) / best_s;
    if (static_cast<long long>(best_s - 1) * kb_per >= num_kb || best_s > 16 || (best_s > 1 && best_s > max_splits_ws)) {
        set_last_error("gemm_skinny_bf16: invalid split factor %d (num_kb=%d, workspace allows %lld)", best_s, num_kb, max_splits_ws);
        return best_s > max_splits_ws ? ESB_ERR_WORKSPACE : ESB_ERR_INVALID_ARG;
    }
    CUtensorMap tw, tx;
    int rc = (w_tiled & 1) ? make_tma_2d_bf16(&tw, W, static_cast<long long>(n_tiles) * num_kb * SK_ROWS, SK_BK, SK_BK, SK_ROWS)
                     : make_tma_2d_bf16(&tw, W, N, K, ldw, SK_ROWS);
    if (rc) return rc;
    rc = make_tma_2d_bf16(&tx, X, M, K, ldx, bt);
    if (rc) return rc;
    SkinnyParams p;
    p.M = M; p.N = N; p.K = K;
    p.C = reinterpret_cast<__nv_bfloat16*>(C); p.ldc = ldc;
    p.res = reinterpret_cast<const __nv_bfloat16*>(residual); p.ldr = ldr;
    p.n_tiles = n_tiles; p.splits = best_s; p.kb_per = kb_per; p.num_kb = num_kb; p.w_tiled = (w_tiled & 1) ? 1 : 0;
    p.counters = reinterpret_cast<int*>(workspace);
    p.partial = workspace ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + SK_COUNTER_BYTES) : nullptr;
    p.w_tiles = reinterpret_cast<const uint8_t*>(W);
    p.scale_src = scale_src; p.scale_parts = scale_parts; p.scale_inv_d = 1.0f / static_cast<float>(K); p.scale_eps = eps;
    p.ssq_out = ssq_out;
    const long long units = static_cast<long long>(n_tiles) * best_s;
    const int grid = static_cast<int>(units < slots ? units : slots);
    const bool pdl = (w_tiled & 2) != 0;          // caller promises W is not written by the kernel launched just before
    return bt == 32 ? launch_skinny<32, 4>(tw, tx, p, grid, pdl, stream) : launch_skinny<64, 4>(tw, tx, p, grid, pdl, stream);
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// esb_gemm_skinny_static_bf16: the same kernel with its split-K workspace in a file-scope device array instead of a caller
// buffer - for callers that cannot allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds SK_STATIC_BYTES until exit.
constexpr long long SK_STATIC_PARTIAL_BYTES = 8ll << 20;
constexpr long long SK_STATIC_BYTES = SK_COUNTER_BYTES + SK_STATIC_PARTIAL_BYTES;        // 16 KiB counters + 8 MiB fp32 partials
constexpr int SK_STATIC_MAX_N = 16384, SK_STATIC_MAX_K = 16384;
constexpr int SK_STATIC_MAX_DEVICES = 64;

}  // namespace esb

// In host code this name is only a handle for cudaGetSymbolAddress - never a device address.
static __device__ __align__(256) unsigned char g_esb_skinny_scratch[esb::SK_STATIC_BYTES];

namespace esb {

namespace {
// device address of g_esb_skinny_scratch on the CURRENT device, resolved once per device
int skinny_static_scratch(void** out) {
    static void* ptrs[SK_STATIC_MAX_DEVICES] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= SK_STATIC_MAX_DEVICES) {
        set_last_error("gemm_skinny_static_bf16: cudaGetDevice -> %d (%s)", dev, cudaGetErrorString(e));
        return ESB_ERR_CUDA;
    }
    void* ptr = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (!ptr) {
        e = cudaGetSymbolAddress(&ptr, g_esb_skinny_scratch);
        if (e != cudaSuccess || !ptr) {
            set_last_error("gemm_skinny_static_bf16: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            (void)cudaGetLastError();
            return ESB_ERR_CUDA;
        }
        __atomic_store_n(&ptrs[dev], ptr, __ATOMIC_RELEASE);      // same value from every thread that races here
    }
    *out = ptr;
    return ESB_OK;
}
}  // namespace

long long gemm_skinny_static_scratch_bytes() { return SK_STATIC_BYTES; }

int gemm_skinny_static_bf16(const void* X, long long ldx, const void* W, long long ldw, void* C, long long ldc, int M, int N, int K,
                            const void* residual, long long ldr, int force_splits, int w_tiled, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return ESB_OK;
    if (M > 64 || N > SK_STATIC_MAX_N || K > SK_STATIC_MAX_K) {
        set_last_error("gemm_skinny_static_bf16: M=%d N=%d K=%d exceeds the static scratch limits (M <= 64, N <= %d, K <= %d); "
                       "use esb_gemm_skinny_bf16 with a caller workspace", M, N, K, SK_STATIC_MAX_N, SK_STATIC_MAX_K);
        return ESB_ERR_UNSUPPORTED;
    }
    void* scratch = nullptr;
    int rc = skinny_static_scratch(&scratch);
    if (rc) return rc;
    // the tile counters must start at zero; the kernel leaves them at zero, but an aborted launch would not - clear them every call
    // (a memset node, no allocation).  The partials need no clearing: every slot is written before it is read.
    cudaError_t e = cudaMemsetAsync(scratch, 0, SK_COUNTER_BYTES, stream);
    if (e != cudaSuccess) {
        set_last_error("gemm_skinny_static_bf16: cudaMemsetAsync: %s", cudaGetErrorString(e));
        return ESB_ERR_CUDA;
    }
    // no programmatic dependent launch (w_tiled bit 1 dropped): the kernel must not start before the memset has finished
    return gemm_skinny_normed_bf16(X, ldx, W, ldw, C, ldc, M, N, K, residual, ldr, scratch, SK_STATIC_BYTES, force_splits, w_tiled & 1,
                                   nullptr, 0, 0.0f, nullptr, stream);
}

}  // namespace esb
