// Normalisation kernels (HBM-bound; fp32 statistics, bf16 I/O, 128-bit accesses).
//
// RMSNorm  : HF LlamaRMSNorm.forward (transformers modeling_llama.py:53-67), reached from the reference through
//            pointllm/model/pointllm.py:173-178.  Order of operations kept: x32 * rsqrt(mean(x32^2)+eps) is cast to
//            the activation dtype (bf16) BEFORE it is multiplied by the (bf16) weight.
// LayerNorm: nn.LayerNorm(eps=1e-5) of the PointBERT blocks (pointbert/point_encoder.py:62,66,142); the per-block
//            `x + pos` (point_encoder.py:96-97) is fused in as an optional prologue that also emits the sum.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"
#include <cstdlib>

namespace esb {

// ESB_RMSNORM_GENERIC=1 (read at every call): the general two-pass kernels also for d = 16 x threads (tests compare the two)
static int rmsnorm_fast_path() {
    const char* e = getenv("ESB_RMSNORM_GENERIC");
    return (e && e[0] == '1') ? 0 : 1;
}

ESB_DEVICE float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    float t = (l < nw) ? red[l] : 0.0f;
    t = warp_sum(t);
    return t;
}

// one CTA per row; d % 8 == 0
__global__ void __launch_bounds__(256) rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                                          __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out,
                                                          int d, long long ldx, long long ldy, float eps, int fast) {
    __shared__ float red[32];
    griddep_launch();
    griddep_wait();
    const long long row = blockIdx.x;
    const uint4* xp = reinterpret_cast<const uint4*>(x + row * ldx);
    const int nv = d >> 3;
    if (fast && nv == 2 * static_cast<int>(blockDim.x)) {
        // d = 16 x threads (4096 with 256 threads): the row stays in registers between the two passes and the weight vectors
        // travel under the reduction; same formulas as the general path below (results agree to fp32 round-off)
        const int t = threadIdx.x, T = blockDim.x;
        const uint4 u[2] = {xp[t], xp[t + T]};
        const uint4* wp = reinterpret_cast<const uint4*>(w);
        const uint4 wv2[2] = {__ldg(wp + t), __ldg(wp + t + T)};
        float ss = 0.0f;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const float2 a = unpack_bf16x2(u[k].x), b = unpack_bf16x2(u[k].y), c = unpack_bf16x2(u[k].z), e = unpack_bf16x2(u[k].w);
            ss += a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y + c.x * c.x + c.y * c.y + e.x * e.x + e.y * e.y;
        }
        ss = block_sum(ss, red);
        const float rstd = rsqrtf(ss / static_cast<float>(d) + eps);
        if (t == 0 && rstd_out) rstd_out[row] = rstd;
        uint4* yp = reinterpret_cast<uint4*>(y + row * ldy);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint32_t xi[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
            const uint32_t wi[4] = {wv2[k].x, wv2[k].y, wv2[k].z, wv2[k].w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 xv = unpack_bf16x2(xi[j]);
                const float2 wv = unpack_bf16x2(wi[j]);
                o[j] = pack_bf16x2(wv.x * bf16_round(xv.x * rstd), wv.y * bf16_round(xv.y * rstd));
            }
            yp[t + k * T] = make_uint4(o[0], o[1], o[2], o[3]);
        }
        return;
    }
    float ss = 0.0f;
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
        const uint4 u = xp[i];
        const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
        ss += a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y + c.x * c.x + c.y * c.y + e.x * e.x + e.y * e.y;
    }
    ss = block_sum(ss, red);
    const float rstd = rsqrtf(ss / static_cast<float>(d) + eps);
    if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
    const uint4* wp = reinterpret_cast<const uint4*>(w);
    uint4* yp = reinterpret_cast<uint4*>(y + row * ldy);
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
        const uint4 u = xp[i];
        const uint4 ww = __ldg(wp + i);
        const uint32_t xi[4] = {u.x, u.y, u.z, u.w};
        const uint32_t wi[4] = {ww.x, ww.y, ww.z, ww.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 xv = unpack_bf16x2(xi[j]);
            const float2 wv = unpack_bf16x2(wi[j]);
            o[j] = pack_bf16x2(wv.x * bf16_round(xv.x * rstd), wv.y * bf16_round(xv.y * rstd));
        }
        yp[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// Backward.  dx = rstd * (g - xhat * mean(g * xhat)), g = dy * w, xhat = x * rstd;  optional  dx += dres;
// optional dw[j] += sum_rows dy[j] * xhat[j] (fp32 atomics, one per column per CTA).  Persistent over rows.
__global__ void __launch_bounds__(256) rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                          const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd,
                                                          const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                                                          float* __restrict__ dw, long long rows, int d, int fast) {
    __shared__ float red[32];
    extern __shared__ float dw_acc[];            // [d] when dw != null
    const int nv = d >> 3;
    if (dw) {
        for (int i = threadIdx.x; i < d; i += blockDim.x) dw_acc[i] = 0.0f;
        __syncthreads();
    }
    const uint4* wp = reinterpret_cast<const uint4*>(w);
    if (fast && nv == 2 * static_cast<int>(blockDim.x)) {
        // d = 16 x threads: weight vectors loaded once per CTA, the row's x / dy stay in registers between the two passes and the
        // NEXT row is requested before this row's reduction; same formulas as the general path (results agree to fp32 round-off)
        const int t = threadIdx.x, T = blockDim.x;
        const uint4 wq[2] = {__ldg(wp + t), __ldg(wp + t + T)};
        long long row = blockIdx.x;
        uint4 xq[2], gq[2];
        float rs = 0.0f;
        if (row < rows) {
            const uint4* xp = reinterpret_cast<const uint4*>(x + row * d);
            const uint4* gp = reinterpret_cast<const uint4*>(dy + row * d);
            xq[0] = xp[t]; xq[1] = xp[t + T]; gq[0] = gp[t]; gq[1] = gp[t + T];
            rs = rstd[row];
        }
        while (row < rows) {
            const long long nrow = row + gridDim.x;
            uint4 nx[2], ng[2], rq[2];
            float nrs = 0.0f;
            rq[0] = rq[1] = make_uint4(0, 0, 0, 0);
            if (dres) {
                const uint4* rp = reinterpret_cast<const uint4*>(dres + row * d);
                rq[0] = rp[t]; rq[1] = rp[t + T];
            }
            if (nrow < rows) {
                const uint4* xp = reinterpret_cast<const uint4*>(x + nrow * d);
                const uint4* gp = reinterpret_cast<const uint4*>(dy + nrow * d);
                nx[0] = xp[t]; nx[1] = xp[t + T]; ng[0] = gp[t]; ng[1] = gp[t + T];
                nrs = rstd[nrow];
            }
            float dot = 0.0f;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const uint32_t xi[4] = {xq[k].x, xq[k].y, xq[k].z, xq[k].w}, gi[4] = {gq[k].x, gq[k].y, gq[k].z, gq[k].w};
                const uint32_t wi[4] = {wq[k].x, wq[k].y, wq[k].z, wq[k].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 xv = unpack_bf16x2(xi[j]), gv = unpack_bf16x2(gi[j]), wv = unpack_bf16x2(wi[j]);
                    dot += gv.x * wv.x * xv.x * rs + gv.y * wv.y * xv.y * rs;
                }
            }
            dot = block_sum(dot, red) / static_cast<float>(d);
            uint4* op = reinterpret_cast<uint4*>(dx + row * d);
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int i = t + k * T;
                const uint32_t xi[4] = {xq[k].x, xq[k].y, xq[k].z, xq[k].w}, gi[4] = {gq[k].x, gq[k].y, gq[k].z, gq[k].w};
                const uint32_t wi[4] = {wq[k].x, wq[k].y, wq[k].z, wq[k].w}, ri[4] = {rq[k].x, rq[k].y, rq[k].z, rq[k].w};
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 xv = unpack_bf16x2(xi[j]), gv = unpack_bf16x2(gi[j]), wv = unpack_bf16x2(wi[j]);
                    const float2 rv = unpack_bf16x2(ri[j]);
                    const float xh0 = xv.x * rs, xh1 = xv.y * rs;
                    const float d0 = rs * (gv.x * wv.x - xh0 * dot) + rv.x;
                    const float d1 = rs * (gv.y * wv.y - xh1 * dot) + rv.y;
                    o[j] = pack_bf16x2(d0, d1);
                    if (dw) {
                        dw_acc[i * 8 + j * 2] += gv.x * xh0;          // column owned by this thread: no race
                        dw_acc[i * 8 + j * 2 + 1] += gv.y * xh1;
                    }
                }
                op[i] = make_uint4(o[0], o[1], o[2], o[3]);
            }
            row = nrow;
            xq[0] = nx[0]; xq[1] = nx[1]; gq[0] = ng[0]; gq[1] = ng[1];
            rs = nrs;
        }
        if (dw) {
            __syncthreads();
            for (int i = threadIdx.x; i < d; i += blockDim.x) atomicAdd(dw + i, dw_acc[i]);
        }
        return;
    }
    for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
        const uint4* xp = reinterpret_cast<const uint4*>(x + row * d);
        const uint4* gp = reinterpret_cast<const uint4*>(dy + row * d);
        const float rs = rstd[row];
        float dot = 0.0f;
        for (int i = threadIdx.x; i < nv; i += blockDim.x) {
            const uint4 xu = xp[i], gu = gp[i], wu = __ldg(wp + i);
            const uint32_t xi[4] = {xu.x, xu.y, xu.z, xu.w}, gi[4] = {gu.x, gu.y, gu.z, gu.w}, wi[4] = {wu.x, wu.y, wu.z, wu.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 xv = unpack_bf16x2(xi[j]), gv = unpack_bf16x2(gi[j]), wv = unpack_bf16x2(wi[j]);
                dot += gv.x * wv.x * xv.x * rs + gv.y * wv.y * xv.y * rs;
            }
        }
        dot = block_sum(dot, red) / static_cast<float>(d);
        uint4* op = reinterpret_cast<uint4*>(dx + row * d);
        const uint4* rp = dres ? reinterpret_cast<const uint4*>(dres + row * d) : nullptr;
        for (int i = threadIdx.x; i < nv; i += blockDim.x) {
            const uint4 xu = xp[i], gu = gp[i], wu = __ldg(wp + i);
            uint4 ru = make_uint4(0, 0, 0, 0);
            if (rp) ru = rp[i];
            const uint32_t xi[4] = {xu.x, xu.y, xu.z, xu.w}, gi[4] = {gu.x, gu.y, gu.z, gu.w}, wi[4] = {wu.x, wu.y, wu.z, wu.w};
            const uint32_t ri[4] = {ru.x, ru.y, ru.z, ru.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 xv = unpack_bf16x2(xi[j]), gv = unpack_bf16x2(gi[j]), wv = unpack_bf16x2(wi[j]);
                const float2 rv = unpack_bf16x2(ri[j]);
                const float xh0 = xv.x * rs, xh1 = xv.y * rs;
                const float d0 = rs * (gv.x * wv.x - xh0 * dot) + rv.x;
                const float d1 = rs * (gv.y * wv.y - xh1 * dot) + rv.y;
                o[j] = pack_bf16x2(d0, d1);
                if (dw) {
                    dw_acc[i * 8 + j * 2] += gv.x * xh0;          // column owned by this thread: no race
                    dw_acc[i * 8 + j * 2 + 1] += gv.y * xh1;
                }
            }
            op[i] = make_uint4(o[0], o[1], o[2], o[3]);
        }
    }
    if (dw) {
        __syncthreads();
        for (int i = threadIdx.x; i < d; i += blockDim.x) atomicAdd(dw + i, dw_acc[i]);
    }
}

// LayerNorm forward, one warp per row.  Optional prologue: s = x + add (add broadcast over `add_period` rows,
// i.e. add row = row % add_period), s written to sum_out (may alias x).  d % 2 == 0.
__global__ void __launch_bounds__(128) layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ add,
                                                            long long add_period, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, __nv_bfloat16* __restrict__ sum_out,
                                                            __nv_bfloat16* __restrict__ y, long long rows, int d, float eps) {
    const long long row = static_cast<long long>(blockIdx.x) * 4 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const int n2 = d >> 1;
    const uint32_t* xp = reinterpret_cast<const uint32_t*>(x + row * d);
    const uint32_t* ap = add ? reinterpret_cast<const uint32_t*>(add + (row % add_period) * d) : nullptr;
    uint32_t* sp = sum_out ? reinterpret_cast<uint32_t*>(sum_out + row * d) : nullptr;
    constexpr int MAXP = 24;                       // up to d = 1536 kept in registers
    float2 v[MAXP];
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < MAXP; ++k) {
        const int i = lane + k * 32;
        if (i < n2) {
            float2 t = unpack_bf16x2(xp[i]);
            if (ap) {
                const float2 a = unpack_bf16x2(ap[i]);
                const uint32_t packed = pack_bf16x2(t.x + a.x, t.y + a.y);   // the sum is a bf16 tensor downstream
                t = unpack_bf16x2(packed);
                if (sp) sp[i] = packed;
            }
            v[k] = t;
            s += t.x + t.y;
        }
    }
    const float mean = warp_sum(s) / static_cast<float>(d);
    float q = 0.0f;
#pragma unroll
    for (int k = 0; k < MAXP; ++k) {
        const int i = lane + k * 32;
        if (i < n2) {
            const float a = v[k].x - mean, b = v[k].y - mean;
            q += a * a + b * b;
        }
    }
    const float rstd = rsqrtf(warp_sum(q) / static_cast<float>(d) + eps);
    uint32_t* yp = reinterpret_cast<uint32_t*>(y + row * d);
#pragma unroll
    for (int k = 0; k < MAXP; ++k) {
        const int i = lane + k * 32;
        if (i < n2) {
            const float2 g = __ldg(reinterpret_cast<const float2*>(gamma) + i);
            const float2 b = __ldg(reinterpret_cast<const float2*>(beta) + i);
            yp[i] = pack_bf16x2((v[k].x - mean) * rstd * g.x + b.x, (v[k].y - mean) * rstd * g.y + b.y);
        }
    }
}

int rmsnorm_fwd(const void* x, long long ldx, const void* w, void* y, long long ldy, float* rstd, long long rows, int d,
                float eps, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if ((d & 7) || (ldx & 7) || (ldy & 7)) { set_last_error("rmsnorm_fwd: d/ld must be multiples of 8 (d=%d)", d); return ESB_ERR_INVALID_ARG; }
    int threads = ((d / 8 + 31) / 32) * 32;
    threads = threads > 256 ? 256 : (threads < 32 ? 32 : threads);
    launch_chain(rmsnorm_fwd_kernel, dim3(static_cast<unsigned>(rows)), dim3(threads), 0, s,
                 static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(w), static_cast<__nv_bfloat16*>(y), rstd, d, ldx, ldy, eps,
                 rmsnorm_fast_path());
    return check_launch("rmsnorm_fwd");
}

int rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dres, void* dx, float* dw,
                long long rows, int d, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if (d & 7) { set_last_error("rmsnorm_bwd: d must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    int threads = ((d / 8 + 31) / 32) * 32;
    threads = threads > 256 ? 256 : (threads < 32 ? 32 : threads);
    const int fast = rmsnorm_fast_path() && (d >> 3) == 2 * threads;
    // register-resident path: 2 CTAs fit an SM (114 registers), each walks its rows with the next row in flight
    const long long cap = static_cast<long long>(device_sm_count()) * (fast ? 2 : 8);
    const int grid = static_cast<int>(rows < cap ? rows : cap);
    const size_t smem = dw ? static_cast<size_t>(d) * 4 : 0;
    rmsnorm_bwd_kernel<<<grid, threads, smem, s>>>(static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
                                                   static_cast<const __nv_bfloat16*>(w), rstd, static_cast<const __nv_bfloat16*>(dres),
                                                   static_cast<__nv_bfloat16*>(dx), dw, rows, d, fast);
    return check_launch("rmsnorm_bwd");
}

int layernorm_fwd(const void* x, const void* add, long long add_period, const float* gamma, const float* beta, void* sum_out,
                  void* y, long long rows, int d, float eps, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if ((d & 1) || d > 1536) { set_last_error("layernorm_fwd: d must be even and <= 1536 (d=%d)", d); return ESB_ERR_INVALID_ARG; }
    if (add && add_period <= 0) add_period = rows;
    layernorm_fwd_kernel<<<static_cast<unsigned>((rows + 3) / 4), 128, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(add), add_period, gamma, beta,
        static_cast<__nv_bfloat16*>(sum_out), static_cast<__nv_bfloat16*>(y), rows, d, eps);
    return check_launch("layernorm_fwd");
}

}  // namespace esb
