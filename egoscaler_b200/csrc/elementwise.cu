// HBM-bound element-wise / gather kernels of the LLaMA trunk and the splice (bf16 I/O, fp32 math, 128-bit accesses).
//
// rope        : HF apply_rotary_pos_emb / rotate_half (modeling_llama.py:138-168), cos/sin tables cast to bf16 like
//               LlamaRotaryEmbedding.forward (:117-135); positions = arange + past (:394-397).
// silu_mul    : HF LlamaMLP act_fn(gate) * up (:171-184).
// gelu        : nn.GELU() of point_proj in training (pointllm/model/pointllm.py:72-79).
// embed_splice: embed_tokens + point-token splice (pointllm/model/pointllm.py:107,134-171) without the Python loop.
#include <cstdlib>
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

static inline int ew_grid(long long items, int block) {
    const long long want = (items + block - 1) / block;
    static const int per_sm = [] { const char* e = getenv("ESB_EW_CTAS_PER_SM"); return e ? atoi(e) : 64; }();   // ncu sweep 8..uncapped: 64 is best (silu_bwd 97.7 -> 88.1 us, fwd 80.2 -> 76.9)
    const long long cap = static_cast<long long>(device_sm_count()) * per_sm;
    const long long g = want < cap ? want : cap;
    return static_cast<int>(g < 1 ? 1 : g);
}

// ------------------------------------------------------------------------------------------------ RoPE
// buf rows = B*L tokens, row stride ld; heads [0, n_heads) start at column head*hd (q and k heads laid out
// contiguously).  One thread handles 8 rotation pairs (i..i+7 and i+hd/2..).  sign=+1 forward, -1 transposed.
// Optional KV-cache write: rows of K (heads [kv_head0, kv_head0+kv_heads)) and V (v_col0..) go to
// kcache/vcache [B, cache_len, kv_heads*hd] at position pos0 + l.
__global__ void __launch_bounds__(256) rope_kernel(__nv_bfloat16* __restrict__ buf, long long ld, int L, int n_heads, int hd,
                                                   const __nv_bfloat16* __restrict__ cos_tab, const __nv_bfloat16* __restrict__ sin_tab,
                                                   int pos0, const int* __restrict__ pos_ptr, float sign, long long rows,
                                                   __nv_bfloat16* __restrict__ kcache, __nv_bfloat16* __restrict__ vcache,
                                                   int kv_head0, int kv_heads, long long v_col0, int cache_len) {
    griddep_launch();
    griddep_wait();
    const int half = hd >> 1;
    const int vec_per_head = half >> 3;
    if (pos_ptr) pos0 += *pos_ptr;                 // device-side position (CUDA-graph replay of the decode step)
    const long long total = rows * n_heads * vec_per_head;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const int vi = static_cast<int>(idx % vec_per_head);
        const long long t = idx / vec_per_head;
        const int head = static_cast<int>(t % n_heads);
        const long long row = t / n_heads;
        const int l = static_cast<int>(row % L);
        const int pos = pos0 + l;
        __nv_bfloat16* p = buf + row * ld + static_cast<long long>(head) * hd + vi * 8;
        const uint4 lo = *reinterpret_cast<const uint4*>(p);
        const uint4 hi = *reinterpret_cast<const uint4*>(p + half);
        const uint4 cu = __ldg(reinterpret_cast<const uint4*>(cos_tab + static_cast<long long>(pos) * half + vi * 8));
        const uint4 su = __ldg(reinterpret_cast<const uint4*>(sin_tab + static_cast<long long>(pos) * half + vi * 8));
        const uint32_t a[4] = {lo.x, lo.y, lo.z, lo.w}, b[4] = {hi.x, hi.y, hi.z, hi.w};
        const uint32_t c[4] = {cu.x, cu.y, cu.z, cu.w}, sn[4] = {su.x, su.y, su.z, su.w};
        uint32_t o1[4], o2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 x1 = unpack_bf16x2(a[j]), x2 = unpack_bf16x2(b[j]), cc = unpack_bf16x2(c[j]);
            float2 ss = unpack_bf16x2(sn[j]);
            ss.x *= sign; ss.y *= sign;
            // bf16 arithmetic like the reference: each product and the sum round to bf16
            const float r1x = bf16_round(x1.x * cc.x) - bf16_round(x2.x * ss.x);
            const float r1y = bf16_round(x1.y * cc.y) - bf16_round(x2.y * ss.y);
            const float r2x = bf16_round(x2.x * cc.x) + bf16_round(x1.x * ss.x);
            const float r2y = bf16_round(x2.y * cc.y) + bf16_round(x1.y * ss.y);
            o1[j] = pack_bf16x2(r1x, r1y);
            o2[j] = pack_bf16x2(r2x, r2y);
        }
        const uint4 w1 = make_uint4(o1[0], o1[1], o1[2], o1[3]), w2 = make_uint4(o2[0], o2[1], o2[2], o2[3]);
        *reinterpret_cast<uint4*>(p) = w1;
        *reinterpret_cast<uint4*>(p + half) = w2;
        if (kcache && head >= kv_head0) {
            const int kh = head - kv_head0;
            const long long b_idx = row / L;
            const long long crow = (b_idx * cache_len + pos) * (static_cast<long long>(kv_heads) * hd) + static_cast<long long>(kh) * hd + vi * 8;
            *reinterpret_cast<uint4*>(kcache + crow) = w1;
            *reinterpret_cast<uint4*>(kcache + crow + half) = w2;
            const __nv_bfloat16* vp = buf + row * ld + v_col0 + static_cast<long long>(kh) * hd + vi * 8;
            *reinterpret_cast<uint4*>(vcache + crow) = *reinterpret_cast<const uint4*>(vp);
            *reinterpret_cast<uint4*>(vcache + crow + half) = *reinterpret_cast<const uint4*>(vp + half);
        }
    }
}

int rope(void* buf, long long ld, long long rows, int L, int n_heads, int hd, const void* cos_tab, const void* sin_tab,
         int pos0, int inverse, void* kcache, void* vcache, int kv_head0, int kv_heads, long long v_col0, int cache_len,
         const int* pos_ptr, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if ((hd & 15) || (ld & 7)) { set_last_error("rope: head_dim must be a multiple of 16 and ld of 8"); return ESB_ERR_INVALID_ARG; }
    const long long total = rows * n_heads * (hd / 16);
    launch_chain(rope_kernel, dim3(ew_grid(total, 256)), dim3(256), 0, s, static_cast<__nv_bfloat16*>(buf), ld, L, n_heads, hd,
                 static_cast<const __nv_bfloat16*>(cos_tab), static_cast<const __nv_bfloat16*>(sin_tab), pos0, pos_ptr,
                 inverse ? -1.0f : 1.0f, rows, static_cast<__nv_bfloat16*>(kcache), static_cast<__nv_bfloat16*>(vcache), kv_head0,
                 kv_heads, v_col0, cache_len);
    return check_launch("rope");
}

// ------------------------------------------------------------------------------------------------ SwiGLU
// gu [rows, 2f] (gate | up)  ->  act [rows, f] = bf16(bf16(silu(g)) * u)
__global__ void __launch_bounds__(256) silu_mul_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act,
                                                           long long rows, int f) {
    griddep_launch();
    griddep_wait();
    const int fv = f >> 3;
    const long long total = rows * fv;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const long long row = idx / fv;
        const int c = static_cast<int>(idx % fv) * 8;
        const uint4 g = *reinterpret_cast<const uint4*>(gu + row * 2 * f + c);
        const uint4 u = *reinterpret_cast<const uint4*>(gu + row * 2 * f + f + c);
        const uint32_t gi[4] = {g.x, g.y, g.z, g.w}, ui[4] = {u.x, u.y, u.z, u.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 gv = unpack_bf16x2(gi[j]), uv = unpack_bf16x2(ui[j]);
            o[j] = pack_bf16x2(bf16_round(silu(gv.x)) * uv.x, bf16_round(silu(gv.y)) * uv.y);
        }
        *reinterpret_cast<uint4*>(act + row * f + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
}
// dgu [rows, 2f]: dgate = dact * u * silu'(g), dup = dact * silu(g)
__global__ void __launch_bounds__(256) silu_mul_bwd_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ dact,
                                                           __nv_bfloat16* __restrict__ dgu, long long rows, int f) {
    const int fv = f >> 3;
    const long long total = rows * fv;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const long long row = idx / fv;
        const int c = static_cast<int>(idx % fv) * 8;
        const uint4 g = *reinterpret_cast<const uint4*>(gu + row * 2 * f + c);
        const uint4 u = *reinterpret_cast<const uint4*>(gu + row * 2 * f + f + c);
        const uint4 d = *reinterpret_cast<const uint4*>(dact + row * f + c);
        const uint32_t gi[4] = {g.x, g.y, g.z, g.w}, ui[4] = {u.x, u.y, u.z, u.w}, di[4] = {d.x, d.y, d.z, d.w};
        uint32_t og[4], ou[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 gv = unpack_bf16x2(gi[j]), uv = unpack_bf16x2(ui[j]), dv = unpack_bf16x2(di[j]);
            const float s0 = 1.0f / (1.0f + __expf(-gv.x)), s1 = 1.0f / (1.0f + __expf(-gv.y));
            og[j] = pack_bf16x2(dv.x * uv.x * s0 * (1.0f + gv.x * (1.0f - s0)), dv.y * uv.y * s1 * (1.0f + gv.y * (1.0f - s1)));
            ou[j] = pack_bf16x2(dv.x * gv.x * s0, dv.y * gv.y * s1);
        }
        *reinterpret_cast<uint4*>(dgu + row * 2 * f + c) = make_uint4(og[0], og[1], og[2], og[3]);
        *reinterpret_cast<uint4*>(dgu + row * 2 * f + f + c) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
    }
}

int silu_mul_fwd(const void* gu, void* act, long long rows, int f, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if (f & 7) { set_last_error("silu_mul: f must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    launch_chain(silu_mul_fwd_kernel, dim3(ew_grid(rows * (f / 8), 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(gu),
                 static_cast<__nv_bfloat16*>(act), rows, f);
    return check_launch("silu_mul_fwd");
}
int silu_mul_bwd(const void* gu, const void* dact, void* dgu, long long rows, int f, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if (f & 7) { set_last_error("silu_mul: f must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    silu_mul_bwd_kernel<<<ew_grid(rows * (f / 8), 256), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(gu), static_cast<const __nv_bfloat16*>(dact),
                                                                    static_cast<__nv_bfloat16*>(dgu), rows, f);
    return check_launch("silu_mul_bwd");
}

// ------------------------------------------------------------------------------------------------ GELU (erf)
// mode 0: y = gelu(z);  mode 1: y = dy * gelu'(z)   (n % 8 == 0)
__global__ void __launch_bounds__(256) gelu_kernel(const __nv_bfloat16* __restrict__ z, const __nv_bfloat16* __restrict__ dy,
                                                   __nv_bfloat16* __restrict__ y, long long nvec, int mode) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const uint4 zu = reinterpret_cast<const uint4*>(z)[i];
        uint4 du = make_uint4(0, 0, 0, 0);
        if (mode) du = reinterpret_cast<const uint4*>(dy)[i];
        const uint32_t zi[4] = {zu.x, zu.y, zu.z, zu.w}, di[4] = {du.x, du.y, du.z, du.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 zv = unpack_bf16x2(zi[j]);
            if (mode) {
                const float2 dv = unpack_bf16x2(di[j]);
                o[j] = pack_bf16x2(dv.x * gelu_erf_grad(zv.x), dv.y * gelu_erf_grad(zv.y));
            } else {
                o[j] = pack_bf16x2(gelu_erf(zv.x), gelu_erf(zv.y));
            }
        }
        reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}
int gelu(const void* z, const void* dy, void* y, long long n, int backward, cudaStream_t s) {
    if (n <= 0) return ESB_OK;
    if (n & 7) { set_last_error("gelu: n must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    gelu_kernel<<<ew_grid(n / 8, 256), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(z), static_cast<const __nv_bfloat16*>(dy),
                                                   static_cast<__nv_bfloat16*>(y), n / 8, backward);
    return check_launch("gelu");
}

}  // namespace esb
