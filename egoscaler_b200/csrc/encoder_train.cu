// Kernels that only the TRAINABLE point encoder needs (--unfreeze_pc_encoder, model_arch.py:36-38,121-122):
// the reference then runs PointBERT in train mode, i.e. BatchNorm1d with BATCH statistics in the mini-PointNet
// (pointbert/dvae.py:194-205), DropPath in the 12 blocks (point_encoder.py:62,74-75,134) and autograd through
// LayerNorm / max-pool / GELU.  Everything here is HBM-bound byte moving (no tensor cores on purpose); the GEMMs of
// the backward reuse gemm_tcgen05.cu.
//
//   layernorm_bwd      nn.LayerNorm backward (eps 1e-5): dx, dgamma, dbeta            point_encoder.py:58,64,142
//   colsum2            out0[c] = sum_r a[r,c], out1[c] = sum_r a[r,c]*b[r,c]          BatchNorm statistics fwd + bwd
//   bn_coeffs          per-channel coefficients from those sums (+ running stats)     dvae.py:196,202
//   affine_cols        out = act(A[c]*x + B[c]*y + C[c])                              BN apply fwd / bwd
//   relu_mask          out = y > 0 ? dy : 0                                           nn.ReLU backward
//   group_maxpool_bwd  torch.max(dim) backward (first arg-max), plain and concat form dvae.py:217-220
//   scale_rows_add     out = x + s[row / rows_per_batch] * y                          DropPath fwd / bwd
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

namespace {

ESB_DEVICE float block_sum256(float v, float* red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.0f;
    const int nw = blockDim.x >> 5;
    for (int w = 0; w < nw; ++w) t += red[w];
    return t;
}

// One warp per row (d <= 1536, even).  x is the LayerNorm INPUT; mean / rstd are recomputed (cheaper than saving them).
// dgamma / dbeta: per-CTA shared accumulation, then one fp32 atomic per column per CTA.
__global__ void __launch_bounds__(128) layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                            const float* __restrict__ gamma, const __nv_bfloat16* __restrict__ dres,
                                                            __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, long long rows, int d, float eps) {
    extern __shared__ float acc[];                  // [2*d]: dgamma | dbeta partials of this CTA
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 2 * d; i += blockDim.x) acc[i] = 0.0f;
    __syncthreads();
    const int nv = d >> 1;
    for (long long row = static_cast<long long>(blockIdx.x) * 4 + warp; row < rows; row += static_cast<long long>(gridDim.x) * 4) {
        const uint32_t* xp = reinterpret_cast<const uint32_t*>(x + row * d);
        const uint32_t* gp = reinterpret_cast<const uint32_t*>(dy + row * d);
        float s = 0.0f, ss = 0.0f;
        for (int i = lane; i < nv; i += 32) {
            const float2 v = unpack_bf16x2(xp[i]);
            s += v.x + v.y;
        }
        const float mean = warp_sum(s) / d;
        for (int i = lane; i < nv; i += 32) {
            const float2 v = unpack_bf16x2(xp[i]);
            ss += (v.x - mean) * (v.x - mean) + (v.y - mean) * (v.y - mean);
        }
        const float rstd = rsqrtf(warp_sum(ss) / d + eps);
        float c1 = 0.0f, c2 = 0.0f;                  // sum(g), sum(g * xhat), g = dy * gamma
        for (int i = lane; i < nv; i += 32) {
            const float2 v = unpack_bf16x2(xp[i]), g = unpack_bf16x2(gp[i]);
            const float g0 = g.x * gamma[2 * i], g1 = g.y * gamma[2 * i + 1];
            const float h0 = (v.x - mean) * rstd, h1 = (v.y - mean) * rstd;
            c1 += g0 + g1;
            c2 += g0 * h0 + g1 * h1;
        }
        c1 = warp_sum(c1) / d;
        c2 = warp_sum(c2) / d;
        uint32_t* op = reinterpret_cast<uint32_t*>(dx + row * d);
        const uint32_t* rp = dres ? reinterpret_cast<const uint32_t*>(dres + row * d) : nullptr;
        for (int i = lane; i < nv; i += 32) {
            const float2 v = unpack_bf16x2(xp[i]), g = unpack_bf16x2(gp[i]);
            const float h0 = (v.x - mean) * rstd, h1 = (v.y - mean) * rstd;
            float o0 = rstd * (g.x * gamma[2 * i] - c1 - h0 * c2);
            float o1 = rstd * (g.y * gamma[2 * i + 1] - c1 - h1 * c2);
            if (rp) { const float2 r = unpack_bf16x2(rp[i]); o0 += r.x; o1 += r.y; }
            op[i] = pack_bf16x2(o0, o1);
            atomicAdd(&acc[2 * i], g.x * h0);
            atomicAdd(&acc[2 * i + 1], g.y * h1);
            atomicAdd(&acc[d + 2 * i], g.x);
            atomicAdd(&acc[d + 2 * i + 1], g.y);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
        if (dgamma) atomicAdd(dgamma + i, acc[i]);
        if (dbeta) atomicAdd(dbeta + i, acc[d + i]);
    }
}

// Column reductions over a row-major bf16 matrix: CTA = 32 columns x 8 row lanes, grid.y slices the rows.  DETERMINISTIC: the
// slices leave their partial sums in a device scratch and the last slice of a column block to arrive adds them up in slice order
// (an fp32 atomicAdd per slice made the train-mode BatchNorm statistics - and through bf16 roundings the whole forward of a
// --unfreeze_pc_encoder step - differ in the last bits from run to run: bench.py dp_check, 3 % on the embedding gradient).
// The scratch is shared by all launches of a GPU: launches on different streams must not overlap (the trainer uses one stream).
constexpr int CS2_MAX_SLICES = 512, CS2_MAX_COLS = 4096;
__device__ float g_cs2_part[2][CS2_MAX_SLICES][CS2_MAX_COLS];       // 16 MiB
__device__ unsigned int g_cs2_count[CS2_MAX_COLS / 32];

__global__ void __launch_bounds__(256) colsum2_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                                      long long lda, long long ldb, long long M, int N, float* __restrict__ out0,
                                                      float* __restrict__ out1) {
    __shared__ float s0[8][33], s1[8][33];
    __shared__ bool s_last;
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + cx;
    float t0 = 0.0f, t1 = 0.0f;
    if (col < N) {
        for (long long r = static_cast<long long>(blockIdx.y) * 8 + ry; r < M; r += static_cast<long long>(gridDim.y) * 8) {
            const float av = __bfloat162float(a[r * lda + col]);
            t0 += av;
            t1 += av * (b ? __bfloat162float(b[r * ldb + col]) : av);
        }
    }
    s0[ry][cx] = t0;
    s1[ry][cx] = t1;
    __syncthreads();
    if (ry == 0 && col < N) {
#pragma unroll
        for (int j = 1; j < 8; ++j) { t0 += s0[j][cx]; t1 += s1[j][cx]; }
        if (gridDim.y == 1) {
            out0[col] += t0;
            if (out1) out1[col] += t1;
        } else {
            g_cs2_part[0][blockIdx.y][col] = t0;
            g_cs2_part[1][blockIdx.y][col] = t1;
        }
    }
    if (gridDim.y == 1) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(&g_cs2_count[blockIdx.x], 1u);
        s_last = prev == gridDim.y - 1;
        if (s_last) g_cs2_count[blockIdx.x] = 0;                    // ready for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // 8 row lanes take the slices round-robin, then a fixed-order sum over the lanes
    float u0 = 0.0f, u1 = 0.0f;
    if (col < N)
        for (int y = ry; y < static_cast<int>(gridDim.y); y += 8) {
            u0 += __ldcg(&g_cs2_part[0][y][col]);
            u1 += __ldcg(&g_cs2_part[1][y][col]);
        }
    s0[ry][cx] = u0;
    s1[ry][cx] = u1;
    __syncthreads();
    if (ry == 0 && col < N) {
#pragma unroll
        for (int j = 1; j < 8; ++j) { u0 += s0[j][cx]; u1 += s1[j][cx]; }
        out0[col] += u0;
        if (out1) out1[col] += u1;
    }
}

// mode 0 (forward): sums = {sum z, sum z^2} over R rows -> A = gamma*rstd, C = beta - mean*A (B unused = 0); also stores
//   mean / rstd for the backward and updates the running statistics (momentum, unbiased variance) like nn.BatchNorm1d.
// mode 1 (backward): sums = {sum dy, sum dy*z}; with the saved mean / rstd:  dz = A*dy + B*z + C,
//   A = gamma*rstd, B = -gamma*rstd^2*S/R, C = (-gamma*rstd*sum_dy + gamma*rstd^2*mean*S)/R, S = rstd*(sum_dyz - mean*sum_dy);
//   dgamma += S, dbeta += sum_dy.
__global__ void bn_coeffs_kernel(int mode, const float* __restrict__ sum0, const float* __restrict__ sum1, long long R, int C,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float momentum,
                                 float* __restrict__ mean_io, float* __restrict__ rstd_io, float* __restrict__ running_mean,
                                 float* __restrict__ running_var, float* __restrict__ A, float* __restrict__ B, float* __restrict__ Cc,
                                 float* __restrict__ dgamma, float* __restrict__ dbeta) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float invR = 1.0f / static_cast<float>(R);
    if (mode == 0) {
        const float mean = sum0[c] * invR;
        float var = sum1[c] * invR - mean * mean;
        var = var < 0.0f ? 0.0f : var;
        const float rstd = rsqrtf(var + eps);
        mean_io[c] = mean;
        rstd_io[c] = rstd;
        const float a = gamma[c] * rstd;
        A[c] = a;
        if (B) B[c] = 0.0f;
        Cc[c] = beta[c] - mean * a;
        if (running_mean) {
            const float unb = R > 1 ? var * static_cast<float>(R) / static_cast<float>(R - 1) : var;
            running_mean[c] = (1.0f - momentum) * running_mean[c] + momentum * mean;
            running_var[c] = (1.0f - momentum) * running_var[c] + momentum * unb;
        }
    } else {
        const float mean = mean_io[c], rstd = rstd_io[c], g = gamma[c];
        const float sdy = sum0[c];
        const float S = rstd * (sum1[c] - mean * sdy);
        A[c] = g * rstd;
        B[c] = -g * rstd * rstd * S * invR;
        Cc[c] = (-g * rstd * sdy + g * rstd * rstd * mean * S) * invR;
        if (dgamma) dgamma[c] += S;
        if (dbeta) dbeta[c] += sdy;
    }
}

__global__ void __launch_bounds__(256) affine_cols_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                                          const float* __restrict__ A, const float* __restrict__ B,
                                                          const float* __restrict__ Cc, __nv_bfloat16* __restrict__ out,
                                                          long long total, int C, int relu) {
    for (long long i = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) * 2; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x * 2) {
        const int c = static_cast<int>(i % C);
        const float2 xv = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + i));
        float o0 = A[c] * xv.x + Cc[c], o1 = A[c + 1] * xv.y + Cc[c + 1];
        if (y) {
            const float2 yv = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(y + i));
            o0 += B[c] * yv.x;
            o1 += B[c + 1] * yv.y;
        }
        if (relu) { o0 = fmaxf(o0, 0.0f); o1 = fmaxf(o1, 0.0f); }
        *reinterpret_cast<uint32_t*>(out + i) = pack_bf16x2(o0, o1);
    }
}

__global__ void __launch_bounds__(256) relu_mask_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                                                        __nv_bfloat16* __restrict__ out, long long total) {
    for (long long i = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) * 2; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x * 2) {
        const float2 g = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(dy + i));
        const float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(y + i));
        *reinterpret_cast<uint32_t*>(out + i) = pack_bf16x2(v.x > 0.0f ? g.x : 0.0f, v.y > 0.0f ? g.y : 0.0f);
    }
}

// f [groups*n, C] (the max-pool input).  plain : df[r,c] = (r == first argmax) ? dpool[g,c] : 0
// concat: dcat [groups*n, 2C] (grad of cat(global.expand, f)):  df[r,c] = dcat[r, C+c] + (r == argmax ? sum_k dcat[k,c] : 0)
__global__ void __launch_bounds__(256) group_maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ f, const __nv_bfloat16* __restrict__ dpool,
                                                                const __nv_bfloat16* __restrict__ dcat, __nv_bfloat16* __restrict__ df,
                                                                long long groups, int n, int C) {
    const long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (idx >= groups * C) return;
    const long long g = idx / C;
    const int c = static_cast<int>(idx % C);
    const __nv_bfloat16* fp = f + g * n * C + c;
    float best = -INFINITY;
    int bi = 0;
    for (int k = 0; k < n; ++k) {
        const float v = __bfloat162float(fp[static_cast<long long>(k) * C]);
        if (v > best) { best = v; bi = k; }
    }
    if (dcat) {
        const __nv_bfloat16* dp = dcat + g * n * 2 * C;
        float tot = 0.0f;
        for (int k = 0; k < n; ++k) tot += __bfloat162float(dp[static_cast<long long>(k) * 2 * C + c]);
        for (int k = 0; k < n; ++k) {
            const float local = __bfloat162float(dp[static_cast<long long>(k) * 2 * C + C + c]);
            df[(g * n + k) * C + c] = __float2bfloat16_rn(local + (k == bi ? tot : 0.0f));
        }
    } else {
        const __nv_bfloat16 gv = dpool[g * C + c];
        for (int k = 0; k < n; ++k) df[(g * n + k) * C + c] = (k == bi) ? gv : __float2bfloat16(0.0f);
    }
}

__global__ void __launch_bounds__(256) scale_rows_add_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                                             const float* __restrict__ s, __nv_bfloat16* __restrict__ out,
                                                             long long rows, int d, long long rows_per_batch) {
    const long long total = rows * d;
    for (long long i = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) * 2; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x * 2) {
        const float sc = s[(i / d) / rows_per_batch];
        const float2 yv = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(y + i));
        float o0 = sc * yv.x, o1 = sc * yv.y;
        if (x) {
            const float2 xv = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + i));
            o0 += xv.x;
            o1 += xv.y;
        }
        *reinterpret_cast<uint32_t*>(out + i) = pack_bf16x2(o0, o1);
    }
}

int ew_grid(long long total_pairs) {
    const long long cap = static_cast<long long>(device_sm_count()) * 16;
    const long long want = (total_pairs + 255) / 256;
    return static_cast<int>(want < 1 ? 1 : (want < cap ? want : cap));
}

}  // namespace

int layernorm_bwd(const void* dy, const void* x, const float* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                  long long rows, int d, float eps, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if ((d & 1) || d > 1536 || !dy || !x || !gamma || !dx) {
        set_last_error("layernorm_bwd: bad arguments (d=%d must be even and <= 1536)", d);
        return ESB_ERR_INVALID_ARG;
    }
    const long long cap = static_cast<long long>(device_sm_count()) * 4;
    const long long want = (rows + 3) / 4;
    const int grid = static_cast<int>(want < cap ? want : cap);
    layernorm_bwd_kernel<<<grid, 128, static_cast<size_t>(2 * d) * 4, s>>>(
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x), gamma, static_cast<const __nv_bfloat16*>(dres),
        static_cast<__nv_bfloat16*>(dx), dgamma, dbeta, rows, d, eps);
    return check_launch("layernorm_bwd");
}

int colsum2_bf16(const void* a, long long lda, const void* b, long long ldb, long long M, int N, float* out0, float* out1,
                 cudaStream_t s) {
    if (M <= 0 || N <= 0) return ESB_OK;
    if (!a || !out0) { set_last_error("colsum2_bf16: null argument"); return ESB_ERR_INVALID_ARG; }
    if (N > CS2_MAX_COLS) { set_last_error("colsum2_bf16: N=%d columns, at most %d", N, CS2_MAX_COLS); return ESB_ERR_UNSUPPORTED; }
    long long slices = (M + 8 * 64 - 1) / (8 * 64);
    if (slices > 512) slices = 512;
    if (slices < 1) slices = 1;
    dim3 grid((N + 31) / 32, static_cast<unsigned>(slices));
    colsum2_kernel<<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b), lda, ldb, M, N, out0,
                                        out1);
    return check_launch("colsum2");
}

int bn_coeffs(int mode, const float* sum0, const float* sum1, long long R, int C, const float* gamma, const float* beta, float eps,
              float momentum, float* mean_io, float* rstd_io, float* running_mean, float* running_var, float* A, float* B, float* Cc,
              float* dgamma, float* dbeta, cudaStream_t s) {
    if (C <= 0) return ESB_OK;
    if (!sum0 || !sum1 || !gamma || !mean_io || !rstd_io || !A || !Cc || (mode == 0 && !beta) || (mode == 1 && !B) || R <= 0) {
        set_last_error("bn_coeffs: bad arguments (mode=%d R=%lld C=%d)", mode, R, C);
        return ESB_ERR_INVALID_ARG;
    }
    bn_coeffs_kernel<<<(C + 127) / 128, 128, 0, s>>>(mode, sum0, sum1, R, C, gamma, beta, eps, momentum, mean_io, rstd_io, running_mean,
                                                     running_var, A, B, Cc, dgamma, dbeta);
    return check_launch("bn_coeffs");
}

int affine_cols(const void* x, const void* y, const float* A, const float* B, const float* Cc, void* out, long long rows, int C,
                int relu, cudaStream_t s) {
    if (rows <= 0 || C <= 0) return ESB_OK;
    if ((C & 1) || !x || !A || !Cc || !out || (y && !B)) { set_last_error("affine_cols: bad arguments (C=%d)", C); return ESB_ERR_INVALID_ARG; }
    const long long total = rows * C;
    affine_cols_kernel<<<ew_grid(total / 2), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(y), A, B,
                                                          Cc, static_cast<__nv_bfloat16*>(out), total, C, relu);
    return check_launch("affine_cols");
}

int relu_mask(const void* dy, const void* y, void* out, long long n, cudaStream_t s) {
    if (n <= 0) return ESB_OK;
    if (n & 1) { set_last_error("relu_mask: n must be even"); return ESB_ERR_INVALID_ARG; }
    relu_mask_kernel<<<ew_grid(n / 2), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y),
                                                    static_cast<__nv_bfloat16*>(out), n);
    return check_launch("relu_mask");
}

int group_maxpool_bwd(const void* f, const void* dpool, const void* dcat, void* df, long long groups, int n, int C, cudaStream_t s) {
    if (groups <= 0) return ESB_OK;
    if (!f || !df || (!dpool == !dcat)) { set_last_error("group_maxpool_bwd: exactly one of dpool / dcat"); return ESB_ERR_INVALID_ARG; }
    const long long total = groups * C;
    group_maxpool_bwd_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(
        static_cast<const __nv_bfloat16*>(f), static_cast<const __nv_bfloat16*>(dpool), static_cast<const __nv_bfloat16*>(dcat),
        static_cast<__nv_bfloat16*>(df), groups, n, C);
    return check_launch("group_maxpool_bwd");
}

int scale_rows_add(const void* x, const void* y, const float* scale, void* out, long long rows, int d, long long rows_per_batch,
                   cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if ((d & 1) || !y || !scale || !out || rows_per_batch <= 0) { set_last_error("scale_rows_add: bad arguments"); return ESB_ERR_INVALID_ARG; }
    scale_rows_add_kernel<<<ew_grid(rows * d / 2), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(y),
                                                               scale, static_cast<__nv_bfloat16*>(out), rows, d, rows_per_batch);
    return check_launch("scale_rows_add");
}

}  // namespace esb
