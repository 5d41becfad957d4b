// Gather / scatter / layout kernels: embedding + point-token splice, row copies, transpose, column sums,
// group max-pool of the mini-PointNet.
//
// embed_splice : pointllm/model/pointllm.py:107 (embed_tokens) + :134-171 (per-sample torch.cat splice), one pass.
// point_layout : the validation scans of pointllm.py:138-164 (counts of start/end/patch tokens, start position,
//                end-follows-start check) done on device; the host raises the reference's ValueErrors from it.
// group_maxpool_concat : pointbert/dvae.py:211-215 (max over the group, expand, concat global|local).
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

static inline int g_grid(long long items, int block) {
    const long long want = (items + block - 1) / block;
    const long long cap = static_cast<long long>(device_sm_count()) * 16;
    const long long g = want < cap ? want : cap;
    return static_cast<int>(g < 1 ? 1 : g);
}

// info[b] = {n_start, n_end, n_patch, first_start_pos, end_ok(1/0), patch_consecutive(1/0), first_patch_pos, 0}
__global__ void point_layout_kernel(const long long* __restrict__ ids, int B, int L, long long patch_id, long long start_id,
                                    long long end_id, int n_point, int* __restrict__ info) {
    const int b = blockIdx.x;
    const long long* row = ids + static_cast<long long>(b) * L;
    __shared__ int s_cnt[3];
    __shared__ int s_first_start, s_first_patch, s_last_patch;
    if (threadIdx.x == 0) { s_cnt[0] = s_cnt[1] = s_cnt[2] = 0; s_first_start = L; s_first_patch = L; s_last_patch = -1; }
    __syncthreads();
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
        const long long t = row[i];
        if (t == start_id) { atomicAdd(&s_cnt[0], 1); atomicMin(&s_first_start, i); }
        if (t == end_id) atomicAdd(&s_cnt[1], 1);
        if (t == patch_id) { atomicAdd(&s_cnt[2], 1); atomicMin(&s_first_patch, i); atomicMax(&s_last_patch, i); }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int* o = info + b * 8;
        o[0] = s_cnt[0]; o[1] = s_cnt[1]; o[2] = s_cnt[2];
        o[3] = s_cnt[0] ? s_first_start : -1;
        int end_ok = 1;
        if (s_cnt[0]) {
            const int e = s_first_start + n_point + 1;
            end_ok = (e < L && row[e] == end_id) ? 1 : 0;
        }
        o[4] = end_ok;
        o[5] = (s_cnt[2] == 0 || (s_last_patch - s_first_patch + 1) == s_cnt[2]) ? 1 : 0;
        o[6] = s_cnt[2] ? s_first_patch : -1;
        o[7] = 0;
    }
}

// x[b, l, :] = (splice_pos[b] >= 0 && splice_pos[b] <= l < splice_pos[b] + n_point) ? feats[b, l - splice_pos[b], :]
//                                                                                  : embed[ids[b, l], :]
__global__ void __launch_bounds__(256) embed_splice_fwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ embed,
                                                               const __nv_bfloat16* __restrict__ feats, const int* __restrict__ splice_pos,
                                                               int n_point, __nv_bfloat16* __restrict__ x, int B, int L, int d, long long vocab) {
    const int dv = d >> 3;
    const long long total = static_cast<long long>(B) * L * dv;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const int v = static_cast<int>(idx % dv);
        const long long tok = idx / dv;
        const int l = static_cast<int>(tok % L);
        const int b = static_cast<int>(tok / L);
        const int sp = (feats && splice_pos) ? splice_pos[b] : -1;
        const uint4* src;
        if (sp >= 0 && l >= sp && l < sp + n_point)
            src = reinterpret_cast<const uint4*>(feats + (static_cast<long long>(b) * n_point + (l - sp)) * d) + v;
        else {
            const long long id = ids[tok];
            if (id < 0 || id >= vocab) {                      // torch raises a device assert; the raw-pointer ABI stays in bounds
                reinterpret_cast<uint4*>(x + tok * d)[v] = make_uint4(0u, 0u, 0u, 0u);
                continue;
            }
            src = reinterpret_cast<const uint4*>(embed + id * d) + v;
        }
        reinterpret_cast<uint4*>(x + tok * d)[v] = __ldg(src);
    }
}

// backward: rows inside the point block go to dfeats (plain copy), all others are scatter-added (fp32 atomics)
// into dembed[ids] unless ids == padding_idx (nn.Embedding(padding_idx) keeps that row's gradient at zero).
__global__ void __launch_bounds__(256) embed_splice_bwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ dx,
                                                               const int* __restrict__ splice_pos, int n_point,
                                                               float* __restrict__ dembed, __nv_bfloat16* __restrict__ dfeats,
                                                               long long padding_idx, int B, int L, int d, long long vocab) {
    const int dv = d >> 3;
    const long long total = static_cast<long long>(B) * L * dv;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const int v = static_cast<int>(idx % dv);
        const long long tok = idx / dv;
        const int l = static_cast<int>(tok % L);
        const int b = static_cast<int>(tok / L);
        const int sp = splice_pos ? splice_pos[b] : -1;
        const uint4 g = reinterpret_cast<const uint4*>(dx + tok * d)[v];
        if (sp >= 0 && l >= sp && l < sp + n_point) {
            if (dfeats) reinterpret_cast<uint4*>(dfeats + (static_cast<long long>(b) * n_point + (l - sp)) * d)[v] = g;
        } else if (dembed) {
            const long long id = ids[tok];
            if (id == padding_idx || id < 0 || id >= vocab) continue;
            float* o = dembed + id * d + v * 8;
            const float2 a = unpack_bf16x2(g.x), bb = unpack_bf16x2(g.y), c = unpack_bf16x2(g.z), e = unpack_bf16x2(g.w);
            atomicAdd(o + 0, a.x); atomicAdd(o + 1, a.y); atomicAdd(o + 2, bb.x); atomicAdd(o + 3, bb.y);
            atomicAdd(o + 4, c.x); atomicAdd(o + 5, c.y); atomicAdd(o + 6, e.x); atomicAdd(o + 7, e.y);
        }
    }
}

// ---- sparse exchange of the embedding gradient (data-parallel training) -------------------------------
// Only the rows of embed_tokens named by this step's token ids receive gradient (<= B * (L - n_point) of 32262 rows), so the
// ranks exchange (token id, gradient row) pairs instead of all-reducing the dense [V, d] fp32 matrix.
// pack: compact the non-point rows of dx [B, L, d] to rows_out [B * rmax, d] (+ their token ids; unused slots carry `pad`).
__global__ void __launch_bounds__(128) embed_rows_pack_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ dx,
                                                              const int* __restrict__ splice_pos, int n_point, int rmax,
                                                              __nv_bfloat16* __restrict__ rows_out, long long* __restrict__ ids_out,
                                                              long long pad, int B, int L, int d) {
    const int r = blockIdx.x % rmax, b = blockIdx.x / rmax;
    const int sp = splice_pos ? splice_pos[b] : -1;
    const int l = (sp >= 0 && r >= sp) ? r + n_point : r;
    const bool valid = l < L;
    if (threadIdx.x == 0) ids_out[blockIdx.x] = valid ? ids[static_cast<long long>(b) * L + l] : pad;
    uint4* dst = reinterpret_cast<uint4*>(rows_out + static_cast<long long>(blockIdx.x) * d);
    const uint4* src = reinterpret_cast<const uint4*>(dx + (static_cast<long long>(b) * L + (valid ? l : 0)) * d);
    for (int v = threadIdx.x; v < (d >> 3); v += blockDim.x) dst[v] = valid ? src[v] : make_uint4(0u, 0u, 0u, 0u);
}
// scatter: dembed[id, :] += sum of the rows k with ids[k] == id, added in INCREASING k.  One CTA column-slice per row j does the
// work for its id iff j is the first occurrence; every rank runs this on the same all-gathered (ids, rows), so all ranks hold
// bit-identical gradients (an atomicAdd scatter would not: its summation order differs from rank to rank).
constexpr int SCATTER_MAX_ROWS = 1 << 16;
__global__ void __launch_bounds__(128) embed_scatter_rows_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ rows,
                                                                 float* __restrict__ dembed, long long pad, long long vocab, int R, int d) {
    __shared__ uint32_t match[SCATTER_MAX_ROWS / 32];
    __shared__ int dup;
    const int j = blockIdx.x;
    const long long id = ids[j];
    if (id == pad || id < 0 || id >= vocab) return;
    const int words = (R + 31) >> 5;
    if (threadIdx.x == 0) dup = 0;
    __syncthreads();
    for (int w = threadIdx.x; w < words; w += blockDim.x) {
        uint32_t bits = 0;
        const int k0 = w << 5;
#pragma unroll 4
        for (int i = 0; i < 32; ++i) {
            const int k = k0 + i;
            if (k < R && ids[k] == id) bits |= 1u << i;
        }
        match[w] = bits;
        if (k0 < j) {                                               // an earlier occurrence owns this id
            const uint32_t below = (j - k0 >= 32) ? 0xffffffffu : ((1u << (j - k0)) - 1u);
            if (bits & below) dup = 1;
        }
    }
    __syncthreads();
    if (dup) return;
    const int c0 = (blockIdx.y * blockDim.x + threadIdx.x) * 8;      // 8 columns per thread
    if (c0 >= d) return;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int w = j >> 5; w < words; ++w) {
        uint32_t bits = match[w];
        while (bits) {
            const int k = (w << 5) + __ffs(bits) - 1;
            bits &= bits - 1;
            const uint4 u = *reinterpret_cast<const uint4*>(rows + static_cast<long long>(k) * d + c0);
            const float2 a = unpack_bf16x2(u.x), b2 = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
            acc[0] += a.x; acc[1] += a.y; acc[2] += b2.x; acc[3] += b2.y; acc[4] += c.x; acc[5] += c.y; acc[6] += e.x; acc[7] += e.y;
        }
    }
    float4* o = reinterpret_cast<float4*>(dembed + id * d + c0);
    float4 o0 = o[0], o1 = o[1];
    o0.x += acc[0]; o0.y += acc[1]; o0.z += acc[2]; o0.w += acc[3];
    o1.x += acc[4]; o1.y += acc[5]; o1.z += acc[6]; o1.w += acc[7];
    o[0] = o0; o[1] = o1;
}

// The same sum with the O(R^2) id scan replaced by a stable sort done by the caller: sorted_ids ascending, order[p] = original row
// index, rows of one id in increasing original index (stable).  CTA p works iff it starts a run of equal ids; with the 8 x 1464
// rows of an 8-GPU step the scanning kernel above took 2.8 ms at the very end of the step.
__global__ void __launch_bounds__(128) embed_scatter_sorted_kernel(const long long* __restrict__ sorted_ids, const long long* __restrict__ order,
                                                                   const __nv_bfloat16* __restrict__ rows, float* __restrict__ dembed,
                                                                   long long pad, long long vocab, int R, int d) {
    const int p = blockIdx.x;
    const long long id = sorted_ids[p];
    if (id == pad || id < 0 || id >= vocab) return;
    if (p > 0 && sorted_ids[p - 1] == id) return;
    const int c0 = (blockIdx.y * blockDim.x + threadIdx.x) * 8;      // 8 columns per thread
    if (c0 >= d) return;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int q = p; q < R && sorted_ids[q] == id; ++q) {
        const long long k = order[q];
        const uint4 u = *reinterpret_cast<const uint4*>(rows + k * d + c0);
        const float2 a = unpack_bf16x2(u.x), b2 = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
        acc[0] += a.x; acc[1] += a.y; acc[2] += b2.x; acc[3] += b2.y; acc[4] += c.x; acc[5] += c.y; acc[6] += e.x; acc[7] += e.y;
    }
    float4* o = reinterpret_cast<float4*>(dembed + id * d + c0);
    float4 o0 = o[0], o1 = o[1];
    o0.x += acc[0]; o0.y += acc[1]; o0.z += acc[2]; o0.w += acc[3];
    o1.x += acc[4]; o1.y += acc[5]; o1.z += acc[6]; o1.w += acc[7];
    o[0] = o0; o[1] = o1;
}

// Batched row copy (16-byte vectors): dst[b, r, :ncols] = src[b, r, :ncols]; strides in elements of `elem_bytes`.
__global__ void __launch_bounds__(256) copy_rows_kernel(const uint8_t* __restrict__ src, long long src_bstride, long long src_ld,
                                                        uint8_t* __restrict__ dst, long long dst_bstride, long long dst_ld,
                                                        int batches, int rows, int row_bytes) {
    const int rv = row_bytes >> 4;
    const long long total = static_cast<long long>(batches) * rows * rv;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const int v = static_cast<int>(idx % rv);
        const long long t = idx / rv;
        const int r = static_cast<int>(t % rows);
        const long long b = t / rows;
        const uint4 u = *reinterpret_cast<const uint4*>(src + b * src_bstride + r * src_ld + v * 16);
        *reinterpret_cast<uint4*>(dst + b * dst_bstride + r * dst_ld + v * 16) = u;
    }
}

// out[c, r] = in[r, c]  (bf16), 32x32 tiles through padded shared memory
__global__ void __launch_bounds__(256) transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in, long long ld_in,
                                                             __nv_bfloat16* __restrict__ out, long long ld_out, int R, int C) {
    __shared__ __nv_bfloat16 tile[32][34];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int i = ty; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + tx;
        tile[i][tx] = (r < R && c < C) ? in[static_cast<long long>(r) * ld_in + c] : __float2bfloat16(0.0f);
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + tx;
        if (c < C && r < R) out[static_cast<long long>(c) * ld_out + r] = tile[tx][i];
    }
}

// out[n] (+)= sum_m in[m, n]   (bf16 in, fp32 out); grid.x over column blocks of 64, grid.y splits rows
__global__ void __launch_bounds__(256) colsum_kernel(const __nv_bfloat16* __restrict__ in, long long ld, int M, int N,
                                                     float* __restrict__ out) {
    __shared__ float part[4][64];
    const int c = blockIdx.x * 64 + (threadIdx.x & 63);
    const int rlane = threadIdx.x >> 6;                 // 0..3
    float acc = 0.0f;
    if (c < N)
        for (int r = blockIdx.y * 4 + rlane; r < M; r += gridDim.y * 4) acc += __bfloat162float(in[static_cast<long long>(r) * ld + c]);
    part[rlane][threadIdx.x & 63] = acc;
    __syncthreads();
    if (rlane == 0 && c < N) atomicAdd(out + c, part[0][threadIdx.x] + part[1][threadIdx.x] + part[2][threadIdx.x] + part[3][threadIdx.x]);
}

// f [G*n, C] -> mode 0: out [G, C] = max over n ; mode 1: out [G*n, 2C] = [max over n (broadcast) | f]
__global__ void __launch_bounds__(128) group_maxpool_kernel(const __nv_bfloat16* __restrict__ f, __nv_bfloat16* __restrict__ out,
                                                            int n, int C, int mode) {
    const long long g = blockIdx.x;
    const int c2 = C >> 1;
    for (int c = threadIdx.x; c < c2; c += blockDim.x) {
        float mx = -INFINITY, my = -INFINITY;
        for (int j = 0; j < n; ++j) {
            const float2 v = unpack_bf16x2(reinterpret_cast<const uint32_t*>(f + (g * n + j) * C)[c]);
            mx = fmaxf(mx, v.x); my = fmaxf(my, v.y);
        }
        const uint32_t m = pack_bf16x2(mx, my);
        if (mode == 0) {
            reinterpret_cast<uint32_t*>(out + g * C)[c] = m;
        } else {
            for (int j = 0; j < n; ++j) {
                uint32_t* o = reinterpret_cast<uint32_t*>(out + (g * n + j) * 2 * C);
                o[c] = m;
                o[c2 + c] = reinterpret_cast<const uint32_t*>(f + (g * n + j) * C)[c];
            }
        }
    }
}

int point_layout(const long long* ids, int B, int L, long long patch_id, long long start_id, long long end_id, int n_point,
                 int* info, cudaStream_t s) {
    if (B <= 0) return ESB_OK;
    point_layout_kernel<<<B, 256, 0, s>>>(ids, B, L, patch_id, start_id, end_id, n_point, info);
    return check_launch("point_layout");
}
int embed_splice_fwd(const long long* ids, const void* embed, const void* feats, const int* splice_pos, int n_point, void* x,
                     int B, int L, int d, long long vocab, cudaStream_t s) {
    if (B <= 0 || L <= 0) return ESB_OK;
    if (d & 7) { set_last_error("embed_splice: d must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    embed_splice_fwd_kernel<<<g_grid(static_cast<long long>(B) * L * (d / 8), 256), 256, 0, s>>>(
        ids, static_cast<const __nv_bfloat16*>(embed), static_cast<const __nv_bfloat16*>(feats), splice_pos, n_point,
        static_cast<__nv_bfloat16*>(x), B, L, d, vocab);
    return check_launch("embed_splice_fwd");
}
int embed_splice_bwd(const long long* ids, const void* dx, const int* splice_pos, int n_point, float* dembed, void* dfeats,
                     long long padding_idx, int B, int L, int d, long long vocab, cudaStream_t s) {
    if (B <= 0 || L <= 0) return ESB_OK;
    if (d & 7) { set_last_error("embed_splice: d must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    embed_splice_bwd_kernel<<<g_grid(static_cast<long long>(B) * L * (d / 8), 256), 256, 0, s>>>(
        ids, static_cast<const __nv_bfloat16*>(dx), splice_pos, n_point, dembed, static_cast<__nv_bfloat16*>(dfeats), padding_idx, B, L, d, vocab);
    return check_launch("embed_splice_bwd");
}
int embed_rows_pack(const long long* ids, const void* dx, const int* splice_pos, int n_point, int rmax, void* rows_out,
                    long long* ids_out, long long pad, int B, int L, int d, cudaStream_t s) {
    if (B <= 0 || rmax <= 0) return ESB_OK;
    if (d & 7) { set_last_error("embed_rows_pack: d must be a multiple of 8"); return ESB_ERR_INVALID_ARG; }
    embed_rows_pack_kernel<<<B * rmax, 128, 0, s>>>(ids, static_cast<const __nv_bfloat16*>(dx), splice_pos, n_point, rmax,
                                                    static_cast<__nv_bfloat16*>(rows_out), ids_out, pad, B, L, d);
    return check_launch("embed_rows_pack");
}
int embed_scatter_rows(const long long* ids, const void* rows, float* dembed, long long pad, long long vocab, int R, int d, cudaStream_t s) {
    if (R <= 0) return ESB_OK;
    if ((d & 7) || R > SCATTER_MAX_ROWS) {
        set_last_error("embed_scatter_rows: d must be a multiple of 8 and R <= %d (R=%d d=%d)", SCATTER_MAX_ROWS, R, d);
        return ESB_ERR_INVALID_ARG;
    }
    dim3 grid(R, (d / 8 + 127) / 128);
    embed_scatter_rows_kernel<<<grid, 128, 0, s>>>(ids, static_cast<const __nv_bfloat16*>(rows), dembed, pad, vocab, R, d);
    return check_launch("embed_scatter_rows");
}
int embed_scatter_sorted(const long long* sorted_ids, const long long* order, const void* rows, float* dembed, long long pad,
                         long long vocab, int R, int d, cudaStream_t s) {
    if (R <= 0) return ESB_OK;
    if (d & 7) { set_last_error("embed_scatter_sorted: d must be a multiple of 8 (d=%d)", d); return ESB_ERR_INVALID_ARG; }
    dim3 grid(R, (d / 8 + 127) / 128);
    embed_scatter_sorted_kernel<<<grid, 128, 0, s>>>(sorted_ids, order, static_cast<const __nv_bfloat16*>(rows), dembed, pad, vocab, R, d);
    return check_launch("embed_scatter_sorted");
}
int copy_rows(const void* src, long long src_bstride_bytes, long long src_ld_bytes, void* dst, long long dst_bstride_bytes,
              long long dst_ld_bytes, int batches, int rows, int row_bytes, cudaStream_t s) {
    if (batches <= 0 || rows <= 0 || row_bytes <= 0) return ESB_OK;
    if ((row_bytes & 15) || (src_ld_bytes & 15) || (dst_ld_bytes & 15) || (src_bstride_bytes & 15) || (dst_bstride_bytes & 15) ||
        (reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst) & 15)) {
        set_last_error("copy_rows: everything must be 16-byte aligned");
        return ESB_ERR_INVALID_ARG;
    }
    copy_rows_kernel<<<g_grid(static_cast<long long>(batches) * rows * (row_bytes / 16), 256), 256, 0, s>>>(
        static_cast<const uint8_t*>(src), src_bstride_bytes, src_ld_bytes, static_cast<uint8_t*>(dst), dst_bstride_bytes, dst_ld_bytes,
        batches, rows, row_bytes);
    return check_launch("copy_rows");
}
int transpose_bf16(const void* in, long long ld_in, void* out, long long ld_out, int R, int C, cudaStream_t s) {
    if (R <= 0 || C <= 0) return ESB_OK;
    dim3 grid((C + 31) / 32, (R + 31) / 32);
    transpose_bf16_kernel<<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(in), ld_in, static_cast<__nv_bfloat16*>(out), ld_out, R, C);
    return check_launch("transpose_bf16");
}
int colsum_bf16(const void* in, long long ld, int M, int N, float* out, cudaStream_t s) {
    if (M <= 0 || N <= 0) return ESB_OK;
    int ysplit = (M + 255) / 256;
    if (ysplit > 64) ysplit = 64;
    dim3 grid((N + 63) / 64, ysplit);
    colsum_kernel<<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(in), ld, M, N, out);
    return check_launch("colsum_bf16");
}
int group_maxpool(const void* f, void* out, long long groups, int n, int C, int concat, cudaStream_t s) {
    if (groups <= 0) return ESB_OK;
    if (C & 1) { set_last_error("group_maxpool: C must be even"); return ESB_ERR_INVALID_ARG; }
    group_maxpool_kernel<<<static_cast<unsigned>(groups), 128, 0, s>>>(static_cast<const __nv_bfloat16*>(f), static_cast<__nv_bfloat16*>(out), n, C, concat);
    return check_launch("group_maxpool");
}

}  // namespace esb
