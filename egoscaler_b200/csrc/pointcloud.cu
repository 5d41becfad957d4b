// Point-cloud front end of the PointBERT encoder: farthest-point sampling and kNN grouping.
//
// fps       : pointbert/misc.py:40-60 (fps) -- 512 sequential argmax iterations; the reference spends ~6 tiny kernels
//             per iteration from a Python loop.  Here: ONE persistent CTA per cloud, points and running distances
//             live in registers, one block-wide (value,index) arg-max per iteration.  Arithmetic replicates the
//             reference's fp32 order ((dx^2 + dy^2) + dz^2, separate roundings, first-max-wins) so the selected
//             indices are exact.  The random start index (misc.py:52, CPU generator) is an input.
// knn_group : pointbert/dvae.py:107-140 (square_distance + topk(k, largest=False)) and :150-187 (Group.forward:
//             gather neighbours, subtract the centre from xyz only).  One CTA per (cloud, centre): distances to
//             shared memory, exact k-th-smallest by 4-pass radix select (no 512x8192 matrix round trip), neighbours
//             emitted in ascending index order, written as zero-padded bf16 rows ready for the first conv GEMM.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

constexpr int FPS_THREADS = 1024;

template <int PPT>
__global__ void __launch_bounds__(FPS_THREADS, 1) fps_kernel(const float* __restrict__ pts, int N, int C, int G,
                                                             const long long* __restrict__ start, float* __restrict__ centers,
                                                             int* __restrict__ center_idx, __nv_bfloat16* __restrict__ centers_pad,
                                                             int pad_ld) {
    __shared__ float s_val[32];
    __shared__ int s_idx[32];
    __shared__ float s_c[3];
    __shared__ int s_far;
    const int b = blockIdx.x;
    const float* P = pts + static_cast<long long>(b) * N * C;
    float x[PPT], y[PPT], z[PPT], dist[PPT];
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
        const int i = threadIdx.x + k * FPS_THREADS;
        if (i < N) { x[k] = P[static_cast<long long>(i) * C]; y[k] = P[static_cast<long long>(i) * C + 1]; z[k] = P[static_cast<long long>(i) * C + 2]; }
        else { x[k] = y[k] = z[k] = 0.0f; }
        dist[k] = 1e10f;
    }
    int far = static_cast<int>(start[b]);
    if (far < 0 || far >= N) far = 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int it = 0; it < G; ++it) {
        if (threadIdx.x == 0) {
            const float cx = P[static_cast<long long>(far) * C], cy = P[static_cast<long long>(far) * C + 1], cz = P[static_cast<long long>(far) * C + 2];
            s_c[0] = cx; s_c[1] = cy; s_c[2] = cz;
            center_idx[b * G + it] = far;
            centers[(static_cast<long long>(b) * G + it) * 3 + 0] = cx;
            centers[(static_cast<long long>(b) * G + it) * 3 + 1] = cy;
            centers[(static_cast<long long>(b) * G + it) * 3 + 2] = cz;
        }
        __syncthreads();
        const float cx = s_c[0], cy = s_c[1], cz = s_c[2];
        if (centers_pad && threadIdx.x < pad_ld)            // zero-padded bf16 copy = A operand of the pos_embed GEMM
            centers_pad[(static_cast<long long>(b) * G + it) * pad_ld + threadIdx.x] =
                __float2bfloat16_rn(threadIdx.x == 0 ? cx : (threadIdx.x == 1 ? cy : (threadIdx.x == 2 ? cz : 0.0f)));
        float best = -1.0f;
        int bi = 0x7fffffff;
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
            const int i = threadIdx.x + k * FPS_THREADS;
            if (i < N) {
                const float dx = __fsub_rn(x[k], cx), dy = __fsub_rn(y[k], cy), dz = __fsub_rn(z[k], cz);
                const float d = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                dist[k] = fminf(dist[k], d);
                if (dist[k] > best) { best = dist[k]; bi = i; }      // ascending i within a thread: first max wins
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = best; s_idx[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            best = s_val[lane]; bi = s_idx[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
            }
            if (lane == 0) s_far = bi;
        }
        __syncthreads();
        far = s_far;
    }
}

int fps(const float* pts, int B, int N, int C, int G, const long long* start, float* centers, int* center_idx, void* centers_pad,
        int pad_ld, cudaStream_t s) {
    if (B <= 0 || G <= 0) return ESB_OK;
    if (C < 3 || N <= 0) { set_last_error("fps: need N > 0 and at least xyz channels"); return ESB_ERR_INVALID_ARG; }
    if (N <= 8 * FPS_THREADS) fps_kernel<8><<<B, FPS_THREADS, 0, s>>>(pts, N, C, G, start, centers, center_idx, static_cast<__nv_bfloat16*>(centers_pad), pad_ld);
    else if (N <= 16 * FPS_THREADS) fps_kernel<16><<<B, FPS_THREADS, 0, s>>>(pts, N, C, G, start, centers, center_idx, static_cast<__nv_bfloat16*>(centers_pad), pad_ld);
    else { set_last_error("fps: N=%d exceeds %d", N, 16 * FPS_THREADS); return ESB_ERR_UNSUPPORTED; }
    return check_launch("fps");
}

// ------------------------------------------------------------------------------------------------ kNN + group
ESB_DEVICE uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int KNN_THREADS = 256;
constexpr int KNN_MAXK = 64;

__global__ void __launch_bounds__(KNN_THREADS) knn_group_kernel(const float* __restrict__ pts, const float* __restrict__ centers,
                                                                int N, int C, int G, int K, int out_ld,
                                                                __nv_bfloat16* __restrict__ nb_out, int* __restrict__ idx_out) {
    extern __shared__ uint32_t keys[];                  // [N]
    __shared__ int hist[256];
    __shared__ uint32_t s_prefix;
    __shared__ int s_remaining;
    __shared__ int s_sel[KNN_MAXK];
    __shared__ int s_sorted[KNN_MAXK];
    __shared__ int s_cnt_less, s_cnt_eq;
    const int g = blockIdx.x, b = blockIdx.y;
    const float* P = pts + static_cast<long long>(b) * N * C;
    const float* cp = centers + (static_cast<long long>(b) * G + g) * 3;
    const float cx = cp[0], cy = cp[1], cz = cp[2];
    const float cc = cx * cx + cy * cy + cz * cz;
    for (int i = threadIdx.x; i < N; i += KNN_THREADS) {
        const float x = P[static_cast<long long>(i) * C], y = P[static_cast<long long>(i) * C + 1], z = P[static_cast<long long>(i) * C + 2];
        const float d = -2.0f * (cx * x + cy * y + cz * z) + cc + (x * x + y * y + z * z);   // dvae.py:137-139
        keys[i] = f2ord(d);
    }
    if (threadIdx.x == 0) { s_prefix = 0; s_remaining = K; s_cnt_less = 0; s_cnt_eq = 0; }
    __syncthreads();
    uint32_t mask = 0;
    for (int pass = 3; pass >= 0; --pass) {
        hist[threadIdx.x] = 0;
        __syncthreads();
        const uint32_t prefix = s_prefix;
        const int shift = pass * 8;
        for (int i = threadIdx.x; i < N; i += KNN_THREADS) {
            const uint32_t key = keys[i];
            if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            const int lane = threadIdx.x;
            int local[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { local[j] = hist[lane * 8 + j]; sum += local[j]; }
            int incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int tv = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += tv;
            }
            const int excl = incl - sum;
            const int rem = s_remaining;
            if (excl < rem && incl >= rem) {            // exactly one lane
                int cum = excl;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (cum + local[j] >= rem) {
                        s_prefix = prefix | (static_cast<uint32_t>(lane * 8 + j) << shift);
                        s_remaining = rem - cum;
                        break;
                    }
                    cum += local[j];
                }
            }
        }
        mask |= 0xFFu << shift;
        __syncthreads();
    }
    const uint32_t thr = s_prefix;
    const int need_eq = s_remaining;                    // how many keys == thr belong to the k smallest
    const int n_less = K - need_eq;
    for (int i = threadIdx.x; i < N; i += KNN_THREADS) {
        const uint32_t key = keys[i];
        if (key < thr) {
            const int slot = atomicAdd(&s_cnt_less, 1);
            s_sel[slot] = i;
        } else if (key == thr) {
            const int slot = atomicAdd(&s_cnt_eq, 1);
            if (slot < need_eq) s_sel[n_less + slot] = i;
        }
    }
    __syncthreads();
    if (threadIdx.x < K) {                              // rank sort -> ascending point index (deterministic order)
        const int mine = s_sel[threadIdx.x];
        int rank = 0;
        for (int j = 0; j < K; ++j) rank += (s_sel[j] < mine);
        s_sorted[rank] = mine;
    }
    __syncthreads();
    const long long row0 = (static_cast<long long>(b) * G + g) * K;
    for (int e = threadIdx.x; e < K * out_ld; e += KNN_THREADS) {
        const int r = e / out_ld, c = e % out_ld;
        float val = 0.0f;
        if (c < C) {
            val = P[static_cast<long long>(s_sorted[r]) * C + c];
            if (c < 3) val -= (c == 0 ? cx : (c == 1 ? cy : cz));
        }
        nb_out[(row0 + r) * out_ld + c] = __float2bfloat16_rn(val);
    }
    if (idx_out && threadIdx.x < K) idx_out[row0 + threadIdx.x] = s_sorted[threadIdx.x];
}

int knn_group(const float* pts, const float* centers, int B, int N, int C, int G, int K, int out_ld, void* nb_out, int* idx_out,
              cudaStream_t s) {
    if (B <= 0 || G <= 0) return ESB_OK;
    if (K < 1 || K > KNN_MAXK || K > N || C < 3 || out_ld < C) { set_last_error("knn_group: bad K/C/out_ld (K=%d C=%d ld=%d)", K, C, out_ld); return ESB_ERR_INVALID_ARG; }
    const size_t smem = static_cast<size_t>(N) * 4;
    if (smem > 200 * 1024) { set_last_error("knn_group: N=%d too large", N); return ESB_ERR_UNSUPPORTED; }
    static bool set = false;
    if (!set) { cudaFuncSetAttribute(knn_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); set = true; }
    dim3 grid(G, B);
    knn_group_kernel<<<grid, KNN_THREADS, smem, s>>>(pts, centers, N, C, G, K, out_ld, static_cast<__nv_bfloat16*>(nb_out), idx_out);
    return check_launch("knn_group");
}

}  // namespace esb
