// Trajectory token <-> continuous 6DoF pose kernels (bit-exact with the reference NumPy code).
//
// Reference: egoscaler/models/pointllm/utils/utils.py
//   discretize_action :13-16   np.digitize(x, np.linspace(-1, 1, num_bins)) - 1
//   token_to_action   :18-21   np.linspace(-1, 1, num_bins)[id]           (left bin edge, float64)
//   rt2_scaler        :23-34   rot * pi ; z -> [d_min, d_max] ; x,y pixel -> metric (f=605.343, c=703.5)
//   simple_scaler     :36-45
//   str_to_float      :47-104  split on <tsep>, first run of 6 <p*> tokens per segment, repeat previous on a miss,
//                              float32 cast (:96) before the scaler
// and dataset.py:126-148 (denorm, do_norm branch), train.py:251-255 (pad by repeating the last frame).
//
// Exactness notes.  np.linspace builds edge i as  i*step + start  in float64 (two roundings, last edge forced to
// `stop`); the kernels rebuild that table with __dmul_rn/__dadd_rn (no FMA contraction) and bracket every input
// against it, so values sitting exactly on an edge, NaN (-> last bin), +-inf and out-of-range inputs land where
// np.digitize puts them.  The float32 scalers use explicit _rn intrinsics so nvcc cannot fuse mul+add.
#include "common.cuh"
#include "esb_internal.h"

namespace esb {

constexpr int kMaxBins = 1024;

ESB_DEVICE void build_edges(double* edges, int num_bins) {
    const double step = __ddiv_rn(2.0, static_cast<double>(num_bins - 1));
    for (int i = threadIdx.x; i < num_bins; i += blockDim.x)
        edges[i] = (i == num_bins - 1) ? 1.0 : __dadd_rn(__dmul_rn(static_cast<double>(i), step), -1.0);
    __syncthreads();
}

// number of edges <= x, minus one  (np.digitize(x, edges, right=False) - 1); NaN sorts last.
ESB_DEVICE int digitize_m1(double x, const double* edges, int num_bins) {
    if (x != x) return num_bins - 1;
    double g = floor((x + 1.0) * (0.5 * static_cast<double>(num_bins - 1)));
    int b = g < -1.0 ? -1 : (g > static_cast<double>(num_bins - 1) ? num_bins - 1 : static_cast<int>(g));
    while (b + 1 < num_bins && edges[b + 1] <= x) ++b;
    while (b >= 0 && edges[b] > x) --b;
    return b;
}

template <typename T>
__global__ void __launch_bounds__(256) traj_quantize_kernel(const T* __restrict__ x, long long n, int num_bins,
                                                            long long id_offset, long long* __restrict__ out) {
    __shared__ double edges[kMaxBins];
    build_edges(edges, num_bins);
    const long long nvec = n >> 2;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    // two 16-byte loads in flight per thread per iteration (the SMs are at full occupancy, so memory-level parallelism has to
    // come from the instruction stream: first B200 measurement of the one-load loop was 0.74 of the HBM peak), streaming stores
    auto quant4 = [&](const T* v, long long i) {
        longlong2 o0, o1;
        o0.x = digitize_m1(static_cast<double>(v[0]), edges, num_bins) + id_offset;
        o0.y = digitize_m1(static_cast<double>(v[1]), edges, num_bins) + id_offset;
        o1.x = digitize_m1(static_cast<double>(v[2]), edges, num_bins) + id_offset;
        o1.y = digitize_m1(static_cast<double>(v[3]), edges, num_bins) + id_offset;
        __stcs(reinterpret_cast<longlong2*>(out) + 2 * i, o0);
        __stcs(reinterpret_cast<longlong2*>(out) + 2 * i + 1, o1);
    };
    auto load4 = [&](T* v, long long i) {
        if constexpr (sizeof(T) == 4) {
            const float4 f = __ldcs(reinterpret_cast<const float4*>(x) + i);
            v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
        } else {
            const double2 a = __ldcs(reinterpret_cast<const double2*>(x) + 2 * i);
            const double2 b = __ldcs(reinterpret_cast<const double2*>(x) + 2 * i + 1);
            v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
        }
    };
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += 2 * stride) {
        T v0[4], v1[4];
        const long long i1 = i + stride;
        load4(v0, i);
        if (i1 < nvec) load4(v1, i1);
        quant4(v0, i);
        if (i1 < nvec) quant4(v1, i1);
    }
    // tail
    for (long long i = (nvec << 2) + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = digitize_m1(static_cast<double>(x[i]), edges, num_bins) + id_offset;
}

// mode: 0 none (left bin edge as float32), 1 rt2_scaler, 2 simple_scaler, 3 denorm(do_norm)
ESB_DEVICE void scale_pose(float* p, int mode, float d_max_minus_min, float d_min) {
    const float kPi = 3.14159274101257324f;          // float32(np.pi)
    if (mode == 1) {
        p[3] = __fmul_rn(kPi, p[3]); p[4] = __fmul_rn(kPi, p[4]); p[5] = __fmul_rn(kPi, p[5]);
        float z = __fadd_rn(__fmul_rn(0.5f, p[2]), 0.5f);
        z = __fadd_rn(__fmul_rn(d_max_minus_min, z), d_min);
        p[2] = z;
        float x = __fadd_rn(__fmul_rn(704.0f, p[0]), 704.0f);
        p[0] = __fdiv_rn(__fmul_rn(__fsub_rn(x, 703.5f), z), 605.343f);
        float y = __fadd_rn(__fmul_rn(704.0f, p[1]), 704.0f);
        p[1] = __fdiv_rn(__fmul_rn(__fsub_rn(y, 703.5f), z), 605.343f);
    } else if (mode == 2) {
#pragma unroll
        for (int j = 3; j < 6; ++j)
            p[j] = __fmul_rn(kPi, __fsub_rn(__fmul_rn(2.0f, __fdiv_rn(p[j], 100.0f)), 1.0f));
        float z = __fdiv_rn(p[2], 100.0f);
        z = __fadd_rn(__fmul_rn(z, d_max_minus_min), d_min);
        p[2] = z;
        p[0] = __fdiv_rn(__fmul_rn(__fsub_rn(p[0], 703.5f), z), 605.343f);
        p[1] = __fdiv_rn(__fmul_rn(__fsub_rn(p[1], 703.5f), z), 605.343f);
    } else if (mode == 3) {
        // egoscaler/configs/dataset.py:1-6 : x,y in [-2,2], z in [0,2.5]
#pragma unroll
        for (int j = 0; j < 3; ++j) p[j] = __fdiv_rn(__fadd_rn(p[j], 1.0f), 2.0f);
        p[0] = __fadd_rn(__fmul_rn(p[0], 4.0f), -2.0f);
        p[1] = __fadd_rn(__fmul_rn(p[1], 4.0f), -2.0f);
        p[2] = __fadd_rn(__fmul_rn(p[2], 2.5f), 0.0f);
        p[3] = __fmul_rn(p[3], kPi); p[4] = __fmul_rn(p[4], kPi); p[5] = __fmul_rn(p[5], kPi);
    }
}

// One thread per pose: 6 ids (int64) -> 6 float32, optional scaler.  Invalid ids give NaN.
__global__ void __launch_bounds__(256) traj_dequantize_kernel(const long long* __restrict__ ids, long long n_poses,
                                                              int num_bins, long long id_offset, int mode,
                                                              float d_max_minus_min, float d_min, float* __restrict__ out) {
    __shared__ double edges[kMaxBins];
    build_edges(edges, num_bins);
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_poses; i += stride) {
        const longlong2* ip = reinterpret_cast<const longlong2*>(ids + 6 * i);
        const longlong2 a = __ldcs(ip), b = __ldcs(ip + 1), c = __ldcs(ip + 2);
        const long long id[6] = {a.x, a.y, b.x, b.y, c.x, c.y};
        float p[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const long long k = id[j] - id_offset;
            p[j] = (k >= 0 && k < num_bins) ? static_cast<float>(edges[k]) : __int_as_float(0x7fc00000);
        }
        scale_pose(p, mode, d_max_minus_min, d_min);
        float2* op = reinterpret_cast<float2*>(out + 6 * i);
        __stcs(op, make_float2(p[0], p[1])); __stcs(op + 1, make_float2(p[2], p[3])); __stcs(op + 2, make_float2(p[4], p[5]));
    }
}

__global__ void __launch_bounds__(256) traj_scale_kernel(float* __restrict__ poses, long long n_poses, int mode,
                                                         float d_max_minus_min, float d_min) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_poses; i += stride) {
        float p[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) p[j] = poses[6 * i + j];
        scale_pose(p, mode, d_max_minus_min, d_min);
#pragma unroll
        for (int j = 0; j < 6; ++j) poses[6 * i + j] = p[j];
    }
}

// Generated token ids -> poses, one thread per sequence (the sequential <tsep>/regex logic of str_to_float).
// ids [B, T]; out [B, max_steps, 6]; count[B] = number of parsed timesteps (before padding).
__global__ void traj_ids_to_poses_kernel(const long long* __restrict__ ids, int B, int T, int num_bins, long long base,
                                         long long tsep, long long eos, int mode, float d_max_minus_min, float d_min,
                                         int max_steps, int pad_repeat, float* __restrict__ out, int* __restrict__ count) {
    __shared__ double edges[kMaxBins];
    build_edges(edges, num_bins);
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const long long* row = ids + static_cast<long long>(b) * T;
    float* o = out + static_cast<long long>(b) * max_steps * 6;
    int n = 0;
    bool have_last = false, found = false, stop = false;
    float last[6], cur[6];
    int run = 0;
    for (int t = 0; t <= T && !stop; ++t) {
        const long long tok = (t < T) ? row[t] : eos;
        const bool end = (t == T) || (tok == eos);
        if (end || tok == tsep) {
            // close the segment
            if (found) {
#pragma unroll
                for (int j = 0; j < 6; ++j) last[j] = cur[j];
                have_last = true;
            }
            if ((found || have_last) && n < max_steps) {
#pragma unroll
                for (int j = 0; j < 6; ++j) o[n * 6 + j] = last[j];
                ++n;
            }
            found = false;
            run = 0;
            if (end) stop = true;
        } else if (!found) {
            if (tok >= base && tok < base + num_bins) {
                cur[run++] = static_cast<float>(edges[tok - base]);
                if (run == 6) found = true;
            } else {
                run = 0;
            }
        }
    }
    for (int s = 0; s < n; ++s) scale_pose(o + s * 6, mode, d_max_minus_min, d_min);
    count[b] = n;
    if (pad_repeat && n > 0)
        for (int s = n; s < max_steps; ++s)
#pragma unroll
            for (int j = 0; j < 6; ++j) o[s * 6 + j] = o[(n - 1) * 6 + j];
}

static int grid_for(long long work_items, int block, int per_sm) {
    const long long want = (work_items + block - 1) / block;
    const long long cap = static_cast<long long>(device_sm_count()) * per_sm;
    long long g = want < cap ? want : cap;
    return static_cast<int>(g < 1 ? 1 : g);
}

int traj_quantize(const void* x, int is_f64, long long n, int num_bins, long long id_offset, long long* out,
                  cudaStream_t s) {
    if (num_bins < 2 || num_bins > kMaxBins) { set_last_error("traj_quantize: num_bins %d out of [2,%d]", num_bins, kMaxBins); return ESB_ERR_INVALID_ARG; }
    if (n <= 0) return ESB_OK;
    if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) { set_last_error("traj_quantize: pointers must be 16-byte aligned"); return ESB_ERR_INVALID_ARG; }
    const int grid = grid_for((n + 3) / 4, 256, 8);
    if (is_f64) traj_quantize_kernel<double><<<grid, 256, 0, s>>>(static_cast<const double*>(x), n, num_bins, id_offset, out);
    else traj_quantize_kernel<float><<<grid, 256, 0, s>>>(static_cast<const float*>(x), n, num_bins, id_offset, out);
    return check_launch("traj_quantize");
}

int traj_dequantize(const long long* ids, long long n_poses, int num_bins, long long id_offset, int mode, double d_max,
                    double d_min, float* out, cudaStream_t s) {
    if (num_bins < 2 || num_bins > kMaxBins || mode < 0 || mode > 3) { set_last_error("traj_dequantize: bad num_bins/mode"); return ESB_ERR_INVALID_ARG; }
    if (n_poses <= 0) return ESB_OK;
    if ((reinterpret_cast<uintptr_t>(ids) & 15) || (reinterpret_cast<uintptr_t>(out) & 7)) { set_last_error("traj_dequantize: misaligned pointers"); return ESB_ERR_INVALID_ARG; }
    // (d_max - d_min) is formed in float64 by the reference and rounded to float32 when it meets the array
    const float dmm = static_cast<float>(d_max - d_min);
    traj_dequantize_kernel<<<grid_for(n_poses, 256, 8), 256, 0, s>>>(ids, n_poses, num_bins, id_offset, mode, dmm,
                                                                     static_cast<float>(d_min), out);
    return check_launch("traj_dequantize");
}

int traj_scale(float* poses, long long n_poses, int mode, double d_max, double d_min, cudaStream_t s) {
    if (mode < 1 || mode > 3) { set_last_error("traj_scale: bad mode %d", mode); return ESB_ERR_INVALID_ARG; }
    if (n_poses <= 0) return ESB_OK;
    traj_scale_kernel<<<grid_for(n_poses, 256, 8), 256, 0, s>>>(poses, n_poses, mode, static_cast<float>(d_max - d_min), static_cast<float>(d_min));
    return check_launch("traj_scale");
}

int traj_ids_to_poses(const long long* ids, int B, int T, int num_bins, long long base, long long tsep, long long eos,
                      int mode, double d_max, double d_min, int max_steps, int pad_repeat, float* out, int* count,
                      cudaStream_t s) {
    if (num_bins < 2 || num_bins > kMaxBins || mode < 0 || mode > 3 || max_steps < 1) { set_last_error("traj_ids_to_poses: bad arguments"); return ESB_ERR_INVALID_ARG; }
    if (B <= 0) return ESB_OK;
    traj_ids_to_poses_kernel<<<(B + 63) / 64, 64, 0, s>>>(ids, B, T, num_bins, base, tsep, eos, mode,
                                                          static_cast<float>(d_max - d_min), static_cast<float>(d_min),
                                                          max_steps, pad_repeat, out, count);
    return check_launch("traj_ids_to_poses");
}

// ------------------------------------------------------------------------------------------------ poses -> ids
// Data-side tokenisation, the inverse of traj_ids_to_poses: normalised poses in [-1, 1] are discretised
// with utils/utils.py:13-16 discretize_action (bit-exact np.digitize) and laid out as the token sequence the parser
// utils/utils.py:47-104 / constant.py expects:  <ts> (6 x <p*> <tsep>) x len  <te> eos pad...   One warp per sequence.
// norm = 1 first applies the inverse of dataset.py:139-146 denorm(do_norm=True): xyz affine [min,max] -> [-1,1], rot / pi
// (the forward normaliser itself is not shipped by the reference).
__global__ void __launch_bounds__(128) traj_poses_to_ids_kernel(const float* __restrict__ poses, const int* __restrict__ lens, int B,
                                                                int T, int num_bins, long long base, long long ts, long long tsep,
                                                                long long te, long long eos, long long pad, int norm,
                                                                const double* __restrict__ range6, int out_ld,
                                                                long long* __restrict__ out, int* __restrict__ out_len) {
    __shared__ double edges[kMaxBins];
    build_edges(edges, num_bins);
    const int b = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= B) return;
    int n = lens ? lens[b] : T;
    n = n < 0 ? 0 : (n > T ? T : n);
    const int need = 1 + 7 * n + 2;
    long long* o = out + static_cast<long long>(b) * out_ld;
    for (int i = lane; i < out_ld; i += 32) {
        long long tok = pad;
        if (i == 0) tok = ts;
        else if (i < 1 + 7 * n) {
            const int step = (i - 1) / 7, j = (i - 1) % 7;
            if (j == 6) tok = tsep;
            else {
                double v = static_cast<double>(poses[(static_cast<long long>(b) * T + step) * 6 + j]);
                if (norm) {
                    if (j < 3) v = 2.0 * (v - range6[2 * j]) / (range6[2 * j + 1] - range6[2 * j]) - 1.0;
                    else v = v / 3.141592653589793;
                    v = static_cast<double>(static_cast<float>(v));          // the stored normalised pose is float32
                }
                tok = base + digitize_m1(v, edges, num_bins);
            }
        } else if (i == 1 + 7 * n) tok = te;
        else if (i == 2 + 7 * n) tok = eos;
        o[i] = tok;
    }
    if (lane == 0 && out_len) out_len[b] = need <= out_ld ? need : out_ld;
}

int traj_poses_to_ids(const float* poses, const int* lens, int B, int T, int num_bins, long long base, long long ts, long long tsep,
                      long long te, long long eos, long long pad, int norm, const double* range6, int out_ld, long long* out,
                      int* out_len, cudaStream_t s) {
    if (B <= 0) return ESB_OK;
    if (!poses || !out || num_bins < 2 || num_bins > kMaxBins || T < 1 || out_ld < 3 + 7 * T || (norm && !range6)) {
        set_last_error("traj_poses_to_ids: bad arguments (T=%d out_ld=%d needs >= %d)", T, out_ld, 3 + 7 * T);
        return ESB_ERR_INVALID_ARG;
    }
    traj_poses_to_ids_kernel<<<(B + 3) / 4, 128, 0, s>>>(poses, lens, B, T, num_bins, base, ts, tsep, te, eos, pad, norm, range6, out_ld,
                                                         out, out_len);
    return check_launch("traj_poses_to_ids");
}

// ------------------------------------------------------------------------------------------------ pre-processing
// egoscaler/models/utils/traj_utils.py: smoothing_traj :42-96 (moving average over the position columns with the reference's
// edge rules, fp32 left-to-right like numpy on float32 rows, no FMA contraction) followed by preprocess_traj :3-40
// (np.linspace(0, T-1, num_steps).astype(int) gather when T >= num_steps, else pad with the last point; mask 1 = real).
ESB_DEVICE float smoothed_pos(const float* __restrict__ src, int D, int n, int j, int d) {
    auto P = [&](int t) { return src[static_cast<long long>(t) * D + d]; };
    auto add = [](float a, float b) { return __fadd_rn(a, b); };
    if (j == 0) {
        if (n >= 3) return __fdiv_rn(add(add(__fmul_rn(3.0f, P(0)), P(1)), P(2)), 5.0f);
        if (n == 2) return __fdiv_rn(add(__fmul_rn(3.0f, P(0)), P(1)), 4.0f);
        return P(0);
    }
    if (j == 1) {
        if (n >= 4) return __fdiv_rn(add(add(add(__fmul_rn(2.0f, P(0)), P(1)), P(2)), P(3)), 5.0f);
        if (n == 3) return __fdiv_rn(add(add(__fmul_rn(2.0f, P(0)), P(1)), P(2)), 4.0f);
        return P(1);
    }
    if (j == n - 2) {                                   // n >= 4 here (j >= 2)
        return __fdiv_rn(add(add(add(P(j - 2), P(j - 1)), P(j)), P(j + 1)), 4.0f);
    }
    if (j == n - 1) {                                   // n >= 3 here
        return __fdiv_rn(add(add(P(j - 2), P(j - 1)), P(j)), 3.0f);
    }
    return __fdiv_rn(add(add(add(add(P(j - 2), P(j - 1)), P(j)), P(j + 1)), P(j + 2)), 5.0f);
}

__global__ void __launch_bounds__(128) traj_preprocess_kernel(const float* __restrict__ traj, const int* __restrict__ lens, int T,
                                                              int D, int num_steps, int smooth, float* __restrict__ out,
                                                              int* __restrict__ mask) {
    const int b = blockIdx.x;
    int n = lens ? lens[b] : T;
    n = n < 1 ? 1 : (n > T ? T : n);
    const float* src = traj + static_cast<long long>(b) * T * D;
    float* dst = out + static_cast<long long>(b) * num_steps * D;
    const double step = num_steps > 1 ? static_cast<double>(n - 1) / static_cast<double>(num_steps - 1) : 0.0;
    for (int e = threadIdx.x; e < num_steps * D; e += blockDim.x) {
        const int s = e / D, d = e - s * D;
        int j;
        bool real = true;
        if (n >= num_steps) {
            j = (num_steps > 1 && s == num_steps - 1) ? n - 1 : static_cast<int>(__dmul_rn(static_cast<double>(s), step));
        } else {
            real = s < n;
            j = real ? s : n - 1;
        }
        dst[e] = (smooth && d < 3) ? smoothed_pos(src, D, n, j, d) : src[static_cast<long long>(j) * D + d];
        if (mask && d == 0) mask[static_cast<long long>(b) * num_steps + s] = real ? 1 : 0;
    }
}

int traj_preprocess(const float* traj, const int* lens, int B, int T, int D, int num_steps, int smooth, float* out, int* mask,
                    cudaStream_t s) {
    if (B <= 0) return ESB_OK;
    if (!traj || !out || T < 1 || D < 1 || num_steps < 1 || (smooth && D < 3)) {
        set_last_error("traj_preprocess: bad arguments (T=%d D=%d num_steps=%d)", T, D, num_steps);
        return ESB_ERR_INVALID_ARG;
    }
    traj_preprocess_kernel<<<B, 128, 0, s>>>(traj, lens, T, D, num_steps, smooth, out, mask);
    return check_launch("traj_preprocess");
}

}  // namespace esb
