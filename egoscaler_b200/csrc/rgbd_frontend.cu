// RGB-D -> coloured point cloud front-end, the caller-side producer of `pcrgbs`.
//
//   rgbd_to_points   egoscaler/data/tools/pcm_tools.py:68-96 get_points_colors (call sites
//                    data/train/7_get_object_trajectory.py:191-200,244-253; same unprojection in
//                    third_party/Depth-Anything-V2/metric_depth/depth.py:55-60): pinhole unprojection
//                    x = (i - c) / fx * z, y = (j - c) / fy * z in fp64 (numpy promotes int64 grid / python float), colours / 255
//                    in fp32 (uint8 ++ float32 depth concatenates to float32, :157), pixels kept where all three colour channels
//                    are non-zero, outside every bbox, and z < d_thres; survivors in row-major pixel order (boolean indexing).
//   pc_sample_norm   pointllm/data/utils.py:146-157 pc_norm on a caller-chosen subset (the sampler down to 8192 points is not
//                    shipped by the reference: indices come from the caller): xyz - mean, / max L2 norm, colours untouched.
//
// HBM-bound byte work: one float4 load per pixel in each of two passes (count, then ordered write) + 24 B per survivor.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

namespace {

constexpr int RF_THREADS = 256;

struct RgbdParams {
    const float4* rgbd;
    long long n_pix;
    int W;
    const int* boxes;     // [n_boxes,4] xmin,ymin,xmax,ymax
    int n_boxes;
    double c, fx, fy;
    int has_thres;
    float thres;
};

ESB_DEVICE bool pixel_valid(const RgbdParams& p, const float4 v, int x, int y) {
    bool ok = (v.x != 0.0f) && (v.y != 0.0f) && (v.z != 0.0f);
    if (p.has_thres) ok = ok && (v.w < p.thres);
    for (int b = 0; b < p.n_boxes && ok; ++b) {
        const int4 bx = *reinterpret_cast<const int4*>(p.boxes + 4 * b);
        if (x >= bx.x && x < bx.z && y >= bx.y && y < bx.w) ok = false;
    }
    return ok;
}

__global__ void __launch_bounds__(RF_THREADS) rgbd_count_kernel(const RgbdParams p, long long n_seg, int* __restrict__ seg_count) {
    for (long long seg = blockIdx.x; seg < n_seg; seg += gridDim.x) {
        const long long i = seg * RF_THREADS + threadIdx.x;
        bool ok = false;
        if (i < p.n_pix) {
            const float4 v = p.rgbd[i];
            ok = pixel_valid(p, v, static_cast<int>(i % p.W), static_cast<int>(i / p.W));
        }
        const int c = __syncthreads_count(ok);
        if (threadIdx.x == 0) seg_count[seg] = c;
    }
}

// exclusive scan of seg_count -> seg_off (int64), total -> *out_count.  One CTA of 1024 threads, chunks of 1024 with a carry.
__global__ void __launch_bounds__(1024) rgbd_scan_kernel(const int* __restrict__ seg_count, long long n_seg,
                                                         long long* __restrict__ seg_off, long long* __restrict__ out_count) {
    __shared__ long long warp_sum[32];
    __shared__ long long carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (long long base = 0; base < n_seg; base += 1024) {
        const long long i = base + tid;
        const long long v = i < n_seg ? seg_count[i] : 0;
        long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_sum[w] = inc;
        __syncthreads();
        if (w == 0) {
            long long s = warp_sum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += t;
            }
            warp_sum[lane] = s;                       // inclusive over warps
        }
        __syncthreads();
        const long long before = carry + (w ? warp_sum[w - 1] : 0) + inc - v;
        if (i < n_seg) seg_off[i] = before;
        __syncthreads();
        if (tid == 1023) carry = before + v;
        __syncthreads();
    }
    if (tid == 0 && out_count) *out_count = carry;
}

__global__ void __launch_bounds__(RF_THREADS) rgbd_write_kernel(const RgbdParams p, long long n_seg,
                                                                const long long* __restrict__ seg_off, float* __restrict__ out,
                                                                long long cap) {
    __shared__ int warp_cnt[RF_THREADS / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (long long seg = blockIdx.x; seg < n_seg; seg += gridDim.x) {
        const long long i = seg * RF_THREADS + threadIdx.x;
        bool ok = false;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        int x = 0, y = 0;
        if (i < p.n_pix) {
            v = p.rgbd[i];
            x = static_cast<int>(i % p.W);
            y = static_cast<int>(i / p.W);
            ok = pixel_valid(p, v, x, y);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) warp_cnt[w] = __popc(bal);
        __syncthreads();
        int before = __popc(bal & ((1u << lane) - 1u));
        for (int j = 0; j < w; ++j) before += warp_cnt[j];
        if (ok) {
            const long long r = seg_off[seg] + before;
            if (r < cap) {
                const double z = static_cast<double>(v.w);
                const double px = __ddiv_rn(static_cast<double>(x) - p.c, p.fx);
                const double py = __ddiv_rn(static_cast<double>(y) - p.c, p.fy);
                float* o = out + r * 6;
                o[0] = static_cast<float>(__dmul_rn(px, z));
                o[1] = static_cast<float>(__dmul_rn(py, z));
                o[2] = v.w;
                o[3] = __fdiv_rn(v.x, 255.0f);
                o[4] = __fdiv_rn(v.y, 255.0f);
                o[5] = __fdiv_rn(v.z, 255.0f);
            }
        }
        __syncthreads();
    }
}

// One CTA per cloud: gather rows idx[s] (or the first S rows), pc_norm in fp64.
__global__ void __launch_bounds__(1024) pc_sample_norm_kernel(const float* __restrict__ pts, long long cloud_stride, long long n,
                                                              const long long* __restrict__ idx, int S, float* __restrict__ out) {
    __shared__ double red[3][32];
    __shared__ double cen[3];
    __shared__ double s_m;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const float* src = pts + static_cast<long long>(b) * cloud_stride;
    const long long* ip = idx ? idx + static_cast<long long>(b) * S : nullptr;
    float* dst = out + static_cast<long long>(b) * S * 6;
    double sx = 0.0, sy = 0.0, sz = 0.0;
    for (int s = tid; s < S; s += 1024) {
        long long r = ip ? ip[s] : s;
        r = r < 0 ? 0 : (r >= n ? n - 1 : r);
        const float* q = src + r * 6;
        sx += q[0]; sy += q[1]; sz += q[2];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o);
    }
    if (lane == 0) { red[0][w] = sx; red[1][w] = sy; red[2][w] = sz; }
    __syncthreads();
    if (tid < 3) {
        double t = 0.0;
        for (int j = 0; j < 32; ++j) t += red[tid][j];
        cen[tid] = t / static_cast<double>(S);
    }
    __syncthreads();
    const double cx = cen[0], cy = cen[1], cz = cen[2];
    double m2 = 0.0;
    for (int s = tid; s < S; s += 1024) {
        long long r = ip ? ip[s] : s;
        r = r < 0 ? 0 : (r >= n ? n - 1 : r);
        const float* q = src + r * 6;
        const double dx = q[0] - cx, dy = q[1] - cy, dz = q[2] - cz;
        m2 = fmax(m2, dx * dx + dy * dy + dz * dz);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) m2 = fmax(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    __syncthreads();
    if (lane == 0) red[0][w] = m2;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int j = 0; j < 32; ++j) t = fmax(t, red[0][j]);
        s_m = sqrt(t);
    }
    __syncthreads();
    const double m = s_m;
    for (int s = tid; s < S; s += 1024) {
        long long r = ip ? ip[s] : s;
        r = r < 0 ? 0 : (r >= n ? n - 1 : r);
        const float* q = src + r * 6;
        float* o = dst + static_cast<long long>(s) * 6;
        o[0] = static_cast<float>((q[0] - cx) / m);
        o[1] = static_cast<float>((q[1] - cy) / m);
        o[2] = static_cast<float>((q[2] - cz) / m);
        o[3] = q[3]; o[4] = q[4]; o[5] = q[5];
    }
}

}  // namespace

long long rgbd_workspace_bytes(int H, int W) {
    const long long n_seg = (static_cast<long long>(H) * W + RF_THREADS - 1) / RF_THREADS;
    return n_seg * 16;                                  // int32 counts (padded to 8 B) + int64 offsets
}

int rgbd_to_points(const float* rgbd, int H, int W, const int* boxes, int n_boxes, double principal, double fx, double fy,
                   int has_thres, double d_thres, float* out, long long cap, long long* out_count, void* workspace,
                   long long workspace_bytes, cudaStream_t s) {
    if (!rgbd || !out || H <= 0 || W <= 0 || n_boxes < 0 || (n_boxes > 0 && !boxes) || !(fx != 0.0) || !(fy != 0.0) || cap < 0) {
        set_last_error("rgbd_to_points: bad arguments (H=%d W=%d n_boxes=%d fx=%g fy=%g)", H, W, n_boxes, fx, fy);
        return ESB_ERR_INVALID_ARG;
    }
    if ((reinterpret_cast<uintptr_t>(rgbd) & 15) || (n_boxes > 0 && (reinterpret_cast<uintptr_t>(boxes) & 15)) ||
        (reinterpret_cast<uintptr_t>(workspace) & 7)) {
        set_last_error("rgbd_to_points: rgbd / boxes must be 16-byte aligned and the workspace 8-byte aligned");
        return ESB_ERR_INVALID_ARG;
    }
    if (!workspace || workspace_bytes < rgbd_workspace_bytes(H, W)) {
        set_last_error("rgbd_to_points: workspace %lld B < %lld B", workspace_bytes, rgbd_workspace_bytes(H, W));
        return ESB_ERR_INVALID_ARG;
    }
    RgbdParams p;
    p.rgbd = reinterpret_cast<const float4*>(rgbd);
    p.n_pix = static_cast<long long>(H) * W;
    p.W = W; p.boxes = boxes; p.n_boxes = n_boxes;
    p.c = principal; p.fx = fx; p.fy = fy; p.has_thres = has_thres;
    p.thres = static_cast<float>(d_thres);
    const long long n_seg = (p.n_pix + RF_THREADS - 1) / RF_THREADS;
    long long* seg_off = static_cast<long long*>(workspace);
    int* seg_count = reinterpret_cast<int*>(seg_off + n_seg);
    const unsigned grid = static_cast<unsigned>(n_seg < 148LL * 16 ? n_seg : 148LL * 16);
    rgbd_count_kernel<<<grid, RF_THREADS, 0, s>>>(p, n_seg, seg_count);
    rgbd_scan_kernel<<<1, 1024, 0, s>>>(seg_count, n_seg, seg_off, out_count);
    rgbd_write_kernel<<<grid, RF_THREADS, 0, s>>>(p, n_seg, seg_off, out, cap);
    return check_launch("rgbd_to_points");
}

int pc_sample_norm(const float* pts, long long cloud_stride, long long n, const long long* idx, int B, int S, float* out,
                   cudaStream_t s) {
    if (B <= 0 || S <= 0) return ESB_OK;
    if (!pts || !out || n <= 0 || (!idx && S > n)) {
        set_last_error("pc_sample_norm: bad arguments (n=%lld S=%d)", n, S);
        return ESB_ERR_INVALID_ARG;
    }
    pc_sample_norm_kernel<<<B, 1024, 0, s>>>(pts, cloud_stride, n, idx, S, out);
    return check_launch("pc_sample_norm");
}

}  // namespace esb
