// Trajectory metrics on device, batched: ADE / FDE / IDE / angular (geodesic) distance per generated trajectory.
//
// Reference: egoscaler/models/utils/metrics.py -- average_displacement_error :37-54, final_displacement_error :7-26,
// initial_displacement_error :28-35, anglar_distance :60-88 (scipy Rotation.from_rotvec(...).as_quat(), float64),
// called per sample in a Python loop after a D2H copy (train.py:240-264, evaluate.py:127-152).  Here the poses stay in
// HBM (esb_traj_ids_to_poses writes them) and one warp scores one sample; the numbers are tiny, the win is removing the
// per-sample host round trip.  Displacements in fp32 like NumPy on float32 arrays, rotations in fp64 like scipy.
//
// Length rule of the reference functions: a generated trajectory longer than the ground truth is truncated, a shorter one
// is padded by repeating its last pose.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

namespace {

__device__ void rotvec_to_quat(const float* v, double* q) {
    const double x = v[0], y = v[1], z = v[2];
    const double a = sqrt(x * x + y * y + z * z);
    double s;
    if (a <= 1e-3) {
        const double a2 = a * a;
        s = 0.5 - a2 / 48.0 + a2 * a2 / 3840.0;
    } else {
        s = sin(a / 2.0) / a;
    }
    q[0] = s * x; q[1] = s * y; q[2] = s * z; q[3] = cos(a / 2.0);
}

// gen [B, Tg, D], gt [B, Tt, D] fp32; gen_len / gt_len [B] (NULL: full); out [B, 4] fp64 = (ADE, FDE, IDE, angular).
// Displacement over columns [c0, c1); rotation vector = columns [r0, r0+3) (r0 < 0: angular = NaN).
__global__ void __launch_bounds__(128) traj_metrics_kernel(const float* __restrict__ gen, const float* __restrict__ gt,
                                                           const int* __restrict__ gen_len, const int* __restrict__ gt_len, int B,
                                                           int Tg, int Tt, int D, int c0, int c1, int r0, double* __restrict__ out) {
    const int b = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= B) return;
    const int lg = gen_len ? gen_len[b] : Tg;
    const int lt = gt_len ? gt_len[b] : Tt;
    double* o = out + static_cast<long long>(b) * 4;
    if (lg <= 0 || lt <= 0) {                            // train.py:247-248: no parsable trajectory -> sample skipped
        if (lane == 0) { o[0] = o[1] = o[2] = o[3] = nan(""); }
        return;
    }
    const float* gp = gen + static_cast<long long>(b) * Tg * D;
    const float* tp = gt + static_cast<long long>(b) * Tt * D;
    float ade = 0.0f;
    double ang = 0.0;
    float fde = 0.0f, ide = 0.0f;
    for (int t = lane; t < lt; t += 32) {
        const float* g = gp + static_cast<long long>(t < lg ? t : lg - 1) * D;
        const float* r = tp + static_cast<long long>(t) * D;
        float ss = 0.0f;
        for (int c = c0; c < c1; ++c) { const float d = r[c] - g[c]; ss += d * d; }
        const float nrm = sqrtf(ss);
        ade += nrm;
        if (t == lt - 1) fde = nrm;
        if (t == 0) ide = nrm;
        if (r0 >= 0) {
            double qa[4], qb[4];
            rotvec_to_quat(g + r0, qa);
            rotvec_to_quat(r + r0, qb);
            double dot = qa[0] * qb[0] + qa[1] * qb[1] + qa[2] * qb[2] + qa[3] * qb[3];
            dot = dot > 1.0 ? 1.0 : (dot < -1.0 ? -1.0 : dot);
            ang += 2.0 * acos(dot);
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        ade += __shfl_xor_sync(0xffffffffu, ade, off);
        fde += __shfl_xor_sync(0xffffffffu, fde, off);
        ide += __shfl_xor_sync(0xffffffffu, ide, off);
        ang += __shfl_xor_sync(0xffffffffu, ang, off);
    }
    if (lane == 0) {
        // IDE of the reference compares gen[0] with gt[0] (no padding involved) -- same value as computed at t = 0
        o[0] = static_cast<double>(ade / static_cast<float>(lt));
        o[1] = static_cast<double>(fde);
        o[2] = static_cast<double>(ide);
        o[3] = r0 >= 0 ? ang / static_cast<double>(lt) : nan("");
    }
}

}  // namespace

int traj_metrics(const float* gen, const float* gt, const int* gen_len, const int* gt_len, int B, int Tg, int Tt, int D, int c0,
                 int c1, int r0, double* out, cudaStream_t s) {
    if (B <= 0) return ESB_OK;
    if (!gen || !gt || !out || Tg <= 0 || Tt <= 0 || D <= 0 || c0 < 0 || c1 > D || c0 >= c1 || r0 + 3 > D) {
        set_last_error("traj_metrics: bad arguments (Tg=%d Tt=%d D=%d cols [%d,%d) rot %d)", Tg, Tt, D, c0, c1, r0);
        return ESB_ERR_INVALID_ARG;
    }
    traj_metrics_kernel<<<(B + 3) / 4, 128, 0, s>>>(gen, gt, gen_len, gt_len, B, Tg, Tt, D, c0, c1, r0, out);
    return check_launch("traj_metrics");
}

}  // namespace esb
