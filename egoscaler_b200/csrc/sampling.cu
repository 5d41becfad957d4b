// Sampling step of generate(): the reference's DEFAULT decode mode (model_arch.py:77-108: do_sample=True, top_k=50,
// top_p=0.95, temperature=1.0, repetition_penalty=1.0; callers train.py:221-231, evaluate.py:114-124).
//
// Arithmetic restated from HF transformers 5.5.0 (un-vendored dependency):
//   generation/utils.py  _sample: scores = processors(input_ids, logits[:, -1].float()); probs = softmax(scores);
//                        next = multinomial(probs, 1)
//   generation/logits_process.py
//     RepetitionPenaltyLogitsProcessor :302  s = where(s < 0, s * penalty, s / penalty) at every id already in the row
//     TemperatureLogitsWarper          :236  s = s / temperature
//     TopKLogitsWarper                 :536  remove s < (k-th largest s)          (ties with the k-th value are kept)
//     TopPLogitsWarper                 :469  ascending sort, cumsum(softmax) <= 1 - top_p removed, largest always kept
// The filter (which ids survive, and their fp32 scores) is exactly comparable with the reference on identical logits;
// the draw is an inverse-CDF walk over the survivors in descending-score order with a caller-supplied uniform u, so it
// is distribution-comparable with torch.multinomial (which uses a different, exponential-race algorithm).
//
// One CTA per row; the row lives in shared memory as fp32 (V = 32262 -> 126 KiB).  k-th largest value by a 4-pass
// 8-bit radix select over order-preserving keys, survivors (<= 2048 incl. ties) compacted and bitonic-sorted by
// (score desc, id desc) - reversed that is the stable ascending order torch.sort gives the top-p warper.
#include "common.cuh"
#include "esb_internal.h"
#include "esb_ops.h"

namespace esb {

namespace {

constexpr int SMP_THREADS = 1024;
constexpr int SMP_CAP = 2048;          // survivors of top-k incl. ties that fit the sort buffer

ESB_DEVICE uint32_t f2key(float f) {   // monotone: larger float -> larger key; NaN handled by caller
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct SampleParams {
    const __nv_bfloat16* logits;
    long long ld;
    int V;
    float temperature, top_p, penalty;
    int top_k;
    const long long* hist;      // [rows, hist_ld] ids seen so far (prompt + generated) or null
    long long hist_ld;
    int hist_len;               // + *pos_ptr when pos_ptr != null
    const int* pos_ptr;
    const float* u;             // uniform [0,1) draws: u[(step) * u_step_stride + row]
    long long u_step_stride;
    const int* step_ptr;        // step = *step_ptr + step_off (0 + step_off when null)
    int step_off;
    float* scores;              // optional processed scores [.., rows, V] fp32 (-inf where filtered)
    long long scores_ld, scores_step_stride;
    int suppress_col;
    long long* out_idx;
    int* n_kept;                // optional [rows]: survivors after top-p
};

__global__ void __launch_bounds__(SMP_THREADS, 1) sample_rows_kernel(const SampleParams p) {
    extern __shared__ float sv[];                         // [V]
    __shared__ int hist_bins[256];
    __shared__ float s_val[SMP_CAP];
    __shared__ int s_idx[SMP_CAP];
    __shared__ int s_count, s_keep, s_pick;
    __shared__ uint32_t s_prefix;
    __shared__ int s_remaining;

    const int row = blockIdx.x;
    const int tid = threadIdx.x;
    const int V = p.V;
    const __nv_bfloat16* lp = p.logits + static_cast<long long>(row) * p.ld;
    const int step = (p.step_ptr ? *p.step_ptr : 0) + p.step_off;

    // ---- 1. fp32 row (+ suppressed column)
    for (int i = tid; i < V; i += SMP_THREADS) {
        float v = __bfloat162float(lp[i]);
        if (i == p.suppress_col) v = -INFINITY;
        sv[i] = v;
    }
    __syncthreads();
    // ---- 2. repetition penalty: idempotent writes computed from the ORIGINAL logit (duplicates in the history are safe)
    if (p.hist && p.penalty != 1.0f) {
        const int hl = p.hist_len + (p.pos_ptr ? *p.pos_ptr : 0);
        const long long* hp = p.hist + static_cast<long long>(row) * p.hist_ld;
        for (int t = tid; t < hl; t += SMP_THREADS) {
            const long long id = hp[t];
            if (id >= 0 && id < V && id != p.suppress_col) {
                const float v0 = __bfloat162float(lp[id]);
                sv[id] = v0 < 0.0f ? v0 * p.penalty : v0 / p.penalty;
            }
        }
        __syncthreads();
    }
    // ---- 3. temperature
    if (p.temperature != 1.0f) {
        for (int i = tid; i < V; i += SMP_THREADS) sv[i] = sv[i] / p.temperature;
        __syncthreads();
    }
    // ---- 4. k-th largest key: MSB-first radix select
    const int k = p.top_k < 1 ? 1 : (p.top_k > V ? V : p.top_k);
    if (tid == 0) { s_prefix = 0u; s_remaining = k; }
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        if (tid < 256) hist_bins[tid] = 0;
        __syncthreads();
        const uint32_t prefix = s_prefix;
        const uint32_t mask_hi = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
        for (int i = tid; i < V; i += SMP_THREADS) {
            const uint32_t key = f2key(sv[i]);
            if ((key & mask_hi) == prefix) atomicAdd(&hist_bins[(key >> shift) & 255u], 1);
        }
        __syncthreads();
        if (tid == 0) {
            int rem = s_remaining;
            int b = 255;
            for (; b > 0; --b) {
                if (hist_bins[b] >= rem) break;
                rem -= hist_bins[b];
            }
            s_prefix = prefix | (static_cast<uint32_t>(b) << shift);
            s_remaining = rem;
        }
        __syncthreads();
    }
    const uint32_t kth_key = s_prefix;
    // ---- 5. compact survivors (key >= kth): strictly-greater ones are < k, ties all kept (HF: remove s < kth)
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int i = tid; i < V; i += SMP_THREADS) {
        const float v = sv[i];
        if (f2key(v) >= kth_key && v > -INFINITY) {
            const int slot = atomicAdd(&s_count, 1);
            if (slot < SMP_CAP) { s_val[slot] = v; s_idx[slot] = i; }
        }
    }
    __syncthreads();
    int n = s_count < SMP_CAP ? s_count : SMP_CAP;
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    for (int i = n + tid; i < n2; i += SMP_THREADS) { s_val[i] = -INFINITY; s_idx[i] = -1; }
    __syncthreads();
    // ---- 6. bitonic sort: (score desc, id desc)
    for (int size = 2; size <= n2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (n2 >> 1); t += SMP_THREADS) {
                const int lo = 2 * t - (t & (stride - 1));
                const int hi = lo + stride;
                const bool desc = ((lo & size) == 0);
                const float a = s_val[lo], b = s_val[hi];
                const int ia = s_idx[lo], ib = s_idx[hi];
                const bool a_first = (a > b) || (a == b && ia > ib);      // a belongs before b in the final order
                if (a_first != desc) { s_val[lo] = b; s_val[hi] = a; s_idx[lo] = ib; s_idx[hi] = ia; }
            }
            __syncthreads();
        }
    }
    // ---- 7. top-p over the survivors + inverse-CDF draw (sequential fp32, ascending like a CPU cumsum)
    if (tid == 0) {
        int keep = n;
        if (n > 0) {
            const float m = s_val[0];
            float Z = 0.0f;
            for (int j = 0; j < n; ++j) Z += expf(s_val[j] - m);
            if (p.top_p < 1.0f) {
                const float thr = 1.0f - p.top_p;
                float c = 0.0f;
                for (int j = n - 1; j >= 1; --j) {        // ascending order = reversed list; index 0 (largest) always kept
                    c += expf(s_val[j] - m) / Z;
                    if (c <= thr) keep = j; else break;
                }
            }
            float Zk = 0.0f;
            for (int j = 0; j < keep; ++j) Zk += expf(s_val[j] - m);
            const float uu = p.u ? p.u[static_cast<long long>(step) * p.u_step_stride + row] : 0.0f;
            const float target = uu * Zk;
            float c = 0.0f;
            int pick = keep - 1;
            for (int j = 0; j < keep; ++j) {
                c += expf(s_val[j] - m);
                if (c > target) { pick = j; break; }
            }
            s_pick = s_idx[pick];
        } else {
            s_pick = 0;
        }
        s_keep = keep;
    }
    __syncthreads();
    const int keep = s_keep;
    if (tid == 0) {
        p.out_idx[row] = s_pick;
        if (p.n_kept) p.n_kept[row] = keep;
    }
    // ---- 8. processed scores (HF output_scores): -inf everywhere except the survivors
    if (p.scores) {
        float* out = p.scores + static_cast<long long>(step) * p.scores_step_stride + static_cast<long long>(row) * p.scores_ld;
        for (int i = tid; i < V; i += SMP_THREADS) sv[i] = -INFINITY;
        __syncthreads();
        for (int j = tid; j < keep; j += SMP_THREADS) sv[s_idx[j]] = s_val[j];
        __syncthreads();
        for (int i = tid; i < V; i += SMP_THREADS) out[i] = sv[i];
    }
}

}  // namespace

int sample_rows(const void* logits, long long ld, long long rows, int V, float temperature, int top_k, float top_p, float penalty,
                const long long* hist, long long hist_ld, int hist_len, const int* pos_ptr, const float* u,
                long long u_step_stride, const int* step_ptr, int step_off, float* scores, long long scores_ld,
                long long scores_step_stride, int suppress_col, long long* out_idx, int* n_kept, cudaStream_t s) {
    if (rows <= 0) return ESB_OK;
    if (!logits || !out_idx || V <= 0 || top_k < 1 || top_k > 1024 || !(temperature > 0.0f) || !(top_p > 0.0f) || !(penalty > 0.0f)) {
        set_last_error("sample_rows: bad arguments (V=%d top_k=%d (1..1024) temperature=%g top_p=%g penalty=%g)", V, top_k,
                       (double)temperature, (double)top_p, (double)penalty);
        return ESB_ERR_INVALID_ARG;
    }
    const size_t smem = static_cast<size_t>(V) * sizeof(float);
    if (smem > 200 * 1024) {
        set_last_error("sample_rows: V=%d does not fit the shared-memory row buffer", V);
        return ESB_ERR_UNSUPPORTED;
    }
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(sample_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) {
            set_last_error("cudaFuncSetAttribute(sample_rows): %s", cudaGetErrorString(e));
            return ESB_ERR_CUDA;
        }
        attr_set = true;
    }
    SampleParams p;
    p.logits = static_cast<const __nv_bfloat16*>(logits); p.ld = ld; p.V = V;
    p.temperature = temperature; p.top_p = top_p; p.penalty = penalty; p.top_k = top_k;
    p.hist = hist; p.hist_ld = hist_ld; p.hist_len = hist_len; p.pos_ptr = pos_ptr;
    p.u = u; p.u_step_stride = u_step_stride; p.step_ptr = step_ptr; p.step_off = step_off;
    p.scores = scores; p.scores_ld = scores_ld; p.scores_step_stride = scores_step_stride;
    p.suppress_col = suppress_col; p.out_idx = out_idx; p.n_kept = n_kept;
    sample_rows_kernel<<<static_cast<unsigned>(rows), SMP_THREADS, smem, s>>>(p);
    return check_launch("sample_rows");
}

}  // namespace esb
