// Internal declarations of the non-GEMM operators (namespace esb); see include/egoscaler_b200.h for semantics.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace esb {

// traj_tokens.cu
int traj_quantize(const void* x, int is_f64, long long n, int num_bins, long long id_offset, long long* out, cudaStream_t s);
int traj_dequantize(const long long* ids, long long n_poses, int num_bins, long long id_offset, int mode, double d_max,
                    double d_min, float* out, cudaStream_t s);
int traj_scale(float* poses, long long n_poses, int mode, double d_max, double d_min, cudaStream_t s);
int traj_ids_to_poses(const long long* ids, int B, int T, int num_bins, long long base, long long tsep, long long eos,
                      int mode, double d_max, double d_min, int max_steps, int pad_repeat, float* out, int* count,
                      cudaStream_t s);

}  // namespace esb
