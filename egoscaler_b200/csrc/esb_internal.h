// Internal (C++) declarations shared by the translation units of libegoscaler_b200.so.
// The public C ABI is include/egoscaler_b200.h; everything here is namespace esb.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>

namespace esb {

void set_last_error(const char* fmt, ...);
const char* last_error();
int check_launch(const char* what);
int device_sm_count();

int make_tma_2d_bf16(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld, int box_rows);

int gemm_bf16(const void* A, long long lda, const void* B, long long ldb, void* C, long long ldc, int M, int N, int K,
              const float* bias, const void* residual, long long ldr, int act, int out_fp32, int accumulate, float alpha,
              int force_bn, cudaStream_t stream);

}  // namespace esb
