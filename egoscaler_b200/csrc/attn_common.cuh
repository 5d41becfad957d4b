// mma.sync (HMMA m16n8k16 bf16) building blocks shared by the attention kernels.
// Attention is ~1.4 % of the model FLOPs; the dense 98 % runs on tcgen05 (gemm_tcgen05.cu).
#pragma once
#include "common.cuh"

namespace esb {

ESB_DEVICE void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz) : "memory");
}
ESB_DEVICE void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
ESB_DEVICE void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

ESB_DEVICE void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
ESB_DEVICE void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
ESB_DEVICE void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Shared-memory tile [rows][HD] bf16.  Power-of-two HD (64, 128): 16-byte chunks XOR-swizzled by (row & 7) so ldmatrix
// is conflict-free.  Other HD (96 = PointBERT base, 1152 / 12 heads): rows padded by one 16-byte chunk instead (an odd
// number of chunks per row spreads 8 consecutive rows over the 8 bank groups).
template <int HD>
constexpr int tile_row_bytes() { return (HD & (HD - 1)) == 0 ? HD * 2 : HD * 2 + 16; }
template <int HD>
ESB_DEVICE uint32_t tile_off(int row, int chunk) {
    if constexpr ((HD & (HD - 1)) == 0) return static_cast<uint32_t>(row * (HD * 2) + ((chunk ^ (row & 7)) << 4));
    else return static_cast<uint32_t>(row * tile_row_bytes<HD>() + (chunk << 4));
}

// Async load of a [ROWS x HD] tile: global row r is base + (row0 + r) * row_stride (elements); rows >= nrows -> zeros.
template <int HD, int ROWS, int THREADS>
ESB_DEVICE void load_tile_async(uint8_t* smem_tile, const __nv_bfloat16* base, long long row_stride, int row0, int nrows) {
    constexpr int CPR = HD / 8;
    constexpr int TOTAL = ROWS * CPR;
#pragma unroll
    for (int i = threadIdx.x; i < TOTAL; i += THREADS) {
        const int r = i / CPR, c = i % CPR;
        const int gr = row0 + r;
        const bool ok = gr < nrows;
        const __nv_bfloat16* src = base + static_cast<long long>(ok ? gr : 0) * row_stride + c * 8;
        cp_async16(smem_tile + tile_off<HD>(r, c), src, ok);
    }
}

}  // namespace esb
