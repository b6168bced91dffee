// Unit-test kernel for the tcgen05 plumbing of tc.cuh: one CTA, one 128-row tile.
//   mode 0 (K-major views):   C[128 x N] = A[128 x K] * B[N x K]^T          (forward / input-gradient shape)
//   mode 1 (MN-major views):  C[128 x N] = X[128 x 128]^T * D[128 x N]      (weight-gradient shape: reduce over rows)
// Operands arrive as fp32 row-major in global memory, are split into bf16 hi/lo tiles exactly like the fused decoder
// does, multiplied with three MMAs per 16-wide reduction step and read back through tcgen05.ld.
#include "args.cuh"
#include "tc.cuh"

namespace cb200 {

__global__ void __launch_bounds__(128) tc_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B, int K, int N, int mode,
                                                          float* __restrict__ C) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int a_cols = mode == 0 ? K : 128, b_rows = mode == 0 ? N : 128, b_cols = mode == 0 ? K : N;
    const uint32_t a_gs = (uint32_t)(a_cols / 8) * 128u, b_gs = (uint32_t)(b_cols / 8) * 128u;
    const uint32_t a_bytes = 16u * a_gs, b_bytes = (uint32_t)((b_rows + 7) / 8) * b_gs;
    const uint32_t a_hi = smem_u32(smem), a_lo = a_hi + a_bytes, b_hi = a_lo + a_bytes, b_lo = b_hi + b_bytes;

    if (warp == 0) tmem_alloc<128>(&tmem_slot);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    for (int c = 0; c < a_cols / 8; ++c) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = A[tid * a_cols + c * 8 + i];
        tile_store8(a_hi, a_lo, a_gs, tid, c, v);
    }
    if (tid < b_rows) {
        for (int c = 0; c < b_cols / 8; ++c) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = B[tid * b_cols + c * 8 + i];
            tile_store8(b_hi, b_lo, b_gs, tid, c, v);
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (tid < 32 && elect_one()) {
        if (mode == 0) {
            umma_split3(tmem, a_hi, a_lo, 128u, a_gs, 256u, b_hi, b_lo, 128u, b_gs, 256u, K / 16, umma_idesc_bf16(128, N, false, false), false);
        } else {
            umma_split3(tmem, a_hi, a_lo, a_gs, 128u, 2u * a_gs, b_hi, b_lo, b_gs, 128u, 2u * b_gs, 128 / 16,
                        umma_idesc_bf16(128, N, true, true), false);
        }
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        tmem_ld_x16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) C[tid * N + c0 + i] = __uint_as_float(v[i]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<128>(tmem);
}

cudaError_t launch_tc_selftest(const float* A, const float* B, int K, int N, int mode, float* C, cudaStream_t st) {
    const int a_cols = mode == 0 ? K : 128, b_rows = mode == 0 ? N : 128, b_cols = mode == 0 ? K : N;
    const size_t smem = 2 * (size_t)16 * (a_cols / 8) * 128 + 2 * (size_t)((b_rows + 7) / 8) * (b_cols / 8) * 128 + 128;
    cudaError_t e = cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    tc_selftest_kernel<<<1, 128, smem, st>>>(A, B, K, N, mode, C);
    return cudaGetLastError();
}

}  // namespace cb200
