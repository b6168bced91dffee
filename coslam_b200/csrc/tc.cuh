// tcgen05 / TMEM helpers for the fused decoder (sm_100a).
//
// Operand tiles live in shared memory in the canonical NO-SWIZZLE ("interleave") UMMA layout for 16-bit elements:
//   a tile is a [rows x cols] bf16 matrix cut into core matrices of 8 rows x 16 bytes (8 elements), each core matrix
//   128 contiguous bytes;   byte offset of element (r, c) = (r/8)*group_stride + (c/8)*128 + (r%8)*16 + (c%8)*2.
// The same bytes serve two views of the matrix (CUTLASS cute/atom/mma_traits_sm100.hpp, "canonical layouts"):
//   K-major  (reduce over c):  leading-byte-offset = 128 (next 8 columns), stride-byte-offset = group_stride (next 8 rows)
//   MN-major (reduce over r):  leading-byte-offset = group_stride (next 8 rows), stride-byte-offset = 128 (next 8 columns)
// fp32 values are split into bf16 hi + bf16 lo (x = hi + lo + O(2^-17 |x|)); a product is issued as three MMAs
// (hi*hi + lo*hi + hi*lo) accumulated in fp32 in TMEM, which keeps the decoder within ~1e-5 of fp32 FMA arithmetic
// while the tensor core, not the LSU/shared-memory pipe, does the contractions.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

#include "tc_ld.cuh"

namespace cb200 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- descriptors ---------------------------------------------------------------------------------------------------
// shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (Blackwell)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46);
}
// instruction descriptor for kind::f16, bf16 x bf16 -> fp32
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) /* D = f32 */ | (1u << 7) /* A = bf16 */ | (1u << 10) /* B = bf16 */ | ((a_mn_major ? 1u : 0u) << 15) |
           ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- TMEM allocation (one warp) ----------------------------------------------------------------------------------------
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "n"(kCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
}
template <int kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols));
}

// ---- fences / barriers -------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// the mbarrier receives one arrival when every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// one lane of a fully active warp (the MMA issue / commit blocks run under `warp == 0 && elect_one()`: a warp-uniform
// branch plus elect.sync, so the compiler emits straight-line UTCHMMA issue instead of a per-instruction election loop)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

// D[tmem] (+)= A[smem] * B[smem]^T ; one thread issues
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// ---- operand tiles -----------------------------------------------------------------------------------------------------
// fp32 pair -> packed bf16x2 (element 0 in the low half) of the round-to-nearest bf16 and of the residual
__device__ __forceinline__ void split_bf16x2(float a, float b, uint32_t& hi, uint32_t& lo) {
    uint32_t h;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(b), "f"(a));
    const float ra = a - __uint_as_float(h << 16);
    const float rb = b - __uint_as_float(h & 0xffff0000u);
    uint32_t l;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(l) : "f"(rb), "f"(ra));
    hi = h;
    lo = l;
}

// write 8 consecutive columns (chunk) of row `r` into the hi and lo tiles (byte addresses in the shared window)
__device__ __forceinline__ void tile_store8(uint32_t hi_base, uint32_t lo_base, uint32_t group_stride, int r, int chunk, const float (&v)[8]) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) split_bf16x2(v[2 * i], v[2 * i + 1], h[i], l[i]);
    const uint32_t off = (uint32_t)(r >> 3) * group_stride + (uint32_t)chunk * 128u + (uint32_t)(r & 7) * 16u;
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(hi_base + off), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3]) : "memory");
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(lo_base + off), "r"(l[0]), "r"(l[1]), "r"(l[2]), "r"(l[3]) : "memory");
}

// one logical fp32 product  D (+)= A * B^T  over `ksteps` x 16 reduction elements, as hi*hi + lo*hi + hi*lo.
//   a_*/b_*: tile byte addresses at the first reduction element; *_kstep: byte advance per 16 reduction elements
//   (2*128 for a K-major view, 2*group_stride for an MN-major view); *_lbo/*_sbo as in the header comment.
__device__ __forceinline__ void umma_split3(uint32_t tmem_d, uint32_t a_hi, uint32_t a_lo, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_kstep,
                                            uint32_t b_hi, uint32_t b_lo, uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep, int ksteps,
                                            uint32_t idesc, bool accumulate_first) {
    uint32_t acc = accumulate_first ? 1u : 0u;
    // base descriptors once, then the start-address field (bits 0..13, units of 16 bytes) advances per reduction step
    uint64_t ah = umma_desc(a_hi, a_lbo, a_sbo), al = umma_desc(a_lo, a_lbo, a_sbo);
    uint64_t bh = umma_desc(b_hi, b_lbo, b_sbo), bl = umma_desc(b_lo, b_lbo, b_sbo);
    const uint64_t a_inc = a_kstep >> 4, b_inc = b_kstep >> 4;
#pragma unroll
    for (int k = 0; k < ksteps; ++k) {
        umma_bf16(tmem_d, ah, bh, idesc, acc);
        umma_bf16(tmem_d, al, bh, idesc, 1u);
        umma_bf16(tmem_d, ah, bl, idesc, 1u);
        acc = 1u;
        ah += a_inc; al += a_inc; bh += b_inc; bl += b_inc;
    }
}

}  // namespace cb200
