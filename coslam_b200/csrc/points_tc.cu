// Per-sample-point kernels, version 2 ("tc"): the decoder contractions run on the 5th-generation tensor cores.
//
// Why (profiles/r1/NOTES.md): in the SIMT version the L1TEX data pipe was the top unit and MORE than half of its
// wavefronts were shared-memory reads of decoder weights (fwd 55 %, bwd 64 %); only the rest were the hash-table
// gathers the path is really about.  Here one thread still owns one point for the gathers, but its 80 (+15) decoder
// inputs are written once into a shared-memory operand tile (bf16 hi/lo split, tc.cuh) and every layer is a
// tcgen05.mma with M = 128 points, accumulators in TMEM, read back with tcgen05.ld in the natural "lane = point"
// layout.  Weights are staged once per persistent CTA.
//
// Tile bookkeeping (bytes; "group stride" = distance between 8-row groups, see tc.cuh):
//   A   [128 x 96]  features: 32 grid | 48 OneBlob | 15 geo | 1 zero      group stride 1536   hi + lo = 48 KB
//   H   [128 x 32]  relu(h1) then relu(h3)                                 group stride  512   hi + lo = 16 KB
//   W1  [32 x 96] (cols 80.. zero)  W2 [16 x 32]  W3 [32 x 64] (col 63 zero)  W4 [16 x 32] (rows 3.. zero)   24 KB
// TMEM columns: h1 0..31 | h2 32..47 | h3 48..79 | rgb 80..95   (128 allocated)
#include "args.cuh"
#include "point_common.cuh"
#include "tc.cuh"

namespace cb200 {

namespace tc {

constexpr int kTile = 128;
constexpr uint32_t kAGS = 12 * 128, kABytes = 16 * kAGS;      // 24576
constexpr uint32_t kHGS = 4 * 128, kHBytes = 16 * kHGS;       // 8192
constexpr uint32_t kW1GS = 12 * 128, kW1Bytes = 4 * kW1GS;    // 6144
constexpr uint32_t kW2GS = 4 * 128, kW2Bytes = 2 * kW2GS;     // 1024
constexpr uint32_t kW3GS = 8 * 128, kW3Bytes = 4 * kW3GS;     // 4096
constexpr uint32_t kW4GS = 4 * 128, kW4Bytes = 2 * kW4GS;     // 1024

constexpr uint32_t kOffAHi = 0, kOffALo = kOffAHi + kABytes;
constexpr uint32_t kOffHHi = kOffALo + kABytes, kOffHLo = kOffHHi + kHBytes;
constexpr uint32_t kOffW1Hi = kOffHLo + kHBytes, kOffW1Lo = kOffW1Hi + kW1Bytes;
constexpr uint32_t kOffW2Hi = kOffW1Lo + kW1Bytes, kOffW2Lo = kOffW2Hi + kW2Bytes;
constexpr uint32_t kOffW3Hi = kOffW2Lo + kW2Bytes, kOffW3Lo = kOffW3Hi + kW3Bytes;
constexpr uint32_t kOffW4Hi = kOffW3Lo + kW3Bytes, kOffW4Lo = kOffW4Hi + kW4Bytes;
constexpr uint32_t kFwdSmem = kOffW4Lo + kW4Bytes;            // 90112

constexpr uint32_t kColH1 = 0, kColH2 = 32, kColH3 = 48, kColRgb = 80;
constexpr int kTmemCols = 128;

// stage one weight matrix [rows_valid x cols_valid] (row-major fp32, global) as a zero-padded [rows x 8*chunks] tile
__device__ __forceinline__ void stage_weight(uint32_t hi, uint32_t lo, uint32_t gs, const float* __restrict__ w, int rows, int chunks,
                                             int rows_valid, int cols_valid) {
    for (int i = threadIdx.x; i < rows * chunks; i += blockDim.x) {
        const int r = i / chunks, c = i % chunks;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int col = c * 8 + e;
            v[e] = (r < rows_valid && col < cols_valid) ? w[r * cols_valid + col] : 0.f;
        }
        tile_store8(hi, lo, gs, r, c, v);
    }
}

__device__ __forceinline__ void stage_all_weights(uint32_t sbase, const PointsArgs& a) {
    stage_weight(sbase + kOffW1Hi, sbase + kOffW1Lo, kW1GS, a.w1, 32, 12, 32, kSdfIn);
    stage_weight(sbase + kOffW2Hi, sbase + kOffW2Lo, kW2GS, a.w2, 16, 4, 16, kHidden);
    stage_weight(sbase + kOffW3Hi, sbase + kOffW3Lo, kW3GS, a.w3, 32, 8, 32, kColIn);
    stage_weight(sbase + kOffW4Hi, sbase + kOffW4Lo, kW4GS, a.w4, 16, 4, 3, kHidden);
}

// hash-grid features of one point -> chunks 0..3 of the A tile.  Four levels (32 gathers) are in flight per thread.
__device__ __forceinline__ void encode_grid_to_tile(const PointsArgs& a, const float x[3], uint32_t a_hi, uint32_t a_lo, int row) {
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
        float2 v[4][8];
        float w[4][3];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            LevelCorners lc;
            level_corners(a.scene, 4 * c + q, x, lc);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[q][k] = ldg_f2(a.table + lc.idx[k]);
            w[q][0] = lc.w[0]; w[q][1] = lc.w[1]; w[q][2] = lc.w[2];
        }
        float feat[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float f0 = 0.f, f1 = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float wc = corner_weight(w[q], k);
                f0 = fmaf(wc, v[q][k].x, f0);
                f1 = fmaf(wc, v[q][k].y, f1);
            }
            feat[2 * q] = f0;
            feat[2 * q + 1] = f1;
        }
        tile_store8(a_hi, a_lo, kAGS, row, c, feat);
    }
}

// OneBlob features -> chunks 4..9 of the A tile
__device__ __forceinline__ void encode_blob_to_tile(const float x[3], uint32_t a_hi, uint32_t a_lo, int row) {
#pragma unroll 1
    for (int d = 0; d < 3; ++d) {
        const float f_first = blob_edge_cdf(0.f, x[d], (float)kBins);
        float f_prev = f_first;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int b = half * 8 + e;
                const float f_next = (b == kBins - 1) ? f_first + 1.0f : blob_edge_cdf((float)(b + 1) / (float)kBins, x[d], (float)kBins);
                v[e] = f_next - f_prev;
                f_prev = f_next;
            }
            tile_store8(a_hi, a_lo, kAGS, row, 4 + d * 2 + half, v);
        }
    }
}

// make this thread's shared-memory writes visible to the tensor core, then CTA barrier
__device__ __forceinline__ void publish_tiles() {
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
}

}  // namespace tc

// ------------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(tc::kTile) points_fwd_tc_kernel(PointsArgs a) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t a_hi = sbase + kOffAHi, a_lo = sbase + kOffALo, h_hi = sbase + kOffHHi, h_lo = sbase + kOffHLo;

    if (warp == 0) tmem_alloc<kTmemCols>(&tmem_slot);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    stage_all_weights(sbase, a);
    publish_tiles();
    const uint32_t tmem = tmem_slot;
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);      // this warp's 32 TMEM lanes
    uint32_t parity = 0;

    const int64_t n_tiles = (a.n_points + kTile - 1) / kTile;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t p_raw = tile * kTile + tid;
        const bool valid = p_raw < a.n_points;
        const int64_t p = valid ? p_raw : a.n_points - 1;
        float x[3], zdummy;
        point_position(a, p, x, zdummy);
        encode_grid_to_tile(a, x, a_hi, a_lo, tid);
        encode_blob_to_tile(x, a_hi, a_lo, tid);

        // ---- SDF net layer 1: h1[128,32] = in1[128,80] W1^T ----------------------------------------------------------
        publish_tiles();
        if (tid == 0) {
            umma_split3(tmem + kColH1, a_hi, a_lo, 128u, kAGS, 256u, sbase + kOffW1Hi, sbase + kOffW1Lo, 128u, kW1GS, 256u, 5,
                        umma_idesc_bf16(128, 32, false, false), false);
            umma_commit(&bar);
        }
        mbar_wait(&bar, parity);
        parity ^= 1u;
        tc_fence_after();
        uint32_t m1 = 0;
        {
            uint32_t r[32];
            tmem_ld_x32(trow + kColH1, r);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float h = __uint_as_float(r[c * 8 + e]);
                    const bool on = h > 0.f;
                    m1 |= on ? (1u << (c * 8 + e)) : 0u;
                    v[e] = on ? h : 0.f;
                }
                tile_store8(h_hi, h_lo, kHGS, tid, c, v);
            }
        }
        // ---- layer 2: h2[128,16] = relu(h1) W2^T ------------------------------------------------------------------------
        publish_tiles();
        if (tid == 0) {
            umma_split3(tmem + kColH2, h_hi, h_lo, 128u, kHGS, 256u, sbase + kOffW2Hi, sbase + kOffW2Lo, 128u, kW2GS, 256u, 2,
                        umma_idesc_bf16(128, 16, false, false), false);
            umma_commit(&bar);
        }
        mbar_wait(&bar, parity);
        parity ^= 1u;
        tc_fence_after();
        float sdf;
        {
            uint32_t r[16];
            tmem_ld_x16(trow + kColH2, r);
            tmem_ld_wait();
            sdf = __uint_as_float(r[0]);
            if (a.geo && valid) {
#pragma unroll
                for (int g = 0; g < kGeo; ++g) a.geo[p * kGeo + g] = __uint_as_float(r[1 + g]);
            }
            if (!a.sdf_only) {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[1 + e]);
                tile_store8(a_hi, a_lo, kAGS, tid, 10, v);
#pragma unroll
                for (int e = 0; e < 7; ++e) v[e] = __uint_as_float(r[9 + e]);
                v[7] = 0.f;
                tile_store8(a_hi, a_lo, kAGS, tid, 11, v);
            }
        }
        float rgb[3] = {0.f, 0.f, 0.f};
        uint32_t m3 = 0;
        if (!a.sdf_only) {
            // ---- colour net layer 1: h3[128,32] = [blob | geo][128,64] W3^T ------------------------------------------------
            publish_tiles();
            if (tid == 0) {
                umma_split3(tmem + kColH3, a_hi + 4 * 128, a_lo + 4 * 128, 128u, kAGS, 256u, sbase + kOffW3Hi, sbase + kOffW3Lo, 128u, kW3GS,
                            256u, 4, umma_idesc_bf16(128, 32, false, false), false);
                umma_commit(&bar);
            }
            mbar_wait(&bar, parity);
            parity ^= 1u;
            tc_fence_after();
            {
                uint32_t r[32];
                tmem_ld_x32(trow + kColH3, r);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float v[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float h = __uint_as_float(r[c * 8 + e]);
                        const bool on = h > 0.f;
                        m3 |= on ? (1u << (c * 8 + e)) : 0u;
                        v[e] = on ? h : 0.f;
                    }
                    tile_store8(h_hi, h_lo, kHGS, tid, c, v);
                }
            }
            // ---- colour net layer 2: rgb[128,16(3)] = relu(h3) W4^T --------------------------------------------------------
            publish_tiles();
            if (tid == 0) {
                umma_split3(tmem + kColRgb, h_hi, h_lo, 128u, kHGS, 256u, sbase + kOffW4Hi, sbase + kOffW4Lo, 128u, kW4GS, 256u, 2,
                            umma_idesc_bf16(128, 16, false, false), false);
                umma_commit(&bar);
            }
            mbar_wait(&bar, parity);
            parity ^= 1u;
            tc_fence_after();
            uint32_t r[4];
            tmem_ld_x4(trow + kColRgb, r);
            tmem_ld_wait();
            rgb[0] = __uint_as_float(r[0]);
            rgb[1] = __uint_as_float(r[1]);
            rgb[2] = __uint_as_float(r[2]);
        }
        if (valid) {
            if (a.raw) reinterpret_cast<float4*>(a.raw)[p] = make_float4(rgb[0], rgb[1], rgb[2], sdf);
            if (a.act_mask) reinterpret_cast<uint2*>(a.act_mask)[p] = make_uint2(m1, m3);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<tc::kTmemCols>(tmem);
}

static int tc_persistent_grid(int64_t n_tiles, int ctas_per_sm) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t g = (int64_t)sms * ctas_per_sm;
    if (g > n_tiles) g = n_tiles;
    if (g < 1) g = 1;
    return (int)g;
}

cudaError_t launch_points_fwd_tc(const PointsArgs& a, cudaStream_t st) {
    const size_t smem = tc::kFwdSmem;
    cudaError_t e = cudaFuncSetAttribute(points_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int64_t n_tiles = (a.n_points + tc::kTile - 1) / tc::kTile;
    points_fwd_tc_kernel<<<tc_persistent_grid(n_tiles, 2), tc::kTile, smem, st>>>(a);
    return cudaGetLastError();
}

}  // namespace cb200
