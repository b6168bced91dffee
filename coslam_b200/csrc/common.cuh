// Shared device helpers of the coslam_b200 kernels (sm_100a).
//
// Rounding contract (hash indices and sample positions are bit-exact):
//   - sample positions follow torch eager arithmetic: every multiply, add, subtract and divide is a separately
//     rounded fp32 operation (model/scene_rep.py:251-258, :186) -> __fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn;
//   - the grid position is tiny-cuda-nn's fused multiply-add pos = fmaf(scale, x, 0.5f) followed by floorf and a
//     (uint32_t)(int) cast.
// Everything that is only tolerance-gated (interpolation, MLP, rendering) may contract to FMA freely.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/coslam_b200.h"

namespace cb200 {

constexpr int kLevels = CB200_MAX_LEVELS;
constexpr int kGridFeats = CB200_GRID_FEATS;   // 32
constexpr int kBlobFeats = CB200_BLOB_FEATS;   // 48
constexpr int kBins = 16;
constexpr int kSdfIn = CB200_SDF_IN;           // 80
constexpr int kHidden = CB200_HIDDEN;          // 32
constexpr int kSdfOut = CB200_SDF_OUT;         // 16
constexpr int kGeo = kSdfOut - 1;              // 15
constexpr int kColIn = CB200_COL_IN;           // 63

// kernel-parameter copy of cb200_scene (passed by value: 4 KB parameter space is plenty)
struct SceneDev {
    uint32_t n_levels;
    uint32_t offset[kLevels + 1];
    uint32_t res[kLevels];
    uint32_t hashed[kLevels];
    float scale[kLevels];
    float bb_lo[3];
    float bb_ext[3];
    uint32_t normalise;
    // optional cell-major copies of the dense levels (cb200_cell_table_build): block of cell (cx, cy, cz) of dense level l =
    // cells + 4 * (cell_offset[l] + cx + res * (cy + res * cz)), 4 float4 = [corner x][corner (y, z)][2 features]
    const float4* cells;
    float4* dcells;
    uint32_t cell_offset[kLevels];
};

// number of cells (one per vertex, so that edge cells exist) of the dense levels in front of level l
__host__ inline uint32_t cell_table_cells(const cb200_grid& g, int upto) {
    uint32_t n = 0;
    for (int l = 0; l < upto && l < (int)g.n_levels; ++l)
        if (!g.hashed[l]) n += g.res[l] * g.res[l] * g.res[l];
    return n;
}

__host__ inline SceneDev make_scene_dev(const cb200_scene& s) {
    SceneDev d;
    d.n_levels = s.grid.n_levels;
    for (int l = 0; l <= kLevels; ++l) d.offset[l] = s.grid.offset[l];
    for (int l = 0; l < kLevels; ++l) {
        d.res[l] = s.grid.res[l];
        d.hashed[l] = s.grid.hashed[l];
        d.scale[l] = s.grid.scale[l];
    }
    for (int i = 0; i < 3; ++i) {
        d.bb_lo[i] = s.bb_lo[i];
        d.bb_ext[i] = s.bb_ext[i];
    }
    d.normalise = s.normalise;
    d.cells = reinterpret_cast<const float4*>(s.cells);
    d.dcells = reinterpret_cast<float4*>(s.dcells);
    for (int l = 0; l < kLevels; ++l) d.cell_offset[l] = cell_table_cells(s.grid, l);
    return d;
}

__host__ inline SceneDev make_scene_dev(const cb200_grid& g) {
    cb200_scene s{};
    s.grid = g;
    s.normalise = 0;
    for (int i = 0; i < 3; ++i) s.bb_ext[i] = 1.f;
    return make_scene_dev(s);
}

// ---- sample position: model/scene_rep.py:258 then :186 --------------------------------------------------------
__device__ __forceinline__ float ray_point(float o, float d, float z) { return __fadd_rn(o, __fmul_rn(d, z)); }
__device__ __forceinline__ float box_normalise(float p, float lo, float ext) { return __fdiv_rn(__fsub_rn(p, lo), ext); }

// ---- tiny-cuda-nn grid position / index -------------------------------------------
__device__ __forceinline__ void grid_pos(float scale, float x, uint32_t& cell, float& w) {
    const float pos = fmaf(scale, x, 0.5f);
    const float fl = floorf(pos);
    w = pos - fl;
    cell = (uint32_t)(int)fl;
}

__device__ __forceinline__ uint32_t grid_index(bool hashed, uint32_t size, uint32_t res, uint32_t px, uint32_t py,
                                               uint32_t pz) {
    if (hashed) {
        // coherent prime hash; a hashed level always has size == 2^log2_hashmap_size
        return (px ^ (py * 2654435761u) ^ (pz * 805459861u)) & (size - 1u);
    }
    uint32_t idx = px + py * res + pz * res * res;
    if (idx >= size) idx %= size;      // only for corners outside the unit cube
    return idx;
}

__device__ __forceinline__ float2 ldg_f2(const float2* p) { return __ldg(p); }

__device__ __forceinline__ void red_add_f2(float2* addr, float a, float b) {
    // one vectorised reduction per table entry (sm_90+): red.global.add.v2.f32
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

// ---- OneBlob (tiny-cuda-nn oneblob.h / common_device.h) ----------------------------------
__device__ __forceinline__ float quartic_cdf(float t, float inv_radius) {
    const float u = t * inv_radius;
    const float u2 = u * u;
    const float u4 = u2 * u2;
    const float v = (15.0f / 16.0f) * u * (1.0f - (2.0f / 3.0f) * u2 + (1.0f / 5.0f) * u4) + 0.5f;
    return fminf(fmaxf(v, 0.0f), 1.0f);
}
__device__ __forceinline__ float quartic_pdf(float t, float inv_radius) {
    const float u = t * inv_radius;
    const float a = fmaxf(1.0f - u * u, 0.0f);
    return (15.0f / 16.0f) * a * a * inv_radius;
}
// periodic (3 images) CDF / kernel at bin edge e for coordinate x
__device__ __forceinline__ float blob_edge_cdf(float e, float x, float nb) {
    const float t = e - x;
    return quartic_cdf(t, nb) + quartic_cdf(t - 1.0f, nb) + quartic_cdf(t + 1.0f, nb);
}
__device__ __forceinline__ float blob_edge_pdf(float e, float x, float nb) {
    const float t = e - x;
    return quartic_pdf(t, nb) + quartic_pdf(t - 1.0f, nb) + quartic_pdf(t + 1.0f, nb);
}

__device__ __forceinline__ float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }
// The rendering / loss-head kernels (tolerance-gated at 1e-3 relative) evaluate five sigmoids per sample: with
// IEEE division and expf they were ~40 % of those kernels' instructions (render_fwd_kernel<2>: 15 division slow-path calls, 24 BSSY).
// ex2.approx + rcp.approx: relative error ~|x| 2^-21 + 2 ulp (2e-5 at |x| = 50), saturates to 0 / 1 like the exact form.
__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }

// ---- Philox4x32-10 (counter-based noise for the non-parity "device noise" mode) --------------------------------
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}
__device__ __forceinline__ float u32_to_unit(uint32_t v) { return (float)(v >> 8) * (1.0f / 16777216.0f); }

}  // namespace cb200
