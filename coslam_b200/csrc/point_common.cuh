// Device helpers shared by the per-point kernels (points.cu: SIMT decoder, points_tc.cu: tcgen05 decoder).
#pragma once
#include "args.cuh"

namespace cb200 {

// position of point p in the unit cube (model/scene_rep.py:258, :186); z_out = its ray parameter (0 for explicit points).
// Split into the global loads and the arithmetic so that a tile loop can issue the loads of its NEXT tile's points early (the chain
// list entry -> ray / sample -> position is three dependent global loads: ~6 % of the forward's warp-stall samples when it sits at
// the top of a tile, profiles/r2/NOTES.md section 19).
struct PointLoads {
    float o[3], d[3], z;      // ray mode: origin, direction, ray parameter; explicit points: o = the point
};

__device__ __forceinline__ void point_loads(const PointsArgs& a, int64_t p, PointLoads& L) {
    if (a.rays_o) {
        const int64_t ray = p / a.n_samples;
        L.z = a.z[p];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            L.o[d] = a.rays_o[ray * 3 + d];
            L.d[d] = a.rays_d[ray * 3 + d];
        }
    } else {
        L.z = 0.f;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            L.o[d] = a.pts[p * 3 + d];
            L.d[d] = 0.f;
        }
    }
}

__device__ __forceinline__ void point_from_loads(const PointsArgs& a, const PointLoads& L, float x[3], float& z_out) {
    float pt[3];
    z_out = L.z;
#pragma unroll
    for (int d = 0; d < 3; ++d) pt[d] = a.rays_o ? ray_point(L.o[d], L.d[d], L.z) : L.o[d];
#pragma unroll
    for (int d = 0; d < 3; ++d) x[d] = a.scene.normalise ? box_normalise(pt[d], a.scene.bb_lo[d], a.scene.bb_ext[d]) : pt[d];
}

__device__ __forceinline__ void point_position(const PointsArgs& a, int64_t p, float x[3], float& z_out) {
    PointLoads L;
    point_loads(a, p, L);
    point_from_loads(a, L, x, z_out);
}

struct LevelCorners {
    uint32_t idx[8];
    float w[3];
};

__device__ __forceinline__ void level_corners(const SceneDev& sc, int l, const float x[3], LevelCorners& lc) {
    const float scale = sc.scale[l];
    const uint32_t res = sc.res[l];
    const uint32_t size = sc.offset[l + 1] - sc.offset[l];
    const bool hashed = sc.hashed[l] != 0;
    uint32_t cell[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) grid_pos(scale, x[d], cell[d], lc.w[d]);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        lc.idx[c] = sc.offset[l] + grid_index(hashed, size, res, cell[0] + (c & 1), cell[1] + ((c >> 1) & 1), cell[2] + (c >> 2));
    }
}

__device__ __forceinline__ float corner_weight(const float w[3], int c) {
    return ((c & 1) ? w[0] : 1.f - w[0]) * ((c & 2) ? w[1] : 1.f - w[1]) * ((c & 4) ? w[2] : 1.f - w[2]);
}

}  // namespace cb200
