// Device helpers shared by the per-point kernels (points.cu: SIMT decoder, points_tc.cu: tcgen05 decoder).
#pragma once
#include "args.cuh"

namespace cb200 {

// position of point p in the unit cube (model/scene_rep.py:258, :186); z_out = its ray parameter (0 for explicit points)
__device__ __forceinline__ void point_position(const PointsArgs& a, int64_t p, float x[3], float& z_out) {
    float pt[3];
    if (a.rays_o) {
        const int64_t ray = p / a.n_samples;
        const float z = a.z[p];
        z_out = z;
#pragma unroll
        for (int d = 0; d < 3; ++d) pt[d] = ray_point(a.rays_o[ray * 3 + d], a.rays_d[ray * 3 + d], z);
    } else {
        z_out = 0.f;
#pragma unroll
        for (int d = 0; d < 3; ++d) pt[d] = a.pts[p * 3 + d];
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) x[d] = a.scene.normalise ? box_normalise(pt[d], a.scene.bb_lo[d], a.scene.bb_ext[d]) : pt[d];
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
