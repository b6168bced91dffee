// Cell-major copies of the DENSE hash-grid levels (include/coslam_b200.h: cb200_cell_table_build / _fold).
//
// A dense level stores its vertices x + y*res + z*res^2 (tiny-cuda-nn grid.h; the first 5 levels of the Replica / TUM scenes, 6 of
// scene0000), so the 8 corners of the cell a point falls into lie in 4 different places of the table: a lane pair of the per-point
// kernels pays 4 requests per level for them (6 reduction requests in the backward, where an odd x splits the pair).  The copy made
// here holds the 8 corner values of every cell in ONE 64-byte block [corner x][corner (y, z)][2 features]: one 32-byte load per lane
// and level in the forward, two 16-byte reductions per lane and level in the backward.  8 x the dense levels' size (7.9 MB for room0),
// rebuilt from the table in front of a forward (~5 us) and folded into the table gradient behind a backward.
#include "args.cuh"

namespace cb200 {

// thread per (cell, corner x): 4 gathers -> one 32-byte store
__global__ void __launch_bounds__(256) cell_table_build_kernel(SceneDev sc, const float2* __restrict__ table, float4* __restrict__ cells,
                                                               uint32_t n_cells) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2u * n_cells) return;
    const uint32_t cell = t >> 1, xb = t & 1u;
    int l = 0;
    while (l + 1 < (int)sc.n_levels && !sc.hashed[l + 1] && cell >= sc.cell_offset[l + 1]) ++l;
    const uint32_t res = sc.res[l], size = sc.offset[l + 1] - sc.offset[l];
    uint32_t c = cell - sc.cell_offset[l];
    const uint32_t cx = c % res;
    c /= res;
    const uint32_t cy = c % res, cz = c / res;
    float2 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = table[sc.offset[l] + grid_index(false, size, res, cx + xb, cy + (j & 1), cz + (j >> 1))];
    float4* dst = cells + (size_t)cell * 4 + xb * 2;
    dst[0] = make_float4(v[0].x, v[0].y, v[1].x, v[1].y);
    dst[1] = make_float4(v[2].x, v[2].y, v[3].x, v[3].y);
}

// thread per (cell, corner x): the non-zero gradients of its 4 corners go to the table gradient
__global__ void __launch_bounds__(256) cell_table_fold_kernel(SceneDev sc, const float4* __restrict__ dcells, float2* __restrict__ dtable,
                                                              uint32_t n_cells) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2u * n_cells) return;
    const uint32_t cell = t >> 1, xb = t & 1u;
    const float4* src = dcells + (size_t)cell * 4 + xb * 2;
    const float4 a = src[0], b = src[1];
    const float2 g[4] = {make_float2(a.x, a.y), make_float2(a.z, a.w), make_float2(b.x, b.y), make_float2(b.z, b.w)};
    if (a.x == 0.f && a.y == 0.f && a.z == 0.f && a.w == 0.f && b.x == 0.f && b.y == 0.f && b.z == 0.f && b.w == 0.f) return;
    int l = 0;
    while (l + 1 < (int)sc.n_levels && !sc.hashed[l + 1] && cell >= sc.cell_offset[l + 1]) ++l;
    const uint32_t res = sc.res[l], size = sc.offset[l + 1] - sc.offset[l];
    uint32_t c = cell - sc.cell_offset[l];
    const uint32_t cx = c % res;
    c /= res;
    const uint32_t cy = c % res, cz = c / res;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (g[j].x != 0.f || g[j].y != 0.f)
            red_add_f2(dtable + sc.offset[l] + grid_index(false, size, res, cx + xb, cy + (j & 1), cz + (j >> 1)), g[j].x, g[j].y);
    }
}

// dense levels must come first (level 0 dense, no dense level behind a hashed one): that is how cell_offset is laid out
static uint32_t dense_cells_or_zero(const cb200_grid& g) {
    uint32_t k = 0;
    while (k < g.n_levels && !g.hashed[k]) ++k;
    for (uint32_t l = k; l < g.n_levels; ++l)
        if (!g.hashed[l]) return 0;
    return cell_table_cells(g, (int)k);
}

}  // namespace cb200

extern "C" {

size_t cb200_cell_table_bytes(const cb200_scene* scene) {
    if (scene == nullptr) return 0;
    return (size_t)cb200::dense_cells_or_zero(scene->grid) * 64;
}

int cb200_cell_table_build(const cb200_scene* scene, const float* table, void* cells, void* stream) {
    if (!scene || !table || !cells) return cb200_set_error(-1, "cell_table_build: scene, table and cells must be given");
    if (scene->grid.n_features != 2) return cb200_set_error(-1, "cell_table_build: written for 2 features per level");
    const uint32_t n = cb200::dense_cells_or_zero(scene->grid);
    if (n == 0) return cb200_set_error(-1, "cell_table_build: no leading dense levels in this grid");
    cb200_scene s = *scene;
    s.cells = nullptr;
    s.dcells = nullptr;
    cb200::cell_table_build_kernel<<<(2 * n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        cb200::make_scene_dev(s), reinterpret_cast<const float2*>(table), reinterpret_cast<float4*>(cells), n);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : cb200_set_error((int)e, "cell_table_build: %s", cudaGetErrorString(e));
}

int cb200_cell_table_fold(const cb200_scene* scene, const void* dcells, float* d_table, void* stream) {
    if (!scene || !dcells || !d_table) return cb200_set_error(-1, "cell_table_fold: scene, dcells and d_table must be given");
    const uint32_t n = cb200::dense_cells_or_zero(scene->grid);
    if (n == 0) return cb200_set_error(-1, "cell_table_fold: no leading dense levels in this grid");
    cb200_scene s = *scene;
    s.cells = nullptr;
    s.dcells = nullptr;
    cb200::cell_table_fold_kernel<<<(2 * n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        cb200::make_scene_dev(s), reinterpret_cast<const float4*>(dcells), reinterpret_cast<float2*>(d_table), n);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : cb200_set_error((int)e, "cell_table_fold: %s", cudaGetErrorString(e));
}

}  // extern "C"
