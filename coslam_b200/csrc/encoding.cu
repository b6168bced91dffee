// Stand-alone encodings behind the tinycudann-compatible lower boundary (model/encodings.py:31-46, 61-71) and the
// feature-grid smoothness term (coslam.py:251-274 -> model/scene_rep.py:126-139 query_sdf(embed=True)).
// The fused per-point kernels in points.cu share the index / interpolation helpers of common.cuh with these.
#include "args.cuh"

namespace cb200 {

// thread (p, l): one level of one point, like tiny-cuda-nn's kernel_grid -- 8 independent gathers per thread and
// blockIdx.y = level keeps each CTA inside one level's slice of the table (L1/L2 locality on the coarse levels).
__global__ void __launch_bounds__(256) hash_fwd_kernel(SceneDev sc, const float* __restrict__ x, const float2* __restrict__ table,
                                                       int64_t n_points, float* __restrict__ out) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int l = blockIdx.y;
    if (p >= n_points) return;
    const float xp[3] = {x[p * 3], x[p * 3 + 1], x[p * 3 + 2]};
    const uint32_t size = sc.offset[l + 1] - sc.offset[l];
    uint32_t cell[3];
    float w[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) grid_pos(sc.scale[l], xp[d], cell[d], w[d]);
    float2 v[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const uint32_t idx = grid_index(sc.hashed[l] != 0, size, sc.res[l], cell[0] + (c & 1), cell[1] + ((c >> 1) & 1), cell[2] + (c >> 2));
        v[c] = ldg_f2(table + sc.offset[l] + idx);
    }
    float f0 = 0.f, f1 = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float wc = ((c & 1) ? w[0] : 1.f - w[0]) * ((c & 2) ? w[1] : 1.f - w[1]) * ((c & 4) ? w[2] : 1.f - w[2]);
        f0 = fmaf(wc, v[c].x, f0);
        f1 = fmaf(wc, v[c].y, f1);
    }
    reinterpret_cast<float2*>(out)[p * sc.n_levels + l] = make_float2(f0, f1);
}

__global__ void __launch_bounds__(256) hash_indices_kernel(SceneDev sc, const float* __restrict__ x, int64_t n_points, uint32_t* __restrict__ idx_out) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int l = blockIdx.y;
    if (p >= n_points) return;
    const uint32_t size = sc.offset[l + 1] - sc.offset[l];
    uint32_t cell[3];
    float w[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) grid_pos(sc.scale[l], x[p * 3 + d], cell[d], w[d]);
#pragma unroll
    for (int c = 0; c < 8; ++c)
        idx_out[(p * sc.n_levels + l) * 8 + c] =
            grid_index(sc.hashed[l] != 0, size, sc.res[l], cell[0] + (c & 1), cell[1] + ((c >> 1) & 1), cell[2] + (c >> 2));
}

// backward: thread per point, loop over levels (dx needs the sum over levels; the scatter is one v2 reduction/corner)
__global__ void __launch_bounds__(256) hash_bwd_kernel(SceneDev sc, const float* __restrict__ x, const float2* __restrict__ table,
                                                       const float* __restrict__ dy, int64_t n_points, float2* dtable, float* __restrict__ dx_out) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_points) return;
    const float xp[3] = {x[p * 3], x[p * 3 + 1], x[p * 3 + 2]};
    float dx[3] = {0.f, 0.f, 0.f};
    const int L = sc.n_levels;
    for (int l = 0; l < L; ++l) {
        const uint32_t size = sc.offset[l + 1] - sc.offset[l];
        uint32_t cell[3];
        float w[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) grid_pos(sc.scale[l], xp[d], cell[d], w[d]);
        const float2 g = reinterpret_cast<const float2*>(dy)[p * L + l];
        uint32_t idx[8];
#pragma unroll
        for (int c = 0; c < 8; ++c)
            idx[c] = sc.offset[l] + grid_index(sc.hashed[l] != 0, size, sc.res[l], cell[0] + (c & 1), cell[1] + ((c >> 1) & 1), cell[2] + (c >> 2));
        if (dx_out) {
            float s[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float2 v = ldg_f2(table + idx[c]);
                s[c] = v.x * g.x + v.y * g.y;
            }
            const float w0 = w[0], w1 = w[1], w2 = w[2];
            dx[0] += sc.scale[l] * ((1.f - w1) * (1.f - w2) * (s[1] - s[0]) + w1 * (1.f - w2) * (s[3] - s[2]) +
                                    (1.f - w1) * w2 * (s[5] - s[4]) + w1 * w2 * (s[7] - s[6]));
            dx[1] += sc.scale[l] * ((1.f - w0) * (1.f - w2) * (s[2] - s[0]) + w0 * (1.f - w2) * (s[3] - s[1]) +
                                    (1.f - w0) * w2 * (s[6] - s[4]) + w0 * w2 * (s[7] - s[5]));
            dx[2] += sc.scale[l] * ((1.f - w0) * (1.f - w1) * (s[4] - s[0]) + w0 * (1.f - w1) * (s[5] - s[1]) +
                                    (1.f - w0) * w1 * (s[6] - s[2]) + w0 * w1 * (s[7] - s[3]));
        }
        if (dtable) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float wc = ((c & 1) ? w[0] : 1.f - w[0]) * ((c & 2) ? w[1] : 1.f - w[1]) * ((c & 4) ? w[2] : 1.f - w[2]);
                red_add_f2(dtable + idx[c], wc * g.x, wc * g.y);
            }
        }
    }
    if (dx_out) {
        dx_out[p * 3 + 0] = dx[0];
        dx_out[p * 3 + 1] = dx[1];
        dx_out[p * 3 + 2] = dx[2];
    }
}

// table gradient only (the smoothness term, coslam.py:251-274: the lattice points are constants): thread (p, l) like the
// forward.  Neighbouring lanes that sit in the same cell with the same y / z weights (a row of an axis-aligned lattice whose
// fastest axis is x, cb200_smooth_points) are summed in the warp first -- sum_p wx wy wz g = wy wz sum_p wx g, so four values
// per lane are reduced, and the first lane of each run issues the eight reductions.  The same holds with z fastest (the
// reference's own lattice order, CoSLAM.smoothness through autograd).  Any other point set has no such runs and takes the
// direct path.
__global__ void __launch_bounds__(256) hash_bwd_table_kernel(SceneDev sc, const float* __restrict__ x, const float* __restrict__ dy,
                                                             int64_t n_points, float2* dtable) {
    // level is the FASTEST block coordinate: the CTAs in flight spread their reductions over all levels' slices of the table
    // instead of all landing on the few thousand entries of one coarse level
    const int l = blockIdx.x % sc.n_levels;
    const int64_t p = (int64_t)(blockIdx.x / sc.n_levels) * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool live = p < n_points;
    const int64_t pc = live ? p : n_points - 1;
    const uint32_t size = sc.offset[l + 1] - sc.offset[l];
    uint32_t cell[3];
    float w[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) grid_pos(sc.scale[l], x[pc * 3 + d], cell[d], w[d]);
    float2 g = reinterpret_cast<const float2*>(dy)[pc * sc.n_levels + l];
    if (!live) g = make_float2(0.f, 0.f);
    bool cell_same = lane != 0;
    cell_same &= __shfl_up_sync(0xffffffffu, cell[0], 1) == cell[0];
    cell_same &= __shfl_up_sync(0xffffffffu, cell[1], 1) == cell[1];
    cell_same &= __shfl_up_sync(0xffffffffu, cell[2], 1) == cell[2];
    bool eq[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) eq[d] = __float_as_uint(__shfl_up_sync(0xffffffffu, w[d], 1)) == __float_as_uint(w[d]);
    // the lattice's fastest axis is the one whose weight changes along a run: x (cb200_smooth_points) or z (the reference's
    // meshgrid order, CoSLAM.smoothness through autograd); the warp takes whichever gives it more merged lanes
    const uint32_t run_x = __ballot_sync(0xffffffffu, cell_same && eq[1] && eq[2]);
    const uint32_t run_z = __ballot_sync(0xffffffffu, cell_same && eq[0] && eq[1]);
    const int axis = __popc(run_z) > __popc(run_x) ? 2 : 0;
    const uint32_t heads = ~(axis == 2 ? run_z : run_x);
    const bool same = !((heads >> lane) & 1u);
    const float wa = axis == 2 ? w[2] : w[0];
    float a0 = (1.f - wa) * g.x, a1 = (1.f - wa) * g.y, b0 = wa * g.x, b1 = wa * g.y;
    if (heads != 0xffffffffu) {
        const uint32_t above = lane == 31 ? 0u : heads >> (lane + 1);
        const int end = above ? lane + __ffs(above) : 32;             // first lane of the next run
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const float t0 = __shfl_down_sync(0xffffffffu, a0, s), t1 = __shfl_down_sync(0xffffffffu, a1, s);
            const float t2 = __shfl_down_sync(0xffffffffu, b0, s), t3 = __shfl_down_sync(0xffffffffu, b1, s);
            if (lane + s < end) { a0 += t0; a1 += t1; b0 += t2; b1 += t3; }
        }
    }
    if (same || !live) return;          // lanes past the end repeat the last point with g = 0: they extend its run or head a run of zeros
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const uint32_t idx = sc.offset[l] + grid_index(sc.hashed[l] != 0, size, sc.res[l], cell[0] + (c & 1), cell[1] + ((c >> 1) & 1), cell[2] + (c >> 2));
        float wo = 1.f;                 // the two weights that are constant along the run
#pragma unroll
        for (int d = 0; d < 3; ++d)
            if (d != axis) wo *= ((c >> d) & 1) ? w[d] : 1.f - w[d];
        const bool hi = (c >> axis) & 1;
        red_add_f2(dtable + idx, wo * (hi ? b0 : a0), wo * (hi ? b1 : a1));
    }
}

// OneBlob: thread per (point, dim)
__global__ void __launch_bounds__(256) oneblob_fwd_kernel(const float* __restrict__ x, int64_t n, int n_bins, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;      // flat (point, dim)
    if (i >= n) return;
    const float xv = x[i];
    const float nb = (float)n_bins;
    const float f_first = blob_edge_cdf(0.f, xv, nb);
    float f_prev = f_first;
    for (int b = 0; b < n_bins; ++b) {
        const float f_next = (b == n_bins - 1) ? f_first + 1.0f : blob_edge_cdf((float)(b + 1) / nb, xv, nb);
        out[i * n_bins + b] = f_next - f_prev;
        f_prev = f_next;
    }
}

__global__ void __launch_bounds__(256) oneblob_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, int64_t n, int n_bins,
                                                          float* __restrict__ dx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float xv = x[i];
    const float nb = (float)n_bins;
    const float k_first = blob_edge_pdf(0.f, xv, nb);
    float k_prev = k_first, acc = 0.f;
    for (int b = 0; b < n_bins; ++b) {
        const float k_next = (b == n_bins - 1) ? k_first : blob_edge_pdf((float)(b + 1) / nb, xv, nb);
        acc += dy[i * n_bins + b] * (k_prev - k_next);
        k_prev = k_next;
    }
    dx[i] = acc;
}

cudaError_t launch_hash_fwd(const SceneDev& sc, const float* x, const float* table, int64_t n, float* out, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    dim3 grid((unsigned)((n + 255) / 256), sc.n_levels);
    hash_fwd_kernel<<<grid, 256, 0, st>>>(sc, x, reinterpret_cast<const float2*>(table), n, out);
    return cudaGetLastError();
}
cudaError_t launch_hash_indices(const SceneDev& sc, const float* x, int64_t n, uint32_t* idx, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    dim3 grid((unsigned)((n + 255) / 256), sc.n_levels);
    hash_indices_kernel<<<grid, 256, 0, st>>>(sc, x, n, idx);
    return cudaGetLastError();
}
cudaError_t launch_hash_bwd(const SceneDev& sc, const float* x, const float* table, const float* dy, int64_t n, float* dtable, float* dx,
                            cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (dtable && !dx) {
        const int64_t blocks = (n + 255) / 256 * sc.n_levels;
        if (blocks > 0x7fffffffLL) return cudaErrorInvalidValue;
        hash_bwd_table_kernel<<<(unsigned)blocks, 256, 0, st>>>(sc, x, dy, n, reinterpret_cast<float2*>(dtable));
        return cudaGetLastError();
    }
    hash_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(sc, x, reinterpret_cast<const float2*>(table), dy, n,
                                                                reinterpret_cast<float2*>(dtable), dx);
    return cudaGetLastError();
}
cudaError_t launch_oneblob_fwd(const float* x, int64_t n_flat, int n_bins, float* out, cudaStream_t st) {
    if (n_flat == 0) return cudaSuccess;
    oneblob_fwd_kernel<<<(unsigned)((n_flat + 255) / 256), 256, 0, st>>>(x, n_flat, n_bins, out);
    return cudaGetLastError();
}
cudaError_t launch_oneblob_bwd(const float* x, const float* dy, int64_t n_flat, int n_bins, float* dx, cudaStream_t st) {
    if (n_flat == 0) return cudaSuccess;
    oneblob_bwd_kernel<<<(unsigned)((n_flat + 255) / 256), 256, 0, st>>>(x, dy, n_flat, n_bins, dx);
    return cudaGetLastError();
}

}  // namespace cb200
