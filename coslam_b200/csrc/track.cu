// Pose / optimiser glue of the tracking and bundle-adjustment loops on the device.
//
// Tracking (coslam.py:502-573 `tracking_render`): everything between two renders of one
// tracking iteration that the reference does with ~100 tiny autograd / optimiser / bookkeeping kernels and two
// `.cpu().item()` syncs, as ONE single-thread kernel:
//   loss = sum_k w_k loss_k (get_loss_from_ret, coslam.py:140-160)
//   best / last pose and loss, patience counter, early-stop flag (coslam.py:552-566)
//   dL/d(rot, trans) from dL/d c2w through at_to_transform_matrix / qt_to_transform_matrix (optimization/utils.py:5-55)
//   Adam on the two pose parameters (torch.optim.Adam defaults, coslam.py:276-283), skipped once the loop has stopped
//   the camera-to-world matrix of the updated parameters, for the next iteration's rays.
// The rotation derivative is taken with forward-mode dual numbers (value + one partial per rotation parameter) through
// the same formulas as the reference's torch code, so there is no hand-derived Jacobian to get wrong.
#include "args.cuh"
#include "track_update.cuh"

namespace cb200 {

namespace {

__global__ void pose_matrices_kernel(const float* rot, const float* trans, long long n, int n_rot, float* c2w) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Dual R[9];
    rotation_dual(rot + i * n_rot, n_rot, R);
    write_c2w(R, trans + i * 3, c2w + i * 16);
}

// Bundle adjustment (coslam.py:285-397): one Adam step of torch.optim.Adam([rot, trans]) for n poses from the accumulated
// dL/d c2w, thread per pose.  Between two pose steps the reference back-propagates every iteration through the SAME
// pose matrices (`retain_graph=True`, coslam.py:361), so summing dL/d c2w over those iterations and applying the (fixed)
// Jacobian once is the same gradient.
struct PoseAdamArgs {
    float *rot, *trans, *m_rot, *v_rot, *m_trans, *v_trans;
    float* d_c2w;       // [n,4,4] accumulated gradient, cleared here when zero_grad
    float* c2w;         // [n,4,4] out: matrices of the updated parameters
    long long n;
    int n_rot, zero_grad;
    float ss_rot, ss_trans, inv_sqrt_bc2;
};

__global__ void pose_adam_kernel(PoseAdamArgs a) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    float* rot = a.rot + i * a.n_rot;
    float* trans = a.trans + i * 3;
    float* g = a.d_c2w + i * 16;
    Dual R[9];
    rotation_dual(rot, a.n_rot, R);
    float g_rot[4] = {0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c)
            for (int k = 0; k < a.n_rot; ++k) g_rot[k] = fmaf(g[r * 4 + c], R[r * 3 + c].d[k], g_rot[k]);
    for (int k = 0; k < a.n_rot; ++k)
        adam_scalar(rot[k], g_rot[k], a.m_rot[i * a.n_rot + k], a.v_rot[i * a.n_rot + k], a.ss_rot, a.inv_sqrt_bc2, true);
    for (int k = 0; k < 3; ++k) adam_scalar(trans[k], g[k * 4 + 3], a.m_trans[i * 3 + k], a.v_trans[i * 3 + k], a.ss_trans, a.inv_sqrt_bc2, true);
    if (a.zero_grad)
        for (int k = 0; k < 16; ++k) g[k] = 0.f;
    rotation_dual(rot, a.n_rot, R);
    write_c2w(R, trans, a.c2w + i * 16);
}

// Smoothness regulariser (coslam.py:251-274): TV = sum of squared forward differences of the lattice features along x, y, z.
// f [n,n,n,F]; d_f = scale * dTV/df (each difference enters the gradient of both of its end points); tv_sum += TV.
// The sum runs over all three axes, so which lattice axis is the fastest one does not matter here (smooth_points_kernel
// makes it x).  Thread per four features, grid-stride, one double atomic per CTA.
__global__ void __launch_bounds__(256) smooth_tv_kernel(const float* __restrict__ f, int n, int F, float scale, float* __restrict__ d_f,
                                                        double* tv_sum) {
    const int F4 = F >> 2;
    const long long total = (long long)n * n * n * F4;
    const float4* f4 = reinterpret_cast<const float4*>(f);
    const long long stride[3] = {(long long)n * n * F4, (long long)n * F4, (long long)F4};
    float local = 0.f;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long p = idx / F4;
        const int c[3] = {(int)(p / ((long long)n * n)), (int)((p / n) % n), (int)(p % n)};
        const float4 v = __ldg(f4 + idx);
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            if (c[ax] + 1 < n) {
                const float4 u = __ldg(f4 + idx + stride[ax]);
                const float dx = u.x - v.x, dy = u.y - v.y, dz = u.z - v.z, dw = u.w - v.w;
                local = fmaf(dx, dx, local); local = fmaf(dy, dy, local); local = fmaf(dz, dz, local); local = fmaf(dw, dw, local);
                g.x -= 2.f * dx; g.y -= 2.f * dy; g.z -= 2.f * dz; g.w -= 2.f * dw;
            }
            if (c[ax] > 0) {
                const float4 u = __ldg(f4 + idx - stride[ax]);
                g.x += 2.f * (v.x - u.x); g.y += 2.f * (v.y - u.y); g.z += 2.f * (v.z - u.z); g.w += 2.f * (v.w - u.w);
            }
        }
        reinterpret_cast<float4*>(d_f)[idx] = make_float4(scale * g.x, scale * g.y, scale * g.z, scale * g.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    __shared__ float part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0 && tv_sum) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += (double)part[w];
        atomicAdd(tv_sum, t);
    }
}

// lattice positions of the regulariser, normalised to the unit cube, with torch's operation order (separately rounded):
//   pts = (coords + jitter) * vox + lo + offset;   x = (pts - lo) / ext          (coslam.py:258-266)
// Point p = (iz * n + iy) * n + ix: x is the fastest lattice axis (the reference's meshgrid has z fastest; the term is a sum
// over the whole lattice, so the order is free).  The dense levels index x fastest, so neighbouring lanes gather and scatter
// neighbouring table entries: 164 us instead of 359 us for the table scatter of the 63^3 lattice (tools/smooth_probe.py).
__global__ void __launch_bounds__(256) smooth_points_kernel(int n, float vox, float3 jitter, float3 offset, float3 lo, float3 ext,
                                                            float* __restrict__ x) {
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= (long long)n * n * n) return;
    const float c[3] = {(float)(p % n), (float)((p / n) % n), (float)(p / ((long long)n * n))};
    const float j[3] = {jitter.x, jitter.y, jitter.z}, o[3] = {offset.x, offset.y, offset.z}, l[3] = {lo.x, lo.y, lo.z},
                e[3] = {ext.x, ext.y, ext.z};
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float pt = __fadd_rn(__fadd_rn(__fmul_rn(__fadd_rn(c[d], j[d]), vox), l[d]), o[d]);
        x[p * 3 + d] = __fdiv_rn(__fsub_rn(pt, l[d]), e[d]);
    }
}

__global__ void track_update_kernel(cb200_track_step s) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    track_update_step(s, s.first_iter != 0);
}

}  // namespace

cudaError_t launch_pose_matrices(const float* rot, const float* trans, long long n, int n_rot, float* c2w, cudaStream_t st) {
    pose_matrices_kernel<<<(unsigned)((n + 63) / 64), 64, 0, st>>>(rot, trans, n, n_rot, c2w);
    return cudaGetLastError();
}

cudaError_t launch_pose_adam(float* rot, float* trans, float* m_rot, float* v_rot, float* m_trans, float* v_trans, float* d_c2w, float* c2w,
                             long long n, int n_rot, float lr_rot, float lr_trans, long long step, int zero_grad, cudaStream_t st) {
    // bias corrections in double like the Python scalars of torch.optim.Adam (betas 0.9 / 0.999), rounded once
    const double bc1 = 1.0 - pow(0.9, (double)step), bc2 = 1.0 - pow(0.999, (double)step);
    PoseAdamArgs a{rot, trans, m_rot, v_rot, m_trans, v_trans, d_c2w, c2w, n, n_rot, zero_grad,
                   (float)((double)lr_rot / bc1), (float)((double)lr_trans / bc1), (float)(1.0 / sqrt(bc2))};
    pose_adam_kernel<<<(unsigned)((n + 63) / 64), 64, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_smooth_tv(const float* f, int n, int n_feat, float scale, float* d_f, double* tv_sum, cudaStream_t st) {
    if (n_feat & 3) return cudaErrorInvalidValue;
    const long long total = (long long)n * n * n * (n_feat >> 2);
    const long long blocks = (total + 255) / 256;
    smooth_tv_kernel<<<(unsigned)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, st>>>(f, n, n_feat, scale, d_f, tv_sum);
    return cudaGetLastError();
}

cudaError_t launch_smooth_points(int n, float vox, const float* jitter, const float* offset, const float* lo, const float* ext, float* x,
                                 cudaStream_t st) {
    const long long total = (long long)n * n * n;
    smooth_points_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(n, vox, make_float3(jitter[0], jitter[1], jitter[2]),
                                                                          make_float3(offset[0], offset[1], offset[2]),
                                                                          make_float3(lo[0], lo[1], lo[2]), make_float3(ext[0], ext[1], ext[2]), x);
    return cudaGetLastError();
}

// Tracking iteration: dL/d c2w of the ONE pose all rays share (assemble.cu's segment sum for a single slot) and, in the CTA that
// finishes last (ticket), the single-thread pose update -- two launches of the iteration in one.  s.d_c2w [16] and *ticket are
// zero on entry; the noise counter of the device Philox stream (may be NULL) advances by one.
__global__ void __launch_bounds__(256) track_pose_update_kernel(const float* kf_rays, const float* d_rays_o, const float* d_rays_d, long long n,
                                                                cb200_track_step s, unsigned* ticket, unsigned long long* noise_counter) {
    __shared__ float red[12][8];
    __shared__ bool last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float v[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) v[k] = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float d[3] = {kf_rays[i * 7 + 0], kf_rays[i * 7 + 1], kf_rays[i * 7 + 2]};
#pragma unroll
        for (int row = 0; row < 3; ++row) {
            const float gd = d_rays_d[i * 3 + row], go = d_rays_o[i * 3 + row];
            v[row * 4 + 0] = fmaf(gd, d[0], v[row * 4 + 0]);
            v[row * 4 + 1] = fmaf(gd, d[1], v[row * 4 + 1]);
            v[row * 4 + 2] = fmaf(gd, d[2], v[row * 4 + 2]);
            v[row * 4 + 3] += go;
        }
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if (lane == 0) red[k][warp] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += red[threadIdx.x][w];
        atomicAdd(const_cast<float*>(s.d_c2w) + threadIdx.x, t);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        if (last) {
            __threadfence();
            track_update_step(s, s.first_iter != 0);
            if (noise_counter) *noise_counter += 1ull;
        }
    }
}

cudaError_t launch_track_pose_update(const float* kf_rays, const float* d_rays_o, const float* d_rays_d, long long n, const cb200_track_step& s,
                                     unsigned* ticket, unsigned long long* noise_counter, cudaStream_t st) {
    long long blocks = (n + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 64) blocks = 64;
    track_pose_update_kernel<<<(unsigned)blocks, 256, 0, st>>>(kf_rays, d_rays_o, d_rays_d, n, s, ticket, noise_counter);
    return cudaGetLastError();
}

__global__ void pose_params_kernel(const float* c2w, long long n, int n_rot, float* rot, float* trans) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float m[16], r[4];
#pragma unroll
    for (int k = 0; k < 16; ++k) m[k] = c2w[i * 16 + k];
    rotation_params_from_matrix(m, n_rot, r);
    for (int k = 0; k < n_rot; ++k) rot[i * n_rot + k] = r[k];
    trans[i * 3 + 0] = m[3]; trans[i * 3 + 1] = m[7]; trans[i * 3 + 2] = m[11];
}

cudaError_t launch_pose_params(const float* c2w, long long n, int n_rot, float* rot, float* trans, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    pose_params_kernel<<<(unsigned)((n + 63) / 64), 64, 0, st>>>(c2w, n, n_rot, rot, trans);
    return cudaGetLastError();
}

__global__ void track_init_kernel(const float* c2w_init, cb200_track_step s) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float c2w[16];
    track_init_state(c2w_init, s, c2w, true);
}

cudaError_t launch_track_init(const float* c2w_init, const cb200_track_step& s, cudaStream_t st) {
    track_init_kernel<<<1, 32, 0, st>>>(c2w_init, s);
    return cudaGetLastError();
}

cudaError_t launch_track_update(const cb200_track_step& s, cudaStream_t st) {
    track_update_kernel<<<1, 32, 0, st>>>(s);
    return cudaGetLastError();
}

}  // namespace cb200
