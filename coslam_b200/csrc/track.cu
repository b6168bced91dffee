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

namespace cb200 {

namespace {

struct Dual {
    float v;
    float d[4];
};

__device__ __forceinline__ Dual dconst(float v) {
    Dual r;
    r.v = v;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = 0.f;
    return r;
}
__device__ __forceinline__ Dual dvar(float v, int i) {
    Dual r = dconst(v);
    r.d[i] = 1.f;
    return r;
}
__device__ __forceinline__ Dual operator+(const Dual& a, const Dual& b) {
    Dual r;
    r.v = a.v + b.v;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] + b.d[i];
    return r;
}
__device__ __forceinline__ Dual operator-(const Dual& a, const Dual& b) {
    Dual r;
    r.v = a.v - b.v;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] - b.d[i];
    return r;
}
__device__ __forceinline__ Dual operator-(const Dual& a) {
    Dual r;
    r.v = -a.v;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = -a.d[i];
    return r;
}
__device__ __forceinline__ Dual operator*(const Dual& a, const Dual& b) {
    Dual r;
    r.v = a.v * b.v;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i];
    return r;
}
__device__ __forceinline__ Dual operator/(const Dual& a, const Dual& b) {
    Dual r;
    const float inv = 1.f / b.v;
    r.v = a.v * inv;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;
    return r;
}
__device__ __forceinline__ Dual dsqrt(const Dual& a) {
    Dual r;
    r.v = sqrtf(a.v);
    const float k = 0.5f / r.v;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * k;
    return r;
}
__device__ __forceinline__ Dual dsin(const Dual& a) {
    Dual r;
    r.v = sinf(a.v);
    const float c = cosf(a.v);
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * c;
    return r;
}
__device__ __forceinline__ Dual dcos(const Dual& a) {
    Dual r;
    r.v = cosf(a.v);
    const float s = -sinf(a.v);
#pragma unroll
    for (int i = 0; i < 4; ++i) r.d[i] = a.d[i] * s;
    return r;
}

// rotation parameters -> R (row-major 3x3 of duals).  n_rot == 3: axis-angle, Rodrigues with theta = |r| (singular at 0 like
// optimization/utils.py:5-21); n_rot == 4: real-first un-normalised quaternion (pytorch3d quaternion_to_matrix).
__device__ void rotation_dual(const float* rot, int n_rot, Dual R[9]) {
    if (n_rot == 3) {
        const Dual r0 = dvar(rot[0], 0), r1 = dvar(rot[1], 1), r2 = dvar(rot[2], 2);
        const Dual theta = dsqrt(r0 * r0 + r1 * r1 + r2 * r2);
        const Dual w0 = r0 / theta, w1 = r1 / theta, w2 = r2 / theta;
        const Dual zero = dconst(0.f);
        const Dual K[9] = {zero, -w2, w1, w2, zero, -w0, -w1, w0, zero};
        const Dual s = dsin(theta), c1 = dconst(1.f) - dcos(theta);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const Dual kk = K[i * 3 + 0] * K[0 * 3 + j] + K[i * 3 + 1] * K[1 * 3 + j] + K[i * 3 + 2] * K[2 * 3 + j];
                R[i * 3 + j] = dconst(i == j ? 1.f : 0.f) + s * K[i * 3 + j] + c1 * kk;
            }
        }
    } else {
        const Dual r = dvar(rot[0], 0), i = dvar(rot[1], 1), j = dvar(rot[2], 2), k = dvar(rot[3], 3);
        const Dual two_s = dconst(2.f) / (r * r + i * i + j * j + k * k);
        const Dual one = dconst(1.f);
        R[0] = one - two_s * (j * j + k * k);
        R[1] = two_s * (i * j - k * r);
        R[2] = two_s * (i * k + j * r);
        R[3] = two_s * (i * j + k * r);
        R[4] = one - two_s * (i * i + k * k);
        R[5] = two_s * (j * k - i * r);
        R[6] = two_s * (i * k - j * r);
        R[7] = two_s * (j * k + i * r);
        R[8] = one - two_s * (i * i + j * j);
    }
}

__device__ void write_c2w(const Dual R[9], const float* trans, float* c2w) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) c2w[i * 4 + j] = R[i * 3 + j].v;
        c2w[i * 4 + 3] = trans[i];
    }
    c2w[12] = 0.f; c2w[13] = 0.f; c2w[14] = 0.f; c2w[15] = 1.f;
}

__global__ void pose_matrices_kernel(const float* rot, const float* trans, long long n, int n_rot, float* c2w) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Dual R[9];
    rotation_dual(rot + i * n_rot, n_rot, R);
    write_c2w(R, trans + i * 3, c2w + i * 16);
}

__device__ __forceinline__ void adam_scalar(float& p, float g, float& m, float& v, float step_size, float inv_sqrt_bc2, bool write) {
    // same arithmetic as adam_step_kernel (optim.cu) with beta1 0.9, beta2 0.999, eps 1e-8, no weight decay
    m = fmaf(1.f - 0.9f, g - m, m);
    v = fmaf((1.f - 0.999f) * g, g, 0.999f * v);
    const float denom = fmaf(sqrtf(v), inv_sqrt_bc2, 1e-8f);
    if (write) p = p - step_size * (m / denom);
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
    // every input first, into registers: the buffers may alias as far as the compiler knows, so loads interleaved with the
    // stores below would each wait for the previous store (a chain of ~20 dependent global round trips)
    const float l0 = s.losses[0], l1 = s.losses[1], l2 = s.losses[2], l3 = s.losses[3];
    bool alive = *s.alive != 0;
    const float best_in = *s.best_loss;
    int64_t patience = *s.patience;
    const int64_t t = *s.step + 1;
    const int nr = s.n_rot;
    float c2w[16], g[12], rot[4], trans[3], ad[14];
#pragma unroll
    for (int i = 0; i < 16; ++i) c2w[i] = s.c2w[i];
#pragma unroll
    for (int i = 0; i < 12; ++i) g[i] = s.d_c2w[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) rot[i] = i < nr ? s.rot[i] : 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) trans[i] = s.trans[i];
#pragma unroll
    for (int i = 0; i < 14; ++i) ad[i] = i < 2 * nr + 6 ? s.adam[i] : 0.f;

    const float lv = s.loss_w[0] * l0 + s.loss_w[1] * l1 + s.loss_w[2] * l2 + s.loss_w[3] * l3;
    const bool keep = alive;                       // the loop was still running when this iteration rendered
    // ---- best / last bookkeeping (coslam.py:552-566) ----------------------------------------------------------------------
    bool new_best;
    if (s.first_iter) {
        new_best = true;
        patience = 1;
    } else {
        new_best = lv < best_in && alive;
        if (s.may_stop && alive) patience = new_best ? 0 : patience + 1;
    }
    if (s.may_stop) alive = alive && (patience <= (int64_t)s.wait_iters);      // coslam.py:565-566: break before the update
    // ---- dL/d(rot, trans) from dL/d c2w -----------------------------------------------------------------------------------
    Dual R[9];
    rotation_dual(rot, nr, R);
    float g_rot[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
            for (int k = 0; k < 4; ++k) g_rot[k] = fmaf(g[i * 4 + j], R[i * 3 + j].d[k], g_rot[k]);      // d[k] = 0 for k >= n_rot
    // ---- Adam (skipped for the parameters once the loop has stopped) --------------------------------------------------------
    double b1t = 1.0, b2t = 1.0;                 // beta^t by repeated multiplication (a tracking loop runs 10-20 steps)
    if (t <= 64) {
        for (int64_t i = 0; i < t; ++i) { b1t *= 0.9; b2t *= 0.999; }
    } else {
        b1t = pow(0.9, (double)t);
        b2t = pow(0.999, (double)t);
    }
    const double bc1 = 1.0 - b1t, bc2 = 1.0 - b2t;
    const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    const float ss_rot = (float)((double)s.lr_rot / bc1), ss_trans = (float)((double)s.lr_trans / bc1);
    float* m_rot = ad;
    float* v_rot = ad + nr;
    float* m_tr = ad + 2 * nr;
    float* v_tr = m_tr + 3;
    for (int k = 0; k < nr; ++k) adam_scalar(rot[k], g_rot[k], m_rot[k], v_rot[k], ss_rot, inv_sqrt_bc2, alive);
    for (int k = 0; k < 3; ++k) adam_scalar(trans[k], g[k * 4 + 3], m_tr[k], v_tr[k], ss_trans, inv_sqrt_bc2, alive);
    // ---- pose of the updated parameters -----------------------------------------------------------------------------------
    rotation_dual(rot, nr, R);
    // ---- all outputs ------------------------------------------------------------------------------------------------------
    if (new_best) {
        *s.best_loss = lv;
        for (int i = 0; i < 16; ++i) s.best_c2w[i] = c2w[i];
    }
    if (!s.may_stop || keep) {
        for (int i = 0; i < 16; ++i) s.last_c2w[i] = c2w[i];
        *s.last_loss = lv;
    }
    *s.patience = patience;
    if (s.may_stop) *s.alive = alive ? 1 : 0;
    if (s.psnr) *s.psnr = -10.f * logf(l0 + 1e-6f) / logf(10.f);      // mse2psnr, model/utils.py:8-12
    if (s.loss) *s.loss = lv;
    *s.step = t;
    for (int k = 0; k < nr; ++k) s.rot[k] = rot[k];
    for (int k = 0; k < 3; ++k) s.trans[k] = trans[k];
    for (int k = 0; k < 2 * nr + 6; ++k) s.adam[k] = ad[k];
    if (s.g_rot_out) for (int k = 0; k < nr; ++k) s.g_rot_out[k] = g_rot[k];
    write_c2w(R, trans, s.c2w);
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

cudaError_t launch_track_update(const cb200_track_step& s, cudaStream_t st) {
    track_update_kernel<<<1, 32, 0, st>>>(s);
    return cudaGetLastError();
}

}  // namespace cb200
