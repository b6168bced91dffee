// Single-thread pose update of a tracking iteration (coslam.py:552-573) and the dual-number pose parametrisation it shares with the
// bundle-adjustment pose kernels (track.cu); also called from the one-launch tracking loop (track_frame.cu).
#pragma once
#include "args.cuh"
#include "render_ray.cuh"

namespace cb200 {

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
__device__ inline void rotation_dual(const float* rot, int n_rot, Dual R[9]) {
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

__device__ inline void write_c2w(const Dual R[9], const float* trans, float* c2w) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) c2w[i * 4 + j] = R[i * 3 + j].v;
        c2w[i * 4 + 3] = trans[i];
    }
    c2w[12] = 0.f; c2w[13] = 0.f; c2w[14] = 0.f; c2w[15] = 1.f;
}

__device__ __forceinline__ void adam_scalar(float& p, float g, float& m, float& v, float step_size, float inv_sqrt_bc2, bool write) {
    // same arithmetic as adam_step_kernel (optim.cu) with beta1 0.9, beta2 0.999, eps 1e-8, no weight decay
    m = fmaf(1.f - 0.9f, g - m, m);
    v = fmaf((1.f - 0.999f) * g, g, 0.999f * v);
    const float denom = fmaf(sqrtf(v), inv_sqrt_bc2, 1e-8f);
    if (write) p = p - step_size * (m / denom);
}

// Rotation parameters of an initial camera-to-world matrix, as the reference derives them on the host with pytorch3d
// (optimization/utils.py:23-28, coslam.py:276-283; restated in coslam_b200/pose.py): matrix_to_quaternion -- the candidate with the
// largest denominator, real part made non-negative -- and, for n_rot == 3, quaternion_to_axis_angle.  m = c2w, row-major 4x4.
__device__ inline void rotation_params_from_matrix(const float* m, int n_rot, float* rot) {
    const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[4], m11 = m[5], m12 = m[6], m20 = m[8], m21 = m[9], m22 = m[10];
    const float t[4] = {__fadd_rn(__fadd_rn(__fadd_rn(1.0f, m00), m11), m22), __fsub_rn(__fsub_rn(__fadd_rn(1.0f, m00), m11), m22),
                        __fsub_rn(__fadd_rn(__fsub_rn(1.0f, m00), m11), m22), __fadd_rn(__fsub_rn(__fsub_rn(1.0f, m00), m11), m22)};
    float q_abs[4];
    int best = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        q_abs[i] = t[i] > 0.f ? sqrtf(t[i]) : 0.f;
        if (q_abs[i] > q_abs[best]) best = i;                  // first maximum, like argmax
    }
    const float sq = __fmul_rn(q_abs[best], q_abs[best]);
    float c[4];
    if (best == 0) { c[0] = sq; c[1] = m21 - m12; c[2] = m02 - m20; c[3] = m10 - m01; }
    else if (best == 1) { c[0] = m21 - m12; c[1] = sq; c[2] = m10 + m01; c[3] = m02 + m20; }
    else if (best == 2) { c[0] = m02 - m20; c[1] = m10 + m01; c[2] = sq; c[3] = m12 + m21; }
    else { c[0] = m10 - m01; c[1] = m20 + m02; c[2] = m21 + m12; c[3] = sq; }
    const float den = __fmul_rn(2.0f, fmaxf(q_abs[best], 0.1f));
    float q[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = __fdiv_rn(c[i], den);
    if (q[0] < 0.f) {
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = -q[i];
    }
    if (n_rot == 4) {
#pragma unroll
        for (int i = 0; i < 4; ++i) rot[i] = q[i];
        return;
    }
    const float norm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(q[1], q[1]), __fmul_rn(q[2], q[2])), __fmul_rn(q[3], q[3])));
    const float half = atan2f(norm, q[0]);
    const float angle = __fmul_rn(2.0f, half);
    const float sc = fabsf(angle) < 1e-6f ? __fsub_rn(0.5f, __fdiv_rn(__fmul_rn(angle, angle), 48.0f)) : __fdiv_rn(sinf(half), angle);
#pragma unroll
    for (int i = 0; i < 3; ++i) rot[i] = __fdiv_rn(q[1 + i], sc);
}

// Fresh per-frame tracking state from an initial camera-to-world matrix (coslam.py:508-521): pose parameters, zero Adam moments,
// step and patience counters, alive flag, and the matrix of the parameters (what iteration 0 renders with).  c2w_out: 16 floats
// (registers / local); write_state: also store everything into the step block's buffers.
__device__ inline void track_init_state(const float* c2w_init, const cb200_track_step& s, float* c2w_out, bool write_state) {
    float m[16], rot[4] = {0.f, 0.f, 0.f, 0.f}, trans[3];
#pragma unroll
    for (int i = 0; i < 16; ++i) m[i] = __ldcg(c2w_init + i);
    rotation_params_from_matrix(m, s.n_rot, rot);
    trans[0] = m[3]; trans[1] = m[7]; trans[2] = m[11];
    Dual R[9];
    rotation_dual(rot, s.n_rot, R);
    write_c2w(R, trans, c2w_out);
    if (write_state) {
        for (int k = 0; k < s.n_rot; ++k) s.rot[k] = rot[k];
        for (int k = 0; k < 3; ++k) s.trans[k] = trans[k];
        for (int k = 0; k < 2 * s.n_rot + 6; ++k) s.adam[k] = 0.f;
        *s.step = 0;
        *s.patience = 0;
        *s.alive = 1;
        for (int i = 0; i < 16; ++i) s.c2w[i] = c2w_out[i];
    }
}

// Everything between two renders of one tracking iteration (see track.cu).  `first_iter` overrides s.first_iter.  Inputs are read
// with ld.global.cg: inside the one-launch loop they were written by other SMs earlier in the same kernel.
__device__ inline void track_update_step(const cb200_track_step& s, bool first_iter) {
    // every input first, into registers: the buffers may alias as far as the compiler knows, so loads interleaved with the
    // stores below would each wait for the previous store (a chain of ~20 dependent global round trips)
    float l0, l1, l2, l3;
    const bool own_losses = s.sums != nullptr;
    if (own_losses) {                              // losses from this iteration's sums and counts (cb200_loss_finish arithmetic)
        double sums[4];
        long long counts[3];
        for (int i = 0; i < 4; ++i) sums[i] = __ldcg(s.sums + i);
        for (int i = 0; i < 3; ++i) counts[i] = (long long)__ldcg(s.counts + i);
        float L[4], coef[4];
        loss_from_sums(sums, counts, s.n_rays_total, s.n_samples, L, coef);
        l0 = L[0]; l1 = L[1]; l2 = L[2]; l3 = L[3];
    } else {
        l0 = __ldcg(s.losses); l1 = __ldcg(s.losses + 1); l2 = __ldcg(s.losses + 2); l3 = __ldcg(s.losses + 3);
    }
    bool alive = __ldcg(s.alive) != 0;
    const float best_in = __ldcg(s.best_loss);
    int64_t patience = __ldcg(s.patience);
    const int64_t t = __ldcg(s.step) + 1;
    const int nr = s.n_rot;
    float c2w[16], g[12], rot[4], trans[3], ad[14];
#pragma unroll
    for (int i = 0; i < 16; ++i) c2w[i] = __ldcg(s.c2w + i);
#pragma unroll
    for (int i = 0; i < 12; ++i) g[i] = __ldcg(s.d_c2w + i);
#pragma unroll
    for (int i = 0; i < 4; ++i) rot[i] = i < nr ? __ldcg(s.rot + i) : 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) trans[i] = __ldcg(s.trans + i);
#pragma unroll
    for (int i = 0; i < 14; ++i) ad[i] = i < 2 * nr + 6 ? __ldcg(s.adam + i) : 0.f;

    const float lv = s.loss_w[0] * l0 + s.loss_w[1] * l1 + s.loss_w[2] * l2 + s.loss_w[3] * l3;
    const bool keep = alive;                       // the loop was still running when this iteration rendered
    // ---- best / last bookkeeping (coslam.py:552-566) ----------------------------------------------------------------------
    bool new_best;
    if (first_iter) {
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
    if (own_losses) {
        float* lo = const_cast<float*>(s.losses);
        lo[0] = l0; lo[1] = l1; lo[2] = l2; lo[3] = l3;
    }
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


}  // namespace cb200
