// Dense Adam step (torch.optim.Adam semantics: L2 weight decay folded into the gradient, bias correction, eps added after
// the square root) for the map parameters -- coslam.py:610-634: decoder {weight_decay 1e-6}, hash table {eps 1e-15},
// betas (0.9, 0.99); also used for the pose parameters (coslam.py:276-283).  Dense on purpose: table entries that received
// no gradient this step still decay m, v and move, exactly like the reference's optimiser.
// One pass over p, g, m, v (28 B per parameter), optionally clearing the gradient for the next step.
#include "args.cuh"

namespace cb200 {

struct AdamArgs {
    float* p;
    float* g;
    float* m;
    float* v;
    int64_t n;
    float lr, beta1, beta2, eps, weight_decay;
    const int64_t* step_dev;      // step counter on the device (graph-captured loops), or NULL
    int64_t step;                 // 1-based step when step_dev is NULL
    int zero_grad;
};

__global__ void __launch_bounds__(256) adam_step_kernel(AdamArgs a) {
    const int64_t t = a.step_dev ? *a.step_dev : a.step;
    // bias corrections in double like the Python scalars of torch.optim.Adam, then rounded once
    const double bc1 = 1.0 - pow((double)a.beta1, (double)t);
    const double bc2 = 1.0 - pow((double)a.beta2, (double)t);
    const float step_size = (float)((double)a.lr / bc1);
    const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    const float one_m_b1 = 1.f - a.beta1, one_m_b2 = 1.f - a.beta2;
    const int64_t n4 = a.n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto update = [&](float& p, float& g, float& m, float& v) {
        const float gr = a.weight_decay != 0.f ? fmaf(a.weight_decay, p, g) : g;
        m = fmaf(one_m_b1, gr - m, m);                        // exp_avg.lerp_(grad, 1 - beta1)
        v = fmaf(one_m_b2 * gr, gr, a.beta2 * v);             // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        const float denom = fmaf(sqrtf(v), inv_sqrt_bc2, a.eps);
        p = p - step_size * (m / denom);
        if (a.zero_grad) g = 0.f;
    };
    const bool aligned = ((reinterpret_cast<uintptr_t>(a.p) | reinterpret_cast<uintptr_t>(a.g) | reinterpret_cast<uintptr_t>(a.m) |
                           reinterpret_cast<uintptr_t>(a.v)) & 15u) == 0;
    if (aligned) {
        for (int64_t i = tid; i < n4; i += stride) {
            float4 p = reinterpret_cast<float4*>(a.p)[i], g = reinterpret_cast<float4*>(a.g)[i];
            float4 m = reinterpret_cast<float4*>(a.m)[i], v = reinterpret_cast<float4*>(a.v)[i];
            update(p.x, g.x, m.x, v.x);
            update(p.y, g.y, m.y, v.y);
            update(p.z, g.z, m.z, v.z);
            update(p.w, g.w, m.w, v.w);
            reinterpret_cast<float4*>(a.p)[i] = p;
            reinterpret_cast<float4*>(a.m)[i] = m;
            reinterpret_cast<float4*>(a.v)[i] = v;
            if (a.zero_grad) reinterpret_cast<float4*>(a.g)[i] = g;
        }
        for (int64_t i = (n4 << 2) + tid; i < a.n; i += stride) update(a.p[i], a.g[i], a.m[i], a.v[i]);
    } else {
        for (int64_t i = tid; i < a.n; i += stride) update(a.p[i], a.g[i], a.m[i], a.v[i]);
    }
}

// The same update for up to 8 parameter tensors in one launch (map optimiser: hash table + four decoder matrices; the four
// matrices alone are launch-bound as separate kernels).  Blocks are assigned to tensors by a prefix table.
constexpr int kAdamMulti = 8;
struct AdamMultiArgs {
    cb200_adam_param p[kAdamMulti];
    int block_end[kAdamMulti];     // exclusive prefix of blocks per tensor
    int n;
    int64_t step;
    int zero_grad;
};

__global__ void __launch_bounds__(256) adam_multi_kernel(AdamMultiArgs a) {
    int t = 0;
    while (t < a.n - 1 && (int)blockIdx.x >= a.block_end[t]) ++t;
    const cb200_adam_param& q = a.p[t];
    const int first = t == 0 ? 0 : a.block_end[t - 1];
    const int64_t nblk = a.block_end[t] - first;
    const double bc1 = 1.0 - pow((double)q.beta1, (double)a.step);
    const double bc2 = 1.0 - pow((double)q.beta2, (double)a.step);
    const float step_size = (float)((double)q.lr / bc1);
    const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    const float one_m_b1 = 1.f - q.beta1, one_m_b2 = 1.f - q.beta2;
    for (int64_t i = ((int64_t)blockIdx.x - first) * blockDim.x + threadIdx.x; i < q.n; i += nblk * blockDim.x) {
        float p = q.param[i], g = q.grad[i], m = q.exp_avg[i], v = q.exp_avg_sq[i];
        const float gr = q.weight_decay != 0.f ? fmaf(q.weight_decay, p, g) : g;
        m = fmaf(one_m_b1, gr - m, m);
        v = fmaf(one_m_b2 * gr, gr, q.beta2 * v);
        const float denom = fmaf(sqrtf(v), inv_sqrt_bc2, q.eps);
        q.param[i] = p - step_size * (m / denom);
        q.exp_avg[i] = m;
        q.exp_avg_sq[i] = v;
        if (a.zero_grad) q.grad[i] = 0.f;
    }
}

cudaError_t launch_adam_multi(const cb200_adam_param* params, int n, int64_t step, int zero_grad, cudaStream_t st) {
    for (int base = 0; base < n; base += kAdamMulti) {
        AdamMultiArgs a{};
        a.n = n - base < kAdamMulti ? n - base : kAdamMulti;
        a.step = step;
        a.zero_grad = zero_grad;
        int blocks = 0;
        for (int t = 0; t < a.n; ++t) {
            a.p[t] = params[base + t];
            int64_t b = (a.p[t].n + 1023) / 1024;                 // 4 elements per thread
            if (b < 1) b = 1;
            if (b > 148 * 8) b = 148 * 8;
            blocks += (int)b;
            a.block_end[t] = blocks;
        }
        adam_multi_kernel<<<blocks, 256, 0, st>>>(a);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
                             float weight_decay, const int64_t* step_dev, int64_t step, int zero_grad, cudaStream_t st) {
    AdamArgs a{p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, step_dev, step, zero_grad};
    int64_t blocks = ((n >> 2) + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 8) blocks = 148 * 8;
    adam_step_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace cb200
