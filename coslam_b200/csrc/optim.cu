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
