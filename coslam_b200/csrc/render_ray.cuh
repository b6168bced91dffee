// Per-ray device functions shared by the per-ray kernels (render.cu) and the one-launch tracking loop (track_frame.cu):
// one warp owns one ray, a ray's samples are strided over the lanes (sample s lives on lane s % 32).
//   sampling            model/scene_rep.py:227-255
//   sdf2weights         model/scene_rep.py:77-97
//   raw2outputs         model/scene_rep.py:99-124
//   loss head           model/scene_rep.py:319-338, model/utils.py:89-161
#pragma once
#include "args.cuh"

namespace cb200 {

constexpr int kRayWarps = 8;            // warps (= rays in flight) per CTA
constexpr int kMaxPerLane = 8;          // samples per lane -> n_samples <= 256

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------------------------
// sampling: one ray.  `row` = S floats of shared memory owned by the warp; lin_u / lin_g / lin_n = shared copies of the linspace
// constants.  Writes z_vals[ray, :] and adds this lane's share of the get_masks counts (model/utils.py:89-115).
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void sample_one_ray(const SampleArgs& a, int64_t ray, int lane, float* row, const float* lin_u,
                                               const float* lin_g, const float* lin_n, uint64_t seed, unsigned& n_front,
                                               unsigned& n_sdf, unsigned& n_valid) {
    const int S = a.n_uniform + a.n_guided;
    const float d = a.target_d[ray];
    const bool nodepth = d <= 0.0f;                      // scene_rep.py:239
    // merge of the two ascending lists == torch.sort of their concatenation (scene_rep.py:243):
    // a uniform sample goes after every guided sample strictly below it, a guided sample after every
    // uniform sample not above it.
    for (int i = lane; i < a.n_uniform; i += 32) {
        const float u = lin_u[i];
        int lo = 0, hi = a.n_guided;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            const float g = nodepth ? lin_n[mid] : __fadd_rn(lin_g[mid], d);
            if (g < u) lo = mid + 1; else hi = mid;
        }
        row[i + lo] = u;
    }
    for (int j = lane; j < a.n_guided; j += 32) {
        const float g = nodepth ? lin_n[j] : __fadd_rn(lin_g[j], d);
        int lo = 0, hi = a.n_uniform;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (lin_u[mid] <= g) lo = mid + 1; else hi = mid;
        }
        row[j + lo] = g;
    }
    __syncwarp();
    const float front_edge = __fsub_rn(d, a.trunc_sc);   // model/utils.py:103
    const float back_edge = __fadd_rn(d, a.trunc_sc);    // model/utils.py:105
    for (int s = lane; s < S; s += 32) {
        float z = row[s];
        if (a.perturb) {                                  // scene_rep.py:251-255
            const float zc = z;
            const float upper = (s + 1 < S) ? __fmul_rn(0.5f, __fadd_rn(row[s + 1], zc)) : zc;
            const float lower = (s > 0) ? __fmul_rn(0.5f, __fadd_rn(zc, row[s - 1])) : zc;
            float u;
            if (a.noise) {
                u = a.noise[ray * S + s];
            } else {
                const uint4 r = philox4x32(make_uint4((uint32_t)ray, (uint32_t)(ray >> 32), (uint32_t)(s >> 2), 0u),
                                           make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
                const uint32_t pick = (s & 3) == 0 ? r.x : (s & 3) == 1 ? r.y : (s & 3) == 2 ? r.z : r.w;
                u = u32_to_unit(pick);
            }
            z = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), u));
        }
        a.z_vals[ray * S + s] = z;
        const bool front = z < front_edge;
        const bool back = z > back_edge;
        n_front += front;
        n_sdf += (!front && !back && d > 0.0f);
    }
    if (lane == 0) n_valid += (d > 0.0f && d < a.depth_trunc);   // scene_rep.py:319
    __syncwarp();
}

// ------------------------------------------------------------------------------------------------------------------
// rendering forward / backward
// ------------------------------------------------------------------------------------------------------------------

// everything the forward and the backward share for one ray, held in registers
template <int kPerLane>
struct RayState {
    float4 raw[kPerLane];
    float z[kPerLane];
    float wn[kPerLane];       // normalised weight
    float bell[kPerLane];     // sigmoid(a) * sigmoid(-a)
    float sa[kPerLane];       // sigmoid(a)
    bool keep[kPerLane];      // truncation mask of sdf2weights
    float inv_den;               // 1 / (sum(bell * keep) + 1e-8)
    float rgb[3], depth, acc;
};

template <int kPerLane>
__device__ __forceinline__ void ray_forward(const RenderArgs& a, int64_t ray, int lane, RayState<kPerLane>& st) {
    const int S = a.n_samples;
    const float4* raw4 = reinterpret_cast<const float4*>(a.raw) + ray * S;
    const float* zrow = a.z_vals + ray * S;
    unsigned first = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < kPerLane; ++k) {
        const int s = lane + 32 * k;
        if (s < S) {
            st.raw[k] = raw4[s];
            st.z[k] = zrow[s];
            if (s + 1 < S) {
                const float nxt = raw4[s + 1].w;
                if (__fmul_rn(nxt, st.raw[k].w) < 0.0f) first = min(first, (unsigned)s);   // scene_rep.py:89-91
            }
        } else {
            st.raw[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            st.z[k] = 0.f;
        }
    }
    first = __reduce_min_sync(0xffffffffu, first);
    if (first == 0xffffffffu) first = 0;                       // argmax of an all-zero mask
    const float z_lim = __fadd_rn(zrow[first], a.trunc_sc);    // scene_rep.py:93-94
    float sum = 0.f;
    const float inv_trunc = 1.0f / a.trunc;
#pragma unroll
    for (int k = 0; k < kPerLane; ++k) {
        const int s = lane + 32 * k;
        const float arg = st.raw[k].w * inv_trunc;
        const float sp = sigmoid_fast(arg), sn = sigmoid_fast(-arg);
        st.sa[k] = sp;
        st.bell[k] = sp * sn;
        st.keep[k] = (s < S) && (st.z[k] < z_lim);
        sum += st.keep[k] ? st.bell[k] : 0.f;
    }
    sum = warp_sum(sum);
    st.inv_den = 1.0f / (sum + 1e-8f);
    float r = 0.f, g = 0.f, b = 0.f, dep = 0.f, acc = 0.f;
#pragma unroll
    for (int k = 0; k < kPerLane; ++k) {
        const float w = st.keep[k] ? st.bell[k] * st.inv_den : 0.f;
        st.wn[k] = w;
        r += w * sigmoid_fast(st.raw[k].x);
        g += w * sigmoid_fast(st.raw[k].y);
        b += w * sigmoid_fast(st.raw[k].z);
        dep += w * st.z[k];
        acc += w;
    }
    st.rgb[0] = warp_sum(r);
    st.rgb[1] = warp_sum(g);
    st.rgb[2] = warp_sum(b);
    st.depth = warp_sum(dep);
    st.acc = warp_sum(acc);
}

// the ray's terms of the four squared-residual sums of the loss head (scene_rep.py:319-338); lane 0 accumulates into part[]
template <int kPerLane>
__device__ __forceinline__ void ray_loss_sums(const RenderArgs& a, int64_t ray, int lane, const RayState<kPerLane>& st, float rgb0,
                                              float rgb1, float rgb2, double (&part)[4]) {
    const int S = a.n_samples;
    const float d = a.target_d[ray];
    const bool valid = d > 0.f && d < a.depth_trunc;
    const float rw = (valid || a.rgb_missing_nz) ? 1.f : 0.f;               // bool-typed rgb_weight, :320-321
    const float front_edge = __fsub_rn(d, a.trunc_sc), back_edge = __fadd_rn(d, a.trunc_sc);
    float fs = 0.f, sd = 0.f;
#pragma unroll
    for (int k = 0; k < kPerLane; ++k) {
        if (lane + 32 * k < S) {
            const float z = st.z[k], sdf = st.raw[k].w;
            const bool front = z < front_edge, back = z > back_edge;
            if (front) { const float e = sdf - 1.f; fs += e * e; }
            if (!front && !back && d > 0.f) { const float e = (z + sdf * a.trunc_sc) - d; sd += e * e; }
        }
    }
    fs = warp_sum(fs);
    sd = warp_sum(sd);
    if (lane == 0) {
        const float e0 = rgb0 * rw - a.target_rgb[ray * 3 + 0] * rw;
        const float e1 = rgb1 * rw - a.target_rgb[ray * 3 + 1] * rw;
        const float e2 = rgb2 * rw - a.target_rgb[ray * 3 + 2] * rw;
        part[0] += (double)(e0 * e0) + (double)(e1 * e1) + (double)(e2 * e2);
        if (valid) { const float e = st.depth - d; part[1] += (double)(e * e); }
        part[2] += (double)sd;
        part[3] += (double)fs;
    }
}

// d loss_k / d sum_k: needs the batch-global mask counts only (model/utils.py:111-115: int64 counts, true division in fp32), not the sums,
// so the backward of a ray can follow its forward without waiting for the other rays' residuals
__device__ __forceinline__ void loss_coef_from_counts(const long long* counts, long long n_rays_total, long long n_samples, float* coef,
                                                      float* sdf_w_out = nullptr, float* fs_w_out = nullptr) {
    const long long n_front = counts[0], n_sdf = counts[1], n_valid = counts[2];
    const float n_tot = (float)(n_front + n_sdf);
    const float fs_w = 1.0f - (float)n_front / n_tot;
    const float sdf_w = 1.0f - (float)n_sdf / n_tot;
    const double inv_rgb = 1.0 / (3.0 * (double)n_rays_total);
    const double inv_dep = 1.0 / (double)n_valid;                  // 0 valid rays -> NaN like mse_loss of empty
    const double inv_smp = 1.0 / ((double)n_rays_total * (double)n_samples);
    coef[0] = (float)inv_rgb;
    coef[1] = n_valid > 0 ? (float)inv_dep : 0.f;
    coef[2] = (float)inv_smp * sdf_w;
    coef[3] = (float)inv_smp * fs_w;
    if (sdf_w_out) *sdf_w_out = sdf_w;
    if (fs_w_out) *fs_w_out = fs_w;
}

// the four losses and d loss_k / d sum_k from the batch-global sums and mask counts (model/utils.py:111-161)
__device__ __forceinline__ void loss_from_sums(const double* sums, const long long* counts, long long n_rays_total, long long n_samples,
                                               float* losses, float* coef) {
    float sdf_w, fs_w;
    loss_coef_from_counts(counts, n_rays_total, n_samples, coef, &sdf_w, &fs_w);
    const long long n_valid = counts[2];
    const double inv_rgb = 1.0 / (3.0 * (double)n_rays_total);
    const double inv_dep = 1.0 / (double)n_valid;
    const double inv_smp = 1.0 / ((double)n_rays_total * (double)n_samples);
    const float mse_rgb = (float)(sums[0] * inv_rgb);
    const float mse_dep = n_valid > 0 ? (float)(sums[1] * inv_dep) : __int_as_float(0x7fc00000);
    const float mse_sdf = (float)(sums[2] * inv_smp);
    const float mse_fs = (float)(sums[3] * inv_smp);
    losses[0] = mse_rgb;
    losses[1] = mse_dep;
    losses[2] = mse_sdf * sdf_w;
    losses[3] = mse_fs * fs_w;
}

// backward of one ray: writes d_raw[ray, :, :]; gl[k] = 2 * upstream gradient of loss k * coef[k].  Returns the bit mask of this
// lane's sample slots whose d_raw is not exactly zero (bit k = sample lane + 32 k).
template <int kPerLane>
__device__ __forceinline__ unsigned ray_backward_from_state(const RenderArgs& a, int64_t ray, int lane, const RayState<kPerLane>& st,
                                                            const float (&gl)[4]) {
    const int S = a.n_samples;
    float d = 0.f, rw = 0.f;
    bool valid = false;
    float g_map[3] = {0.f, 0.f, 0.f}, g_dep = 0.f;
    if (a.has_loss) {
        d = a.target_d[ray];
        valid = d > 0.f && d < a.depth_trunc;
        rw = (valid || a.rgb_missing_nz) ? 1.f : 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float m = st.rgb[c] + (a.white_bkgd ? 1.f - st.acc : 0.f);
            g_map[c] = gl[0] * rw * (m * rw - a.target_rgb[ray * 3 + c] * rw);
        }
        if (valid) g_dep = gl[1] * (st.depth - d);
    }
    if (a.g_rgb) {
#pragma unroll
        for (int c = 0; c < 3; ++c) g_map[c] += a.g_rgb[ray * 3 + c];
    }
    if (a.g_depth) g_dep += a.g_depth[ray];
    const float g_acc = a.white_bkgd ? -(g_map[0] + g_map[1] + g_map[2]) : 0.f;

    // d wn_s and the normalisation term  sum_t d wn_t * wn_t
    float gwn[kPerLane];
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < kPerLane; ++k) {
        const float cr = sigmoid_fast(st.raw[k].x), cg = sigmoid_fast(st.raw[k].y), cb = sigmoid_fast(st.raw[k].z);
        gwn[k] = g_map[0] * cr + g_map[1] * cg + g_map[2] * cb + g_dep * st.z[k] + g_acc;
        dot += gwn[k] * st.wn[k];
    }
    dot = warp_sum(dot);
    const float front_edge = __fsub_rn(d, a.trunc_sc), back_edge = __fadd_rn(d, a.trunc_sc);
    float4* out = reinterpret_cast<float4*>(a.d_raw) + ray * S;
    const float inv_trunc = 1.0f / a.trunc;
    unsigned live = 0;
#pragma unroll
    for (int k = 0; k < kPerLane; ++k) {
        const int s = lane + 32 * k;
        if (s < S) {
            const float cr = sigmoid_fast(st.raw[k].x), cg = sigmoid_fast(st.raw[k].y), cb = sigmoid_fast(st.raw[k].z);
            float4 o;
            o.x = g_map[0] * st.wn[k] * cr * (1.f - cr);
            o.y = g_map[1] * st.wn[k] * cg * (1.f - cg);
            o.z = g_map[2] * st.wn[k] * cb * (1.f - cb);
            // weights = bell*keep / (sum + eps);  bell = s(a) s(-a), a = sdf / trunc
            float gs = 0.f;
            if (st.keep[k]) {
                const float g_bell = (gwn[k] - dot) * st.inv_den;
                gs = g_bell * st.bell[k] * (1.f - 2.f * st.sa[k]) * inv_trunc;
            }
            if (a.has_loss) {
                const float z = st.z[k], sdf = st.raw[k].w;
                const bool front = z < front_edge, back = z > back_edge;
                if (front) gs += gl[3] * (sdf - 1.f);
                if (!front && !back && d > 0.f) gs += gl[2] * ((z + sdf * a.trunc_sc) - d) * a.trunc_sc;
            }
            o.w = gs;
            out[s] = o;
            if (!(o.x == 0.f && o.y == 0.f && o.z == 0.f && o.w == 0.f)) live |= 1u << k;      // NaN counts as live
        }
    }
    return live;
}

template <int kPerLane>
__device__ __forceinline__ unsigned ray_backward(const RenderArgs& a, int64_t ray, int lane, const float (&gl)[4]) {
    RayState<kPerLane> st;
    ray_forward<kPerLane>(a, ray, lane, st);
    return ray_backward_from_state<kPerLane>(a, ray, lane, st, gl);
}

// appends the ray's samples with non-zero d_raw (bit j of `live` = sample lane + 32 j) to a point list, ascending and adjacent
template <int kPerLane>
__device__ __forceinline__ void ray_append_live(unsigned live, int64_t ray, int n_samples, int lane, uint32_t* list, unsigned long long* count) {
    unsigned total = 0, mine[kPerLane];
#pragma unroll
    for (int j = 0; j < kPerLane; ++j) {             // rank in ascending sample order s = lane + 32 j
        const unsigned b = __ballot_sync(0xffffffffu, (live >> j) & 1u);
        mine[j] = total + __popc(b & ((1u << lane) - 1u));
        total += __popc(b);
    }
    unsigned long long base = 0;
    if (lane == 0 && total) base = atomicAdd(count, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
    for (int j = 0; j < kPerLane; ++j)
        if ((live >> j) & 1u) list[base + mine[j]] = (uint32_t)(ray * n_samples + lane + 32 * j);
}

}  // namespace cb200
