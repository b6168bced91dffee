// Per-ray kernels of the Co-SLAM hot path: depth-guided sampling, SDF -> weights volume rendering, the loss head
// and their backward.  One warp owns one ray; a ray's samples are strided over the lanes (sample s lives on lane
// s % 32), so all per-ray reductions are warp shuffles and all global accesses are coalesced along the sample axis.
//
// Reference behaviour restated here (file:line relative to the reference tree):
//   sampling            model/scene_rep.py:227-255
//   sdf2weights         model/scene_rep.py:77-97
//   raw2outputs         model/scene_rep.py:99-124
//   loss head           model/scene_rep.py:319-338, model/utils.py:89-161
#include "render_ray.cuh"

namespace cb200 {

// ------------------------------------------------------------------------------------------------------------------
// sampling
// ------------------------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kRayWarps * 32) sample_rays_kernel(SampleArgs a) {
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = a.n_uniform + a.n_guided;
    float* row = smem + warp * S;
    float* lin_u = smem + kRayWarps * S;            // CTA copies of the linspace constants
    float* lin_g = lin_u + a.n_uniform;
    float* lin_n = lin_g + a.n_guided;
    for (int i = threadIdx.x; i < a.n_uniform; i += blockDim.x) lin_u[i] = a.lin_uniform[i];
    for (int i = threadIdx.x; i < a.n_guided; i += blockDim.x) {
        lin_g[i] = a.lin_guided[i];
        lin_n[i] = a.lin_nodepth[i];
    }
    __syncthreads();

    unsigned n_front = 0, n_sdf = 0, n_valid = 0;
    const uint64_t seed = a.seed + (a.seed_dev ? *a.seed_dev : 0ull);
    for (int64_t ray = (int64_t)blockIdx.x * kRayWarps + warp; ray < a.n_rays; ray += (int64_t)gridDim.x * kRayWarps)
        sample_one_ray(a, ray, lane, row, lin_u, lin_g, lin_n, seed, n_front, n_sdf, n_valid);
    if (a.counts) {
        __shared__ unsigned red[3][kRayWarps];
        n_front = __reduce_add_sync(0xffffffffu, n_front);
        n_sdf = __reduce_add_sync(0xffffffffu, n_sdf);
        n_valid = __reduce_add_sync(0xffffffffu, n_valid);
        if (lane == 0) { red[0][warp] = n_front; red[1][warp] = n_sdf; red[2][warp] = n_valid; }
        __syncthreads();
        if (threadIdx.x < 3) {
            unsigned long long t = 0;
            for (int w = 0; w < kRayWarps; ++w) t += red[threadIdx.x][w];
            if (t) atomicAdd(a.counts + threadIdx.x, t);
        }
    }
}

// Tracking iteration: rays of the frame's sampled pixels from the current pose (coslam.py:546-547; assemble.cu's arithmetic for one
// pose) and their depth-guided samples in one launch -- the warp that samples a ray also rotates its direction.
struct TrackSampleArgs {
    SampleArgs s;                  // target_d = the [N] buffer written here
    const float* kf_rays;          // [N,7] camera-frame direction, rgb, depth
    const float* c2w;              // [16]
    float *rays_o, *rays_d, *target_rgb, *target_d;
};

__global__ void __launch_bounds__(kRayWarps * 32) track_rays_sample_kernel(TrackSampleArgs t) {
    extern __shared__ float smem[];
    const SampleArgs& a = t.s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = a.n_uniform + a.n_guided;
    float* row = smem + warp * S;
    float* lin_u = smem + kRayWarps * S;
    float* lin_g = lin_u + a.n_uniform;
    float* lin_n = lin_g + a.n_guided;
    for (int i = threadIdx.x; i < a.n_uniform; i += blockDim.x) lin_u[i] = a.lin_uniform[i];
    for (int i = threadIdx.x; i < a.n_guided; i += blockDim.x) {
        lin_g[i] = a.lin_guided[i];
        lin_n[i] = a.lin_nodepth[i];
    }
    float T[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) T[i] = t.c2w[i];
    __syncthreads();
    unsigned n_front = 0, n_sdf = 0, n_valid = 0;
    const uint64_t seed = a.seed + (a.seed_dev ? *a.seed_dev : 0ull);
    for (int64_t ray = (int64_t)blockIdx.x * kRayWarps + warp; ray < a.n_rays; ray += (int64_t)gridDim.x * kRayWarps) {
        const float* r = t.kf_rays + ray * 7;
        if (lane < 3) {
            const float d0 = r[0], d1 = r[1], d2 = r[2];
            // products rounded separately, summed left to right: what torch.sum over the last axis of 3 does
            t.rays_d[ray * 3 + lane] = __fadd_rn(__fadd_rn(__fmul_rn(d0, T[lane * 4 + 0]), __fmul_rn(d1, T[lane * 4 + 1])), __fmul_rn(d2, T[lane * 4 + 2]));
            t.rays_o[ray * 3 + lane] = T[lane * 4 + 3];
            t.target_rgb[ray * 3 + lane] = r[3 + lane];
        }
        if (lane == 0) t.target_d[ray] = r[6];
        __syncwarp();
        sample_one_ray(a, ray, lane, row, lin_u, lin_g, lin_n, seed, n_front, n_sdf, n_valid);
    }
    if (a.counts) {
        __shared__ unsigned red[3][kRayWarps];
        n_front = __reduce_add_sync(0xffffffffu, n_front);
        n_sdf = __reduce_add_sync(0xffffffffu, n_sdf);
        n_valid = __reduce_add_sync(0xffffffffu, n_valid);
        if (lane == 0) { red[0][warp] = n_front; red[1][warp] = n_sdf; red[2][warp] = n_valid; }
        __syncthreads();
        if (threadIdx.x < 3) {
            unsigned long long c = 0;
            for (int w = 0; w < kRayWarps; ++w) c += red[threadIdx.x][w];
            if (c) atomicAdd(a.counts + threadIdx.x, c);
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// rendering forward / backward (per-ray bodies: render_ray.cuh)
// ------------------------------------------------------------------------------------------------------------------

// kPerLane = ceil(n_samples / 32): sample slots per lane, so a 43-sample ray does 2 slots of work, not kMaxPerLane
template <int kPerLane>
__global__ void __launch_bounds__(kRayWarps * 32) render_fwd_kernel(RenderArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = a.n_samples;
    double part[4] = {0., 0., 0., 0.};     // lane 0 of each warp accumulates its rays' squared residuals
    for (int64_t ray = (int64_t)blockIdx.x * kRayWarps + warp; ray < a.n_rays; ray += (int64_t)gridDim.x * kRayWarps) {
        RayState<kPerLane> st;
        ray_forward<kPerLane>(a, ray, lane, st);
        float var = 0.f;
#pragma unroll
        for (int k = 0; k < kPerLane; ++k) {
            const float dz = st.z[k] - st.depth;
            var += st.wn[k] * dz * dz;
            if (a.weights && lane + 32 * k < S) a.weights[ray * S + lane + 32 * k] = st.wn[k];
        }
        var = warp_sum(var);
        float rgb0 = st.rgb[0], rgb1 = st.rgb[1], rgb2 = st.rgb[2];
        if (a.white_bkgd) { rgb0 += 1.f - st.acc; rgb1 += 1.f - st.acc; rgb2 += 1.f - st.acc; }
        if (lane == 0) {
            if (a.rgb) { a.rgb[ray * 3 + 0] = rgb0; a.rgb[ray * 3 + 1] = rgb1; a.rgb[ray * 3 + 2] = rgb2; }
            if (a.depth) a.depth[ray] = st.depth;
            if (a.depth_var) a.depth_var[ray] = var;
            if (a.acc) a.acc[ray] = st.acc;
            if (a.disp) a.disp[ray] = 1.0f / fmaxf(1e-10f, st.depth / st.acc);      // scene_rep.py:117
        }
        if (a.loss_sums) ray_loss_sums<kPerLane>(a, ray, lane, st, rgb0, rgb1, rgb2, part);
    }
    if (a.loss_sums) {
        __shared__ double red[4][kRayWarps];
        if (lane == 0) for (int i = 0; i < 4; ++i) red[i][warp] = part[i];
        __syncthreads();
        if (threadIdx.x < 4) {
            double t = 0.;
            for (int w = 0; w < kRayWarps; ++w) t += red[threadIdx.x][w];
            atomicAdd(a.loss_sums + threadIdx.x, t);
        }
    }
}

__global__ void loss_finish_kernel(const double* sums, const long long* counts, long long n_rays_total,
                                   long long n_samples, float* losses, float* coef) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    loss_from_sums(sums, counts, n_rays_total, n_samples, losses, coef);
}

template <int kPerLane>
__global__ void __launch_bounds__(kRayWarps * 32) render_bwd_kernel(RenderArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float gl[4] = {0.f, 0.f, 0.f, 0.f};
    if (a.has_loss) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (a.g_loss[i]) gl[i] = 2.0f * a.g_loss[i][0] * a.coef[i];        // d(sum of squares) = 2 e de
    }
    for (int64_t ray = (int64_t)blockIdx.x * kRayWarps + warp; ray < a.n_rays; ray += (int64_t)gridDim.x * kRayWarps)
        ray_backward<kPerLane>(a, ray, lane, gl);
}

// Loss head forward AND backward of a ray in one pass (tracking / bundle-adjustment loops): the backward coefficients need only the
// batch-global mask counts, which the sampling kernel has already produced, so a ray's d_raw can be written right after its
// rendering -- no loss_finish launch and no second ray_forward in between.  Also accumulates the four squared-residual sums (for the
// reported losses) and, when a list is given, appends the ray's samples with non-zero d_raw (the points the per-point backward must
// visit) -- what compact_live_kernel would find.  gl = 2 * loss weight * d loss / d sum.
struct RenderLossBwdArgs {
    RenderArgs r;
    const unsigned long long* counts;      // [3] batch-global n_front, n_sdf, n_valid
    long long n_rays_total;
    uint32_t* list;                        // may be NULL
    unsigned long long* list_count;
    float* coef_out;                       // [4] optional
};

template <int kPerLane>
__global__ void __launch_bounds__(kRayWarps * 32) render_loss_bwd_kernel(RenderLossBwdArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float coef[4], gl[4];
    {
        long long counts[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) counts[i] = (long long)a.counts[i];
        loss_coef_from_counts(counts, a.n_rays_total, a.r.n_samples, coef);
#pragma unroll
        for (int i = 0; i < 4; ++i) gl[i] = a.r.g_loss[i] ? 2.0f * a.r.g_loss[i][0] * coef[i] : 0.f;
        if (a.coef_out && blockIdx.x == 0 && threadIdx.x < 4) a.coef_out[threadIdx.x] = coef[threadIdx.x];
    }
    double part[4] = {0., 0., 0., 0.};
    for (int64_t ray = (int64_t)blockIdx.x * kRayWarps + warp; ray < a.r.n_rays; ray += (int64_t)gridDim.x * kRayWarps) {
        RayState<kPerLane> st;
        ray_forward<kPerLane>(a.r, ray, lane, st);
        float rgb0 = st.rgb[0], rgb1 = st.rgb[1], rgb2 = st.rgb[2];
        if (a.r.white_bkgd) { rgb0 += 1.f - st.acc; rgb1 += 1.f - st.acc; rgb2 += 1.f - st.acc; }
        if (a.r.loss_sums) ray_loss_sums<kPerLane>(a.r, ray, lane, st, rgb0, rgb1, rgb2, part);
        const unsigned live = ray_backward_from_state<kPerLane>(a.r, ray, lane, st, gl);
        if (a.list) ray_append_live<kPerLane>(live, ray, a.r.n_samples, lane, a.list, a.list_count);
    }
    if (a.r.loss_sums) {
        __shared__ double red[4][kRayWarps];
        if (lane == 0) for (int i = 0; i < 4; ++i) red[i][warp] = part[i];
        __syncthreads();
        if (threadIdx.x < 4) {
            double t = 0.;
            for (int w = 0; w < kRayWarps; ++w) t += red[threadIdx.x][w];
            atomicAdd(a.r.loss_sums + threadIdx.x, t);
        }
    }
}

static inline int ray_grid(int64_t n_rays) {
    int64_t blocks = (n_rays + kRayWarps - 1) / kRayWarps;
    const int64_t cap = 148 * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// ------------------------------------------------------------------------------------------------------------------
// Importance resampling (model/scene_rep.py:265-272 + sample_pdf, model/utils.py:32-71), one warp per ray:
//   bins = mid-points of the coarse samples, pdf from weights[1:-1] + 1e-5, inverse-CDF samples, then the sorted union of
//   coarse and new samples.  Also the mask counts of get_masks on the merged samples (the loss head uses the final z).
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kRayWarps * 32) importance_sample_kernel(ImportanceArgs a) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = a.n_samples, I = a.n_importance, nb = S - 1;      // nb bins, nb - 1 pdf entries, nb cdf entries
    float* zrow = smem + warp * (2 * S + I);
    float* cdf = zrow + S;                                           // [nb]
    float* samp = cdf + S;                                           // [I]  (one float of padding after cdf)
    unsigned n_front = 0, n_sdf = 0, n_valid = 0;
    if (a.new_count && blockIdx.x == 0 && threadIdx.x == 0) *a.new_count = (unsigned long long)a.n_rays * (unsigned long long)I;
    for (int64_t ray = (int64_t)blockIdx.x * kRayWarps + warp; ray < a.n_rays; ray += (int64_t)gridDim.x * kRayWarps) {
        for (int s = lane; s < S; s += 32) zrow[s] = a.z_vals[ray * S + s];
        float part = 0.f;
        for (int k = lane; k < nb - 1; k += 32) part += a.weights[ray * S + k + 1] + 1e-5f;      // utils.py:44
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        // utils.py:45-47: pdf in parallel (staged in the cdf row), then the cumulative sum in index order by one lane -- from
        // shared memory: the serial loop used to wait for a global load of weights[k + 1] per step
        for (int k = lane; k < nb - 1; k += 32) cdf[k + 1] = __fdiv_rn(a.weights[ray * S + k + 1] + 1e-5f, part);
        __syncwarp();
        if (lane == 0) {
            float c = 0.f;
            cdf[0] = 0.f;
#pragma unroll 4
            for (int k = 0; k < nb - 1; ++k) {
                c = __fadd_rn(c, cdf[k + 1]);
                cdf[k + 1] = c;
            }
        }
        __syncwarp();
        for (int i = lane; i < I; i += 32) {
            const float u = a.u[ray * a.u_stride + i];
            int lo = 0, hi = nb;                                     // searchsorted(cdf, u, right=True): #{cdf <= u}
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
            }
            const int below = max(lo - 1, 0), above = min(lo, nb - 1);
            const float cb = cdf[below], ca = cdf[above];
            const float bb = __fmul_rn(0.5f, __fadd_rn(zrow[below + 1], zrow[below]));
            const float ba = __fmul_rn(0.5f, __fadd_rn(zrow[above + 1], zrow[above]));
            float denom = __fsub_rn(ca, cb);
            if (denom < 1e-5f) denom = 1.f;
            const float t = __fdiv_rn(__fsub_rn(u, cb), denom);
            samp[i] = __fadd_rn(bb, __fmul_rn(t, __fsub_rn(ba, bb)));
        }
        __syncwarp();
        const float d = a.target_d ? a.target_d[ray] : 0.f;
        const float front_edge = __fsub_rn(d, a.trunc_sc), back_edge = __fadd_rn(d, a.trunc_sc);
        float mean = 0.f;
        // rank of every element in the sorted union (coarse samples first on ties)
        for (int e = lane; e < S + I; e += 32) {
            float v;
            int pos;
            if (e < S) {
                v = zrow[e];
                pos = e;
#pragma unroll 8
                for (int j = 0; j < I; ++j) pos += samp[j] < v;
                if (a.rank) a.rank[ray * S + e] = pos;
            } else {
                const int i = e - S;
                v = samp[i];
                mean += v;
                pos = 0;
#pragma unroll 8
                for (int j = 0; j < I; ++j) pos += (samp[j] < v) || (samp[j] == v && j < i);
                int lo = 0, hi = S;                                  // #{z <= v}
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (zrow[mid] <= v) lo = mid + 1; else hi = mid;
                }
                pos += lo;
            }
            a.z_out[ray * (S + I) + pos] = v;
            if (a.new_list && e >= S) a.new_list[ray * I + (e - S)] = (uint32_t)(ray * (S + I) + pos);
            if (a.target_d) {
                const bool front = v < front_edge, back = v > back_edge;
                n_front += front;
                n_sdf += (!front && !back && d > 0.0f);
            }
        }
        if (a.z_std) {                                               // scene_rep.py:292, unbiased=False
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mean += __shfl_xor_sync(0xffffffffu, mean, o);
            mean /= (float)I;
            float var = 0.f;
            for (int i = lane; i < I; i += 32) var += (samp[i] - mean) * (samp[i] - mean);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
            if (lane == 0) a.z_std[ray] = sqrtf(var / (float)I);
        }
        if (lane == 0 && a.target_d) n_valid += (d > 0.0f && d < a.depth_trunc);
        __syncwarp();
    }
    if (a.counts) {
        __shared__ unsigned red[3][kRayWarps];
        n_front = __reduce_add_sync(0xffffffffu, n_front);
        n_sdf = __reduce_add_sync(0xffffffffu, n_sdf);
        n_valid = __reduce_add_sync(0xffffffffu, n_valid);
        if (lane == 0) { red[0][warp] = n_front; red[1][warp] = n_sdf; red[2][warp] = n_valid; }
        __syncthreads();
        if (threadIdx.x < 3) {
            unsigned long long t = 0;
            for (int w = 0; w < kRayWarps; ++w) t += red[threadIdx.x][w];
            if (t) atomicAdd(a.counts + threadIdx.x, t);
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Lazy forward (training): which sample points does the step need at all?
//   * the losses read the SDF of the samples in front of the truncation band and inside it: z <= d + T  (get_masks,
//     model/utils.py:103-109) -- a PREFIX of the ascending samples of a ray;
//   * rendering weights are non-zero only below z[first sign change] + T (sdf2weights, scene_rep.py:89-95) -- a prefix too.
// Pass 1 evaluates the loss prefix (+ one sample); `lazy_resolve_kernel` then looks for the first sign change inside it: if
// there is one, only the samples below its cut-off are still missing, otherwise the rest of the ray is.  Everything not
// listed keeps raw = 0, which the rendering kernels mask out exactly as the reference's `mask` does, so the step's outputs
// and gradients are unchanged; the points behind the band are simply never encoded or decoded.
// ------------------------------------------------------------------------------------------------------------------
struct LazyArgs {
    const float* z_vals;       // [N,S]
    const float* target_d;     // [N]
    const float* raw;          // [N,S,4]   (resolve only)
    int64_t n_rays;
    int n_samples;
    float trunc_sc;
    int* k_done;               // [N] samples evaluated so far per ray (written by prefix, read + written by resolve)
    uint32_t* list;            // point indices to evaluate next
    unsigned long long* count;
};

// Appends the points [lo, hi) of each warp's ray to the list.  One reservation (atomicAdd) per CTA and loop trip instead of
// one per ray: every warp of the CTA calls this in every trip (n = 0 when it has nothing to add).
__device__ __forceinline__ void lazy_append(const LazyArgs& a, int64_t ray, int lo, int hi, int lane, int warp) {
    __shared__ int warp_n[kRayWarps];
    __shared__ unsigned long long cta_base;
    const int n = max(hi - lo, 0);
    if (lane == 0) warp_n[warp] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
#pragma unroll
        for (int w = 0; w < kRayWarps; ++w) tot += warp_n[w];
        cta_base = tot ? atomicAdd(a.count, (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    unsigned long long base = cta_base;
    for (int w = 0; w < warp; ++w) base += warp_n[w];
    for (int i = lane; i < n; i += 32) a.list[base + i] = (uint32_t)(ray * a.n_samples + lo + i);
    __syncthreads();                                  // warp_n / cta_base are rewritten by the next trip
}

__global__ void __launch_bounds__(kRayWarps * 32) lazy_prefix_kernel(LazyArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = a.n_samples;
    // CTA-uniform trip count: lazy_append synchronises the CTA
    for (int64_t ray0 = (int64_t)blockIdx.x * kRayWarps; ray0 < a.n_rays; ray0 += (int64_t)gridDim.x * kRayWarps) {
        const int64_t ray = ray0 + warp;
        int k = 0;
        if (ray < a.n_rays) {
            const float d = a.target_d[ray];
            k = S;
            if (d > 0.0f) {
                const float band_end = __fadd_rn(d, a.trunc_sc);
                int c = 0;
                for (int s = lane; s < S; s += 32) c += a.z_vals[ray * S + s] <= band_end;
                c = __reduce_add_sync(0xffffffffu, c);
                k = min(S, c + 1);
            }
            if (lane == 0) a.k_done[ray] = k;
        }
        lazy_append(a, ray, 0, k, lane, warp);
    }
}

__global__ void __launch_bounds__(kRayWarps * 32) lazy_resolve_kernel(LazyArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = a.n_samples;
    for (int64_t ray0 = (int64_t)blockIdx.x * kRayWarps; ray0 < a.n_rays; ray0 += (int64_t)gridDim.x * kRayWarps) {
        const int64_t ray = ray0 + warp;
        int k = S, need = S;
        if (ray < a.n_rays) k = a.k_done[ray];
        if (k < S) {                                             // warp-uniform
            const float4* raw4 = reinterpret_cast<const float4*>(a.raw) + ray * S;
            unsigned first = 0xffffffffu;
            for (int s = lane; s + 1 < k; s += 32) {
                if (__fmul_rn(raw4[s + 1].w, raw4[s].w) < 0.0f) first = min(first, (unsigned)s);      // scene_rep.py:89-91
            }
            first = __reduce_min_sync(0xffffffffu, first);
            if (first != 0xffffffffu) {
                const float z_lim = __fadd_rn(a.z_vals[ray * S + first], a.trunc_sc);                 // scene_rep.py:93-94
                int c = 0;
                for (int s = lane; s < S; s += 32) c += a.z_vals[ray * S + s] < z_lim;
                need = __reduce_add_sync(0xffffffffu, c);
            }
            if (need > k && lane == 0) a.k_done[ray] = need;
        }
        lazy_append(a, ray, k, need, lane, warp);                 // empty when nothing is missing (need <= k)
    }
}

cudaError_t launch_lazy_prefix(const float* z_vals, const float* target_d, int64_t n_rays, int n_samples, float trunc_sc, int* k_done,
                               uint32_t* list, unsigned long long* count, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(count, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    LazyArgs a{z_vals, target_d, nullptr, n_rays, n_samples, trunc_sc, k_done, list, count};
    lazy_prefix_kernel<<<ray_grid(n_rays), kRayWarps * 32, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_lazy_resolve(const float* z_vals, const float* target_d, const float* raw, int64_t n_rays, int n_samples, float trunc_sc,
                                int* k_done, uint32_t* list, unsigned long long* count, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(count, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    LazyArgs a{z_vals, target_d, raw, n_rays, n_samples, trunc_sc, k_done, list, count};
    lazy_resolve_kernel<<<ray_grid(n_rays), kRayWarps * 32, 0, st>>>(a);
    return cudaGetLastError();
}

// The coarse samples of the hierarchical pass are a subset of the merged samples (same positions, hence the same network outputs):
// copies (accumulate = 0) their raw / activation masks into the merged rows, or adds (accumulate = 1) the coarse pass's d_raw to the
// merged rows' d_raw so that ONE per-point backward serves both passes.  Thread per coarse sample.
__global__ void __launch_bounds__(256) merge_coarse_kernel(const int32_t* __restrict__ rank, int64_t n_rays, int S, int F, const float4* __restrict__ src,
                                                           float4* dst, const uint2* __restrict__ src_mask, uint2* dst_mask, int accumulate) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays * S) return;
    const int64_t ray = i / S;
    const int64_t j = ray * F + rank[i];
    if (accumulate) {
        const float4 a = src[i];
        float4 b = dst[j];
        b.x += a.x; b.y += a.y; b.z += a.z; b.w += a.w;
        dst[j] = b;
    } else {
        dst[j] = src[i];
        if (src_mask) dst_mask[j] = src_mask[i];
    }
}

cudaError_t launch_merge_coarse(const int32_t* rank, int64_t n_rays, int n_coarse, int n_fine, const float* src4, float* dst4, const uint32_t* src_mask,
                                uint32_t* dst_mask, int accumulate, cudaStream_t st) {
    const int64_t total = n_rays * n_coarse;
    if (total <= 0) return cudaSuccess;
    merge_coarse_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(rank, n_rays, n_coarse, n_fine, reinterpret_cast<const float4*>(src4),
                                                                         reinterpret_cast<float4*>(dst4), reinterpret_cast<const uint2*>(src_mask),
                                                                         reinterpret_cast<uint2*>(dst_mask), accumulate);
    return cudaGetLastError();
}

cudaError_t launch_importance_sample(const ImportanceArgs& a, cudaStream_t st) {
    const size_t smem = sizeof(float) * (size_t)kRayWarps * (size_t)(2 * a.n_samples + a.n_importance);
    importance_sample_kernel<<<ray_grid(a.n_rays), kRayWarps * 32, smem, st>>>(a);
    return cudaGetLastError();
}
cudaError_t launch_track_rays_sample(const SampleArgs& a, const float* kf_rays, const float* c2w, float* rays_o, float* rays_d, float* target_rgb,
                                     float* target_d, cudaStream_t st) {
    const int S = a.n_uniform + a.n_guided;
    const size_t smem = sizeof(float) * (size_t)(kRayWarps * S + a.n_uniform + 2 * a.n_guided);
    TrackSampleArgs t{a, kf_rays, c2w, rays_o, rays_d, target_rgb, target_d};
    t.s.target_d = target_d;
    track_rays_sample_kernel<<<ray_grid(a.n_rays), kRayWarps * 32, smem, st>>>(t);
    return cudaGetLastError();
}
cudaError_t launch_sample_rays(const SampleArgs& a, cudaStream_t st) {
    const int S = a.n_uniform + a.n_guided;
    const size_t smem = sizeof(float) * (size_t)(kRayWarps * S + a.n_uniform + 2 * a.n_guided);
    sample_rays_kernel<<<ray_grid(a.n_rays), kRayWarps * 32, smem, st>>>(a);
    return cudaGetLastError();
}
#define CB200_DISPATCH_PER_LANE(kernel)                                                                    \
    do {                                                                                                   \
        const int per_lane = (a.n_samples + 31) / 32;                                                      \
        const int grid = ray_grid(a.n_rays);                                                               \
        if (per_lane <= 1) kernel<1><<<grid, kRayWarps * 32, 0, st>>>(a);                                  \
        else if (per_lane == 2) kernel<2><<<grid, kRayWarps * 32, 0, st>>>(a);                             \
        else if (per_lane <= 4) kernel<4><<<grid, kRayWarps * 32, 0, st>>>(a);                             \
        else kernel<kMaxPerLane><<<grid, kRayWarps * 32, 0, st>>>(a);                                      \
    } while (0)

#define CB200_DISPATCH_PER_LANE_R(kernel, args, r)                                                         \
    do {                                                                                                   \
        const int per_lane = ((r).n_samples + 31) / 32;                                                    \
        const int grid = ray_grid((r).n_rays);                                                             \
        if (per_lane <= 1) kernel<1><<<grid, kRayWarps * 32, 0, st>>>(args);                               \
        else if (per_lane == 2) kernel<2><<<grid, kRayWarps * 32, 0, st>>>(args);                          \
        else if (per_lane <= 4) kernel<4><<<grid, kRayWarps * 32, 0, st>>>(args);                          \
        else kernel<kMaxPerLane><<<grid, kRayWarps * 32, 0, st>>>(args);                                   \
    } while (0)

cudaError_t launch_render_fwd(const RenderArgs& a, cudaStream_t st) {
    CB200_DISPATCH_PER_LANE(render_fwd_kernel);
    return cudaGetLastError();
}
cudaError_t launch_render_bwd(const RenderArgs& a, cudaStream_t st) {
    CB200_DISPATCH_PER_LANE(render_bwd_kernel);
    return cudaGetLastError();
}
cudaError_t launch_render_loss_bwd(const RenderArgs& r, const unsigned long long* counts, long long n_rays_total, uint32_t* list,
                                   unsigned long long* list_count, float* coef_out, int count_is_zero, cudaStream_t st) {
    if (list_count && !count_is_zero) {
        cudaError_t e = cudaMemsetAsync(list_count, 0, sizeof(unsigned long long), st);
        if (e != cudaSuccess) return e;
    }
    RenderLossBwdArgs a{r, counts, n_rays_total, list, list_count, coef_out};
    CB200_DISPATCH_PER_LANE_R(render_loss_bwd_kernel, a, r);
    return cudaGetLastError();
}
cudaError_t launch_loss_finish(const double* sums, const long long* counts, long long n_rays_total, long long n_samples,
                               float* losses, float* coef, cudaStream_t st) {
    loss_finish_kernel<<<1, 32, 0, st>>>(sums, counts, n_rays_total, n_samples, losses, coef);
    return cudaGetLastError();
}

}  // namespace cb200
