// One tracked frame (coslam.py:502-573 `tracking_render`) as ONE cooperative, persistent launch.
//
// The graph-captured loop (tracking.py) runs 11 kernels + 2 memsets per iteration; at 1024 rays their ~1.5 us boundaries are a
// third of the iteration and the per-point kernels pay a fresh prologue (TMEM allocation, 24 KB of weight staging) per launch.
// Here every CTA (2 per SM, all co-resident) stays alive for the whole frame; the phases of an iteration are separated by
// grid-wide barriers, and the decoder weights -- constant while tracking -- are staged into shared memory once per frame:
//
//   P0  rays from the current pose (optimization/utils.py get_camera_rays equivalent of coslam.py:546-547), depth-guided
//       sampling (scene_rep.py:227-255), mask counts                                                   warp per ray
//   P1  per-point forward: encodings + decoder (points_tc.cuh fwd_tile_loop)                           128-point tiles
//   P2  sdf2weights / raw2outputs / squared-residual sums of the four losses, and -- d loss / d sum needs the mask counts only,
//       not the sums -- the rendering + loss backward -> d_raw and the list of the points with non-zero d_raw (the ray's live
//       samples stay adjacent and ascending)                                                           warp per ray
//   P4  per-point backward to the ray gradients only (bwd_tile_loop<false>)                            tiles over the list
//   P5  dL/d c2w = segment sum of the ray gradients; the LAST CTA to finish (ticket) runs the single-thread pose update
//       (track_update.cuh: losses from the sums, weighted loss, best / last bookkeeping, patience, dual-number pose Jacobian, Adam, next c2w)
//       and clears the iteration's accumulators
//
// A ray is owned by the same warp in P0, P2, P3 and P5, so per-ray data (targets, z_vals) never needs a barrier of its own.
// Data written by other SMs is only read after a barrier whose acquire fence invalidates L1, and never through ld.global.nc
// (the hash table and the weights, which ARE read that way, do not change while tracking).
// The loop leaves early once the patience counter trips (coslam.py:565-566); the graph version keeps rendering and ignores the result.
#include "points_tc.cuh"
#include "render_ray.cuh"
#include "track_update.cuh"

namespace cb200 {

namespace {

__device__ __forceinline__ void grid_sync(unsigned* counter, unsigned& target, unsigned n_ctas) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += n_ctas;
        __threadfence();
        atomicAdd(counter, 1u);
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

template <int kDense, int kPerLane>
__global__ void __launch_bounds__(tc::kThreads, 2) track_frame_kernel(TrackFrameArgs t) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    __shared__ double red_d[4][kRayWarps];
    __shared__ unsigned red_u[3][kRayWarps];
    __shared__ float red_f[12][kRayWarps];
    __shared__ float sh_c2w[16];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned n_ctas = gridDim.x;
    const int64_t N = t.smp.n_rays;
    const int S = t.rnd.n_samples;
    const int64_t ray_first = (int64_t)warp * n_ctas + blockIdx.x, ray_step = (int64_t)kRayWarps * n_ctas;

    if (warp == 0) tmem_alloc<kBwdTmemCols>(&tmem_slot);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    stage_all_weights(smem_u32(smem), t.pts);
    if (tid < 96) reinterpret_cast<float*>(smem + kBOffW4f)[tid] = t.pts.w4[tid];
    publish_tiles();
    const uint32_t tmem = tmem_slot;
    uint32_t parity = 0;
    unsigned epoch = 0;
    bool first_tile = true;                       // unused without weight gradients
    const uint64_t seed0 = t.smp.seed + (t.seed_dev ? *t.seed_dev : 0ull);
    float* row = reinterpret_cast<float*>(smem) + warp * S;                  // scratch in the (idle) A tile region
    float* lin_u = reinterpret_cast<float*>(smem) + kRayWarps * S;
    float* lin_g = lin_u + t.smp.n_uniform;
    float* lin_n = lin_g + t.smp.n_guided;
#ifdef CB200_TRACK_CLOCKS      // phase times of CTA 0 (ns, summed over the iterations) into bytes 256.. of the workspace header: 6 phases, then the 6 barrier waits
    unsigned long long* clk = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(t.barrier) + 256);
    unsigned long long t_prev = 0;
    auto now = [] { unsigned long long v; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v)); return v; };
    if (blockIdx.x == 0 && tid == 0) t_prev = now();
#define TF_CLK(i) do { if (blockIdx.x == 0 && tid == 0) { const unsigned long long t_now = now(); clk[i] += t_now - t_prev; t_prev = t_now; } } while (0)
#else
#define TF_CLK(i)
#endif
    int k = 0;
    for (; k < t.n_iters; ++k) {
        // ---- P0 ---------------------------------------------------------------------------------------------------------
        for (int i = tid; i < t.smp.n_uniform; i += blockDim.x) lin_u[i] = t.smp.lin_uniform[i];
        for (int i = tid; i < t.smp.n_guided; i += blockDim.x) {
            lin_g[i] = t.smp.lin_guided[i];
            lin_n[i] = t.smp.lin_nodepth[i];
        }
        __syncthreads();
        {
            unsigned n_front = 0, n_sdf = 0, n_valid = 0;
            float T[16];
            if (k == 0 && t.c2w_init) {
                // every CTA derives the same parameters and matrix from the initial pose (identical code, identical bits);
                // CTA 0 also stores the fresh per-frame state, which is first read by the pose update behind five barriers
                if (tid == 0) track_init_state(t.c2w_init, t.step, sh_c2w, blockIdx.x == 0);
                __syncthreads();
#pragma unroll
                for (int i = 0; i < 12; ++i) T[i] = sh_c2w[i];
            } else {
#pragma unroll
                for (int i = 0; i < 12; ++i) T[i] = __ldcg(t.step.c2w + i);
            }
            for (int64_t ray = ray_first; ray < N; ray += ray_step) {
                const float* r = t.kf_rays + ray * 7;
                if (lane < 3) {
                    const float d0 = r[0], d1 = r[1], d2 = r[2];
                    // products rounded separately, summed left to right (assemble.cu, torch.sum over the last axis of 3)
                    const_cast<float*>(t.pts.rays_d)[ray * 3 + lane] =
                        __fadd_rn(__fadd_rn(__fmul_rn(d0, T[lane * 4 + 0]), __fmul_rn(d1, T[lane * 4 + 1])), __fmul_rn(d2, T[lane * 4 + 2]));
                    const_cast<float*>(t.pts.rays_o)[ray * 3 + lane] = T[lane * 4 + 3];
                    t.pts.d_rays_o[ray * 3 + lane] = 0.f;
                    t.pts.d_rays_d[ray * 3 + lane] = 0.f;
                    if (k == 0) const_cast<float*>(t.rnd.target_rgb)[ray * 3 + lane] = r[3 + lane];
                }
                if (k == 0) {
                    if (lane == 0) const_cast<float*>(t.rnd.target_d)[ray] = r[6];
                    __syncwarp();
                }
                sample_one_ray(t.smp, ray, lane, row, lin_u, lin_g, lin_n, seed0 + (uint64_t)k, n_front, n_sdf, n_valid);
            }
            n_front = __reduce_add_sync(0xffffffffu, n_front);
            n_sdf = __reduce_add_sync(0xffffffffu, n_sdf);
            n_valid = __reduce_add_sync(0xffffffffu, n_valid);
            if (lane == 0) { red_u[0][warp] = n_front; red_u[1][warp] = n_sdf; red_u[2][warp] = n_valid; }
            __syncthreads();
            if (tid < 3) {
                unsigned long long c = 0;
                for (int w = 0; w < kRayWarps; ++w) c += red_u[tid][w];
                if (c) atomicAdd(t.counts + tid, c);
            }
        }
        TF_CLK(0);
        grid_sync(t.barrier, epoch, n_ctas);
        TF_CLK(6);
        // ---- P1 ---------------------------------------------------------------------------------------------------------
        fwd_tile_loop<kDense>(t.pts, smem, bar, tmem, parity, nullptr, t.pts.n_points, blockIdx.x, n_ctas);
        TF_CLK(1);
        grid_sync(t.barrier, epoch, n_ctas);
        TF_CLK(7);
        // ---- P2 + P3: rendering, loss sums, d_raw and the live-point list of every ray (d loss / d sum needs the counts only) ------
        {
            float gl[4];
            {
                long long counts[3];
                float coef[4];
#pragma unroll
                for (int i = 0; i < 3; ++i) counts[i] = (long long)__ldcg(t.counts + i);
                loss_coef_from_counts(counts, N, S, coef);
#pragma unroll
                for (int i = 0; i < 4; ++i) gl[i] = 2.0f * t.step.loss_w[i] * coef[i];
            }
            double part[4] = {0., 0., 0., 0.};
            for (int64_t ray = ray_first; ray < N; ray += ray_step) {
                RayState<kPerLane> st;
                ray_forward<kPerLane>(t.rnd, ray, lane, st);
                float rgb0 = st.rgb[0], rgb1 = st.rgb[1], rgb2 = st.rgb[2];
                if (t.rnd.white_bkgd) { rgb0 += 1.f - st.acc; rgb1 += 1.f - st.acc; rgb2 += 1.f - st.acc; }
                ray_loss_sums<kPerLane>(t.rnd, ray, lane, st, rgb0, rgb1, rgb2, part);
                const unsigned live = ray_backward_from_state<kPerLane>(t.rnd, ray, lane, st, gl);
                ray_append_live<kPerLane>(live, ray, S, lane, t.live_list, t.live_count);
            }
            if (lane == 0) for (int i = 0; i < 4; ++i) red_d[i][warp] = part[i];
            __syncthreads();
            if (tid < 4) {
                double sum = 0.;
                for (int w = 0; w < kRayWarps; ++w) sum += red_d[tid][w];
                if (sum != 0.) atomicAdd(t.sums + tid, sum);
            }
        }
        TF_CLK(2);
        grid_sync(t.barrier, epoch, n_ctas);
        TF_CLK(8);
        // ---- P4 ---------------------------------------------------------------------------------------------------------
        {
            const int64_t n_live = (int64_t)__ldcg(t.live_count);
            bwd_tile_loop<false>(t.pts, smem, bar, tmem, parity, first_tile, t.live_list, n_live, blockIdx.x, n_ctas);
        }
        TF_CLK(4);
        grid_sync(t.barrier, epoch, n_ctas);
        TF_CLK(10);
        // ---- P5 ---------------------------------------------------------------------------------------------------------
        {
            float v = 0.f;                                       // lane l < 12: element (l >> 2, l & 3) of dL/d c2w[:3, :4]
            for (int64_t ray = ray_first; ray < N; ray += ray_step) {
                if (lane < 12) {
                    const int r3 = lane >> 2, c = lane & 3;
                    v += c < 3 ? t.pts.d_rays_d[ray * 3 + r3] * t.kf_rays[ray * 7 + c] : t.pts.d_rays_o[ray * 3 + r3];
                }
            }
            if (lane < 12) red_f[lane][warp] = v;
            __syncthreads();
            if (tid < 12) {
                float s = 0.f;
                for (int w = 0; w < kRayWarps; ++w) s += red_f[tid][w];
                if (s != 0.f) atomicAdd(t.d_c2w + tid, s);
            }
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                const unsigned done = atomicAdd(t.ticket, 1u);
                if (done == (unsigned)(k + 1) * n_ctas - 1u) {   // every CTA's sums have landed
                    __threadfence();
                    track_update_step(t.step, k == 0);
                    for (int i = 0; i < 3; ++i) t.counts[i] = 0ull;
                    for (int i = 0; i < 4; ++i) t.sums[i] = 0.;
                    for (int i = 0; i < 16; ++i) t.d_c2w[i] = 0.f;
                    *t.live_count = 0ull;
                }
            }
        }
        TF_CLK(5);
        grid_sync(t.barrier, epoch, n_ctas);
        TF_CLK(11);
        if (t.step.may_stop && __ldcg(t.step.alive) == 0) {      // coslam.py:565-566 (uniform: everyone reads the same flag)
            ++k;
            break;
        }
    }
    if (blockIdx.x == 0 && tid == 0 && t.seed_dev) *t.seed_dev += (unsigned long long)k;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<kBwdTmemCols>(tmem);
}

template <int kDense, int kPerLane>
cudaError_t launch_as(const TrackFrameArgs& t, cudaStream_t st) {
    const size_t smem = tc::kBwdSmem;
    // the co-resident grid of this kernel on each device: set up once (attribute + occupancy query), then only the launch
    static int grid_of_device[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    int grid = dev >= 0 && dev < 64 ? grid_of_device[dev] : 0;
    if (grid == 0) {
        cudaError_t e = cudaFuncSetAttribute(track_frame_kernel<kDense, kPerLane>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, track_frame_kernel<kDense, kPerLane>, tc::kThreads, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorCooperativeLaunchTooLarge;
        if (per_sm > 2) per_sm = 2;
        grid = sms * per_sm;
        if (dev >= 0 && dev < 64) grid_of_device[dev] = grid;
    }
    void* params[] = {const_cast<TrackFrameArgs*>(&t)};
    return cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(track_frame_kernel<kDense, kPerLane>), dim3((unsigned)grid),
                                       dim3(tc::kThreads), params, smem, st);
}

int dense_levels(const SceneDev& sc) {
    if (sc.n_levels != (uint32_t)kLevels) return -1;
    int k = 0;
    while (k < kLevels && !sc.hashed[k]) ++k;
    for (int l = k; l < kLevels; ++l)
        if (!sc.hashed[l]) return -1;
    return k;
}

}  // namespace

// returns cudaErrorNotSupported for a (dense levels, samples per ray) combination that is not instantiated
cudaError_t launch_track_frame(const TrackFrameArgs& t, cudaStream_t st) {
    const int dense = dense_levels(t.pts.scene), per_lane = (t.rnd.n_samples + 31) / 32;
    if (dense == 5 && per_lane == 2) return launch_as<5, 2>(t, st);      // room0 / every Replica scene: 32 + 11 samples
    if (dense == 6 && per_lane == 4) return launch_as<6, 4>(t, st);      // scene0000: 96 + 21
    if (dense == 5 && per_lane == 3) return launch_as<5, 3>(t, st);      // fr1_desk without importance samples: 64 + 21
    return cudaErrorNotSupported;
}

}  // namespace cb200
