// Kernels and launchers of the per-point tensor-core path; the tile loops live in points_tc.cuh.
#include <cstdlib>

#include "points_tc.cuh"

namespace cb200 {

template <int kDense, bool kJac>
__global__ void __launch_bounds__(tc::kThreads, 2) points_fwd_tc_kernel(PointsArgs a) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc<kTmemCols>(&tmem_slot);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    stage_all_weights(smem_u32(smem), a);
    publish_tiles();
    const uint32_t tmem = tmem_slot;
    uint32_t parity = 0;
    const int64_t n_slots = a.live_count ? (int64_t)*a.live_count : a.n_points;
    fwd_tile_loop<kDense, kJac>(a, smem, bar, tmem, parity, a.live_idx, n_slots, blockIdx.x, gridDim.x);
    if (warp == 0) tmem_dealloc<tc::kTmemCols>(tmem);
}


template <bool kWgrad, int kKept>
__global__ void __launch_bounds__(tc::kThreads, 2) points_bwd_tc_kernel(PointsArgs a) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int half = tid >> 7;
    if (warp == 0) tmem_alloc<kBwdTmemCols>(&tmem_slot);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    stage_all_weights(smem_u32(smem), a);
    if (tid < 96) reinterpret_cast<float*>(smem + kBOffW4f)[tid] = a.w4[tid];
    publish_tiles();
    const uint32_t tmem = tmem_slot;
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t parity = 0;
    bool first_tile = true;
    const int64_t n_slots = a.live_count ? (int64_t)*a.live_count : a.n_points;
    bwd_tile_loop<kWgrad, kKept>(a, smem, bar, tmem, parity, first_tile, a.live_idx, n_slots, blockIdx.x, gridDim.x);

    if (kWgrad && !first_tile && half == 0) {
        // lane = row of the transposed weight gradients: dW1^T [80 x 32], dW3^T [63 x 32], dW2^T [32 x 16], dW4^T [32 x 3]
        tc_fence_after();
        uint32_t r[32];
        tmem_ld_x32(trow + kColDW1, r);
        tmem_ld_wait();
        if (tid < kSdfIn) {
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(a.dw1 + j * kSdfIn + tid, __uint_as_float(r[j]));
        }
        tmem_ld_x32(trow + kColDW3, r);
        tmem_ld_wait();
        if (tid < kColIn) {
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(a.dw3 + j * kColIn + tid, __uint_as_float(r[j]));
        }
        tmem_ld_x32(trow + kColDW2, r);      // columns 64..79 = dW2^T, 80..95 = dW4^T
        tmem_ld_wait();
        if (tid < kHidden) {
#pragma unroll
            for (int i = 0; i < kSdfOut; ++i) atomicAdd(a.dw2 + i * kHidden + tid, __uint_as_float(r[i]));
#pragma unroll
            for (int c = 0; c < 3; ++c) atomicAdd(a.dw4 + c * kHidden + tid, __uint_as_float(r[16 + c]));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<tc::kBwdTmemCols>(tmem);
}

// Order-preserving (within a 1024-point block) compaction of the points whose output gradient is non-zero.
__global__ void __launch_bounds__(1024) compact_live_kernel(const float4* __restrict__ d_raw, int64_t n_points, unsigned long long* count,
                                                            uint32_t* __restrict__ idx) {
    __shared__ uint32_t warp_off[32];
    __shared__ unsigned long long base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t p = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    bool live = false;
    if (p < n_points) {
        const float4 g = d_raw[p];
        live = !(g.x == 0.f && g.y == 0.f && g.z == 0.f && g.w == 0.f);      // NaN counts as live
    }
    const uint32_t b = __ballot_sync(0xffffffffu, live);
    if (lane == 0) warp_off[warp] = __popc(b);
    __syncthreads();
    if (warp == 0) {
        const uint32_t v = warp_off[lane];
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        warp_off[lane] = incl - v;
        if (lane == 31 && incl) base = atomicAdd(count, (unsigned long long)incl);
    }
    __syncthreads();
    if (live) idx[base + warp_off[warp] + __popc(b & ((1u << lane) - 1u))] = (uint32_t)p;
}

cudaError_t launch_compact_live(const float* d_raw, int64_t n_points, unsigned long long* count, uint32_t* idx, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(count, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    compact_live_kernel<<<(unsigned)((n_points + 1023) / 1024), 1024, 0, st>>>(reinterpret_cast<const float4*>(d_raw), n_points, count, idx);
    return cudaGetLastError();
}

// per-device host state looked up once: these launchers run ~20 times per bundle-adjustment iteration
static int sm_count() {
    static int sms_of[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (sms_of[dev] == 0) {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        sms_of[dev] = sms;
    }
    return sms_of[dev];
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device); kernels of one signature share a C++ type, so the
// flag is keyed by the function's address
static cudaError_t allow_dynamic_smem(const void* kernel, size_t smem) {
    constexpr int kMax = 32;
    static const void* seen[64][kMax];
    static int n_seen[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64) {
        for (int i = 0; i < n_seen[dev]; ++i)
            if (seen[dev][i] == kernel) return cudaSuccess;
    }
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess && dev >= 0 && dev < 64 && n_seen[dev] < kMax) seen[dev][n_seen[dev]++] = kernel;
    return e;
}

// for the other translation units of the tensor-core path (points_ws.cu)
int cb200_sm_count() { return sm_count(); }
cudaError_t cb200_allow_dynamic_smem(const void* kernel, size_t smem) { return allow_dynamic_smem(kernel, smem); }
cudaError_t launch_points_fwd_ws(const PointsArgs& a, int dense, cudaStream_t st);      // points_ws.cu

static int tc_persistent_grid(int64_t n_tiles, int ctas_per_sm) {
    const int sms = sm_count();
    int64_t g = (int64_t)sms * ctas_per_sm;
    if (g > n_tiles) g = n_tiles;
    if (g < 1) g = 1;
    return (int)g;
}

template <bool kWgrad, int kKept>
static cudaError_t launch_points_bwd_tc_as(const PointsArgs& a, cudaStream_t st) {
    const size_t smem = tc::kBwdSmem;
    const int64_t n_tiles = (a.n_points + tc::kTile - 1) / tc::kTile;
    cudaError_t e = allow_dynamic_smem(reinterpret_cast<const void*>(points_bwd_tc_kernel<kWgrad, kKept>), smem);
    if (e != cudaSuccess) return e;
    points_bwd_tc_kernel<kWgrad, kKept><<<tc_persistent_grid(n_tiles, 2), tc::kThreads, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_points_bwd_tc(const PointsArgs& a, cudaStream_t st) {
    if (a.dw1 == nullptr) return launch_points_bwd_tc_as<false, 0>(a, st);
    // kept rows replace the second gather: features alone when no position gradient is asked for, features + their derivatives else
    const bool want_pos = a.d_pts != nullptr || a.d_rays_o != nullptr;
    if (a.feat != nullptr && !want_pos) return launch_points_bwd_tc_as<true, 1>(a, st);
    if (a.feat != nullptr && a.feat_stride == 4 * kGridFeats) return launch_points_bwd_tc_as<true, 2>(a, st);
    return launch_points_bwd_tc_as<true, 0>(a, st);
}

// number of leading dense levels when the scene is "dense levels first, hashed levels after" over all 16 levels, else -1
static int dense_prefix(const SceneDev& sc) {
    if (sc.n_levels != (uint32_t)kLevels) return -1;
    int k = 0;
    while (k < kLevels && !sc.hashed[k]) ++k;
    for (int l = k; l < kLevels; ++l)
        if (!sc.hashed[l]) return -1;
    return k;
}

template <int kDense, bool kJac>
static cudaError_t launch_points_fwd_tc_as(const PointsArgs& a, cudaStream_t st) {
    const size_t smem = tc::kFwdSmem;
    cudaError_t e = allow_dynamic_smem(reinterpret_cast<const void*>(points_fwd_tc_kernel<kDense, kJac>), smem);
    if (e != cudaSuccess) return e;
    const int64_t n_tiles = (a.n_points + tc::kTile - 1) / tc::kTile;
    points_fwd_tc_kernel<kDense, kJac><<<tc_persistent_grid(n_tiles, 2), tc::kThreads, smem, st>>>(a);
    return cudaGetLastError();
}

template <bool kJac>
static cudaError_t launch_points_fwd_tc_jac(const PointsArgs& a, cudaStream_t st) {
    switch (dense_prefix(a.scene)) {          // 5: room0 / fr1_desk (and every Replica / TUM scene), 6: scene0000
        case 5: return launch_points_fwd_tc_as<5, kJac>(a, st);
        case 6: return launch_points_fwd_tc_as<6, kJac>(a, st);
        default: return launch_points_fwd_tc_as<-1, kJac>(a, st);
    }
}

// Which forward runs: < 0 the one-role kernel above, >= 0 the warp-specialised kernel (points_ws.cu) for batches of at least that many
// points.  Process-wide (cb200_set_forward_pipeline); the first query takes COSLAM_B200_FWD_WS / COSLAM_B200_FWD_WS_MIN from the environment.
static long long g_fwd_ws_min = -2;
long long fwd_ws_min_points() {
    if (g_fwd_ws_min == -2) {
        const char* on = getenv("COSLAM_B200_FWD_WS");
        const char* mn = getenv("COSLAM_B200_FWD_WS_MIN");
        g_fwd_ws_min = (on && atoi(on) > 0) ? (mn ? atoll(mn) : 0) : -1;
    }
    return g_fwd_ws_min;
}
void set_fwd_ws_min_points(long long v) { g_fwd_ws_min = v < 0 ? -1 : v; }

cudaError_t launch_points_fwd_tc(const PointsArgs& a, cudaStream_t st) {
    const long long ws_min = fwd_ws_min_points();
    if (ws_min >= 0 && a.n_points >= ws_min) return launch_points_fwd_ws(a, dense_prefix(a.scene), st);
    return (a.feat != nullptr && a.feat_stride == 4 * kGridFeats) ? launch_points_fwd_tc_jac<true>(a, st) : launch_points_fwd_tc_jac<false>(a, st);
}

}  // namespace cb200
