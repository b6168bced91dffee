// Warp-specialised per-point forward (tcgen05 decoder): the table gathers and the decoder's MMA rounds run in DIFFERENT warps.
//
// Why (profiles/r1/NOTES.md "forward with the four MMA rounds removed": 3.13 vs 4.11 ms; profiles/r2/NOTES.md, DESIGN.md section 7
// "next (1)"): in points_fwd_tc_kernel every warp first gathers (13-15 k cycles of a 22 k-cycle tile) and then sits through four
// MMA rounds with a CTA barrier each; the second resident CTA covers part of that, but the gathers -- bound by how many warps have
// requests in flight, not by any unit -- stop for ~30 % of every warp's time.  Here one 768-thread CTA per SM holds two independent
// halves, each a pipeline of
//     8 gather warps (lane pair per point, as in points_tc.cuh)  -->  4 epilogue warps (thread = point = TMEM lane)
// coupled by mbarriers only: the gather warps never wait for the tensor core (the hash-grid feature tile is double-buffered and
// released by the commit of layer 1, the OneBlob / geo tile by the commit of the colour net's first layer), the epilogue warps
// never gather.  Same arithmetic, same operand tiles (bf16 hi/lo split), same outputs as points_fwd_tc_kernel.
//
// Shared memory per half (80 KB):   G0, G1 [128 x 32] hash-grid features (2 x 16 KB) | BG [128 x 64] OneBlob 48 + geo 15 + 0 (32 KB)
//                                   | H [128 x 32] relu(h1), then relu(h3) (16 KB);          both halves + weights 24 KB = 184 KB
// TMEM per half (128 columns):      h1 0..31 | h2 32..47 | h3 48..79 | rgb 80..95
// mbarriers per half:               full[b] (256 gather arrivals: G[b] and BG written) | g_free[b] (commit of layer 1)
//                                   | bg_free (commit of the last MMA that reads BG) | mma (every round)
#include "points_tc.cuh"

namespace cb200 {

namespace ws {

constexpr int kGatherWarps = 16, kEpiWarps = 8;
constexpr int kThreads = 32 * (kGatherWarps + kEpiWarps);      // 768
constexpr uint32_t kGGS = 4 * 128, kGBytes = 16 * kGGS;        // [128 x 32]: 8192 per hi / lo
constexpr uint32_t kBGGS = 8 * 128, kBGBytes = 16 * kBGGS;     // [128 x 64]: 16384 per hi / lo
constexpr uint32_t kOffBGHi = 4 * kGBytes, kOffBGLo = kOffBGHi + kBGBytes;
constexpr uint32_t kOffHHi = kOffBGLo + kBGBytes, kOffHLo = kOffHHi + tc::kHBytes;
constexpr uint32_t kHalfBytes = kOffHLo + tc::kHBytes;         // 81920
constexpr uint32_t kOffW = 2 * kHalfBytes;
constexpr uint32_t kSmem = kOffW + (tc::kFwdSmem - tc::kOffW1Hi);      // 188416
constexpr int kTmemCols = 256;
enum { kBarFull = 0, kBarGFree = 2, kBarBGFree = 4, kBarMma = 5, kBars = 6 };

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// barrier among the 128 epilogue threads of one half (id 1 / 2; id 0 is __syncthreads)
__device__ __forceinline__ void epi_publish(int h) {
    fence_async_smem();
    tc_fence_before();
    asm volatile("bar.sync %0, 128;" ::"r"(1 + h) : "memory");
    tc_fence_after();
}

}  // namespace ws

template <int kDense, bool kJac>
__device__ __forceinline__ void ws_gather_role(const PointsArgs& a, uint32_t hbase, uint64_t* B, const uint32_t* list, int64_t n_slots,
                                               int64_t n_tiles, int64_t tile_begin, int64_t tile_step, int gw, int lane) {
    using namespace tc;
    using namespace ws;
    const int row_g = gw * 16 + (lane >> 1), xb = lane & 1;
    const uint32_t bg_hi = hbase + kOffBGHi, bg_lo = hbase + kOffBGLo;
    int k = 0;
#ifdef CB200_WS_CLOCKS
    long long t_wait = 0, t_grid = 0, t_blob = 0, t0 = clock64();
#define WSC(acc) do { const long long t1 = clock64(); acc += t1 - t0; t0 = t1; } while (0)
#else
#define WSC(acc)
#endif
    for (int64_t tile = tile_begin; tile < n_tiles; tile += tile_step, ++k) {
        const int b = k & 1;
        const uint32_t g_hi = hbase + (uint32_t)b * 2u * kGBytes, g_lo = g_hi + kGBytes;
        const int64_t pg_raw = tile * kTile + row_g;
        const int64_t pg_slot = pg_raw < n_slots ? pg_raw : n_slots - 1;      // padding slots repeat the last point
        const int64_t pg = list ? (int64_t)list[pg_slot] : pg_slot;
        float x[3], zdummy;
        point_position(a, pg, x, zdummy);
        float* const keep = (kJac || a.feat) ? a.feat + pg * a.feat_stride : nullptr;
        WSC(t_grid);
        if (k >= 2) mbar_wait(&B[kBarGFree + b], (uint32_t)(((k >> 1) + 1) & 1));      // layer 1 of tile k - 2 has read G[b]
        WSC(t_wait);
#ifdef CB200_PHASE_CLOCKS
        long long sub_unused[3] = {0, 0, 0};      // the one-role kernel's per-chunk clocks (tools/phase_probe.py); this kernel has its own
#pragma unroll
        for (int c = 0; c < 4; ++c) encode_grid_chunk<kDense, kJac, kGGS>(a, x, xb, c, g_hi, g_lo, row_g, keep, sub_unused);
#else
#pragma unroll
        for (int c = 0; c < 4; ++c) encode_grid_chunk<kDense, kJac, kGGS>(a, x, xb, c, g_hi, g_lo, row_g, keep);
#endif
        WSC(t_grid);
        if (k >= 1) mbar_wait(&B[kBarBGFree], (uint32_t)((k - 1) & 1));                  // tile k - 1 is done with BG
        WSC(t_wait);
#pragma unroll
        for (int d = 0; d < 3; ++d) {             // OneBlob: thread xb does bins 8 xb .. 8 xb + 7 of every coordinate
            float v[8], unused[8];
            blob_half<false>(x[d], xb, v, unused);
            tile_store8(bg_hi, bg_lo, kBGGS, row_g, 2 * d + xb, v);
        }
        fence_async_smem();
        mbar_arrive(&B[kBarFull + b]);
        WSC(t_blob);
    }
#ifdef CB200_WS_CLOCKS
    if ((blockIdx.x == 0 || blockIdx.x == 100) && lane == 0 && (gw == 0 || gw == 5) && k > 0)
        printf("ws gather cta %d gw %d tiles %d: grid %lld wait %lld blob+arrive %lld (cycles per tile)\n", blockIdx.x, gw, k, t_grid / k, t_wait / k, t_blob / k);
#endif
#undef WSC
}

__device__ __forceinline__ void ws_epilogue_role(const PointsArgs& a, uint32_t hbase, uint32_t wbase, uint64_t* B, uint32_t tcol,
                                                 const uint32_t* list, int64_t n_slots, int64_t n_tiles, int64_t tile_begin,
                                                 int64_t tile_step, int h, int q, int lane) {
    using namespace tc;
    using namespace ws;
    const int row = q * 32 + lane;
    const uint32_t bg_hi = hbase + kOffBGHi, bg_lo = hbase + kOffBGLo, h_hi = hbase + ws::kOffHHi, h_lo = hbase + ws::kOffHLo;
    // stage_all_weights was pointed at wbase - kOffW1Hi, so the forward layout's weight offsets apply relative to that
    const uint32_t w0 = wbase - tc::kOffW1Hi;
    const uint32_t trow = tcol + ((uint32_t)(q * 32) << 16);
    uint32_t par = 0;
#define WS_WAIT()                      \
    do {                               \
        mbar_wait(&B[kBarMma], par);   \
        par ^= 1u;                     \
        tc_fence_after();              \
    } while (0)
    int k = 0;
#ifdef CB200_WS_CLOCKS
    long long t_full = 0, t_work = 0, t0 = clock64();
#define WSC(acc) do { const long long t1 = clock64(); acc += t1 - t0; t0 = t1; } while (0)
#else
#define WSC(acc)
#endif
    for (int64_t tile = tile_begin; tile < n_tiles; tile += tile_step, ++k) {
        const int b = k & 1;
        const uint32_t g_hi = hbase + (uint32_t)b * 2u * kGBytes, g_lo = g_hi + kGBytes;
        const int64_t p_raw = tile * kTile + row;
        const bool valid = p_raw < n_slots;
        const int64_t p_slot = valid ? p_raw : n_slots - 1;
        const int64_t p = list ? (int64_t)list[p_slot] : p_slot;
        WSC(t_work);
        mbar_wait(&B[kBarFull + b], (uint32_t)((k >> 1) & 1));
        tc_fence_after();
        WSC(t_full);
        // ---- SDF net layer 1: h1 = [grid 32 | blob 48] W1^T ---------------------------------------------------------------------
        if (q == 0 && elect_one()) {
            umma_split3(tcol + kColH1, g_hi, g_lo, 128u, kGGS, 256u, w0 + kOffW1Hi, w0 + kOffW1Lo, 128u, kW1GS, 256u, 2,
                        umma_idesc_bf16(128, 32, false, false), false);
            umma_split3(tcol + kColH1, bg_hi, bg_lo, 128u, kBGGS, 256u, w0 + kOffW1Hi + 4 * 128, w0 + kOffW1Lo + 4 * 128, 128u, kW1GS, 256u, 3,
                        umma_idesc_bf16(128, 32, false, false), true);
            umma_commit(&B[kBarGFree + b]);
            if (a.sdf_only) umma_commit(&B[kBarBGFree]);
            umma_commit(&B[kBarMma]);
        }
        WS_WAIT();
        uint32_t m1;
        {
            uint32_t r0[16], r1[16];
            tmem_ld_x16(trow + kColH1, r0);
            tmem_ld_x16(trow + kColH1 + 16, r1);
            tmem_ld_wait();
            m1 = relu16_to_tile(r0, h_hi, h_lo, row, 0);
            m1 |= relu16_to_tile(r1, h_hi, h_lo, row, 1) << 16;
        }
        // ---- layer 2: h2 = relu(h1) W2^T -----------------------------------------------------------------------------------------
        epi_publish(h);
        if (q == 0 && elect_one()) {
            umma_split3(tcol + kColH2, h_hi, h_lo, 128u, kHGS, 256u, w0 + kOffW2Hi, w0 + kOffW2Lo, 128u, kW2GS, 256u, 2,
                        umma_idesc_bf16(128, 16, false, false), false);
            umma_commit(&B[kBarMma]);
        }
        WS_WAIT();
        float sdf;
        {
            uint32_t r[16];
            tmem_ld_x16(trow + kColH2, r);
            tmem_ld_wait();
            sdf = __uint_as_float(r[0]);
            if (a.geo && valid) {
#pragma unroll
                for (int g = 0; g < kGeo; ++g) a.geo[p * kGeo + g] = __uint_as_float(r[1 + g]);
            }
            if (!a.sdf_only) {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[1 + e]);
                tile_store8(bg_hi, bg_lo, kBGGS, row, 6, v);
#pragma unroll
                for (int e = 0; e < 7; ++e) v[e] = __uint_as_float(r[9 + e]);
                v[7] = 0.f;
                tile_store8(bg_hi, bg_lo, kBGGS, row, 7, v);
            }
        }
        float rgb[3] = {0.f, 0.f, 0.f};
        uint32_t m3 = 0;
        if (!a.sdf_only) {
            // ---- colour net layer 1: h3 = [blob | geo] W3^T ----------------------------------------------------------------------
            epi_publish(h);
            if (q == 0 && elect_one()) {
                umma_split3(tcol + kColH3, bg_hi, bg_lo, 128u, kBGGS, 256u, w0 + kOffW3Hi, w0 + kOffW3Lo, 128u, kW3GS, 256u, 4,
                            umma_idesc_bf16(128, 32, false, false), false);
                umma_commit(&B[kBarBGFree]);
                umma_commit(&B[kBarMma]);
            }
            WS_WAIT();
            {
                uint32_t r0[16], r1[16];
                tmem_ld_x16(trow + kColH3, r0);
                tmem_ld_x16(trow + kColH3 + 16, r1);
                tmem_ld_wait();
                m3 = relu16_to_tile(r0, h_hi, h_lo, row, 0);
                m3 |= relu16_to_tile(r1, h_hi, h_lo, row, 1) << 16;
            }
            // ---- colour net layer 2: rgb = relu(h3) W4^T -------------------------------------------------------------------------
            epi_publish(h);
            if (q == 0 && elect_one()) {
                umma_split3(tcol + kColRgb, h_hi, h_lo, 128u, kHGS, 256u, w0 + kOffW4Hi, w0 + kOffW4Lo, 128u, kW4GS, 256u, 2,
                            umma_idesc_bf16(128, 16, false, false), false);
                umma_commit(&B[kBarMma]);
            }
            WS_WAIT();
            uint32_t r[4];
            tmem_ld_x4(trow + kColRgb, r);
            tmem_ld_wait();
            rgb[0] = __uint_as_float(r[0]);
            rgb[1] = __uint_as_float(r[1]);
            rgb[2] = __uint_as_float(r[2]);
        }
        if (valid) {
            if (a.raw) reinterpret_cast<float4*>(a.raw)[p] = make_float4(rgb[0], rgb[1], rgb[2], sdf);
            if (a.act_mask) reinterpret_cast<uint2*>(a.act_mask)[p] = make_uint2(m1, m3);      // bit j = column j of h1 / h3
        }
        // no barrier at the end of a tile: every TMEM / H region is read in front of a barrier that precedes the MMA rewriting it
    }
#ifdef CB200_WS_CLOCKS
    WSC(t_work);
    if ((blockIdx.x == 0 || blockIdx.x == 100) && lane == 0 && k > 0)
        printf("ws epilogue cta %d half %d q %d tiles %d: wait for tile %lld work %lld (cycles per tile)\n", blockIdx.x, h, q, k, t_full / k, t_work / k);
#endif
#undef WSC
#undef WS_WAIT
}

template <int kDense, bool kJac>
__global__ void __launch_bounds__(ws::kThreads, 1) points_fwd_ws_kernel(PointsArgs a) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bars[2][ws::kBars];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == ws::kGatherWarps) tmem_alloc<ws::kTmemCols>(&tmem_slot);
    if (tid == 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            mbar_init(&bars[h][ws::kBarFull], 256);
            mbar_init(&bars[h][ws::kBarFull + 1], 256);
            mbar_init(&bars[h][ws::kBarGFree], 1);
            mbar_init(&bars[h][ws::kBarGFree + 1], 1);
            mbar_init(&bars[h][ws::kBarBGFree], 1);
            mbar_init(&bars[h][ws::kBarMma], 1);
        }
        mbar_fence_init();
    }
    const uint32_t sbase = smem_u32(smem);
    const uint32_t wbase = sbase + ws::kOffW;
    stage_all_weights(wbase - kOffW1Hi, a);
    publish_tiles();
    const uint32_t tmem = tmem_slot;
    const int64_t n_slots = a.live_count ? (int64_t)*a.live_count : a.n_points;
    const int64_t n_tiles = (n_slots + kTile - 1) / kTile;
    if (warp < ws::kGatherWarps) {
        const int h = warp >> 3;
        ws_gather_role<kDense, kJac>(a, sbase + (uint32_t)h * ws::kHalfBytes, bars[h], a.live_idx, n_slots, n_tiles,
                                     (int64_t)blockIdx.x * 2 + h, (int64_t)gridDim.x * 2, warp & 7, lane);
    } else {
        const int ew = warp - ws::kGatherWarps, h = ew >> 2;
        ws_epilogue_role(a, sbase + (uint32_t)h * ws::kHalfBytes, wbase, bars[h], tmem + (uint32_t)h * 128u, a.live_idx, n_slots, n_tiles,
                         (int64_t)blockIdx.x * 2 + h, (int64_t)gridDim.x * 2, h, ew & 3, lane);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == ws::kGatherWarps) tmem_dealloc<ws::kTmemCols>(tmem);
}

// defined in points_tc.cu
int cb200_sm_count();
cudaError_t cb200_allow_dynamic_smem(const void* kernel, size_t smem);

template <int kDense, bool kJac>
static cudaError_t launch_points_fwd_ws_as(const PointsArgs& a, cudaStream_t st) {
    cudaError_t e = cb200_allow_dynamic_smem(reinterpret_cast<const void*>(points_fwd_ws_kernel<kDense, kJac>), ws::kSmem);
    if (e != cudaSuccess) return e;
    const int64_t n_tiles = (a.n_points + tc::kTile - 1) / tc::kTile;
    int64_t grid = (n_tiles + 1) / 2;
    if (grid > cb200_sm_count()) grid = cb200_sm_count();
    if (grid < 1) grid = 1;
    points_fwd_ws_kernel<kDense, kJac><<<(unsigned)grid, ws::kThreads, ws::kSmem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_points_fwd_ws(const PointsArgs& a, int dense, cudaStream_t st) {
    const bool jac = a.feat != nullptr && a.feat_stride == 4 * kGridFeats;
    switch (dense) {
        case 5: return jac ? launch_points_fwd_ws_as<5, true>(a, st) : launch_points_fwd_ws_as<5, false>(a, st);
        case 6: return jac ? launch_points_fwd_ws_as<6, true>(a, st) : launch_points_fwd_ws_as<6, false>(a, st);
        default: return jac ? launch_points_fwd_ws_as<-1, true>(a, st) : launch_points_fwd_ws_as<-1, false>(a, st);
    }
}

}  // namespace cb200
