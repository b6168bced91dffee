// Per-sample-point kernels, version 2 ("tc"): the decoder contractions run on the 5th-generation tensor cores.
//
// Why (profiles/r1/NOTES.md): in the SIMT version the L1TEX data pipe was the top unit and MORE than half of its
// wavefronts were shared-memory reads of decoder weights (fwd 55 %, bwd 64 %); only the rest were the hash-table
// gathers the path is really about.  Here one thread still owns one point for the gathers, but its 80 (+15) decoder
// inputs are written once into a shared-memory operand tile (bf16 hi/lo split, tc.cuh) and every layer is a
// tcgen05.mma with M = 128 points, accumulators in TMEM, read back with tcgen05.ld in the natural "lane = point"
// layout.  Weights are staged once per persistent CTA.
//
// Tile bookkeeping (bytes; "group stride" = distance between 8-row groups, see tc.cuh):
//   A   [128 x 96]  features: 32 grid | 48 OneBlob | 15 geo | 1 zero      group stride 1536   hi + lo = 48 KB
//   H   [128 x 32]  relu(h1) then relu(h3)                                 group stride  512   hi + lo = 16 KB
//   W1  [32 x 96] (cols 80.. zero)  W2 [16 x 32]  W3 [32 x 64] (col 63 zero)  W4 [16 x 32] (rows 3.. zero)   24 KB
// TMEM columns: h1 0..31 | h2 32..47 | h3 48..79 | rgb 80..95   (128 allocated)
#pragma once
#include <cstdio>
#include <type_traits>

#include "args.cuh"
#include "point_common.cuh"
#include "tc.cuh"

namespace cb200 {

namespace tc {

constexpr int kTile = 128;
constexpr uint32_t kAGS = 12 * 128, kABytes = 16 * kAGS;      // 24576
constexpr uint32_t kHGS = 4 * 128, kHBytes = 16 * kHGS;       // 8192
constexpr uint32_t kW1GS = 12 * 128, kW1Bytes = 4 * kW1GS;    // 6144
constexpr uint32_t kW2GS = 4 * 128, kW2Bytes = 2 * kW2GS;     // 1024
constexpr uint32_t kW3GS = 8 * 128, kW3Bytes = 4 * kW3GS;     // 4096
constexpr uint32_t kW4GS = 4 * 128, kW4Bytes = 2 * kW4GS;     // 1024

constexpr uint32_t kOffAHi = 0, kOffALo = kOffAHi + kABytes;
constexpr uint32_t kOffHHi = kOffALo + kABytes, kOffHLo = kOffHHi + kHBytes;
constexpr uint32_t kOffW1Hi = kOffHLo + kHBytes, kOffW1Lo = kOffW1Hi + kW1Bytes;
constexpr uint32_t kOffW2Hi = kOffW1Lo + kW1Bytes, kOffW2Lo = kOffW2Hi + kW2Bytes;
constexpr uint32_t kOffW3Hi = kOffW2Lo + kW2Bytes, kOffW3Lo = kOffW3Hi + kW3Bytes;
constexpr uint32_t kOffW4Hi = kOffW3Lo + kW3Bytes, kOffW4Lo = kOffW4Hi + kW4Bytes;
constexpr uint32_t kFwdSmem = kOffW4Lo + kW4Bytes;            // 90112

constexpr uint32_t kColH1 = 0, kColH2 = 32, kColH3 = 48, kColRgb = 80;
constexpr int kTmemCols = 128;

// stage one weight matrix [rows_valid x cols_valid] (row-major fp32, global) as a zero-padded [rows x 8*chunks] tile
__device__ __forceinline__ void stage_weight(uint32_t hi, uint32_t lo, uint32_t gs, const float* __restrict__ w, int rows, int chunks,
                                             int rows_valid, int cols_valid) {
    for (int i = threadIdx.x; i < rows * chunks; i += blockDim.x) {
        const int r = i / chunks, c = i % chunks;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int col = c * 8 + e;
            v[e] = (r < rows_valid && col < cols_valid) ? w[r * cols_valid + col] : 0.f;
        }
        tile_store8(hi, lo, gs, r, c, v);
    }
}

__device__ __forceinline__ void stage_all_weights(uint32_t sbase, const PointsArgs& a) {
    stage_weight(sbase + kOffW1Hi, sbase + kOffW1Lo, kW1GS, a.w1, 32, 12, 32, kSdfIn);
    stage_weight(sbase + kOffW2Hi, sbase + kOffW2Lo, kW2GS, a.w2, 16, 4, 16, kHidden);
    stage_weight(sbase + kOffW3Hi, sbase + kOffW3Lo, kW3GS, a.w3, 32, 8, 32, kColIn);
    stage_weight(sbase + kOffW4Hi, sbase + kOffW4Lo, kW4GS, a.w4, 16, 4, 3, kHidden);
}

// Thread roles (256 threads per 128-point tile, warp w = (quadrant q = w & 3, hsel = w >> 2), lane l):
//   * epilogue role: row_e = 32 q + l -- the TMEM lane this thread reads with tcgen05.ld (warps q and q + 4 address the same 32
//     lanes) -- and 16 of each 32-column accumulator row, selected by half = hsel;
//   * gather role: row_g = 32 q + 16 hsel + (l >> 1), xb = l & 1: a LANE PAIR shares one point and thread xb owns the four
//     corners at x = cell.x + xb of every level.  The two x-neighbours of a corner are adjacent table entries on the dense
//     levels and, on the hashed levels, share a 128-byte line 15 times out of 16 (the hash leaves x un-multiplied), so one
//     warp-wide LDG / RED touches ~16 lines instead of 32.  profiles/r1/NOTES.md: the L1TEX data pipe replays a scattered
//     request once per line and was the saturated unit (~2 cycles per wavefront) -- halving the lines per request is what
//     this layout is for.  Pair partial sums are combined with shfl.xor 1; TMEM values (lane = row_e) reach the pair that owns
//     the row with one shfl from lane 16 hsel + (l >> 1) of the same warp.
constexpr int kThreads = 256;

struct HalfCorners {
    uint32_t idx[4];      // (y, z) = (j & 1, j >> 1) corners at this thread's x
    float wx, w1, w2;
};

// kDense >= 0: levels below kDense are dense-indexed and the rest hashed (known when the kernel is compiled: no per-level branch,
// one index formula per unrolled level); kDense < 0: read the flag of every level from the scene (any geometry).
template <int kDense>
__device__ __forceinline__ void level_corners_x(const SceneDev& sc, int l, const float x[3], int xb, HalfCorners& hc) {
    const float scale = sc.scale[l];
    const uint32_t res = sc.res[l];
    const uint32_t size = sc.offset[l + 1] - sc.offset[l];
    const bool hashed = kDense >= 0 ? (l >= kDense) : (sc.hashed[l] != 0);
    uint32_t cell[3];
    float w0;
    grid_pos(scale, x[0], cell[0], w0);
    grid_pos(scale, x[1], cell[1], hc.w1);
    grid_pos(scale, x[2], cell[2], hc.w2);
    hc.wx = xb ? w0 : 1.f - w0;
#pragma unroll
    for (int j = 0; j < 4; ++j) hc.idx[j] = sc.offset[l] + grid_index(hashed, size, res, cell[0] + xb, cell[1] + (j & 1), cell[2] + (j >> 1));
#ifdef CB200_EXP_COALESCED      // experiment: same arithmetic, but the warp's addresses fall into one or two lines
#pragma unroll
    for (int j = 0; j < 4; ++j) hc.idx[j] = sc.offset[l] + ((hc.idx[j] & 1u) | ((threadIdx.x & 31u) << 1)) + 64u * j;
#endif
}

__device__ __forceinline__ float yz_weight(float w1, float w2, int j) { return ((j & 1) ? w1 : 1.f - w1) * ((j & 2) ? w2 : 1.f - w2); }

// 32 bytes = this thread's four (y, z) corners of a cell block of the cell-major copy (cells.cu): one 256-bit load
__device__ __forceinline__ void ldg_corners4(const float4* p, float2 (&v)[4]) {
    asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(v[0].x), "=f"(v[0].y), "=f"(v[1].x), "=f"(v[1].y), "=f"(v[2].x), "=f"(v[2].y), "=f"(v[3].x), "=f"(v[3].y)
                 : "l"(p));
}
__device__ __forceinline__ void red_add_f4(float4* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// index (in float4) of thread xb's half of the block of cell (cx, cy, cz) of dense level l; false: the cell lies outside the level
__device__ __forceinline__ bool cell_block(const SceneDev& sc, int l, const uint32_t (&cell)[3], int xb, uint32_t& at) {
    const uint32_t res = sc.res[l];
    at = (sc.cell_offset[l] + cell[0] + res * (cell[1] + res * cell[2])) * 4u + 2u * (uint32_t)xb;
    return cell[0] < res && cell[1] < res && cell[2] < res;
}

// the four corner values at this thread's x of level l, and the interpolation weights: dense levels come from the cell-major copy
// when the scene carries one (identical values), everything else from the table
template <int kDense>
__device__ __forceinline__ void gather_level_x(const PointsArgs& a, int l, const float x[3], int xb, float2 (&v)[4], float& wx, float& w1,
                                               float& w2) {
    const bool dense = kDense >= 0 ? (l < kDense) : (a.scene.hashed[l] == 0);
    if (dense && a.scene.cells != nullptr) {
        const float scale = a.scene.scale[l];
        uint32_t cell[3], at;
        float w0;
        grid_pos(scale, x[0], cell[0], w0);
        grid_pos(scale, x[1], cell[1], w1);
        grid_pos(scale, x[2], cell[2], w2);
        wx = xb ? w0 : 1.f - w0;
        if (cell_block(a.scene, l, cell, xb, at)) {
            ldg_corners4(a.scene.cells + at, v);
        } else {                                  // outside the unit cube: the table's wrapped indices
            const uint32_t res = a.scene.res[l], size = a.scene.offset[l + 1] - a.scene.offset[l];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                v[j] = ldg_f2(a.table + a.scene.offset[l] + grid_index(false, size, res, cell[0] + xb, cell[1] + (j & 1), cell[2] + (j >> 1)));
        }
        return;
    }
    HalfCorners hc;
    level_corners_x<kDense>(a.scene, l, x, xb, hc);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = ldg_f2(a.table + hc.idx[j]);
    wx = hc.wx; w1 = hc.w1; w2 = hc.w2;
}

// 4 consecutive columns (sub = 0 / 1: first / second half of a chunk) of row r into the hi and lo tiles
__device__ __forceinline__ void tile_store4(uint32_t hi_base, uint32_t lo_base, uint32_t group_stride, int r, int chunk, int sub, const float (&v)[4]) {
    uint32_t h[2], l[2];
    split_bf16x2(v[0], v[1], h[0], l[0]);
    split_bf16x2(v[2], v[3], h[1], l[1]);
    const uint32_t off = (uint32_t)(r >> 3) * group_stride + (uint32_t)chunk * 128u + (uint32_t)(r & 7) * 16u + (uint32_t)sub * 8u;
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(hi_base + off), "r"(h[0]), "r"(h[1]) : "memory");
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(lo_base + off), "r"(l[0]), "r"(l[1]) : "memory");
}

// the pair's per-corner partial features of one chunk (4 levels x 2) -> complete features; thread xb keeps columns 4 xb .. 4 xb + 3
// keep (optional): where this thread's four fp32 features go in the point's row of PointsArgs::feat (streamed out for the backward)
template <uint32_t kGS = kAGS>
__device__ __forceinline__ void pair_combine_store(const float (&part)[8], int xb, uint32_t a_hi, uint32_t a_lo, int row, int c,
                                                   float* keep = nullptr) {
    float mine[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float send = xb ? part[i] : part[4 + i];
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        mine[i] = (xb ? part[4 + i] : part[i]) + recv;
    }
    tile_store4(a_hi, a_lo, kGS, row, c, xb, mine);
    if (keep) __stcs(reinterpret_cast<float4*>(keep), make_float4(mine[0], mine[1], mine[2], mine[3]));
}

// where thread xb's four columns of chunk c start in a kept row of `stride` floats (32: features only; 128: see encode_grid_chunk)
__device__ __forceinline__ int kept_offset(int stride, int c, int xb) { return stride == kGridFeats ? 8 * c + 4 * xb : 32 * c + 4 * xb; }

// the same exchange for one derivative of the chunk's features: thread xb streams columns 4 xb .. 4 xb + 3 to dst[8 c + 4 xb ..]
__device__ __forceinline__ void pair_combine_keep(const float (&part)[8], int xb, float* dst) {
    float mine[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float send = xb ? part[i] : part[4 + i];
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        mine[i] = (xb ? part[4 + i] : part[i]) + recv;
    }
    __stcs(reinterpret_cast<float4*>(dst), make_float4(mine[0], mine[1], mine[2], mine[3]));
}

// one chunk (4 levels, 8 features) of hash-grid features -> A tile; 16 gathers in flight per thread.
// kJac: the kept row holds the features' derivatives along the normalised position as well, one 128-byte line per chunk --
// [chunk c][feat | d/dx | d/dy | d/dz][thread xb][columns 8c + 4xb .. + 3] -- so that every store (and every load of the backward) of a
// lane pair is one whole 32-byte sector; they come from the corner values that are in registers anyway.
// kGS: group stride of the tile the features go to (the warp-specialised forward keeps them in a [128 x 32] tile of its own)
// between(): independent work issued while the chunk's 16 gathers are in flight (the forward puts one coordinate's OneBlob there)
template <int kDense, bool kJac, uint32_t kGS, typename Between>
__device__ __forceinline__ void encode_grid_chunk_with(const PointsArgs& a, const float x[3], int xb, int c, uint32_t a_hi, uint32_t a_lo, int row,
                                                       float* keep, Between&& between
#ifdef CB200_PHASE_CLOCKS
                                                       , long long* sub
#endif
) {
#ifdef CB200_PHASE_CLOCKS
    long long t0 = clock64();
#endif
    float2 v[4][4];
    float wx[4], w1[4], w2[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) gather_level_x<kDense>(a, 4 * c + q, x, xb, v[q], wx[q], w1[q], w2[q]);
    between();
#ifdef CB200_PHASE_CLOCKS
    long long t1 = clock64();
    {
        float dummy = v[0][0].x + v[1][1].x + v[2][2].x + v[3][3].y;      // wait for (most of) the loads
        asm volatile("" ::"f"(dummy));
    }
    long long t2 = clock64();
#endif
    float part[8];
    float jx[8], jy[8], jz[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float f0 = 0.f, f1 = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float wc = wx[q] * yz_weight(w1[q], w2[q], j);
            f0 = fmaf(wc, v[q][j].x, f0);
            f1 = fmaf(wc, v[q][j].y, f1);
        }
        part[2 * q] = f0;
        part[2 * q + 1] = f1;
        if (kJac) {
            // d/dx: (value at x+1) - (value at x), bilinear in (y, z): this thread contributes its side with its sign;
            // d/dy, d/dz: this thread's x weight times the differences along y / z
            const float scale = a.scene.scale[4 * c + q];
            const float sx = xb ? scale : -scale, sw = scale * wx[q];
            // y-interpolated values at z0 / z1 (lo, hi) and the y differences there; value at this x = lo + w2 (hi - lo)
            const float dy0x = v[q][1].x - v[q][0].x, dy1x = v[q][3].x - v[q][2].x;
            const float dy0y = v[q][1].y - v[q][0].y, dy1y = v[q][3].y - v[q][2].y;
            const float lox = fmaf(w1[q], dy0x, v[q][0].x), hix = fmaf(w1[q], dy1x, v[q][2].x);
            const float loy = fmaf(w1[q], dy0y, v[q][0].y), hiy = fmaf(w1[q], dy1y, v[q][2].y);
            jx[2 * q] = sx * fmaf(w2[q], hix - lox, lox);
            jx[2 * q + 1] = sx * fmaf(w2[q], hiy - loy, loy);
            jy[2 * q] = sw * fmaf(w2[q], dy1x - dy0x, dy0x);
            jy[2 * q + 1] = sw * fmaf(w2[q], dy1y - dy0y, dy0y);
            jz[2 * q] = sw * (hix - lox);
            jz[2 * q + 1] = sw * (hiy - loy);
        }
    }
    float* const mine_at = keep ? keep + kept_offset(kJac ? 4 * kGridFeats : kGridFeats, c, xb) : nullptr;
    pair_combine_store<kGS>(part, xb, a_hi, a_lo, row, c, mine_at);
    if (kJac) {
        pair_combine_keep(jx, xb, mine_at + 8);
        pair_combine_keep(jy, xb, mine_at + 16);
        pair_combine_keep(jz, xb, mine_at + 24);
    }
#ifdef CB200_PHASE_CLOCKS
    long long t3 = clock64();
    sub[0] += t1 - t0; sub[1] += t2 - t1; sub[2] += t3 - t2;
#endif
}

template <int kDense, bool kJac, uint32_t kGS = kAGS>
__device__ __forceinline__ void encode_grid_chunk(const PointsArgs& a, const float x[3], int xb, int c, uint32_t a_hi, uint32_t a_lo, int row,
                                                  float* keep
#ifdef CB200_PHASE_CLOCKS
                                                  , long long* sub
#endif
) {
#ifdef CB200_PHASE_CLOCKS
    encode_grid_chunk_with<kDense, kJac, kGS>(a, x, xb, c, a_hi, a_lo, row, keep, [] {}, sub);
#else
    encode_grid_chunk_with<kDense, kJac, kGS>(a, x, xb, c, a_hi, a_lo, row, keep, [] {});
#endif
}

// OneBlob bins 8*hb .. 8*hb+7 of one coordinate: values (and optionally d value / d x) from the 9 bin-edge CDFs.
// kImages: evaluate the two periodic images of the kernel as well.  For 0 <= x <= 15/16 they are saturated at every edge used
// (edges 0/16 .. 15/16; the 16th is edge 0 plus one): cdf(t - 1) == 0, cdf(t + 1) == 1 and both pdf images == 0 exactly, so
// (cdf(t) + 0) + 1 is the same fp32 value the three-image sum produces.
template <bool kDeriv, bool kImages>
__device__ __forceinline__ void blob_half_impl(float xd, int hb, float (&val)[8], float (&dval)[8]) {
    const float nb = (float)kBins;
    auto cdf = [&](float e) { return kImages ? blob_edge_cdf(e, xd, nb) : quartic_cdf(e - xd, nb) + 1.0f; };
    auto pdf = [&](float e) { return kImages ? blob_edge_pdf(e, xd, nb) : quartic_pdf(e - xd, nb); };
    float f_prev = cdf((float)(8 * hb) / nb);
    float k_prev = kDeriv ? pdf((float)(8 * hb) / nb) : 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const int b = 8 * hb + e + 1;                    // right edge index
        // the last right edge wraps: F(1 - x) = F(0 - x) + 1, identical kernel value
        const float edge = (b == kBins) ? 0.f : (float)b / nb;
        const float f_next = cdf(edge) + ((b == kBins) ? 1.0f : 0.f);
        val[e] = f_next - f_prev;
        f_prev = f_next;
        if (kDeriv) {
            const float k_next = pdf(edge);
            dval[e] = k_prev - k_next;
            k_prev = k_next;
        }
    }
}

template <bool kDeriv>
__device__ __forceinline__ void blob_half(float xd, int hb, float (&val)[8], float (&dval)[8]) {
    if (xd >= 0.f && xd <= 15.0f / 16.0f) blob_half_impl<kDeriv, false>(xd, hb, val, dval);
    else blob_half_impl<kDeriv, true>(xd, hb, val, dval);
}

// make this thread's shared-memory writes visible to the tensor core, then CTA barrier
__device__ __forceinline__ void publish_tiles() {
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
}

// relu + sign bits of 16 accumulator columns -> two chunks of a [128 x 32] tile
__device__ __forceinline__ uint32_t relu16_to_tile(const uint32_t (&r)[16], uint32_t hi, uint32_t lo, int row, int half) {
    uint32_t bits = 0;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float h = __uint_as_float(r[c * 8 + e]);
            const bool on = h > 0.f;
            bits |= on ? (1u << (c * 8 + e)) : 0u;
            v[e] = on ? h : 0.f;
        }
        tile_store8(hi, lo, kHGS, row, 2 * half + c, v);
    }
    return bits;
}

// masked copy (given sign bits) of 16 accumulator columns -> two chunks of a [128 x 32] tile
__device__ __forceinline__ void mask16_to_tile(const uint32_t (&r)[16], uint32_t bits, uint32_t hi, uint32_t lo, int row, int half) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = ((bits >> (c * 8 + e)) & 1u) ? __uint_as_float(r[c * 8 + e]) : 0.f;
        tile_store8(hi, lo, kHGS, row, 2 * half + c, v);
    }
}

}  // namespace tc

// ------------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------------
// The persistent tile loop of the forward: tiles tile_begin, tile_begin + tile_step, ... of the n_slots points (or list slots).
// Expects the weights staged (stage_all_weights), `bar` initialised with count 1, at least kTmemCols TMEM columns at `tmem`;
// `parity` is the phase the next wait on `bar` expects and is left up to date.  Called by points_fwd_tc_kernel and by the
// one-launch tracking loop (track_frame_kernel).
template <int kDense, bool kJac = false>
__device__ __forceinline__ void fwd_tile_loop(const PointsArgs& a, unsigned char* smem, uint64_t& bar, uint32_t tmem, uint32_t& parity,
                                              const uint32_t* list, int64_t n_slots, int64_t tile_begin, int64_t tile_step) {
    using namespace tc;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = tid & (kTile - 1), half = tid >> 7;                                  // epilogue role
    const int row_g = (warp & 3) * 32 + (warp >> 2) * 16 + ((tid & 31) >> 1), xb = tid & 1;  // gather role
    const uint32_t sbase = smem_u32(smem);
    const uint32_t a_hi = sbase + kOffAHi, a_lo = sbase + kOffALo, h_hi = sbase + kOffHHi, h_lo = sbase + kOffHLo;
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);      // this warp's 32 TMEM lanes
    uint16_t* mask16 = reinterpret_cast<uint16_t*>(a.act_mask);

#define TC_WAIT()                 \
    do {                          \
        mbar_wait(&bar, parity);  \
        parity ^= 1u;             \
        tc_fence_after();         \
    } while (0)

    // with a point list (lazy forward, render.cu) the tiles run over the listed points only: slot i -> point live_idx[i]
    const int64_t n_tiles = (n_slots + kTile - 1) / kTile;
#ifdef CB200_PHASE_CLOCKS
    long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, t_prev = clock64(), sub[3] = {0, 0, 0};
    int n_done = 0;
#define PH(i) do { const long long t_now = clock64(); ph[i] += t_now - t_prev; t_prev = t_now; } while (0)
#else
#define PH(i)
#endif
    // gather-role point of a tile: list slot -> point -> (ray, sample) -> position is a chain of dependent global loads.  The first
    // tile's is resolved here; every later one is started while the previous tile sits in its MMA rounds (index after the first MMA
    // has been issued, ray / sample values after the second) and finished at the end of that tile.
    auto gather_point_of = [&](int64_t t) -> int64_t {
        const int64_t raw = t * kTile + row_g;
        const int64_t slot = raw < n_slots ? raw : n_slots - 1;      // padding slots repeat the last point: same values
        return list ? (int64_t)list[slot] : slot;
    };
    int64_t pg = 0;
    float x[3] = {0.f, 0.f, 0.f};
    if (tile_begin < n_tiles) {
        float zdummy;
        pg = gather_point_of(tile_begin);
        point_position(a, pg, x, zdummy);
    }
    for (int64_t tile = tile_begin; tile < n_tiles; tile += tile_step) {
        const int64_t p_raw = tile * kTile + row;
        const bool valid = p_raw < n_slots;
        const int64_t p_slot = valid ? p_raw : n_slots - 1;
        const int64_t p = list ? (int64_t)list[p_slot] : p_slot;
        const bool has_next = tile + tile_step < n_tiles;
        int64_t pg_next = 0;
        PointLoads next_loads;
        PH(7);
        {
            float* const keep = (kJac || a.feat) ? a.feat + pg * a.feat_stride : nullptr;
            // OneBlob (thread xb does bins 8 xb .. 8 xb + 7 of every coordinate) is pure arithmetic: one coordinate goes into the shadow of
            // each of the level chunks 1..3, between the issue of the chunk's 16 gathers and their first use (forward 2.69 -> 2.56 ms at 2^18
            // rays).  Not when the derivative rows are kept as well: that chunk body has no register to spare (3.41 -> 3.51 ms, a spill).
            auto blob_coordinate = [&](int d) {
                float v[8], unused[8];
                blob_half<false>(x[d], xb, v, unused);
                tile_store8(a_hi, a_lo, kAGS, row_g, 4 + 2 * d + xb, v);
            };
#pragma unroll
            for (int c = 0; c < 4; ++c) {
#ifdef CB200_PHASE_CLOCKS
                encode_grid_chunk_with<kDense, kJac, kAGS>(a, x, xb, c, a_hi, a_lo, row_g, keep, [&] { if (!kJac && c >= 1) blob_coordinate(c - 1); }, sub);
#else
                encode_grid_chunk_with<kDense, kJac, kAGS>(a, x, xb, c, a_hi, a_lo, row_g, keep, [&] { if (!kJac && c >= 1) blob_coordinate(c - 1); });
#endif
            }
            PH(0);
            if (kJac) {
#pragma unroll
                for (int d = 0; d < 3; ++d) blob_coordinate(d);
            }
            PH(1);
        }
        // ---- SDF net layer 1: h1[128,32] = in1[128,80] W1^T ----------------------------------------------------------
        publish_tiles();
        PH(2);
        if (warp == 0 && elect_one()) {
            umma_split3(tmem + kColH1, a_hi, a_lo, 128u, kAGS, 256u, sbase + kOffW1Hi, sbase + kOffW1Lo, 128u, kW1GS, 256u, 5,
                        umma_idesc_bf16(128, 32, false, false), false);
            umma_commit(&bar);
        }
        if (has_next) pg_next = gather_point_of(tile + tile_step);
        TC_WAIT();
        PH(3);
        uint32_t m1;
        {
            uint32_t r[16];
            tmem_ld_x16(trow + kColH1 + 16 * half, r);
            tmem_ld_wait();
            m1 = relu16_to_tile(r, h_hi, h_lo, row, half);
        }
        // ---- layer 2: h2[128,16] = relu(h1) W2^T ------------------------------------------------------------------------
        publish_tiles();
        if (warp == 0 && elect_one()) {
            umma_split3(tmem + kColH2, h_hi, h_lo, 128u, kHGS, 256u, sbase + kOffW2Hi, sbase + kOffW2Lo, 128u, kW2GS, 256u, 2,
                        umma_idesc_bf16(128, 16, false, false), false);
            umma_commit(&bar);
        }
        if (has_next) point_loads(a, pg_next, next_loads);
        TC_WAIT();
        float sdf;
        {
            uint32_t r[16];
            tmem_ld_x16(trow + kColH2, r);
            tmem_ld_wait();
            sdf = __uint_as_float(r[0]);
            if (a.geo && valid && half == 0) {
#pragma unroll
                for (int g = 0; g < kGeo; ++g) a.geo[p * kGeo + g] = __uint_as_float(r[1 + g]);
            }
            if (!a.sdf_only) {
                float v[8];
                if (half == 0) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[1 + e]);
                } else {
#pragma unroll
                    for (int e = 0; e < 7; ++e) v[e] = __uint_as_float(r[9 + e]);
                    v[7] = 0.f;
                }
                tile_store8(a_hi, a_lo, kAGS, row, 10 + half, v);
            }
        }
        float rgb[3] = {0.f, 0.f, 0.f};
        uint32_t m3 = 0;
        if (!a.sdf_only) {
            // ---- colour net layer 1: h3[128,32] = [blob | geo][128,64] W3^T ------------------------------------------------
            publish_tiles();
            if (warp == 0 && elect_one()) {
                umma_split3(tmem + kColH3, a_hi + 4 * 128, a_lo + 4 * 128, 128u, kAGS, 256u, sbase + kOffW3Hi, sbase + kOffW3Lo, 128u, kW3GS,
                            256u, 4, umma_idesc_bf16(128, 32, false, false), false);
                umma_commit(&bar);
            }
            TC_WAIT();
            {
                uint32_t r[16];
                tmem_ld_x16(trow + kColH3 + 16 * half, r);
                tmem_ld_wait();
                m3 = relu16_to_tile(r, h_hi, h_lo, row, half);
            }
            // ---- colour net layer 2: rgb[128,16(3)] = relu(h3) W4^T --------------------------------------------------------
            publish_tiles();
            if (warp == 0 && elect_one()) {
                umma_split3(tmem + kColRgb, h_hi, h_lo, 128u, kHGS, 256u, sbase + kOffW4Hi, sbase + kOffW4Lo, 128u, kW4GS, 256u, 2,
                            umma_idesc_bf16(128, 16, false, false), false);
                umma_commit(&bar);
            }
            TC_WAIT();
            if (half == 0) {
                uint32_t r[4];
                tmem_ld_x4(trow + kColRgb, r);
                tmem_ld_wait();
                rgb[0] = __uint_as_float(r[0]);
                rgb[1] = __uint_as_float(r[1]);
                rgb[2] = __uint_as_float(r[2]);
            }
        }
        PH(4);
        if (valid) {
            if (a.raw && half == 0) reinterpret_cast<float4*>(a.raw)[p] = make_float4(rgb[0], rgb[1], rgb[2], sdf);
            if (a.act_mask) {                     // [P,2] uint32 seen as [P,4] uint16: m1 low, m1 high, m3 low, m3 high
                mask16[p * 4 + half] = (uint16_t)m1;
                mask16[p * 4 + 2 + half] = (uint16_t)m3;
            }
        }
        if (has_next) {
            float zdummy;
            point_from_loads(a, next_loads, x, zdummy);
            pg = pg_next;
        }
        tc_fence_before();
        __syncthreads();
        PH(5);
#ifdef CB200_PHASE_CLOCKS
        ++n_done;
#endif
    }
#ifdef CB200_PHASE_CLOCKS
    if ((blockIdx.x == 0 || blockIdx.x == 100) && (tid == 0 || tid == 200))
        printf("fwd cta %d tid %d tiles %d: gather %lld blob %lld publish %lld mma1wait %lld rounds2-4 %lld stores+sync %lld head %lld | per chunk: index+issue %lld wait %lld consume %lld (cycles per tile)\n",
               blockIdx.x, tid, n_done, ph[0] / n_done, ph[1] / n_done, ph[2] / n_done, ph[3] / n_done, ph[4] / n_done, ph[5] / n_done, ph[7] / n_done, sub[0] / n_done / 4, sub[1] / n_done / 4, sub[2] / n_done / 4);
#endif
#undef PH
#undef TC_WAIT
}

// ------------------------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------------------------
// Per 128-point tile:
//   (1) input-gradient chain on the tensor core, from d_raw and the saved ReLU masks only (no features needed):
//         d_h3 = relu'(h3) . (d_rgb W4)                      SIMT (K = 3)
//         d_in2 [128,64] = d_h3 W3        -> d_geo            round "b"
//         d_h1 = relu'(h1) . (d_h2 W2)                        round "c"
//         d_in1 [128,80] = d_h1 W1  (+= d_h3 W3[:, :48] on the OneBlob columns)   round "d"
//   (2) ONE gather pass per point: dL/dx from the corner values and d_embed (read back from TMEM 8 columns at a time),
//       the table scatter (red.global.add.v2.f32 per corner) and -- for the weight gradients -- the features, written
//       into the A tile;
//   (3) kWgrad: forward recompute on the tensor core (h1, h2, h3) and the four weight-gradient products, which read the
//       SAME activation / gradient tiles through MN-major descriptors (reduction over the 128 points) and accumulate in
//       TMEM for the whole persistent loop; one atomicAdd per weight per CTA at the end.
// Shared memory (kWgrad): A 48 K | HG (d_h1 tile, then relu(h1), then relu(h3)) 16 K | G3 (d_h3) 16 K | G2 (d_h2, then d_rgb) 8 K
//                         | weights 24 K  = 112 K  -> 2 CTAs/SM.
// TMEM (256 columns): dW1^T 0 | dW3^T 32 | dW2^T 64 | dW4^T 80 | d_in1 96..175 | d_in2 176..239 | d_h1 224..255 (after d_in2 died)
//                     forward recompute reuses 176.. : h1 176 | h2 208 | h3 224
namespace tc {

constexpr uint32_t kG2GS = 2 * 128, kG2Bytes = 16 * kG2GS;    // [128 x 16] tile: 4096 per half

constexpr uint32_t kBOffHG = kOffHHi;                          // reuse the forward layout: A | H(=HG) | weights ...
constexpr uint32_t kBOffG3Hi = kFwdSmem, kBOffG3Lo = kBOffG3Hi + kHBytes;
constexpr uint32_t kBOffG2Hi = kBOffG3Lo + kHBytes, kBOffG2Lo = kBOffG2Hi + kG2Bytes;
constexpr uint32_t kBOffW4f = kBOffG2Lo + kG2Bytes;            // fp32 copy of W4 [3][32]
constexpr uint32_t kBwdSmem = kBOffW4f + 3 * 32 * 4;           // 90112 + 16384 + 8192 + 384 = 115072

constexpr uint32_t kColDW1 = 0, kColDW3 = 32, kColDW2 = 64, kColDW4 = 80;
constexpr uint32_t kColDin1 = 96, kColDin2 = 176, kColDh1 = 224;
constexpr uint32_t kColRH1 = 176, kColRH2 = 208, kColRH3 = 224;
constexpr int kBwdTmemCols = 256;

}  // namespace tc

// The persistent tile loop of the backward (same contract as fwd_tile_loop; kBwdTmemCols columns; additionally expects the fp32
// copy of W4 at kBOffW4f).  first_tile: no weight-gradient product has been accumulated in TMEM yet (kWgrad).
// kKept (with kWgrad): the hash-grid features come from PointsArgs::feat (kept by the forward) instead of a second gather of the table.
// 1: rows of 32 features, no position gradient (it needs the corner values); 2: rows of [feat | d feat/dx | d feat/dy | d feat/dz],
// position gradient = d embed . those derivatives.
template <bool kWgrad, int kKept = 0>
__device__ __forceinline__ void bwd_tile_loop(const PointsArgs& a, unsigned char* smem, uint64_t& bar, uint32_t tmem, uint32_t& parity,
                                              bool& first_tile, const uint32_t* list, int64_t n_slots, int64_t tile_begin,
                                              int64_t tile_step) {
    using namespace tc;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row = tid & (kTile - 1), half = tid >> 7;                                // epilogue role
    const int row_g = (warp & 3) * 32 + half * 16 + (lane >> 1), xb = lane & 1;          // gather role
    const int src_lane = half * 16 + (lane >> 1);                                      // lane of this warp that reads row_g from TMEM
    const uint32_t sbase = smem_u32(smem);
    const uint32_t a_hi = sbase + kOffAHi, a_lo = sbase + kOffALo;
    const uint32_t hg_hi = sbase + kOffHHi, hg_lo = sbase + kOffHLo;
    const uint32_t g3_hi = sbase + kBOffG3Hi, g3_lo = sbase + kBOffG3Lo;
    const uint32_t g2_hi = sbase + kBOffG2Hi, g2_lo = sbase + kBOffG2Lo;
    float* sW4f = reinterpret_cast<float*>(smem + kBOffW4f);
    const uint16_t* mask16 = reinterpret_cast<const uint16_t*>(a.act_mask);
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    float2* const dtab = a.dtable;

#define TC_WAIT()                 \
    do {                          \
        mbar_wait(&bar, parity);  \
        parity ^= 1u;             \
        tc_fence_after();         \
    } while (0)

    // with a compacted list the tiles run over the live points only (slot i -> point live_idx[i])
    const int64_t n_tiles = (n_slots + kTile - 1) / kTile;
    // The two pointer chases of a tile are taken off its critical path (~5 % of this kernel's warp-stall samples sat on them):
    // the epilogue-role point of the NEXT tile is read from the list while this tile runs, and this tile's gather-role point is
    // resolved during the MMA rounds of step (1) -- index at the top, ray / sample values after round b -- instead of in front of step (2).
    auto slot_of = [&](int64_t t, int r) -> int64_t {
        const int64_t raw = t * kTile + r;
        return raw < n_slots ? raw : n_slots - 1;
    };
    // (one register: list entries are 32-bit; without a list the point is its slot and needs no load)
    uint32_t p_ahead = (list && tile_begin < n_tiles) ? list[slot_of(tile_begin, row)] : 0u;
    for (int64_t tile = tile_begin; tile < n_tiles; tile += tile_step) {
        const int64_t p_raw = tile * kTile + row;
        const bool valid = p_raw < n_slots;
        const int64_t p = list ? (int64_t)p_ahead : slot_of(tile, row);
        const int64_t pg_raw = tile * kTile + row_g;
        const bool valid_g = pg_raw < n_slots;
        const uint32_t pg_entry = list ? list[slot_of(tile, row_g)] : 0u;

        // ---- (1) input-gradient chain (epilogue role: 16 columns per thread) --------------------------------------------
        float4 g = reinterpret_cast<const float4*>(a.d_raw)[p];
        const uint32_t m1 = mask16[p * 4 + half];
        uint32_t m3 = mask16[p * 4 + 2 + half];
        if (!valid) g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.sdf_only) { g.x = g.y = g.z = 0.f; m3 = 0u; }
        // Points whose output gradient is exactly zero (samples behind the truncation band carry no loss and no rendering
        // weight: typically a third of a batch, contiguous along each ray) contribute exact zeros to every gradient: their
        // gathers and reductions are skipped.  Bit i of `dead` = row 32 q + i of this warp's TMEM quadrant.
        const uint32_t dead = __ballot_sync(0xffffffffu, a.d_geo == nullptr && g.x == 0.f && g.y == 0.f && g.z == 0.f && g.w == 0.f);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int j = half * 16 + c * 8 + e;
                const float t = sW4f[j] * g.x + sW4f[32 + j] * g.y + sW4f[64 + j] * g.z;
                v[e] = ((m3 >> (c * 8 + e)) & 1u) ? t : 0.f;
            }
            tile_store8(g3_hi, g3_lo, kHGS, row, 2 * half + c, v);
        }
        publish_tiles();
        if (warp == 0 && elect_one()) {      // round b: d_in2 = d_h3 W3   (B = W3 tile read MN-major: reduce over its rows)
            umma_split3(tmem + kColDin2, g3_hi, g3_lo, 128u, kHGS, 256u, sbase + kOffW3Hi, sbase + kOffW3Lo, kW3GS, 128u, 2 * kW3GS, 2,
                        umma_idesc_bf16(128, 64, false, true), false);
            umma_commit(&bar);
        }
        if (list && tile + tile_step < n_tiles) p_ahead = list[slot_of(tile + tile_step, row)];
        TC_WAIT();
        {
            uint32_t r[16];
            tmem_ld_x16(trow + kColDin2 + 48, r);      // columns 48..62 = d geo
            tmem_ld_wait();
            float v[8];
            if (half == 0) {                           // d_h2 columns 0..7 = [d sdf, d geo 0..6]
                v[0] = g.w;
#pragma unroll
                for (int e = 1; e < 8; ++e) {
                    float t = __uint_as_float(r[e - 1]);
                    if (a.d_geo && valid) t += a.d_geo[p * kGeo + (e - 1)];
                    v[e] = t;
                }
            } else {                                   // columns 8..15 = d geo 7..14
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float t = __uint_as_float(r[7 + e]);
                    if (a.d_geo && valid) t += a.d_geo[p * kGeo + 7 + e];
                    v[e] = t;
                }
            }
            tile_store8(g2_hi, g2_lo, kG2GS, row, half, v);
        }
        publish_tiles();
        if (warp == 0 && elect_one()) {      // round c: d_h1 (pre-mask) = d_h2 W2
            umma_split3(tmem + kColDh1, g2_hi, g2_lo, 128u, kG2GS, 256u, sbase + kOffW2Hi, sbase + kOffW2Lo, kW2GS, 128u, 2 * kW2GS, 1,
                        umma_idesc_bf16(128, 32, false, true), false);
            umma_commit(&bar);
        }
        const int64_t pg = list ? (int64_t)pg_entry : slot_of(tile, row_g);
        PointLoads g_loads;
        point_loads(a, pg, g_loads);                  // consumed behind round d: two MMA rounds of latency cover them
        TC_WAIT();
        {
            uint32_t r[16];
            tmem_ld_x16(trow + kColDh1 + 16 * half, r);
            tmem_ld_wait();
            mask16_to_tile(r, m1, hg_hi, hg_lo, row, half);
        }
        publish_tiles();
        if (warp == 0 && elect_one()) {      // round d: d_in1 = d_h1 W1, then the OneBlob columns += d_h3 W3[:, :48]
            umma_split3(tmem + kColDin1, hg_hi, hg_lo, 128u, kHGS, 256u, sbase + kOffW1Hi, sbase + kOffW1Lo, kW1GS, 128u, 2 * kW1GS, 2,
                        umma_idesc_bf16(128, 80, false, true), false);
            umma_split3(tmem + kColDin1 + 32, g3_hi, g3_lo, 128u, kHGS, 256u, sbase + kOffW3Hi, sbase + kOffW3Lo, kW3GS, 128u, 2 * kW3GS, 2,
                        umma_idesc_bf16(128, 48, false, true), true);
            umma_commit(&bar);
        }
        TC_WAIT();

        // ---- (2) one gather pass (gather role: lane pair per point): dL/dx, table scatter, features ---------------------
        const bool live = valid_g && !((dead >> src_lane) & 1u);
        float x[3], zray;
        point_from_loads(a, g_loads, x, zray);
        float dx[3] = {0.f, 0.f, 0.f};
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t de_raw[8];
            tmem_ld_x8(trow + kColDin1 + 8 * c, de_raw);        // d embed of levels 4c .. 4c+3, row = this lane's TMEM lane
            if (kKept) {
                // table scatter only: thread xb takes columns 8c + 4xb .. + 3 of the point's kept row
                const float* const krow = a.feat + pg * a.feat_stride + kept_offset(kKept == 2 ? 4 * kGridFeats : a.feat_stride, c, xb);
                float4 kept = make_float4(0.f, 0.f, 0.f, 0.f), jac[3];
                if (live) kept = __ldcs(reinterpret_cast<const float4*>(krow));
                if (kKept == 2) {
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        jac[d] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (live) jac[d] = __ldcs(reinterpret_cast<const float4*>(krow + 8 * (d + 1)));
                    }
                }
                float dm[4] = {0.f, 0.f, 0.f, 0.f};      // d embed of this thread's columns (levels 4c + 2xb, 4c + 2xb + 1)
                uint32_t idx[4][4];
                float wx[4], w1[4], w2[4];
                uint32_t in_cells = 0;      // bit q: level 4c + q reduces into the cell-major gradient block idx[q][0] (dense level, copy present)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    HalfCorners hc;
                    level_corners_x<-1>(a.scene, 4 * c + q, x, xb, hc);
#pragma unroll
                    for (int j = 0; j < 4; ++j) idx[q][j] = hc.idx[j];
                    wx[q] = hc.wx; w1[q] = hc.w1; w2[q] = hc.w2;
                    if (a.scene.dcells != nullptr && a.scene.hashed[4 * c + q] == 0) {
                        const float scale = a.scene.scale[4 * c + q];
                        uint32_t cell[3], at;
                        float unused;
#pragma unroll
                        for (int d = 0; d < 3; ++d) grid_pos(scale, x[d], cell[d], unused);
                        if (cell_block(a.scene, 4 * c + q, cell, xb, at)) {
                            idx[q][0] = at;
                            in_cells |= 1u << q;
                        }
                    }
                }
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float de0 = __uint_as_float(__shfl_sync(0xffffffffu, de_raw[2 * q], src_lane));
                    const float de1 = __uint_as_float(__shfl_sync(0xffffffffu, de_raw[2 * q + 1], src_lane));
                    if ((in_cells >> q) & 1u) {
                        if (dtab && live) {
                            const float c0 = wx[q] * yz_weight(w1[q], w2[q], 0), c1 = wx[q] * yz_weight(w1[q], w2[q], 1);
                            const float c2 = wx[q] * yz_weight(w1[q], w2[q], 2), c3 = wx[q] * yz_weight(w1[q], w2[q], 3);
                            red_add_f4(a.scene.dcells + idx[q][0], c0 * de0, c0 * de1, c1 * de0, c1 * de1);
                            red_add_f4(a.scene.dcells + idx[q][0] + 1, c2 * de0, c2 * de1, c3 * de0, c3 * de1);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float wc = wx[q] * yz_weight(w1[q], w2[q], j);
                            if (dtab && live) red_add_f2(dtab + idx[q][j], wc * de0, wc * de1);
                        }
                    }
                    if (kKept == 2 && (q >> 1) == xb) {
                        dm[2 * (q & 1)] = de0;
                        dm[2 * (q & 1) + 1] = de1;
                    }
                }
                if (kKept == 2) {
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                        dx[d] = fmaf(dm[0], jac[d].x, fmaf(dm[1], jac[d].y, fmaf(dm[2], jac[d].z, fmaf(dm[3], jac[d].w, dx[d]))));
                }
                const float mine[4] = {kept.x, kept.y, kept.z, kept.w};
                tile_store4(a_hi, a_lo, kAGS, row_g, c, xb, mine);
                continue;
            }
            float2 v[4][4];
            uint32_t idx[4][4];
            float wx[4], w1[4], w2[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                HalfCorners hc;
                level_corners_x<-1>(a.scene, 4 * c + q, x, xb, hc);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    idx[q][j] = hc.idx[j];
                    // same-box A/B (profiles/r2/red_cost_model.md): the mapping backward is 1.5 % faster with the predicated gather
                    // and a per-lane branch around every RED, the position-only backward 3 % faster without either
                    if (kWgrad) v[q][j] = live ? ldg_f2(a.table + hc.idx[j]) : make_float2(0.f, 0.f);
                    else v[q][j] = ldg_f2(a.table + hc.idx[j]);
                }
                wx[q] = hc.wx; w1[q] = hc.w1; w2[q] = hc.w2;
            }
            tmem_ld_wait();
            float part[8];
            // Without weight gradients (tracking, or a table-only backward): dead and padding lanes have d_embed == 0 exactly (their
            // d_raw is zero, so every product of the gradient chain is), hence their reductions would add zeros and the decision to
            // reduce is taken per WARP (uniform branch); on a compacted list only the last tile has such lanes.
            const bool red_warp = dtab != nullptr && __any_sync(0xffffffffu, live);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float de0 = __uint_as_float(__shfl_sync(0xffffffffu, de_raw[2 * q], src_lane));
                const float de1 = __uint_as_float(__shfl_sync(0xffffffffu, de_raw[2 * q + 1], src_lane));
                float s[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) s[j] = v[q][j].x * de0 + v[q][j].y * de1;
                const float scale = a.scene.scale[4 * c + q];
                // d/dx: (value at x+1) - (value at x), bilinear in (y, z): this thread contributes its side with its sign
                float t = 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) t = fmaf(yz_weight(w1[q], w2[q], j), s[j], t);
                const float gx = xb ? t : -t;
                const float gy = wx[q] * ((1.f - w2[q]) * (s[1] - s[0]) + w2[q] * (s[3] - s[2]));
                const float gz = wx[q] * ((1.f - w1[q]) * (s[2] - s[0]) + w1[q] * (s[3] - s[1]));
                dx[0] = fmaf(scale, gx, dx[0]);
                dx[1] = fmaf(scale, gy, dx[1]);
                dx[2] = fmaf(scale, gz, dx[2]);
                float f0 = 0.f, f1 = 0.f;
                if (kWgrad) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float wc = wx[q] * yz_weight(w1[q], w2[q], j);
                        if (a.dtable && live) red_add_f2(a.dtable + idx[q][j], wc * de0, wc * de1);
                        f0 = fmaf(wc, v[q][j].x, f0);
                        f1 = fmaf(wc, v[q][j].y, f1);
                    }
                } else if (red_warp) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float wc = wx[q] * yz_weight(w1[q], w2[q], j);
                        red_add_f2(dtab + idx[q][j], wc * de0, wc * de1);
                    }
                }
                part[2 * q] = f0;
                part[2 * q + 1] = f1;
            }
            if (kWgrad) pair_combine_store(part, xb, a_hi, a_lo, row_g, c);
        }
        // OneBlob: thread xb owns bins 8 xb .. 8 xb + 7 of every coordinate (gradient columns 32 + 16 d + 8 xb .. of d_in1)
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            float val[8], dval[8];
            if (kKept == 1) {
                blob_half<false>(x[d], xb, val, dval);
                tile_store8(a_hi, a_lo, kAGS, row_g, 4 + 2 * d + xb, val);
                continue;
            }
            uint32_t gb[16];
            tmem_ld_x16(trow + kColDin1 + 32 + 16 * d, gb);
            blob_half<true>(x[d], xb, val, dval);
            tmem_ld_wait();
            float acc = 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const uint32_t lo8 = __shfl_sync(0xffffffffu, gb[e], src_lane);
                const uint32_t hi8 = __shfl_sync(0xffffffffu, gb[8 + e], src_lane);
                acc = fmaf(__uint_as_float(xb ? hi8 : lo8), dval[e], acc);
            }
            dx[d] += acc;
            if (kWgrad) tile_store8(a_hi, a_lo, kAGS, row_g, 4 + 2 * d + xb, val);
        }
#pragma unroll
        for (int d = 0; d < 3; ++d) dx[d] += __shfl_xor_sync(0xffffffffu, dx[d], 1);      // both threads of the pair: full dL/dx

        // ---- position gradient (pair leader) ----------------------------------------------------------------------------------
        if (kKept != 1) {
            float dp[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) dp[d] = valid_g ? (a.scene.normalise ? dx[d] / a.scene.bb_ext[d] : dx[d]) : 0.f;
            if (a.d_pts && valid_g && xb == 0) {
#pragma unroll
                for (int d = 0; d < 3; ++d) a.d_pts[pg * 3 + d] = dp[d];
            }
            if (a.d_rays_o) {
                // segmented sum over the 16 consecutive points of this warp (even lanes; odd lanes mirror them)
                // A run = adjacent slots of one ray.  The merge is bounded by the run's end (head-flag ballot), not by equal ray
                // ids alone: the compacted list appends 1024-point blocks in arbitrary order, so a ray cut by a block boundary
                // can reappear in the same warp window behind another block's points (X, Y, X) and must not be summed twice.
                const int64_t ray = valid_g ? pg / a.n_samples : -1;
                const int64_t rprev = __shfl_up_sync(0xffffffffu, ray, 2);
                const bool head = lane < 2 || rprev != ray;
                const uint32_t heads = __ballot_sync(0xffffffffu, head && xb == 0);
                const uint32_t above = heads & ~((2u << (lane & ~1)) - 1u);
                const int run_end = above ? __ffs(above) - 1 : 32;
                float v6[6] = {dp[0], dp[1], dp[2], dp[0] * zray, dp[1] * zray, dp[2] * zray};
#pragma unroll
                for (int o = 2; o < 32; o <<= 1) {
                    const bool take = ((lane & ~1) + o < run_end);
#pragma unroll
                    for (int i = 0; i < 6; ++i) {
                        const float t = __shfl_down_sync(0xffffffffu, v6[i], o);
                        if (take) v6[i] += t;
                    }
                }
                if (valid_g && xb == 0 && head) {
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        atomicAdd(a.d_rays_o + ray * 3 + d, v6[d]);
                        atomicAdd(a.d_rays_d + ray * 3 + d, v6[3 + d]);
                    }
                }
            }
        }

        // ---- (3) forward recompute + weight gradients (epilogue role) ----------------------------------------------------------
        if (kWgrad) {
            publish_tiles();
            if (warp == 0 && elect_one()) {      // h1 = in1 W1^T ;  dW1^T += in1^T d_h1   (A tile and HG tile read MN-major, reduce over the points)
                umma_split3(tmem + kColRH1, a_hi, a_lo, 128u, kAGS, 256u, sbase + kOffW1Hi, sbase + kOffW1Lo, 128u, kW1GS, 256u, 5,
                            umma_idesc_bf16(128, 32, false, false), false);
                umma_split3(tmem + kColDW1, a_hi, a_lo, kAGS, 128u, 2 * kAGS, hg_hi, hg_lo, kHGS, 128u, 2 * kHGS, 8,
                            umma_idesc_bf16(128, 32, true, true), !first_tile);
                umma_commit(&bar);
            }
            TC_WAIT();
            {
                uint32_t r[16];
                tmem_ld_x16(trow + kColRH1 + 16 * half, r);
                tmem_ld_wait();
                mask16_to_tile(r, m1, hg_hi, hg_lo, row, half);       // HG now holds relu(h1)
            }
            publish_tiles();
            if (warp == 0 && elect_one()) {      // h2 = relu(h1) W2^T ;  dW2^T += relu(h1)^T d_h2
                umma_split3(tmem + kColRH2, hg_hi, hg_lo, 128u, kHGS, 256u, sbase + kOffW2Hi, sbase + kOffW2Lo, 128u, kW2GS, 256u, 2,
                            umma_idesc_bf16(128, 16, false, false), false);
                umma_split3(tmem + kColDW2, hg_hi, hg_lo, kHGS, 128u, 2 * kHGS, g2_hi, g2_lo, kG2GS, 128u, 2 * kG2GS, 8,
                            umma_idesc_bf16(128, 16, true, true), !first_tile);
                umma_commit(&bar);
            }
            TC_WAIT();
            {
                uint32_t r[16];
                tmem_ld_x16(trow + kColRH2, r);
                tmem_ld_wait();
                float v[8];
                if (half == 0) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[1 + e]);
                } else {
#pragma unroll
                    for (int e = 0; e < 7; ++e) v[e] = __uint_as_float(r[9 + e]);
                    v[7] = 0.f;
                }
                tile_store8(a_hi, a_lo, kAGS, row, 10 + half, v);
                // the d_h2 tile is dead now (dW2 completed): reuse it for d_rgb, the B operand of dW4
                float u[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (half == 0) { u[0] = g.x; u[1] = g.y; u[2] = g.z; }
                tile_store8(g2_hi, g2_lo, kG2GS, row, half, u);
            }
            publish_tiles();
            if (warp == 0 && elect_one()) {      // h3 = [blob | geo] W3^T ;  dW3^T += [blob | geo]^T d_h3
                umma_split3(tmem + kColRH3, a_hi + 4 * 128, a_lo + 4 * 128, 128u, kAGS, 256u, sbase + kOffW3Hi, sbase + kOffW3Lo, 128u, kW3GS,
                            256u, 4, umma_idesc_bf16(128, 32, false, false), false);
                umma_split3(tmem + kColDW3, a_hi + 4 * 128, a_lo + 4 * 128, kAGS, 128u, 2 * kAGS, g3_hi, g3_lo, kHGS, 128u, 2 * kHGS, 8,
                            umma_idesc_bf16(128, 32, true, true), !first_tile);
                umma_commit(&bar);
            }
            TC_WAIT();
            {
                uint32_t r[16];
                tmem_ld_x16(trow + kColRH3 + 16 * half, r);
                tmem_ld_wait();
                mask16_to_tile(r, m3, hg_hi, hg_lo, row, half);       // HG now holds relu(h3)
            }
            publish_tiles();
            if (warp == 0 && elect_one()) {      // dW4^T += relu(h3)^T d_rgb
                umma_split3(tmem + kColDW4, hg_hi, hg_lo, kHGS, 128u, 2 * kHGS, g2_hi, g2_lo, kG2GS, 128u, 2 * kG2GS, 8,
                            umma_idesc_bf16(128, 16, true, true), !first_tile);
            }
            // Not waited for here: the tensor core executes in issue order and the next commit (round b of the next tile, or
            // the one after the loop) covers these MMAs too.  Until that wait returns, the next tile only writes G3, which
            // this product does not read (HG and G2 are rewritten after the waits of rounds b / c).
            first_tile = false;
        }
        tc_fence_before();
        __syncthreads();         // every lane has read its d_in1 columns before the next tile's MMAs overwrite TMEM
    }
    if (kWgrad && !first_tile) {                 // the last tile's dW4 product
        if (warp == 0 && elect_one()) umma_commit(&bar);
        TC_WAIT();
    }
#undef TC_WAIT
}

}  // namespace cb200
