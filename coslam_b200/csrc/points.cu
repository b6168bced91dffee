// Per-sample-point kernels: position -> hash-grid + OneBlob encodings -> SDF / colour decoders, forward and backward.
// (model/scene_rep.py:155-191 run_network/query_color_sdf, model/encodings.py:31-46,61-71, model/decoder.py:165-180)
//
// Version 1 ("simt"): one thread owns one point; 128 points form a CTA tile.
//   * hash gathers: 8 independent float2 loads per level per thread, table read through the read-only path;
//   * decoder: weights transposed in shared memory ([in][out], 16-byte rows) and broadcast to the warp with LDS.128,
//     per-thread activations either in registers (hidden vectors) or in a [feature][thread] shared tile
//     (conflict-free: consecutive threads hit consecutive banks);
//   * backward: the ReLU sign bits saved by the forward make the input gradients computable BEFORE the gathers, so
//     one gather pass serves the feature recompute (needed for the weight gradients), dL/dx and the table scatter
//     (one red.global.add.v2.f32 per corner);
//   * weight gradients: the [feature][thread] tiles double as the operands of a CTA-level [K x 128] x [128 x J]
//     product; each thread keeps 41 accumulators for the whole persistent loop and issues its atomics once.
#include "args.cuh"
#include "point_common.cuh"

namespace cb200 {

constexpr int kTile = 128;   // points per CTA tile == threads per CTA

// transposed weight block in shared memory
constexpr int kW1T = 0;                        // [80][32]
constexpr int kW2T = kW1T + kSdfIn * kHidden;  // [32][16]
constexpr int kW3T = kW2T + kHidden * kSdfOut; // [64][32] (row 63 zero)
constexpr int kW4T = kW3T + 64 * kHidden;      // [32][4]  (col 3 zero)
constexpr int kWFloats = kW4T + kHidden * 4;   // 5248

// rows of the activation tile sX[row][thread]
constexpr int kXE = 0;            // 32 grid features        } in1 = rows 0..79
constexpr int kXB = 32;           // 48 OneBlob features     } in2 = rows 32..94
constexpr int kXG = 80;           // 15 geometric features   }
constexpr int kXH1 = 95;          // 32 relu(h1)
constexpr int kXH3 = 127;         // 32 relu(h3)
constexpr int kXRows = 159;
// rows of the gradient tile sD[row][thread]
constexpr int kD1 = 0;            // 32 d h1 (pre-activation)
constexpr int kD2 = 32;           // 16 d h2
constexpr int kD3 = 48;           // 32 d h3 (pre-activation)
constexpr int kD4 = 80;           // 3  d rgb logits
constexpr int kDRows = 83;


__device__ __forceinline__ void load_weights_transposed(float* sW, const PointsArgs& a) {
    for (int i = threadIdx.x; i < kSdfIn * kHidden; i += blockDim.x) {
        const int k = i / kHidden, j = i % kHidden;
        sW[kW1T + i] = a.w1[j * kSdfIn + k];
    }
    for (int i = threadIdx.x; i < kHidden * kSdfOut; i += blockDim.x) {
        const int j = i / kSdfOut, o = i % kSdfOut;
        sW[kW2T + i] = a.w2[o * kHidden + j];
    }
    for (int i = threadIdx.x; i < 64 * kHidden; i += blockDim.x) {
        const int k = i / kHidden, j = i % kHidden;
        sW[kW3T + i] = k < kColIn ? a.w3[j * kColIn + k] : 0.f;
    }
    for (int i = threadIdx.x; i < kHidden * 4; i += blockDim.x) {
        const int j = i / 4, c = i % 4;
        sW[kW4T + i] = c < 3 ? a.w4[c * kHidden + j] : 0.f;
    }
}

// acc[j] += W[row][j] * v  for the 32 outputs of a hidden layer; W row is 32 contiguous floats in shared memory
__device__ __forceinline__ void axpy32(float (&acc)[kHidden], const float* wrow, float v) {
    const float4* w4 = reinterpret_cast<const float4*>(wrow);
#pragma unroll
    for (int q = 0; q < kHidden / 4; ++q) {
        const float4 w = w4[q];
        acc[4 * q + 0] = fmaf(w.x, v, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(w.y, v, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(w.z, v, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(w.w, v, acc[4 * q + 3]);
    }
}
// sum_j W[row][j] * g[j]
__device__ __forceinline__ float dot32(const float* wrow, const float (&g)[kHidden]) {
    const float4* w4 = reinterpret_cast<const float4*>(wrow);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int q = 0; q < kHidden / 4; ++q) {
        const float4 w = w4[q];
        s0 = fmaf(w.x, g[4 * q + 0], s0);
        s1 = fmaf(w.y, g[4 * q + 1], s1);
        s2 = fmaf(w.z, g[4 * q + 2], s2);
        s3 = fmaf(w.w, g[4 * q + 3], s3);
    }
    return (s0 + s1) + (s2 + s3);
}

// ------------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTile) points_fwd_kernel(PointsArgs a) {
    extern __shared__ __align__(16) float smem[];
    float* sW = smem;
    float* sIn2 = smem + kWFloats;           // [63][kTile]: OneBlob (48) then geo (15)
    load_weights_transposed(sW, a);
    __syncthreads();
    const int tid = threadIdx.x;
    const int64_t n_tiles = (a.n_points + kTile - 1) / kTile;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t p = tile * kTile + tid;
        if (p >= a.n_points) continue;
        float x[3], zdummy;
        point_position(a, p, x, zdummy);

        float h1[kHidden];
#pragma unroll
        for (int j = 0; j < kHidden; ++j) h1[j] = 0.f;

        // ---- hash grid: 16 levels x 8 corners, features folded straight into the first hidden layer
#pragma unroll 2
        for (int l = 0; l < (int)a.scene.n_levels; ++l) {
            LevelCorners lc;
            level_corners(a.scene, l, x, lc);
            float2 v[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = ldg_f2(a.table + lc.idx[c]);
            float f0 = 0.f, f1 = 0.f;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float wc = corner_weight(lc.w, c);
                f0 = fmaf(wc, v[c].x, f0);
                f1 = fmaf(wc, v[c].y, f1);
            }
            axpy32(h1, sW + kW1T + (2 * l) * kHidden, f0);
            axpy32(h1, sW + kW1T + (2 * l + 1) * kHidden, f1);
        }
        // ---- OneBlob: 3 x 16 bins (difference of the periodic quartic CDF at the bin edges)
        for (int d = 0; d < 3; ++d) {
            const float f_first = blob_edge_cdf(0.f, x[d], (float)kBins);
            float f_prev = f_first;
            for (int b = 0; b < kBins; ++b) {
                const float f_next = (b == kBins - 1) ? f_first + 1.0f : blob_edge_cdf((float)(b + 1) / (float)kBins, x[d], (float)kBins);
                const float v = f_next - f_prev;
                f_prev = f_next;
                sIn2[(d * kBins + b) * kTile + tid] = v;
                axpy32(h1, sW + kW1T + (kGridFeats + d * kBins + b) * kHidden, v);
            }
        }
        // ---- SDF net layer 2
        uint32_t m1 = 0;
        float h2[kSdfOut];
#pragma unroll
        for (int i = 0; i < kSdfOut; ++i) h2[i] = 0.f;
#pragma unroll
        for (int j = 0; j < kHidden; ++j) {
            const bool on = h1[j] > 0.f;
            m1 |= on ? (1u << j) : 0u;
            const float r = on ? h1[j] : 0.f;
            const float4* w4 = reinterpret_cast<const float4*>(sW + kW2T + j * kSdfOut);
#pragma unroll
            for (int q = 0; q < kSdfOut / 4; ++q) {
                const float4 w = w4[q];
                h2[4 * q + 0] = fmaf(w.x, r, h2[4 * q + 0]);
                h2[4 * q + 1] = fmaf(w.y, r, h2[4 * q + 1]);
                h2[4 * q + 2] = fmaf(w.z, r, h2[4 * q + 2]);
                h2[4 * q + 3] = fmaf(w.w, r, h2[4 * q + 3]);
            }
        }
        if (a.geo) {
#pragma unroll
            for (int g = 0; g < kGeo; ++g) a.geo[p * kGeo + g] = h2[1 + g];
        }
        float rgb[3] = {0.f, 0.f, 0.f};
        uint32_t m3 = 0;
        if (!a.sdf_only) {
#pragma unroll
            for (int g = 0; g < kGeo; ++g) sIn2[(kBlobFeats + g) * kTile + tid] = h2[1 + g];
            float h3[kHidden];
#pragma unroll
            for (int j = 0; j < kHidden; ++j) h3[j] = 0.f;
#pragma unroll 3
            for (int k = 0; k < kColIn; ++k) axpy32(h3, sW + kW3T + k * kHidden, sIn2[k * kTile + tid]);
#pragma unroll
            for (int j = 0; j < kHidden; ++j) {
                const bool on = h3[j] > 0.f;
                m3 |= on ? (1u << j) : 0u;
                const float r = on ? h3[j] : 0.f;
                const float4 w = *reinterpret_cast<const float4*>(sW + kW4T + j * 4);
                rgb[0] = fmaf(w.x, r, rgb[0]);
                rgb[1] = fmaf(w.y, r, rgb[1]);
                rgb[2] = fmaf(w.z, r, rgb[2]);
            }
        }
        if (a.raw) reinterpret_cast<float4*>(a.raw)[p] = make_float4(rgb[0], rgb[1], rgb[2], h2[0]);
        if (a.act_mask) reinterpret_cast<uint2*>(a.act_mask)[p] = make_uint2(m1, m3);
    }
}

// ------------------------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------------------------
template <bool kWgrad>
__global__ void __launch_bounds__(kTile) points_bwd_kernel(PointsArgs a) {
    extern __shared__ __align__(16) float smem[];
    float* sW = smem;
    float* sE = sW + kWFloats;                       // [32][kTile] d embed
    float* sX = sE + kGridFeats * kTile;             // [kXRows][kTile]   (kWgrad only)
    float* sD = sX + kXRows * kTile;                 // [kDRows][kTile]   (kWgrad only)
    load_weights_transposed(sW, a);
    __syncthreads();
    const int tid = threadIdx.x, lane = tid & 31;

    // weight-gradient accumulators of this thread (see header comment)
    float acc1[4][5], acc3[4][4], acc2[4], acc4 = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int k = 0; k < 5; ++k) acc1[i][k] = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) acc3[i][k] = 0.f;
        acc2[i] = 0.f;
    }

    const int64_t n_tiles = (a.n_points + kTile - 1) / kTile;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t p_raw = tile * kTile + tid;
        const bool valid = p_raw < a.n_points;
        const int64_t p = valid ? p_raw : a.n_points - 1;     // padding threads shadow the last point, contribute zeros
        float x[3], zray;
        point_position(a, p, x, zray);

        // ---- (i) input gradients of the two MLPs from the saved ReLU masks -------------------------------------
        float4 g = reinterpret_cast<const float4*>(a.d_raw)[p];
        uint2 m = reinterpret_cast<const uint2*>(a.act_mask)[p];
        if (!valid) g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.sdf_only) { g.x = g.y = g.z = 0.f; m.y = 0u; }
        float dh3[kHidden], dh1[kHidden], dh2[kSdfOut];
#pragma unroll
        for (int j = 0; j < kHidden; ++j) {
            const float4 w = *reinterpret_cast<const float4*>(sW + kW4T + j * 4);
            const float t = w.x * g.x + w.y * g.y + w.z * g.z;
            dh3[j] = ((m.y >> j) & 1u) ? t : 0.f;
        }
        dh2[0] = g.w;
#pragma unroll
        for (int q = 0; q < kGeo; ++q) {
            float t = dot32(sW + kW3T + (kBlobFeats + q) * kHidden, dh3);
            if (a.d_geo && valid) t += a.d_geo[p * kGeo + q];
            dh2[1 + q] = t;
        }
#pragma unroll
        for (int j = 0; j < kHidden; ++j) {
            const float4* w4 = reinterpret_cast<const float4*>(sW + kW2T + j * kSdfOut);
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < kSdfOut / 4; ++q) {
                const float4 w = w4[q];
                t += w.x * dh2[4 * q] + w.y * dh2[4 * q + 1] + w.z * dh2[4 * q + 2] + w.w * dh2[4 * q + 3];
            }
            dh1[j] = ((m.x >> j) & 1u) ? t : 0.f;
        }
        if (kWgrad) {
#pragma unroll
            for (int j = 0; j < kHidden; ++j) {
                sD[(kD1 + j) * kTile + tid] = dh1[j];
                sD[(kD3 + j) * kTile + tid] = dh3[j];
            }
#pragma unroll
            for (int i = 0; i < kSdfOut; ++i) sD[(kD2 + i) * kTile + tid] = dh2[i];
            sD[(kD4 + 0) * kTile + tid] = g.x;
            sD[(kD4 + 1) * kTile + tid] = g.y;
            sD[(kD4 + 2) * kTile + tid] = g.z;
        }
        // d embed (grid features)
#pragma unroll 4
        for (int k = 0; k < kGridFeats; ++k) sE[k * kTile + tid] = dot32(sW + kW1T + k * kHidden, dh1);

        // ---- (ii) OneBlob forward value (for the weight gradients) and backward into dx --------------------------
        float dx[3] = {0.f, 0.f, 0.f};
        for (int d = 0; d < 3; ++d) {
            const float f_first = blob_edge_cdf(0.f, x[d], (float)kBins);
            const float k_first = blob_edge_pdf(0.f, x[d], (float)kBins);
            float f_prev = f_first, k_prev = k_first;
            float acc = 0.f;
            for (int b = 0; b < kBins; ++b) {
                const float e = (float)(b + 1) / (float)kBins;
                const float f_next = (b == kBins - 1) ? f_first + 1.0f : blob_edge_cdf(e, x[d], (float)kBins);
                const float k_next = (b == kBins - 1) ? k_first : blob_edge_pdf(e, x[d], (float)kBins);
                const int kb = d * kBins + b;
                const float gy = dot32(sW + kW3T + kb * kHidden, dh3) + dot32(sW + kW1T + (kGridFeats + kb) * kHidden, dh1);
                acc += gy * (k_prev - k_next);              // d/dx [F(e_{b+1} - x) - F(e_b - x)]
                if (kWgrad) sX[(kXB + kb) * kTile + tid] = f_next - f_prev;
                f_prev = f_next;
                k_prev = k_next;
            }
            dx[d] = acc;
        }

        // ---- (iii) one gather pass: features, dL/dx and the table scatter ---------------------------------------
        float h1[kHidden];
#pragma unroll
        for (int j = 0; j < kHidden; ++j) h1[j] = 0.f;
#pragma unroll 2
        for (int l = 0; l < (int)a.scene.n_levels; ++l) {
            LevelCorners lc;
            level_corners(a.scene, l, x, lc);
            float2 v[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = ldg_f2(a.table + lc.idx[c]);
            const float de0 = sE[(2 * l) * kTile + tid], de1 = sE[(2 * l + 1) * kTile + tid];
            float s[8];        // corner value projected on the feature gradient
#pragma unroll
            for (int c = 0; c < 8; ++c) s[c] = v[c].x * de0 + v[c].y * de1;
            const float w0 = lc.w[0], w1 = lc.w[1], w2 = lc.w[2];
            const float scale = a.scene.scale[l];
            // derivative of the trilinear interpolant along each axis
            const float gx = (1.f - w1) * (1.f - w2) * (s[1] - s[0]) + w1 * (1.f - w2) * (s[3] - s[2]) +
                             (1.f - w1) * w2 * (s[5] - s[4]) + w1 * w2 * (s[7] - s[6]);
            const float gy = (1.f - w0) * (1.f - w2) * (s[2] - s[0]) + w0 * (1.f - w2) * (s[3] - s[1]) +
                             (1.f - w0) * w2 * (s[6] - s[4]) + w0 * w2 * (s[7] - s[5]);
            const float gz = (1.f - w0) * (1.f - w1) * (s[4] - s[0]) + w0 * (1.f - w1) * (s[5] - s[1]) +
                             (1.f - w0) * w1 * (s[6] - s[2]) + w0 * w1 * (s[7] - s[3]);
            dx[0] = fmaf(scale, gx, dx[0]);
            dx[1] = fmaf(scale, gy, dx[1]);
            dx[2] = fmaf(scale, gz, dx[2]);
            if (a.dtable && valid) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float wc = corner_weight(lc.w, c);
                    red_add_f2(a.dtable + lc.idx[c], wc * de0, wc * de1);
                }
            }
            if (kWgrad) {
                float f0 = 0.f, f1 = 0.f;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float wc = corner_weight(lc.w, c);
                    f0 = fmaf(wc, v[c].x, f0);
                    f1 = fmaf(wc, v[c].y, f1);
                }
                sX[(kXE + 2 * l) * kTile + tid] = f0;
                sX[(kXE + 2 * l + 1) * kTile + tid] = f1;
                axpy32(h1, sW + kW1T + (2 * l) * kHidden, f0);
                axpy32(h1, sW + kW1T + (2 * l + 1) * kHidden, f1);
            }
        }

        // ---- (iv)+(v) activations for the weight gradients, then the CTA-level products ---------------------------
        if (kWgrad) {
#pragma unroll 4
            for (int kb = 0; kb < kBlobFeats; ++kb) axpy32(h1, sW + kW1T + (kGridFeats + kb) * kHidden, sX[(kXB + kb) * kTile + tid]);
            float h2[kSdfOut];
#pragma unroll
            for (int i = 0; i < kSdfOut; ++i) h2[i] = 0.f;
#pragma unroll
            for (int j = 0; j < kHidden; ++j) {
                const float r = ((m.x >> j) & 1u) ? h1[j] : 0.f;
                sX[(kXH1 + j) * kTile + tid] = r;
                const float4* w4 = reinterpret_cast<const float4*>(sW + kW2T + j * kSdfOut);
#pragma unroll
                for (int q = 0; q < kSdfOut / 4; ++q) {
                    const float4 w = w4[q];
                    h2[4 * q + 0] = fmaf(w.x, r, h2[4 * q + 0]);
                    h2[4 * q + 1] = fmaf(w.y, r, h2[4 * q + 1]);
                    h2[4 * q + 2] = fmaf(w.z, r, h2[4 * q + 2]);
                    h2[4 * q + 3] = fmaf(w.w, r, h2[4 * q + 3]);
                }
            }
#pragma unroll
            for (int q = 0; q < kGeo; ++q) sX[(kXG + q) * kTile + tid] = h2[1 + q];
            float h3[kHidden];
#pragma unroll
            for (int j = 0; j < kHidden; ++j) h3[j] = 0.f;
#pragma unroll 3
            for (int k = 0; k < kColIn; ++k) axpy32(h3, sW + kW3T + k * kHidden, sX[(kXB + k) * kTile + tid]);
#pragma unroll
            for (int j = 0; j < kHidden; ++j) sX[(kXH3 + j) * kTile + tid] = ((m.y >> j) & 1u) ? h3[j] : 0.f;

            __syncthreads();
            {
                const int tj = tid & 7, tk = tid >> 3;          // dW1: rows 4*tj.., cols 5*tk.. ; dW3: cols 4*tk..
                const int ti = tid & 3, tc = tid >> 2;          // dW2: rows 4*ti.., col tc
                const int c4 = tid >> 5, j4 = tid & 31;         // dW4: row c4 (<3), col j4
                const float* d1 = sD + (kD1 + 4 * tj) * kTile;
                const float* d3 = sD + (kD3 + 4 * tj) * kTile;
                const float* d2 = sD + (kD2 + 4 * ti) * kTile;
                const float* d4 = sD + (kD4 + (c4 < 3 ? c4 : 0)) * kTile;
                const float* x1 = sX + (kXE + 5 * tk) * kTile;
                const float* x3 = sX + (kXB + 4 * tk) * kTile;
                const float* x2 = sX + (kXH1 + tc) * kTile;
                const float* x4 = sX + (kXH3 + j4) * kTile;
#pragma unroll 2
                for (int it = 0; it < kTile; ++it) {
                    const int q = (it + lane) & (kTile - 1);   // lane skew: every lane reads its own bank
                    float dv[4], xv[5];
#pragma unroll
                    for (int i = 0; i < 4; ++i) dv[i] = d1[i * kTile + q];
#pragma unroll
                    for (int k = 0; k < 5; ++k) xv[k] = x1[k * kTile + q];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int k = 0; k < 5; ++k) acc1[i][k] = fmaf(dv[i], xv[k], acc1[i][k]);
#pragma unroll
                    for (int i = 0; i < 4; ++i) dv[i] = d3[i * kTile + q];
#pragma unroll
                    for (int k = 0; k < 4; ++k) xv[k] = x3[k * kTile + q];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int k = 0; k < 4; ++k) acc3[i][k] = fmaf(dv[i], xv[k], acc3[i][k]);
                    const float xh1 = x2[q];
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc2[i] = fmaf(d2[i * kTile + q], xh1, acc2[i]);
                    acc4 = fmaf(d4[q], x4[q], acc4);
                }
            }
            __syncthreads();
        }

        // ---- (vi) position gradient -------------------------------------------------------------------------------
        float dp[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) dp[d] = valid ? (a.scene.normalise ? dx[d] / a.scene.bb_ext[d] : dx[d]) : 0.f;
        if (a.d_pts && valid) {
#pragma unroll
            for (int d = 0; d < 3; ++d) a.d_pts[p * 3 + d] = dp[d];
        }
        if (a.d_rays_o) {
            // segmented sum over the lanes that belong to the same ray (points are ray-major, so segments are contiguous)
            const int64_t ray = valid ? p / a.n_samples : -1;
            float v6[6] = {dp[0], dp[1], dp[2], dp[0] * zray, dp[1] * zray, dp[2] * zray};
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int64_t r2 = __shfl_down_sync(0xffffffffu, ray, o);
                const bool take = (lane + o < 32) && (r2 == ray);
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const float t = __shfl_down_sync(0xffffffffu, v6[i], o);
                    if (take) v6[i] += t;
                }
            }
            const int64_t rprev = __shfl_up_sync(0xffffffffu, ray, 1);
            if (valid && (lane == 0 || rprev != ray)) {
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    atomicAdd(a.d_rays_o + ray * 3 + d, v6[d]);
                    atomicAdd(a.d_rays_d + ray * 3 + d, v6[3 + d]);
                }
            }
        }
    }

    if (kWgrad) {
        const int tj = tid & 7, tk = tid >> 3, ti = tid & 3, tc = tid >> 2, c4 = tid >> 5, j4 = tid & 31;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int k = 0; k < 5; ++k) atomicAdd(a.dw1 + (4 * tj + i) * kSdfIn + 5 * tk + k, acc1[i][k]);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (4 * tk + k < kColIn) atomicAdd(a.dw3 + (4 * tj + i) * kColIn + 4 * tk + k, acc3[i][k]);
            atomicAdd(a.dw2 + (4 * ti + i) * kHidden + tc, acc2[i]);
        }
        if (c4 < 3) atomicAdd(a.dw4 + c4 * kHidden + j4, acc4);
    }
}

// ------------------------------------------------------------------------------------------------------------------
// bit-exact gate helper: the positions exactly as the kernels above compute them
// ------------------------------------------------------------------------------------------------------------------
__global__ void debug_positions_kernel(PointsArgs a, float* pts_out, float* x_out) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n_points) return;
    const int64_t ray = p / a.n_samples;
    const float z = a.z[p];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float pt = ray_point(a.rays_o[ray * 3 + d], a.rays_d[ray * 3 + d], z);
        pts_out[p * 3 + d] = pt;
        x_out[p * 3 + d] = a.scene.normalise ? box_normalise(pt, a.scene.bb_lo[d], a.scene.bb_ext[d]) : pt;
    }
}

static int persistent_grid(int64_t n_tiles, int ctas_per_sm) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t g = (int64_t)sms * ctas_per_sm;
    if (g > n_tiles) g = n_tiles;
    if (g < 1) g = 1;
    return (int)g;
}

cudaError_t launch_points_fwd(const PointsArgs& a, cudaStream_t st) {
    const size_t smem = sizeof(float) * (size_t)(kWFloats + kColIn * kTile);
    cudaError_t e = cudaFuncSetAttribute(points_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int64_t n_tiles = (a.n_points + kTile - 1) / kTile;
    points_fwd_kernel<<<persistent_grid(n_tiles, 4), kTile, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_points_bwd(const PointsArgs& a, cudaStream_t st) {
    const bool wgrad = a.dw1 != nullptr;
    const int64_t n_tiles = (a.n_points + kTile - 1) / kTile;
    if (wgrad) {
        const size_t smem = sizeof(float) * (size_t)(kWFloats + (kGridFeats + kXRows + kDRows) * kTile);
        cudaError_t e = cudaFuncSetAttribute(points_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        points_bwd_kernel<true><<<persistent_grid(n_tiles, 1), kTile, smem, st>>>(a);
    } else {
        const size_t smem = sizeof(float) * (size_t)(kWFloats + kGridFeats * kTile);
        cudaError_t e = cudaFuncSetAttribute(points_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        points_bwd_kernel<false><<<persistent_grid(n_tiles, 6), kTile, smem, st>>>(a);
    }
    return cudaGetLastError();
}

cudaError_t launch_debug_positions(const PointsArgs& a, float* pts_out, float* x_out, cudaStream_t st) {
    const int threads = 256;
    const int64_t blocks = (a.n_points + threads - 1) / threads;
    debug_positions_kernel<<<(unsigned)blocks, threads, 0, st>>>(a, pts_out, x_out);
    return cudaGetLastError();
}

}  // namespace cb200
