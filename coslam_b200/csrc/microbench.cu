// Measurement hook (not on the product path): how fast can this GPU gather / reduce float2 table entries at the fused
// kernels' own access pattern when NOTHING else is done?  The table (6.7 - 34.7 MB) lives in L2, so the HBM roofline the
// bench line reports is not the physical ceiling of points_fwd / points_bwd; this kernel measures the one that is
// ("report against measured L2 gather bandwidth instead -- state both").
//
// Pattern = points_tc.cu's: a LANE PAIR per point, thread xb owns the four (y, z) corners at x = cell.x + xb of each of the
// 16 levels, i.e. 64 independent LDG.64 (and / or 64 RED.v2.f32) per thread and point.  Corner indices are NOT computed
// here: they are read (coalesced, 32 B per point and level) from the [P, 16, 8] list cb200_hash_indices wrote for real
// sample points, so the lines touched are exactly the product kernels' and the only arithmetic is the level offset add.
#include "args.cuh"

namespace cb200 {

// level_mask: levels that take part (bit l); replicas (power of two) > 1: the reductions of levels below n_repl_levels go to copy
// (smid % replicas) of those levels' rows, stored behind the table gradient (dtable + offset[n_levels] + r * offset[n_repl_levels])
template <bool kGather, bool kScatter>
__global__ void __launch_bounds__(256) gather_peak_kernel(SceneDev sc, const uint32_t* __restrict__ idx, int64_t n_points,
                                                          const float2* __restrict__ table, float2* dtable, float* sink, uint32_t level_mask,
                                                          uint32_t replicas, uint32_t n_repl_levels) {
    const int xb = threadIdx.x & 1;
    float acc = 0.f;
    uint32_t smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    const uint32_t repl_off = sc.offset[kLevels] + (smid & (replicas - 1u)) * sc.offset[n_repl_levels];
    for (int64_t p = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 1; p < n_points; p += (int64_t)gridDim.x * 128) {
        const uint4* row = reinterpret_cast<const uint4*>(idx + p * (int64_t)(kLevels * 8));
#pragma unroll
        for (int l = 0; l < kLevels; ++l) {
            if (!((level_mask >> l) & 1u)) continue;
            // corners c = 0..7 with x bit = c & 1: this thread takes c = xb, xb + 2, xb + 4, xb + 6
            const uint4 lo = __ldg(row + 2 * l), hi = __ldg(row + 2 * l + 1);
            const uint32_t e[4] = {xb ? lo.y : lo.x, xb ? lo.w : lo.z, xb ? hi.y : hi.x, xb ? hi.w : hi.z};
            const uint32_t off = sc.offset[l];
            const uint32_t doff = (replicas > 1u && (uint32_t)l < n_repl_levels) ? repl_off + off : off;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (kGather) {
                    const float2 v = ldg_f2(table + off + e[j]);
                    acc += v.x + v.y;
                }
                if (kScatter) red_add_f2(dtable + doff + e[j], 1e-12f, 1e-12f);
            }
        }
    }
    if (kGather && acc == 1.2345e30f) *sink = acc;      // keeps the loads alive
}

cudaError_t launch_gather_peak(const SceneDev& sc, const uint32_t* idx, int64_t n_points, const float* table, float* dtable, int mode,
                               float* sink, cudaStream_t st) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t blocks = (n_points + 127) / 128;
    if (blocks > (int64_t)sms * 8) blocks = (int64_t)sms * 8;
    if (blocks < 1) return cudaSuccess;
    const float2* t = reinterpret_cast<const float2*>(table);
    float2* dt = reinterpret_cast<float2*>(dtable);
    // mode: bits 0-1 gather / reduce; bits 4-19 level mask (0 = all levels); bits 20-23 log2(replicas); bits 24-28 replicated levels
    uint32_t level_mask = ((uint32_t)mode >> 4) & 0xffffu;
    if (!level_mask) level_mask = 0xffffu;
    const uint32_t replicas = 1u << (((uint32_t)mode >> 20) & 15u), n_repl = ((uint32_t)mode >> 24) & 31u;
    mode &= 3;
    if (mode == 1) gather_peak_kernel<true, false><<<(unsigned)blocks, 256, 0, st>>>(sc, idx, n_points, t, dt, sink, level_mask, replicas, n_repl);
    else if (mode == 2) gather_peak_kernel<false, true><<<(unsigned)blocks, 256, 0, st>>>(sc, idx, n_points, t, dt, sink, level_mask, replicas, n_repl);
    else gather_peak_kernel<true, true><<<(unsigned)blocks, 256, 0, st>>>(sc, idx, n_points, t, dt, sink, level_mask, replicas, n_repl);
    return cudaGetLastError();
}

}  // namespace cb200
