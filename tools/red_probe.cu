// Cost model probe for the table scatter / gather of points_bwd (not on the product path; stand-alone binary):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/red_probe tools/red_probe.cu && gpurun_out/red_probe
// Every mode issues the same number of per-lane operations against a table that lives in L2 and reports lane-ops/s and the
// SM cycles one lane-op costs (SMs x clock / rate).  Question behind it (profiles/r2/l2_gather_peak.md): RED.v2.f32 alone runs
// at 187 G lane-ops/s at the kernel's pattern -- what is that a limit OF (lanes, 16-byte quanta, sectors, lines, the L2), and
// which re-arrangement of the same reductions is cheaper?
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ void red2(float2* p, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red4(float2* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red1(float* p, float a) { asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory"); }
__device__ __forceinline__ float2 ld2(const float2* p) {
    float2 v;
    asm volatile("ld.global.nc.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ld4(const float2* p) {
    float4 v;
    asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

enum Mode {
    RED2_SPREAD = 0,      // every lane its own random entry
    RED4_SPREAD,          // every lane its own random aligned PAIR of entries (16 B)
    RED2_PAIR,            // lanes 2k / 2k+1 take the two entries of one random aligned pair (the kernel's x-neighbours, x even)
    RED4_HALF,            // even lanes only: one v4 per pair (odd lanes idle)
    RED1_SPREAD,          // scalar
    RED_MIX,              // per lane and op: aligned (one v4) or not (two v2 at e, e ^ 3), 50 / 50
    LD2_SPREAD, LD4_SPREAD, LD2_PAIR,
    SMEM_ATOM,            // atomicAdd(float) on a 39 KB shared table, random entries
    RED2_PAIR_RUN2,       // RED2_PAIR with every pair index used by two adjacent lane pairs (same address twice in a warp)
    RED2_HOT,             // RED2_PAIR restricted to a 4 920-entry table (level 0)
    RED2_HOT_REPL,        // ... with 16 replicas of it selected by SM
    RED2_PAIR_ROW4,       // RED2_PAIR where a pair's 4 ops of one "level" go to 4 entries 17 / 289 apart (dense level rows)
    RED2_SPREAD_QUAD,     // lanes 4k..4k+3 take the 4 entries of one aligned 32-byte sector
    N_MODES
};
static const char* kNames[N_MODES] = {"red.v2 spread", "red.v4 spread", "red.v2 lane-pair in one 16 B", "red.v4 even lanes only",
                                      "red.f32 spread", "mix 50% v4 / 50% 2 x v2", "ld.v2 spread", "ld.v4 spread",
                                      "ld.v2 lane-pair in one 16 B", "shared atomicAdd f32 (39 KB)", "red.v2 lane-pair, runs of 2 pairs",
                                      "red.v2 lane-pair, 4 920-entry table", "... 16 replicas by SM", "red.v2 lane-pair, dense rows",
                                      "red.v2 lane-quad in one sector"};

template <int kMode>
__global__ void __launch_bounds__(1024) probe(float2* t, uint32_t mask, int iters, float* sink) {
    __shared__ float stab[9840];
    if (kMode == SMEM_ATOM) {
        for (int i = threadIdx.x; i < 9840; i += 1024) stab[i] = 0.f;
        __syncthreads();
    }
    const uint32_t gtid = blockIdx.x * 1024 + threadIdx.x;
    const uint32_t xb = threadIdx.x & 1;
    uint32_t smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    float acc = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const uint32_t seq = (uint32_t)(it * 16 + j);
            const uint32_t r_lane = mix(gtid * 0x9e3779b1u + seq);
            const uint32_t r_pair = mix((gtid >> 1) * 0x9e3779b1u + seq);
            if (kMode == RED2_SPREAD) red2(t + (r_lane & mask), 1e-12f, 1e-12f);
            if (kMode == RED4_SPREAD) red4(t + (r_lane & mask & ~1u), 1e-12f, 1e-12f, 1e-12f, 1e-12f);
            if (kMode == RED2_PAIR) red2(t + ((r_pair & mask & ~1u) | xb), 1e-12f, 1e-12f);
            if (kMode == RED4_HALF) { if (!xb) red4(t + (r_pair & mask & ~1u), 1e-12f, 1e-12f, 1e-12f, 1e-12f); }
            if (kMode == RED1_SPREAD) red1(reinterpret_cast<float*>(t) + (r_lane & (2 * mask + 1)), 1e-12f);
            if (kMode == RED_MIX) {
                const uint32_t e = r_lane & mask;
                if (r_lane >> 31) red4(t + (e & ~1u), 1e-12f, 1e-12f, 1e-12f, 1e-12f);
                else { red2(t + e, 1e-12f, 1e-12f); red2(t + (e ^ 3u), 1e-12f, 1e-12f); }
            }
            if (kMode == LD2_SPREAD) { const float2 v = ld2(t + (r_lane & mask)); acc += v.x + v.y; }
            if (kMode == LD4_SPREAD) { const float4 v = ld4(t + (r_lane & mask & ~1u)); acc += v.x + v.y + v.z + v.w; }
            if (kMode == LD2_PAIR) { const float2 v = ld2(t + ((r_pair & mask & ~1u) | xb)); acc += v.x + v.y; }
            if (kMode == SMEM_ATOM) atomicAdd(&stab[r_lane % 9840u], 1e-12f);
            if (kMode == RED2_PAIR_RUN2) red2(t + ((mix((gtid >> 2) * 0x9e3779b1u + seq) & mask & ~1u) | xb), 1e-12f, 1e-12f);
            if (kMode == RED2_HOT) red2(t + (((r_pair % 4920u) & ~1u) | xb), 1e-12f, 1e-12f);
            if (kMode == RED2_HOT_REPL) red2(t + (smid & 15u) * 4920u + (((r_pair % 4920u) & ~1u) | xb), 1e-12f, 1e-12f);
            if (kMode == RED2_PAIR_ROW4) {
                const uint32_t base = mix((gtid >> 1) * 0x9e3779b1u + (seq >> 2)) % 4500u;
                const uint32_t e = base + ((seq & 1u) ? 17u : 0u) + ((seq & 2u) ? 289u : 0u) + xb;
                red2(t + e, 1e-12f, 1e-12f);
            }
            if (kMode == RED2_SPREAD_QUAD) red2(t + ((mix((gtid >> 2) * 0x9e3779b1u + seq) & mask & ~3u) | (threadIdx.x & 3u)), 1e-12f, 1e-12f);
        }
    }
    if (kMode == SMEM_ATOM) {
        __syncthreads();
        acc = stab[threadIdx.x];
    }
    if (acc == 1.2345e30f) *sink = acc;
}

template <int kMode>
static void run(float2* t, uint32_t mask, float* sink, int blocks, int sms, double clk_ghz) {
    const int iters = 64;
    probe<kMode><<<blocks, 1024>>>(t, mask, iters, sink);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0));
    const int reps = 5;
    for (int r = 0; r < reps; ++r) probe<kMode><<<blocks, 1024>>>(t, mask, iters, sink);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= reps;
    // "slots" = (lane, j) pairs; RED4_HALF performs half as many instructions but covers the same entries as RED2_PAIR
    const double slots = (double)blocks * 1024 * iters * 16;
    const double rate = slots / (ms * 1e-3);
    const int active_sms = blocks < sms ? blocks : sms;
    printf("%-40s blocks %5d  %8.3f ms  %7.1f G slots/s  %6.3f SM-cycles per slot\n", kNames[kMode], blocks, ms, rate / 1e9,
           active_sms * clk_ghz * 1e9 / rate);
}

int main() {
    int sms = 0, khz = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    const double clk = khz / 1e6;
    printf("SMs %d, clock %.3f GHz\n", sms, clk);
    const uint32_t entries = 1u << 20;      // 8 MB of float2: L2-resident like the 6.7 MB room0 table
    float2* t;
    float* sink;
    CK(cudaMalloc(&t, (size_t)entries * sizeof(float2)));
    CK(cudaMemset(t, 0, (size_t)entries * sizeof(float2)));
    CK(cudaMalloc(&sink, 4));
    const uint32_t mask = entries - 1;
    for (int frac = 1; frac <= 4; frac *= 2) {     // one 1024-thread CTA per SM on all / half / a quarter of the SMs: does the rate follow the SM count (SM-side limit) or not (L2)?
        const int blocks = sms / frac;
        printf("--- %d CTAs of 1024 threads, one per SM\n", blocks);
        run<RED2_SPREAD>(t, mask, sink, blocks, sms, clk);
        run<RED4_SPREAD>(t, mask, sink, blocks, sms, clk);
        run<RED2_PAIR>(t, mask, sink, blocks, sms, clk);
        run<RED4_HALF>(t, mask, sink, blocks, sms, clk);
        run<RED1_SPREAD>(t, mask, sink, blocks, sms, clk);
        run<RED_MIX>(t, mask, sink, blocks, sms, clk);
        run<RED2_PAIR_RUN2>(t, mask, sink, blocks, sms, clk);
        run<RED2_SPREAD_QUAD>(t, mask, sink, blocks, sms, clk);
        run<RED2_HOT>(t, mask, sink, blocks, sms, clk);
        run<RED2_HOT_REPL>(t, mask, sink, blocks, sms, clk);
        run<RED2_PAIR_ROW4>(t, mask, sink, blocks, sms, clk);
        run<LD2_SPREAD>(t, mask, sink, blocks, sms, clk);
        run<LD4_SPREAD>(t, mask, sink, blocks, sms, clk);
        run<LD2_PAIR>(t, mask, sink, blocks, sms, clk);
        run<SMEM_ATOM>(t, mask, sink, blocks, sms, clk);
    }
    return 0;
}
