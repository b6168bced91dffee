// Host routine: the fp32 stream of torch's default CPU generator, bit for bit, without torch.
//
// The reference draws its stratified jitter with `torch.rand(z_vals.shape)` on the CPU generator and uploads it every forward
// (model/scene_rep.py:255); the drop-in's default noise mode does the same so that both consume the SAME stream.  ATen produces
// that tensor one element at a time through a virtual generator call (~3.5-12 ns per float: 0.4 ms for the 99 072 floats of a
// bundle-adjustment iteration, more than the GPU needs for the whole iteration).  This is the same arithmetic in tight, vectorised loops:
//   engine        at::mt19937 (ATen/core/MT19937RNGEngine.h): `if (--left == 0) next_state(); y = state[next++]; temper(y)`
//   distribution  at::uniform_real_distribution<float> (ATen/core/TransformationHelper.h): (y & (2^24 - 1)) * 2^-24
//   kernel        cpu_serial_kernel over the contiguous output (ATen/native/cpu/DistributionTemplates.h): element order
// on the caller's copy of the generator state (`torch.get_rng_state()`, layout below), which the caller writes back with
// `torch.set_rng_state()`, so every later draw of the process is unchanged.  tests/test_oracle.py pins it against torch.rand.
#include <stdint.h>
#include <string.h>

namespace {

constexpr int kN = 624, kM = 397;

// CPUGeneratorImplStateLegacy (ATen/CPUGeneratorImpl.cpp), the first 5 048 of the 5 056 bytes torch.get_rng_state() returns
struct TorchCpuState {
    uint64_t the_initial_seed;
    int32_t left;
    int32_t seeded;
    uint64_t next;
    uint64_t state[kN];      // the 32-bit state words, one per 64-bit slot
    double normal_x, normal_y, normal_rho;
    int32_t normal_is_valid;
    int32_t pad;
};
static_assert(sizeof(TorchCpuState) == 5048, "layout of torch's legacy CPU generator state");

inline uint32_t twist(uint32_t u, uint32_t v) { return (((u & 0x80000000u) | (v & 0x7fffffffu)) >> 1) ^ ((v & 1u) ? 0x9908b0dfu : 0u); }

// Both loops below are plain element-wise integer arithmetic; the compiler's vector clones are picked once at load time by the
// CPU's capabilities (ifunc), the results are the same integers either way: 1.5 ns per float with SSE2, 0.7 with AVX2 / AVX-512.
#define CB200_HOST_CLONES __attribute__((target_clones("avx512f", "avx2", "default")))

CB200_HOST_CLONES
void next_state(uint32_t* s) {
    // at::mt19937::next_state; neither loop carries a dependency shorter than 227 elements
    int p = 0;
#pragma GCC ivdep
    for (; p < kN - kM; ++p) s[p] = s[p + kM] ^ twist(s[p], s[p + 1]);
#pragma GCC ivdep
    for (; p < kN - 1; ++p) s[p] = s[p + kM - kN] ^ twist(s[p], s[p + 1]);
    s[kN - 1] = s[kM - 1] ^ twist(s[kN - 1], s[0]);
}

inline float to_unit(uint32_t y) {
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return (float)(y & 0xffffffu) * (1.0f / 16777216.0f);      // exact: a 24-bit integer times 2^-24
}

CB200_HOST_CLONES
void temper_block(const uint32_t* src, float* dst, int64_t n) {
    for (int64_t i = 0; i < n; ++i) dst[i] = to_unit(src[i]);
}

}  // namespace

// state: the 5 056 bytes of torch.get_rng_state() (read and updated); out: n floats.  Returns 0, or -1 for a state this routine
// does not understand (not seeded, counters out of range).
extern "C" int cb200_host_torch_rand(uint8_t* state, int64_t n, float* out) {
    if (!state || n < 0 || (n > 0 && !out)) return -1;
    TorchCpuState st;
    memcpy(&st, state, sizeof(st));
    if (!st.seeded || st.left < 1 || st.left > kN || st.next > (uint64_t)kN) return -1;
    // between two outputs the engine keeps next == N - left + 1 (k outputs into a block: next = k, left = N - k + 1); a freshly seeded one has left 1, next 0
    if (!(st.left == 1 && st.next == 0) && (uint64_t)(kN - st.left + 1) != st.next) return -1;
    uint32_t s[kN + 1];
    for (int i = 0; i < kN; ++i) s[i] = (uint32_t)st.state[i];
    int64_t left = st.left, next = (int64_t)st.next;
    int64_t done = 0;
    while (done < n) {
        // the engine's `if (--left == 0) next_state()` happens in front of every output: `avail` outputs remain in this block
        int64_t avail = left - 1;
        if (avail == 0) {
            next_state(s);
            left = kN + 1;      // as if the decrement for the first output of the new block were still to come
            next = 0;
            avail = kN;
        }
        const int64_t take = avail < n - done ? avail : n - done;
        const uint32_t* src = s + next;
        float* dst = out + done;
        temper_block(src, dst, take);
        next += take;
        left -= take;
        done += take;
    }
    // engine invariants after k outputs of a block: left = N - k (k >= 1), next = k; `left` above counts one too many after a regeneration
    if (left == kN + 1) left = 1;      // unreachable (a regeneration is always followed by >= 1 output); kept for clarity
    st.left = (int32_t)left;
    st.next = (uint64_t)next;
    for (int i = 0; i < kN; ++i) st.state[i] = (uint64_t)s[i];
    memcpy(state, &st, sizeof(st));
    return 0;
}
