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
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

int cb200_set_error(int code, const char* fmt, ...);

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
// The engine loop shared by the entry points below.  out == nullptr: advance the state as if the n floats had been drawn (the twist
// alone: a fifth of the time).  ckpt != nullptr: additionally record the 624 state words in front of regeneration 0, C, 2C, ... (the
// starting points of the device-side segments); *head = outputs served by the block that was current at entry, *head_next = its read index.
static int host_torch_rand(uint8_t* state, int64_t n, float* out, uint32_t* ckpt = nullptr, int64_t C = 0, int64_t cap = 0,
                           int64_t* n_seg = nullptr, int64_t* head = nullptr, int32_t* head_next = nullptr) {
    if (!state || n < 0) return -1;
    TorchCpuState st;
    memcpy(&st, state, sizeof(st));
    if (!st.seeded || st.left < 1 || st.left > kN || st.next > (uint64_t)kN) return -1;
    // between two outputs the engine keeps next == N - left + 1 (k outputs into a block: next = k, left = N - k + 1); a freshly seeded one has left 1, next 0
    if (!(st.left == 1 && st.next == 0) && (uint64_t)(kN - st.left + 1) != st.next) return -1;
    uint32_t s[kN + 1];
    for (int i = 0; i < kN; ++i) s[i] = (uint32_t)st.state[i];
    int64_t left = st.left, next = (int64_t)st.next;
    int64_t done = 0, regen = 0;
    if (head) *head = (left - 1) < n ? (left - 1) : n;
    if (head_next) *head_next = (int32_t)next;
    if (ckpt) {
        if (C < 1 || cap < 1) return -1;
        memcpy(ckpt, s, kN * sizeof(uint32_t));      // segment 0 starts from the state at entry (it also serves the head)
    }
    while (done < n) {
        // the engine's `if (--left == 0) next_state()` happens in front of every output: `avail` outputs remain in this block
        int64_t avail = left - 1;
        if (avail == 0) {
            if (ckpt && regen % C == 0 && regen > 0) {
                if (regen / C >= cap) return -2;
                memcpy(ckpt + (regen / C) * kN, s, kN * sizeof(uint32_t));
            }
            ++regen;
            next_state(s);
            left = kN + 1;      // as if the decrement for the first output of the new block were still to come
            next = 0;
            avail = kN;
        }
        const int64_t take = avail < n - done ? avail : n - done;
        if (out) temper_block(s + next, out + done, take);
        next += take;
        left -= take;
        done += take;
    }
    if (n_seg) *n_seg = regen == 0 ? 1 : (regen + C - 1) / C;
    st.left = (int32_t)left;
    st.next = (uint64_t)next;
    for (int i = 0; i < kN; ++i) st.state[i] = (uint64_t)s[i];
    memcpy(state, &st, sizeof(st));
    return 0;
}

extern "C" int cb200_host_torch_rand(uint8_t* state, int64_t n, float* out) {
    if (n > 0 && !out) return -1;
    return host_torch_rand(state, n, out);
}

extern "C" int cb200_host_torch_rand_skip(uint8_t* state, int64_t n) { return host_torch_rand(state, n, nullptr); }

extern "C" int cb200_host_torch_rand_segments(uint8_t* state, int64_t n, int64_t blocks_per_segment, uint32_t* checkpoints, int64_t capacity,
                                              int64_t* n_segments, int64_t* head, int32_t* head_next) {
    if (!checkpoints || !n_segments || !head || !head_next) return -1;
    return host_torch_rand(state, n, nullptr, checkpoints, blocks_per_segment, capacity, n_segments, head, head_next);
}

// ------------------------------------------------------------------------------------------------------------------
// The same stream produced ON THE DEVICE (large batches): the values never cross PCIe.  The Mersenne-Twister recurrence
// x[k + 624] = x[k + 397] ^ twist(x[k], x[k + 1]) is sequential (227 words wide at most), so the host still walks it -- the twist
// alone, 0.11 ns per word -- and records the state every `blocks_per_segment` blocks; one CTA per segment then regenerates its
// blocks from that checkpoint (three barrier-separated phases per block: 227 | 227 | 170 words), every thread tempering and
// storing the word it has just produced.  A 2^22-ray batch (1.8e8 floats): ~20 ms of host time, against 97 ms of host generation
// plus a 720 MB upload.
// ------------------------------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ uint32_t twist_dev(uint32_t u, uint32_t v) {
    return (((u & 0x80000000u) | (v & 0x7fffffffu)) >> 1) ^ ((v & 1u) ? 0x9908b0dfu : 0u);
}
__device__ __forceinline__ float to_unit_dev(uint32_t y) {
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return (float)(y & 0xffffffu) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(256) torch_rand_segments_kernel(const uint32_t* __restrict__ ckpt, int64_t C, int64_t head, int head_next,
                                                                  int64_t n, float* __restrict__ out_all) {
    __shared__ uint32_t buf[2][kN];
    const int tid = threadIdx.x;
    const int64_t seg = blockIdx.x;
    for (int i = tid; i < kN; i += 256) buf[0][i] = ckpt[seg * kN + i];
    __syncthreads();
    if (seg == 0) {                                            // what was left of the block that was current at entry
        for (int64_t i = tid; i < head; i += 256) out_all[i] = to_unit_dev(buf[0][head_next + i]);
    }
    int cur = 0;
    for (int64_t i = 0; i < C; ++i) {
        const int64_t base = head + (seg * C + i) * kN;
        if (base >= n) break;
        const uint32_t* A = buf[cur];
        uint32_t* B = buf[cur ^ 1];
        const int64_t take = (n - base) < kN ? (n - base) : kN;
        float* out = out_all + base;
        if (tid < kN - kM) {                                   // words 0 .. 226: old words only
            const uint32_t v = A[tid + kM] ^ twist_dev(A[tid], A[tid + 1]);
            B[tid] = v;
            if (tid < take) out[tid] = to_unit_dev(v);
        }
        __syncthreads();
        if (tid < kN - kM) {                                   // words 227 .. 453: new words 0 .. 226
            const int p = tid + (kN - kM);
            const uint32_t v = B[p - (kN - kM)] ^ twist_dev(A[p], A[p + 1]);
            B[p] = v;
            if (p < take) out[p] = to_unit_dev(v);
        }
        __syncthreads();
        if (tid < kN - 2 * (kN - kM)) {                        // words 454 .. 623: new words 227 .. 396 (and new word 0 for the last)
            const int p = tid + 2 * (kN - kM);
            const uint32_t v = B[p - (kN - kM)] ^ twist_dev(A[p], p == kN - 1 ? B[0] : A[p + 1]);
            B[p] = v;
            if (p < take) out[p] = to_unit_dev(v);
        }
        __syncthreads();
        cur ^= 1;
    }
}

}  // namespace

// checkpoints: DEVICE copy of what cb200_host_torch_rand_segments wrote; out: n floats on the device.  One launch on `stream`.
extern "C" int cb200_torch_rand_device(const uint32_t* checkpoints, int64_t n_segments, int64_t blocks_per_segment, int64_t head,
                                       int32_t head_next, int64_t n, float* out, void* stream) {
    if (!checkpoints || n < 0 || (n > 0 && !out) || n_segments < 1 || blocks_per_segment < 1 || head < 0 || head > n || head_next < 0 ||
        head_next + head > kN)
        return cb200_set_error(-1, "torch_rand_device: bad arguments");
    if (n == 0) return 0;
    torch_rand_segments_kernel<<<(unsigned)n_segments, 256, 0, (cudaStream_t)stream>>>(checkpoints, blocks_per_segment, head, head_next, n, out);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : cb200_set_error((int)e, "torch_rand_device: %s", cudaGetErrorString(e));
}
