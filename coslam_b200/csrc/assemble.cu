// Ray assembly for mapping / bundle adjustment (coslam.py:336-355, model/keyframe.py:96-106): gather the sampled 7-float
// keyframe rays (camera-frame direction, rgb, depth) by index, gather each ray's pose by keyframe slot (negative slot =
// counted from the end, slot -1 = current frame, like the Python indexing of `poses_all[ids_all]`), and rotate:
//   rays_d = sum_k dir_cam[k] * R[:, k]   (same summation order as torch.sum(dir[..., None, :] * R, -1)),  rays_o = t.
// Backward: dL/dposes[slot][:3, :3] += d_rays_d (outer) dir_cam,  dL/dposes[slot][:3, 3] += d_rays_o  -- a segment sum by
// keyframe slot, accumulated per CTA in shared memory (few keyframes, thousands of rays) and flushed once.
#include "args.cuh"

namespace cb200 {

struct AssembleArgs {
    const float* kf_rays;        // [M,7]
    const long long* ray_idx;    // [N] rows of kf_rays
    const long long* pose_slot;  // [N]
    const float* poses;          // [K,4,4] row-major
    long long n, n_poses;
    float* rays_o;
    float* rays_d;
    float* target_rgb;
    float* target_d;
    const float* d_rays_o;       // backward
    const float* d_rays_d;
    float* d_poses;              // [K,4,4] +=
    // optional second source: rays i >= n_first come from (kf_rays2, ray_idx2[i - n_first], pose_slot2[i - n_first]) -- the
    // current frame's rays behind the keyframe rays of a bundle-adjustment batch (coslam.py:339-344) in ONE launch
    long long n_first;
    const float* kf_rays2;
    const long long* ray_idx2;
    const long long* pose_slot2;
};

__device__ __forceinline__ const float* assemble_source(const AssembleArgs& a, long long i, long long& slot) {
    const bool second = i >= a.n_first;
    const long long j = second ? i - a.n_first : i;
    slot = (second ? a.pose_slot2 : a.pose_slot)[j];
    if (slot < 0) slot += a.n_poses;
    return (second ? a.kf_rays2 : a.kf_rays) + (second ? a.ray_idx2 : a.ray_idx)[j] * 7;
}

__global__ void __launch_bounds__(256) assemble_rays_fwd_kernel(AssembleArgs a) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    long long slot;
    const float* r = assemble_source(a, i, slot);
    const float* T = a.poses + slot * 16;
    const float d0 = r[0], d1 = r[1], d2 = r[2];
#pragma unroll
    for (int row = 0; row < 3; ++row) {
        // products rounded separately, summed left to right: what torch.sum over the last axis of 3 does
        a.rays_d[i * 3 + row] = __fadd_rn(__fadd_rn(__fmul_rn(d0, T[row * 4 + 0]), __fmul_rn(d1, T[row * 4 + 1])), __fmul_rn(d2, T[row * 4 + 2]));
        a.rays_o[i * 3 + row] = T[row * 4 + 3];
        a.target_rgb[i * 3 + row] = r[3 + row];
    }
    a.target_d[i] = r[6];
}

constexpr int kMaxSharedPoses = 1024;     // 1024 x 12 floats = 48 KB

// One ray's contribution is 12 sums; consecutive rays usually share their keyframe (tracking: ONE pose for the whole batch),
// so a thread first accumulates its own rays in registers while the slot does not change, then a warp whose lanes all ended
// on the same slot combines them with shuffles and issues one set of 12 atomics (a batch of 1024 same-slot rays would
// otherwise serialise 12 288 atomics on 12 addresses).
__global__ void __launch_bounds__(256) assemble_rays_bwd_kernel(AssembleArgs a, int use_shared) {
    extern __shared__ float acc[];         // [n_poses][12] when use_shared
    if (use_shared) {
        for (long long k = threadIdx.x; k < a.n_poses * 12; k += blockDim.x) acc[k] = 0.f;
        __syncthreads();
    }
    float v[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) v[k] = 0.f;
    long long cur = -1;
    auto emit = [&](long long slot, const float (&w)[12]) {
        float* dst = use_shared ? acc + slot * 12 : a.d_poses + slot * 16;
#pragma unroll
        for (int k = 0; k < 12; ++k) atomicAdd(dst + k, w[k]);           // rows of 4: [slot][row][0..3] is contiguous in both layouts
    };
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (long long)gridDim.x * blockDim.x) {
        long long slot;
        const float* r = assemble_source(a, i, slot);
        if (slot != cur) {
            if (cur >= 0) emit(cur, v);
#pragma unroll
            for (int k = 0; k < 12; ++k) v[k] = 0.f;
            cur = slot;
        }
        const float d[3] = {r[0], r[1], r[2]};
#pragma unroll
        for (int row = 0; row < 3; ++row) {
            const float gd = a.d_rays_d[i * 3 + row], go = a.d_rays_o[i * 3 + row];
            v[row * 4 + 0] = fmaf(gd, d[0], v[row * 4 + 0]);
            v[row * 4 + 1] = fmaf(gd, d[1], v[row * 4 + 1]);
            v[row * 4 + 2] = fmaf(gd, d[2], v[row * 4 + 2]);
            v[row * 4 + 3] += go;
        }
    }
    // lanes without rays (cur < 0) hold zeros and may join any group
    const long long lead = __shfl_sync(0xffffffffu, cur, __ffs(__ballot_sync(0xffffffffu, cur >= 0) | 0x80000000u) - 1);
    const bool uniform = __all_sync(0xffffffffu, cur < 0 || cur == lead);
    if (uniform) {
#pragma unroll
        for (int k = 0; k < 12; ++k) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        }
        if ((threadIdx.x & 31) == 0 && lead >= 0) emit(lead, v);
    } else if (cur >= 0) {
        emit(cur, v);
    }
    if (use_shared) {
        __syncthreads();
        for (long long k = threadIdx.x; k < a.n_poses * 12; k += blockDim.x) {
            const float w = acc[k];
            if (w != 0.f) atomicAdd(a.d_poses + (k / 12) * 16 + (k % 12), w);
        }
    }
}

static cudaError_t run_assemble_fwd(const AssembleArgs& a, cudaStream_t st) {
    if (a.n <= 0) return cudaSuccess;
    assemble_rays_fwd_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, st>>>(a);
    return cudaGetLastError();
}

static cudaError_t run_assemble_bwd(const AssembleArgs& a, cudaStream_t st) {
    if (a.n <= 0) return cudaSuccess;
    const int use_shared = a.n_poses <= kMaxSharedPoses;
    const size_t smem = use_shared ? sizeof(float) * 12 * (size_t)a.n_poses : 0;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(assemble_rays_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    long long blocks = (a.n + 255) / 256;                      // one ray per thread up to 4 CTAs per SM, grid-stride beyond
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 4) blocks = 148 * 4;
    assemble_rays_bwd_kernel<<<(unsigned)blocks, 256, smem, st>>>(a, use_shared);
    return cudaGetLastError();
}

cudaError_t launch_assemble_rays_fwd(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, const float* poses,
                                     long long n, long long n_poses, float* rays_o, float* rays_d, float* target_rgb, float* target_d,
                                     cudaStream_t st) {
    AssembleArgs a{kf_rays, ray_idx, pose_slot, poses, n, n_poses, rays_o, rays_d, target_rgb, target_d, nullptr, nullptr, nullptr,
                   n, nullptr, nullptr, nullptr};
    return run_assemble_fwd(a, st);
}

cudaError_t launch_assemble_rays_bwd(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, const float* d_rays_o,
                                     const float* d_rays_d, long long n, long long n_poses, float* d_poses, cudaStream_t st) {
    AssembleArgs a{kf_rays, ray_idx, pose_slot, nullptr, n, n_poses, nullptr, nullptr, nullptr, nullptr, d_rays_o, d_rays_d, d_poses,
                   n, nullptr, nullptr, nullptr};
    return run_assemble_bwd(a, st);
}

// keyframe rays [0, n_first) and current-frame rays [n_first, n) of one BA batch in one launch (csrc/runtime.cu)
cudaError_t launch_assemble_rays_fwd2(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, long long n_first,
                                      const float* cur_rays, const long long* idx_cur, const long long* slot_cur, long long n_cur,
                                      const float* poses, long long n_poses, float* rays_o, float* rays_d, float* target_rgb,
                                      float* target_d, cudaStream_t st) {
    AssembleArgs a{kf_rays, ray_idx, pose_slot, poses, n_first + n_cur, n_poses, rays_o, rays_d, target_rgb, target_d, nullptr, nullptr,
                   nullptr, n_first, cur_rays, idx_cur, slot_cur};
    return run_assemble_fwd(a, st);
}

cudaError_t launch_assemble_rays_bwd2(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, long long n_first,
                                      const float* cur_rays, const long long* idx_cur, const long long* slot_cur, long long n_cur,
                                      const float* d_rays_o, const float* d_rays_d, long long n_poses, float* d_poses, cudaStream_t st) {
    AssembleArgs a{kf_rays, ray_idx, pose_slot, nullptr, n_first + n_cur, n_poses, nullptr, nullptr, nullptr, nullptr, d_rays_o, d_rays_d,
                   d_poses, n_first, cur_rays, idx_cur, slot_cur};
    return run_assemble_bwd(a, st);
}

}  // namespace cb200
