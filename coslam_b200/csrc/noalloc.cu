// cb200_points_bwd_noalloc: the fused per-point backward for callers that cannot allocate device memory (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the
// reference.
//
// cb200_points_bwd needs scratch in ray mode: the live-point compaction in front of points_bwd_tc_kernel writes a counter
// and a list of up to n_rays * n_samples point indices (cb200_points_bwd_workspace bytes).  Here that scratch is a
// file-scope device array of a fixed, modest size; every GPU that loads the library holds it (4 MiB + 256 B) until exit.
//
//   * `g_scratch` below is, in HOST code, only a stand-in for the symbol: its device address is fetched once per device
//     with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels receive.
//   * All calls share the array and it keeps the previous call's contents: the counter is cleared on the caller's stream
//     at the start of every call (launch_compact_live), the list is written before it is read, up to the new counter.
//   * Calls that overlap on the GPU (two streams, two host threads on one device) RACE on the array: serialise them.
#include <mutex>

#include "args.cuh"

namespace {

constexpr size_t kScratchBytes = 256 + sizeof(uint32_t) * (size_t)CB200_NOALLOC_MAX_POINTS;
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
void* g_scratch_ptr[kMaxDevices] = {};
std::mutex g_scratch_mutex;

// device address of g_scratch on the CURRENT device (looked up once per device); nullptr + *err on failure
void* scratch_for_current_device(cudaError_t* err, int* dev_out) {
    int dev = 0;
    *err = cudaGetDevice(&dev);
    *dev_out = dev;
    if (*err != cudaSuccess) return nullptr;
    if (dev < 0 || dev >= kMaxDevices) {
        *err = cudaErrorInvalidDevice;
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        *err = cudaGetSymbolAddress(&p, g_scratch);
        if (*err != cudaSuccess) return nullptr;
        g_scratch_ptr[dev] = p;
    }
    return g_scratch_ptr[dev];
}

}  // namespace

extern "C" {

int64_t cb200_points_bwd_noalloc_max_points(void) { return (int64_t)CB200_NOALLOC_MAX_POINTS; }

int cb200_points_bwd_noalloc(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* rays_o, const float* rays_d,
                             const float* z_vals, int64_t n_rays, int64_t n_samples, const float* d_raw, const uint32_t* act_mask,
                             float* dtable, const cb200_mlp_grad* d_mlp, float* d_rays_o, float* d_rays_d, int sdf_only, void* stream) {
    if (!(rays_o && rays_d && z_vals)) return cb200_set_error(-1, "points_bwd_noalloc: serves ray mode (rays_o, rays_d, z_vals must be given)");
    if (n_rays < 0 || n_samples < 1) return cb200_set_error(-1, "points_bwd_noalloc: bad sizes");
    if (n_rays > (int64_t)CB200_NOALLOC_MAX_POINTS / n_samples)
        return cb200_set_error(-1, "points_bwd_noalloc: %lld rays x %lld samples exceed the static scratch (at most %lld points per call); "
                                   "use cb200_points_bwd with a caller-owned workspace", (long long)n_rays, (long long)n_samples,
                               (long long)CB200_NOALLOC_MAX_POINTS);
    cudaError_t err = cudaSuccess;
    int dev = 0;
    void* scratch = scratch_for_current_device(&err, &dev);
    if (scratch == nullptr)
        return cb200_set_error((int)err, "points_bwd_noalloc: cudaGetSymbolAddress of the static scratch failed on device %d: %s", dev,
                               cudaGetErrorString(err));
    return cb200_points_bwd(scene, table, mlp, nullptr, rays_o, rays_d, z_vals, n_rays, n_samples, d_raw, nullptr, act_mask, dtable, d_mlp,
                            nullptr, d_rays_o, d_rays_d, sdf_only, scratch, kScratchBytes, stream);
}

}  // extern "C"
