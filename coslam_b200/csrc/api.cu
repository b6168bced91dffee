// extern "C" surface of libcoslam_b200.so (include/coslam_b200.h): argument validation, host-side grid geometry and
// kernel launches.  No allocation, no synchronisation, no default-stream work.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>

#include "args.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_result(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    return fail((int)e, "%s: %s", what, cudaGetErrorString(e));
}

#define CB_REQUIRE(cond, ...)                  \
    do {                                       \
        if (!(cond)) return fail(-1, __VA_ARGS__); \
    } while (0)

int g_decoder = 1;      // 0 = SIMT decoder (points.cu), 1 = tcgen05 decoder (points_tc.cu)

int check_grid(const cb200_grid* g) {
    CB_REQUIRE(g != nullptr, "grid is NULL");
    CB_REQUIRE(g->n_levels >= 1 && g->n_levels <= CB200_MAX_LEVELS, "n_levels %u not in [1,%d]", g->n_levels, CB200_MAX_LEVELS);
    CB_REQUIRE(g->n_features == 2, "n_features_per_level %u unsupported (kernels are written for 2)", g->n_features);
    return 0;
}

int check_fused_scene(const cb200_scene* s) {
    CB_REQUIRE(s != nullptr, "scene is NULL");
    if (int rc = check_grid(&s->grid)) return rc;
    CB_REQUIRE(s->grid.n_levels == CB200_MAX_LEVELS, "fused decoder expects %d grid levels (got %u)", CB200_MAX_LEVELS, s->grid.n_levels);
    CB_REQUIRE(s->n_bins == 16, "fused decoder expects OneBlob n_bins == 16 (got %u)", s->n_bins);
    return 0;
}

}  // namespace

// error reporting for entry points that live in other translation units (noalloc.cu)
int cb200_set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

extern "C" {

int cb200_abi_version(void) { return CB200_ABI_VERSION; }

const char* cb200_last_error(void) { return g_err; }

int cb200_set_decoder(int kind) {
    CB_REQUIRE(kind == 0 || kind == 1, "decoder kind must be 0 (simt) or 1 (tcgen05)");
    g_decoder = kind;
    return 0;
}

int cb200_get_decoder(void) { return g_decoder; }

int cb200_set_forward_pipeline(long long min_points) {
    cb200::set_fwd_ws_min_points(min_points);
    return 0;
}

long long cb200_get_forward_pipeline(void) { return cb200::fwd_ws_min_points(); }

int cb200_device_sm_count(void) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    return sms;
}

// tiny-cuda-nn grid.h: grid_scale / grid_resolution and the offset table of the GridEncodingTemplated constructor
// (SURVEY.md section 9.1-9.2).  exp2f/log2f are taken correctly rounded (double evaluation rounded once).
int cb200_grid_meta(const cb200_grid_cfg* cfg, cb200_grid* grid) {
    CB_REQUIRE(cfg && grid, "cfg/grid is NULL");
    CB_REQUIRE(cfg->n_levels >= 1 && cfg->n_levels <= CB200_MAX_LEVELS, "n_levels %u not in [1,%d]", cfg->n_levels, CB200_MAX_LEVELS);
    CB_REQUIRE(cfg->n_features_per_level == 2, "n_features_per_level %u unsupported", cfg->n_features_per_level);
    CB_REQUIRE(cfg->log2_hashmap_size >= 3 && cfg->log2_hashmap_size <= 28, "log2_hashmap_size %u out of range", cfg->log2_hashmap_size);
    CB_REQUIRE(cfg->per_level_scale > 0.f && cfg->base_resolution >= 1, "bad base_resolution / per_level_scale");
    grid->n_levels = cfg->n_levels;
    grid->n_features = cfg->n_features_per_level;
    const float log2_pls = (float)log2((double)cfg->per_level_scale);
    const uint32_t cap = 1u << cfg->log2_hashmap_size;
    uint64_t offset = 0;
    for (uint32_t l = 0; l < CB200_MAX_LEVELS; ++l) {
        if (l >= cfg->n_levels) {
            grid->scale[l] = 0.f; grid->res[l] = 1; grid->hashed[l] = 0; grid->offset[l + 1] = (uint32_t)offset;
            continue;
        }
        const float arg = (float)l * log2_pls;
        const float e = (float)exp2((double)arg);
        const float scale = e * (float)cfg->base_resolution - 1.0f;
        const uint32_t res = (uint32_t)ceilf(scale) + 1u;
        const uint32_t max_params = UINT32_MAX / 2;
        uint32_t n = powf((float)res, 3.0f) > (float)max_params ? max_params : res * res * res;
        n = (n + 7u) / 8u * 8u;
        if (n > cap) n = cap;
        // grid_index: the dense stride walk stops as soon as stride > size; a level is hashed iff it stopped early
        uint32_t stride = 1;
        for (int d = 0; d < 3 && stride <= n; ++d) stride *= res;
        grid->scale[l] = scale;
        grid->res[l] = res;
        grid->hashed[l] = n < stride ? 1u : 0u;
        grid->offset[l] = (uint32_t)offset;
        offset += n;
        CB_REQUIRE(offset < (1ull << 31), "grid too large");
        grid->offset[l + 1] = (uint32_t)offset;
    }
    return 0;
}

int cb200_hash_fwd(const cb200_grid* grid, const float* x, const float* table, int64_t n_points, float* out, void* stream) {
    if (int rc = check_grid(grid)) return rc;
    CB_REQUIRE(n_points >= 0 && (n_points == 0 || (x && table && out)), "hash_fwd: NULL buffer");
    return cuda_result(cb200::launch_hash_fwd(cb200::make_scene_dev(*grid), x, table, n_points, out, (cudaStream_t)stream), "hash_fwd");
}

int cb200_hash_bwd(const cb200_grid* grid, const float* x, const float* table, const float* dy, int64_t n_points, float* dtable, float* dx,
                   void* stream) {
    if (int rc = check_grid(grid)) return rc;
    CB_REQUIRE(n_points >= 0 && (n_points == 0 || (x && dy)), "hash_bwd: NULL buffer");
    CB_REQUIRE(dx == nullptr || table != nullptr, "hash_bwd: dx needs the table");
    return cuda_result(cb200::launch_hash_bwd(cb200::make_scene_dev(*grid), x, table, dy, n_points, dtable, dx, (cudaStream_t)stream), "hash_bwd");
}

int cb200_hash_indices(const cb200_grid* grid, const float* x, int64_t n_points, uint32_t* idx, void* stream) {
    if (int rc = check_grid(grid)) return rc;
    CB_REQUIRE(n_points >= 0 && (n_points == 0 || (x && idx)), "hash_indices: NULL buffer");
    return cuda_result(cb200::launch_hash_indices(cb200::make_scene_dev(*grid), x, n_points, idx, (cudaStream_t)stream), "hash_indices");
}

int cb200_debug_gather_peak(const cb200_grid* grid, const uint32_t* idx, int64_t n_points, const float* table, float* dtable, int mode,
                            float* sink, void* stream) {
    if (int rc = check_grid(grid)) return rc;
    CB_REQUIRE(grid->n_levels == CB200_MAX_LEVELS, "gather_peak: expects %d levels", CB200_MAX_LEVELS);
    CB_REQUIRE((mode & 3) >= 1, "gather_peak: mode bits 0-1 must be 1 (gather), 2 (reduce) or 3 (both)");
    CB_REQUIRE(n_points >= 0 && idx && sink && ((mode & 1) == 0 || table) && ((mode & 2) == 0 || dtable), "gather_peak: NULL buffer");
    return cuda_result(cb200::launch_gather_peak(cb200::make_scene_dev(*grid), idx, n_points, table, dtable, mode, sink, (cudaStream_t)stream),
                       "gather_peak");
}

int cb200_oneblob_fwd(const float* x, int64_t n_points, uint32_t n_dims, uint32_t n_bins, float* out, void* stream) {
    CB_REQUIRE(n_points >= 0 && n_dims >= 1 && n_bins >= 1, "oneblob_fwd: bad sizes");
    CB_REQUIRE(n_points == 0 || (x && out), "oneblob_fwd: NULL buffer");
    return cuda_result(cb200::launch_oneblob_fwd(x, n_points * n_dims, (int)n_bins, out, (cudaStream_t)stream), "oneblob_fwd");
}

int cb200_oneblob_bwd(const float* x, const float* dy, int64_t n_points, uint32_t n_dims, uint32_t n_bins, float* dx, void* stream) {
    CB_REQUIRE(n_points >= 0 && n_dims >= 1 && n_bins >= 1, "oneblob_bwd: bad sizes");
    CB_REQUIRE(n_points == 0 || (x && dy && dx), "oneblob_bwd: NULL buffer");
    return cuda_result(cb200::launch_oneblob_bwd(x, dy, n_points * n_dims, (int)n_bins, dx, (cudaStream_t)stream), "oneblob_bwd");
}

int cb200_assemble_rays_fwd(const float* kf_rays, const int64_t* ray_idx, const int64_t* pose_slot, const float* poses, int64_t n_rays,
                            int64_t n_poses, float* rays_o, float* rays_d, float* target_rgb, float* target_d, void* stream) {
    CB_REQUIRE(n_rays >= 0 && n_poses >= 1, "assemble_rays: bad sizes");
    if (n_rays == 0) return 0;
    CB_REQUIRE(kf_rays && ray_idx && pose_slot && poses && rays_o && rays_d && target_rgb && target_d, "assemble_rays_fwd: NULL buffer");
    return cuda_result(cb200::launch_assemble_rays_fwd(kf_rays, reinterpret_cast<const long long*>(ray_idx),
                                                       reinterpret_cast<const long long*>(pose_slot), poses, n_rays, n_poses, rays_o, rays_d,
                                                       target_rgb, target_d, (cudaStream_t)stream), "assemble_rays_fwd");
}

int cb200_assemble_rays_bwd(const float* kf_rays, const int64_t* ray_idx, const int64_t* pose_slot, const float* d_rays_o,
                            const float* d_rays_d, int64_t n_rays, int64_t n_poses, float* d_poses, void* stream) {
    CB_REQUIRE(n_rays >= 0 && n_poses >= 1, "assemble_rays: bad sizes");
    if (n_rays == 0) return 0;
    CB_REQUIRE(kf_rays && ray_idx && pose_slot && d_rays_o && d_rays_d && d_poses, "assemble_rays_bwd: NULL buffer");
    return cuda_result(cb200::launch_assemble_rays_bwd(kf_rays, reinterpret_cast<const long long*>(ray_idx),
                                                       reinterpret_cast<const long long*>(pose_slot), d_rays_o, d_rays_d, n_rays, n_poses,
                                                       d_poses, (cudaStream_t)stream), "assemble_rays_bwd");
}

int cb200_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1, float beta2, float eps,
                    float weight_decay, const int64_t* step_dev, int64_t step, int zero_grad, void* stream) {
    CB_REQUIRE(n >= 0, "adam_step: negative size");
    if (n == 0) return 0;
    CB_REQUIRE(param && grad && exp_avg && exp_avg_sq, "adam_step: NULL buffer");
    CB_REQUIRE(step_dev != nullptr || step >= 1, "adam_step: step is 1-based");
    CB_REQUIRE(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f && eps >= 0.f, "adam_step: bad hyper-parameters");
    return cuda_result(cb200::launch_adam_step(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step_dev, step,
                                               zero_grad, (cudaStream_t)stream), "adam_step");
}

int cb200_track_pose(const float* rot, const float* trans, int n_rot, float* c2w, void* stream) {
    CB_REQUIRE(rot && trans && c2w, "track_pose: NULL buffer");
    CB_REQUIRE(n_rot == 3 || n_rot == 4, "track_pose: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    return cuda_result(cb200::launch_pose_matrices(rot, trans, 1, n_rot, c2w, (cudaStream_t)stream), "track_pose");
}

int cb200_pose_matrices(const float* rot, const float* trans, int64_t n_poses, int n_rot, float* c2w, void* stream) {
    CB_REQUIRE(n_poses >= 0, "pose_matrices: negative count");
    if (n_poses == 0) return 0;
    CB_REQUIRE(rot && trans && c2w, "pose_matrices: NULL buffer");
    CB_REQUIRE(n_rot == 3 || n_rot == 4, "pose_matrices: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    return cuda_result(cb200::launch_pose_matrices(rot, trans, n_poses, n_rot, c2w, (cudaStream_t)stream), "pose_matrices");
}

int cb200_pose_adam(float* rot, float* trans, float* exp_avg_rot, float* exp_avg_sq_rot, float* exp_avg_trans, float* exp_avg_sq_trans,
                    float* d_c2w, float* c2w, int64_t n_poses, int n_rot, float lr_rot, float lr_trans, int64_t step, int zero_grad,
                    void* stream) {
    CB_REQUIRE(n_poses >= 0, "pose_adam: negative count");
    if (n_poses == 0) return 0;
    CB_REQUIRE(rot && trans && exp_avg_rot && exp_avg_sq_rot && exp_avg_trans && exp_avg_sq_trans && d_c2w && c2w, "pose_adam: NULL buffer");
    CB_REQUIRE(n_rot == 3 || n_rot == 4, "pose_adam: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    CB_REQUIRE(step >= 1, "pose_adam: step is 1-based");
    return cuda_result(cb200::launch_pose_adam(rot, trans, exp_avg_rot, exp_avg_sq_rot, exp_avg_trans, exp_avg_sq_trans, d_c2w, c2w, n_poses,
                                               n_rot, lr_rot, lr_trans, step, zero_grad, (cudaStream_t)stream), "pose_adam");
}

int cb200_smooth_points(int64_t n_side, float voxel, const float* jitter, const float* offset, const float* bb_lo, const float* bb_ext,
                        float* x, void* stream) {
    CB_REQUIRE(n_side >= 1 && n_side <= 1024, "smooth_points: lattice side %lld out of range", (long long)n_side);
    CB_REQUIRE(jitter && offset && bb_lo && bb_ext && x, "smooth_points: NULL buffer");
    return cuda_result(cb200::launch_smooth_points((int)n_side, voxel, jitter, offset, bb_lo, bb_ext, x, (cudaStream_t)stream), "smooth_points");
}

int cb200_smooth_tv(const float* f, int64_t n_side, int64_t n_feat, float grad_scale, float* d_f, double* tv_sum, void* stream) {
    CB_REQUIRE(n_side >= 1 && n_side <= 1024 && n_feat >= 4 && n_feat <= 1024 && n_feat % 4 == 0, "smooth_tv: bad lattice shape");
    CB_REQUIRE(f && d_f, "smooth_tv: NULL buffer");
    return cuda_result(cb200::launch_smooth_tv(f, (int)n_side, (int)n_feat, grad_scale, d_f, tv_sum, (cudaStream_t)stream), "smooth_tv");
}

int cb200_pose_params(const float* c2w, int64_t n_poses, int32_t n_rot, float* rot, float* trans, void* stream) {
    CB_REQUIRE(n_poses >= 0 && (n_poses == 0 || (c2w && rot && trans)), "pose_params: NULL buffer");
    CB_REQUIRE(n_rot == 3 || n_rot == 4, "pose_params: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    return cuda_result(cb200::launch_pose_params(c2w, n_poses, n_rot, rot, trans, (cudaStream_t)stream), "pose_params");
}

int cb200_track_rays_sample(const cb200_render_cfg* cfg, const float* kf_rays, const float* c2w, const float* lin_uniform,
                            const float* lin_guided, const float* lin_nodepth, const float* noise, uint64_t seed, const uint64_t* seed_dev,
                            int64_t n_rays, float* rays_o, float* rays_d, float* target_rgb, float* target_d, float* z_vals, int64_t* counts,
                            void* stream) {
    CB_REQUIRE(cfg != nullptr, "render cfg is NULL");
    const int S = (int)(cfg->n_uniform + cfg->n_guided);
    CB_REQUIRE(cfg->n_guided >= 1 && S <= 256, "n_samples_d + n_range_d = %d must be in [1,256] with n_range_d >= 1", S);
    CB_REQUIRE(n_rays >= 0, "negative ray count");
    if (n_rays == 0) return 0;
    CB_REQUIRE(kf_rays && c2w && rays_o && rays_d && target_rgb && target_d && z_vals && lin_guided && lin_nodepth &&
                   (cfg->n_uniform == 0 || lin_uniform),
               "track_rays_sample: NULL buffer");
    cb200::SampleArgs a{};
    a.lin_uniform = lin_uniform; a.lin_guided = lin_guided; a.lin_nodepth = lin_nodepth;
    a.noise = noise; a.seed = seed; a.seed_dev = reinterpret_cast<const unsigned long long*>(seed_dev); a.n_rays = n_rays; a.z_vals = z_vals;
    a.counts = reinterpret_cast<unsigned long long*>(counts);
    a.n_uniform = (int)cfg->n_uniform; a.n_guided = (int)cfg->n_guided; a.perturb = (int)cfg->perturb;
    a.trunc_sc = cfg->trunc_sc; a.depth_trunc = cfg->depth_trunc;
    return cuda_result(cb200::launch_track_rays_sample(a, kf_rays, c2w, rays_o, rays_d, target_rgb, target_d, (cudaStream_t)stream),
                       "track_rays_sample");
}

int cb200_track_pose_update(const float* kf_rays, const float* d_rays_o, const float* d_rays_d, int64_t n_rays, const cb200_track_step* s,
                            uint32_t* ticket, uint64_t* noise_counter, void* stream) {
    CB_REQUIRE(s != nullptr && kf_rays && d_rays_o && d_rays_d && ticket && n_rays >= 1, "track_pose_update: NULL argument");
    CB_REQUIRE(s->rot && s->trans && s->adam && s->step && s->losses && s->d_c2w && s->c2w && s->best_loss && s->best_c2w && s->last_loss &&
                   s->last_c2w && s->patience && s->alive,
               "track_pose_update: NULL buffer in the step block");
    CB_REQUIRE(s->n_rot == 3 || s->n_rot == 4, "track_pose_update: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    return cuda_result(cb200::launch_track_pose_update(kf_rays, d_rays_o, d_rays_d, n_rays, *s, ticket,
                                                       reinterpret_cast<unsigned long long*>(noise_counter), (cudaStream_t)stream),
                       "track_pose_update");
}

int cb200_track_init(const float* c2w_init, const cb200_track_step* s, void* stream) {
    CB_REQUIRE(c2w_init != nullptr && s != nullptr, "track_init: NULL argument");
    CB_REQUIRE(s->rot && s->trans && s->adam && s->step && s->c2w && s->patience && s->alive, "track_init: NULL buffer");
    CB_REQUIRE(s->n_rot == 3 || s->n_rot == 4, "track_init: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    return cuda_result(cb200::launch_track_init(c2w_init, *s, (cudaStream_t)stream), "track_init");
}

int cb200_track_update(const cb200_track_step* s, void* stream) {
    CB_REQUIRE(s != nullptr, "track_update: NULL step");
    CB_REQUIRE(s->rot && s->trans && s->adam && s->step && s->losses && s->d_c2w && s->c2w && s->best_loss && s->best_c2w && s->last_loss &&
                   s->last_c2w && s->patience && s->alive,
               "track_update: NULL buffer");
    CB_REQUIRE(s->n_rot == 3 || s->n_rot == 4, "track_update: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    return cuda_result(cb200::launch_track_update(*s, (cudaStream_t)stream), "track_update");
}

static int fill_points_args(cb200::PointsArgs& a, const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts,
                            const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples);
static int fill_render_args(cb200::RenderArgs& a, const cb200_render_cfg* cfg, const float* raw, const float* z_vals,
                            const float* target_rgb, const float* target_d, int64_t n_rays, int64_t n_samples);

// workspace of cb200_track_frame: 512-byte header (barrier, ticket, list count, mask counts, loss sums, dL/d c2w) + per-ray and
// per-sample arrays, each starting on a 256-byte boundary
namespace {
struct TrackFrameLayout {
    size_t rays_o, rays_d, target_rgb, target_d, d_rays_o, d_rays_d, z, raw, mask, d_raw, list, total;
};
TrackFrameLayout track_frame_layout(int64_t n, int64_t s) {
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    TrackFrameLayout l;
    size_t off = 512;
    const size_t N = (size_t)(n > 0 ? n : 0), P = N * (size_t)(s > 0 ? s : 0);
    l.rays_o = off; off = up(off + 12 * N);
    l.rays_d = off; off = up(off + 12 * N);
    l.target_rgb = off; off = up(off + 12 * N);
    l.target_d = off; off = up(off + 4 * N);
    l.d_rays_o = off; off = up(off + 12 * N);
    l.d_rays_d = off; off = up(off + 12 * N);
    l.z = off; off = up(off + 4 * P);
    l.raw = off; off = up(off + 16 * P);
    l.mask = off; off = up(off + 8 * P);
    l.d_raw = off; off = up(off + 16 * P);
    l.list = off; off = up(off + 4 * P);
    l.total = off;
    return l;
}
}  // namespace

size_t cb200_track_frame_workspace(int64_t n_rays, int64_t n_samples) { return track_frame_layout(n_rays, n_samples).total; }

int cb200_track_frame(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const cb200_render_cfg* cfg, const float* lin_uniform,
                      const float* lin_guided, const float* lin_nodepth, uint64_t seed, uint64_t* seed_dev, const float* kf_rays,
                      const float* c2w_init, int64_t n_rays, int32_t n_iters, const cb200_track_step* step, void* workspace,
                      size_t workspace_bytes, void* stream) {
    CB_REQUIRE(cfg != nullptr && step != nullptr, "track_frame: NULL cfg / step");
    CB_REQUIRE(g_decoder == 1, "track_frame: needs the tensor-core decoder");
    const int64_t S = (int64_t)cfg->n_uniform + cfg->n_guided;
    CB_REQUIRE(cfg->n_guided >= 1 && S <= 256, "n_samples_d + n_range_d = %lld must be in [1,256] with n_range_d >= 1", (long long)S);
    CB_REQUIRE(n_rays >= 1 && n_iters >= 1 && n_rays * S < (1ll << 31), "track_frame: bad sizes");
    CB_REQUIRE(kf_rays && lin_guided && lin_nodepth && (cfg->n_uniform == 0 || lin_uniform), "track_frame: NULL buffer");
    CB_REQUIRE(step->rot && step->trans && step->adam && step->step && step->losses && step->c2w && step->best_loss && step->best_c2w &&
                   step->last_loss && step->last_c2w && step->patience && step->alive,
               "track_frame: NULL buffer in the step block");
    CB_REQUIRE(step->n_rot == 3 || step->n_rot == 4, "track_frame: n_rot must be 3 (axis-angle) or 4 (quaternion)");
    const TrackFrameLayout l = track_frame_layout(n_rays, S);
    CB_REQUIRE(workspace != nullptr && workspace_bytes >= l.total && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
               "track_frame: workspace missing, not 256-byte aligned or smaller than cb200_track_frame_workspace (%zu bytes)", l.total);
    char* w = static_cast<char*>(workspace);
    auto f = [&](size_t off) { return reinterpret_cast<float*>(w + off); };
    cb200::TrackFrameArgs t{};
    if (int rc = fill_points_args(t.pts, scene, table, mlp, nullptr, f(l.rays_o), f(l.rays_d), f(l.z), n_rays, S)) return rc;
    t.pts.raw = f(l.raw);
    t.pts.act_mask = reinterpret_cast<uint32_t*>(w + l.mask);
    t.pts.d_raw = f(l.d_raw);
    t.pts.d_rays_o = f(l.d_rays_o);
    t.pts.d_rays_d = f(l.d_rays_d);
    t.smp.target_d = f(l.target_d); t.smp.lin_uniform = lin_uniform; t.smp.lin_guided = lin_guided; t.smp.lin_nodepth = lin_nodepth;
    t.smp.seed = seed; t.smp.n_rays = n_rays; t.smp.z_vals = f(l.z);
    t.smp.n_uniform = (int)cfg->n_uniform; t.smp.n_guided = (int)cfg->n_guided; t.smp.perturb = (int)cfg->perturb;
    t.smp.trunc_sc = cfg->trunc_sc; t.smp.depth_trunc = cfg->depth_trunc;
    if (int rc = fill_render_args(t.rnd, cfg, f(l.raw), f(l.z), f(l.target_rgb), f(l.target_d), n_rays, S)) return rc;
    t.rnd.has_loss = 1;
    t.rnd.d_raw = f(l.d_raw);
    t.barrier = reinterpret_cast<unsigned*>(w);
    t.ticket = reinterpret_cast<unsigned*>(w + 4);
    t.live_count = reinterpret_cast<unsigned long long*>(w + 8);
    t.counts = reinterpret_cast<unsigned long long*>(w + 16);
    t.sums = reinterpret_cast<double*>(w + 64);
    t.d_c2w = f(128);
    t.losses = const_cast<float*>(step->losses);
    t.live_list = reinterpret_cast<uint32_t*>(w + l.list);
    t.kf_rays = kf_rays;
    t.c2w_init = c2w_init;
    t.seed_dev = reinterpret_cast<unsigned long long*>(seed_dev);
    t.n_iters = n_iters;
    t.step = *step;
    t.step.d_c2w = t.d_c2w;
    t.step.sums = t.sums;                       // the pose update takes the losses from this iteration's sums and counts
    t.step.counts = reinterpret_cast<const int64_t*>(t.counts);
    t.step.n_rays_total = n_rays;
    t.step.n_samples = S;
    if (int rc = cuda_result(cudaMemsetAsync(w, 0, 512, (cudaStream_t)stream), "track_frame: clear header")) return rc;
    const cudaError_t e = cb200::launch_track_frame(t, (cudaStream_t)stream);
    CB_REQUIRE(e != cudaErrorNotSupported, "track_frame: no instantiation for this grid (dense levels) / samples-per-ray combination");
    return cuda_result(e, "track_frame");
}

int cb200_adam_multi(const cb200_adam_param* params, int64_t n_params, int64_t step, int zero_grad, void* stream) {
    CB_REQUIRE(n_params >= 0 && n_params <= 64, "adam_multi: bad tensor count");
    if (n_params == 0) return 0;
    CB_REQUIRE(params != nullptr && step >= 1, "adam_multi: NULL list or step < 1");
    for (int64_t i = 0; i < n_params; ++i) {
        const cb200_adam_param& q = params[i];
        CB_REQUIRE(q.n >= 0 && (q.n == 0 || (q.param && q.grad && q.exp_avg && q.exp_avg_sq)), "adam_multi: NULL buffer in tensor %lld", (long long)i);
        CB_REQUIRE(q.beta1 >= 0.f && q.beta1 < 1.f && q.beta2 >= 0.f && q.beta2 < 1.f && q.eps >= 0.f, "adam_multi: bad hyper-parameters");
    }
    return cuda_result(cb200::launch_adam_multi(params, (int)n_params, step, zero_grad, (cudaStream_t)stream), "adam_multi");
}

int cb200_sample_rays(const cb200_render_cfg* cfg, const float* target_d, const float* lin_uniform, const float* lin_guided,
                      const float* lin_nodepth, const float* noise, uint64_t seed, const uint64_t* seed_dev, int64_t n_rays, float* z_vals,
                      int64_t* counts, void* stream) {
    CB_REQUIRE(cfg != nullptr, "render cfg is NULL");
    const int S = (int)(cfg->n_uniform + cfg->n_guided);
    CB_REQUIRE(cfg->n_guided >= 1 && S <= 256, "n_samples_d + n_range_d = %d must be in [1,256] with n_range_d >= 1", S);
    CB_REQUIRE(n_rays >= 0, "negative ray count");
    if (n_rays == 0) return 0;
    CB_REQUIRE(target_d && lin_guided && lin_nodepth && z_vals && (cfg->n_uniform == 0 || lin_uniform), "sample_rays: NULL buffer");
    cb200::SampleArgs a{};
    a.target_d = target_d; a.lin_uniform = lin_uniform; a.lin_guided = lin_guided; a.lin_nodepth = lin_nodepth;
    a.noise = noise; a.seed = seed; a.seed_dev = reinterpret_cast<const unsigned long long*>(seed_dev); a.n_rays = n_rays; a.z_vals = z_vals;
    a.counts = reinterpret_cast<unsigned long long*>(counts);
    a.n_uniform = (int)cfg->n_uniform; a.n_guided = (int)cfg->n_guided; a.perturb = (int)cfg->perturb;
    a.trunc_sc = cfg->trunc_sc; a.depth_trunc = cfg->depth_trunc;
    return cuda_result(cb200::launch_sample_rays(a, (cudaStream_t)stream), "sample_rays");
}

static int importance_sample_impl(const cb200_render_cfg* cfg, const float* z_vals, const float* weights, const float* u, int64_t u_stride,
                                  const float* target_d, int64_t n_rays, int64_t n_samples, int64_t n_importance, float* z_out, float* z_std,
                                  int64_t* counts, int32_t* rank, uint32_t* new_list, uint64_t* new_count, void* stream) {
    CB_REQUIRE(cfg != nullptr, "render cfg is NULL");
    CB_REQUIRE(n_samples >= 3 && n_importance >= 1 && n_samples + n_importance <= 256,
               "importance_sample: need 3 <= n_samples, 1 <= n_importance, n_samples + n_importance <= 256 (got %lld + %lld)",
               (long long)n_samples, (long long)n_importance);
    CB_REQUIRE(n_rays >= 0, "negative ray count");
    CB_REQUIRE(u_stride == 0 || u_stride == n_importance, "importance_sample: u_stride must be 0 or n_importance");
    CB_REQUIRE(counts == nullptr || target_d != nullptr, "importance_sample: counts need target_d");
    CB_REQUIRE((new_list == nullptr) == (new_count == nullptr), "importance_sample: new_list and new_count go together");
    CB_REQUIRE(new_list == nullptr || n_rays * (n_samples + n_importance) < (1ll << 32), "importance_sample: the point list indexes with 32 bits");
    if (n_rays == 0) return 0;
    CB_REQUIRE(z_vals && weights && u && z_out, "importance_sample: NULL buffer");
    cb200::ImportanceArgs a{};
    a.z_vals = z_vals; a.weights = weights; a.u = u; a.u_stride = u_stride; a.target_d = target_d; a.n_rays = n_rays;
    a.n_samples = (int)n_samples; a.n_importance = (int)n_importance; a.z_out = z_out; a.z_std = z_std;
    a.counts = reinterpret_cast<unsigned long long*>(counts);
    a.trunc_sc = cfg->trunc_sc; a.depth_trunc = cfg->depth_trunc;
    a.rank = rank; a.new_list = new_list; a.new_count = reinterpret_cast<unsigned long long*>(new_count);
    return cuda_result(cb200::launch_importance_sample(a, (cudaStream_t)stream), "importance_sample");
}

int cb200_importance_sample(const cb200_render_cfg* cfg, const float* z_vals, const float* weights, const float* u, int64_t u_stride,
                            const float* target_d, int64_t n_rays, int64_t n_samples, int64_t n_importance, float* z_out, float* z_std,
                            int64_t* counts, void* stream) {
    return importance_sample_impl(cfg, z_vals, weights, u, u_stride, target_d, n_rays, n_samples, n_importance, z_out, z_std, counts, nullptr,
                                  nullptr, nullptr, stream);
}

int cb200_importance_sample_merge(const cb200_render_cfg* cfg, const float* z_vals, const float* weights, const float* u, int64_t u_stride,
                                  const float* target_d, int64_t n_rays, int64_t n_samples, int64_t n_importance, float* z_out, float* z_std,
                                  int64_t* counts, int32_t* rank, uint32_t* new_list, uint64_t* new_count, void* stream) {
    CB_REQUIRE(rank != nullptr, "importance_sample_merge: rank is NULL");
    return importance_sample_impl(cfg, z_vals, weights, u, u_stride, target_d, n_rays, n_samples, n_importance, z_out, z_std, counts, rank,
                                  new_list, new_count, stream);
}

int cb200_merge_coarse(const int32_t* rank, int64_t n_rays, int64_t n_coarse, int64_t n_fine, const float* src, float* dst,
                       const uint32_t* src_mask, uint32_t* dst_mask, int accumulate, void* stream) {
    CB_REQUIRE(n_rays >= 0 && n_coarse >= 1 && n_fine >= n_coarse && n_fine <= 256, "merge_coarse: bad sizes");
    if (n_rays == 0) return 0;
    CB_REQUIRE(rank && src && dst, "merge_coarse: NULL buffer");
    CB_REQUIRE((src_mask == nullptr) == (dst_mask == nullptr), "merge_coarse: the two masks go together");
    return cuda_result(cb200::launch_merge_coarse(rank, n_rays, (int)n_coarse, (int)n_fine, src, dst, src_mask, dst_mask, accumulate,
                                                  (cudaStream_t)stream),
                       "merge_coarse");
}

static int fill_points_args(cb200::PointsArgs& a, const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts,
                            const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples) {
    if (int rc = check_fused_scene(scene)) return rc;
    CB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples < (1 << 30), "points: bad sizes");
    a.n_points = n_rays * n_samples;
    if (a.n_points == 0) return 0;
    CB_REQUIRE(table && mlp && mlp->w_sdf0 && mlp->w_sdf1 && mlp->w_col0 && mlp->w_col1, "points: NULL table / weights");
    CB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples < (1 << 30), "points: bad sizes");
    CB_REQUIRE((pts != nullptr) != (rays_o != nullptr), "points: pass either pts or (rays_o, rays_d, z_vals)");
    CB_REQUIRE(rays_o == nullptr || (rays_d && z_vals), "points: ray mode needs rays_d and z_vals");
    a.scene = cb200::make_scene_dev(*scene);
    a.table = reinterpret_cast<const float2*>(table);
    a.w1 = mlp->w_sdf0; a.w2 = mlp->w_sdf1; a.w3 = mlp->w_col0; a.w4 = mlp->w_col1;
    a.pts = pts; a.rays_o = rays_o; a.rays_d = rays_d; a.z = z_vals;
    a.n_points = n_rays * n_samples;
    a.n_samples = (int)n_samples;
    return 0;
}

static int points_fwd_impl(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts, const float* rays_o,
                           const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, float* raw, float* geo,
                           uint32_t* act_mask, int sdf_only, const uint32_t* point_list, const uint64_t* point_count, float* feat,
                           int feat_stride, void* stream) {
    cb200::PointsArgs a{};
    if (int rc = fill_points_args(a, scene, table, mlp, pts, rays_o, rays_d, z_vals, n_rays, n_samples)) return rc;
    if (a.n_points == 0) return 0;
    a.raw = raw; a.geo = geo; a.act_mask = act_mask; a.sdf_only = sdf_only;
    CB_REQUIRE(feat == nullptr || g_decoder == 1, "points_fwd_keep: the kept features need the tensor-core decoder");
    a.feat = feat;
    a.feat_stride = feat_stride;
    CB_REQUIRE((point_list == nullptr) == (point_count == nullptr), "points_fwd: point_list and point_count go together");
    CB_REQUIRE(point_list == nullptr || g_decoder == 1, "points_fwd: point lists need the tensor-core decoder");
    CB_REQUIRE(point_list == nullptr || a.n_points < (1ll << 32), "points_fwd: point lists index with 32 bits");
    a.live_idx = point_list;
    a.live_count = reinterpret_cast<const unsigned long long*>(point_count);
    if (g_decoder == 1) return cuda_result(cb200::launch_points_fwd_tc(a, (cudaStream_t)stream), "points_fwd_tc");
    return cuda_result(cb200::launch_points_fwd(a, (cudaStream_t)stream), "points_fwd");
}

int cb200_points_fwd(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts, const float* rays_o,
                     const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, float* raw, float* geo,
                     uint32_t* act_mask, int sdf_only, const uint32_t* point_list, const uint64_t* point_count, void* stream) {
    return points_fwd_impl(scene, table, mlp, pts, rays_o, rays_d, z_vals, n_rays, n_samples, raw, geo, act_mask, sdf_only, point_list,
                           point_count, nullptr, 0, stream);
}

int cb200_points_fwd_keep(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts, const float* rays_o,
                          const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, float* raw, float* geo,
                          uint32_t* act_mask, int sdf_only, const uint32_t* point_list, const uint64_t* point_count, float* feat,
                          int with_derivatives, void* stream) {
    CB_REQUIRE(feat != nullptr || n_rays == 0, "points_fwd_keep: feat is NULL");
    return points_fwd_impl(scene, table, mlp, pts, rays_o, rays_d, z_vals, n_rays, n_samples, raw, geo, act_mask, sdf_only, point_list,
                           point_count, feat, with_derivatives ? 4 * CB200_GRID_FEATS : CB200_GRID_FEATS, stream);
}

size_t cb200_lazy_workspace(int64_t n_rays, int64_t n_samples) {
    if (n_rays < 0 || n_samples < 0) return 0;
    return 16 + sizeof(int) * (size_t)n_rays + sizeof(uint32_t) * (size_t)n_rays * (size_t)n_samples;
}

static int lazy_split(void* workspace, size_t bytes, int64_t n_rays, int64_t n_samples, unsigned long long** count, int** k_done,
                      uint32_t** list) {
    CB_REQUIRE(workspace != nullptr && bytes >= cb200_lazy_workspace(n_rays, n_samples), "lazy: workspace missing or too small");
    CB_REQUIRE(n_rays * n_samples < (1ll << 32), "lazy: point lists index with 32 bits");
    *count = reinterpret_cast<unsigned long long*>(workspace);
    *k_done = reinterpret_cast<int*>(*count + 2);
    *list = reinterpret_cast<uint32_t*>(*k_done + n_rays);
    return 0;
}

int cb200_lazy_prefix(const cb200_render_cfg* cfg, const float* z_vals, const float* target_d, int64_t n_rays, int64_t n_samples,
                      void* workspace, size_t workspace_bytes, void* stream) {
    CB_REQUIRE(cfg != nullptr, "render cfg is NULL");
    CB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples <= 256, "lazy_prefix: bad sizes");
    if (n_rays == 0) return 0;
    CB_REQUIRE(z_vals && target_d, "lazy_prefix: NULL buffer");
    unsigned long long* count; int* k_done; uint32_t* list;
    if (int rc = lazy_split(workspace, workspace_bytes, n_rays, n_samples, &count, &k_done, &list)) return rc;
    return cuda_result(cb200::launch_lazy_prefix(z_vals, target_d, n_rays, (int)n_samples, cfg->trunc_sc, k_done, list, count,
                                                 (cudaStream_t)stream), "lazy_prefix");
}

int cb200_lazy_resolve(const cb200_render_cfg* cfg, const float* z_vals, const float* target_d, const float* raw, int64_t n_rays,
                       int64_t n_samples, void* workspace, size_t workspace_bytes, void* stream) {
    CB_REQUIRE(cfg != nullptr, "render cfg is NULL");
    CB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples <= 256, "lazy_resolve: bad sizes");
    if (n_rays == 0) return 0;
    CB_REQUIRE(z_vals && target_d && raw, "lazy_resolve: NULL buffer");
    unsigned long long* count; int* k_done; uint32_t* list;
    if (int rc = lazy_split(workspace, workspace_bytes, n_rays, n_samples, &count, &k_done, &list)) return rc;
    return cuda_result(cb200::launch_lazy_resolve(z_vals, target_d, raw, n_rays, (int)n_samples, cfg->trunc_sc, k_done, list, count,
                                                  (cudaStream_t)stream), "lazy_resolve");
}

size_t cb200_points_bwd_workspace(int64_t n_points) { return 16 + sizeof(uint32_t) * (size_t)(n_points > 0 ? n_points : 0); }

static int points_bwd_impl(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts, const float* rays_o,
                           const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, const float* d_raw,
                           const float* d_geo, const uint32_t* act_mask, float* dtable, const cb200_mlp_grad* d_mlp, float* d_pts,
                           float* d_rays_o, float* d_rays_d, int sdf_only, void* workspace, size_t workspace_bytes, const float* feat,
                           int feat_stride, void* stream) {
    cb200::PointsArgs a{};
    if (int rc = fill_points_args(a, scene, table, mlp, pts, rays_o, rays_d, z_vals, n_rays, n_samples)) return rc;
    if (a.n_points == 0) return 0;
    a.feat = const_cast<float*>(feat);
    a.feat_stride = feat_stride;
    CB_REQUIRE(d_raw && act_mask, "points_bwd: d_raw / act_mask is NULL");
    CB_REQUIRE((d_rays_o == nullptr) == (d_rays_d == nullptr), "points_bwd: d_rays_o and d_rays_d go together");
    CB_REQUIRE(d_rays_o == nullptr || rays_o != nullptr, "points_bwd: ray gradients need ray mode");
    a.d_raw = d_raw; a.d_geo = d_geo; a.act_mask = const_cast<uint32_t*>(act_mask); a.sdf_only = sdf_only;
    a.dtable = reinterpret_cast<float2*>(dtable);
    if (d_mlp) {
        CB_REQUIRE(d_mlp->w_sdf0 && d_mlp->w_sdf1 && d_mlp->w_col0 && d_mlp->w_col1, "points_bwd: partial weight-gradient set");
        a.dw1 = d_mlp->w_sdf0; a.dw2 = d_mlp->w_sdf1; a.dw3 = d_mlp->w_col0; a.dw4 = d_mlp->w_col1;
    }
    a.d_pts = d_pts; a.d_rays_o = d_rays_o; a.d_rays_d = d_rays_d;
    if (g_decoder == 1) {
        if (workspace && workspace_bytes == 0 && rays_o && d_geo == nullptr && d_pts == nullptr) {
            // the list is there already (cb200_render_loss_bwd wrote it while it produced d_raw)
            unsigned long long* count = reinterpret_cast<unsigned long long*>(workspace);
            a.live_idx = reinterpret_cast<uint32_t*>(count + 2);
            a.live_count = count;
        } else if (workspace && rays_o && d_geo == nullptr && d_pts == nullptr) {
            CB_REQUIRE(workspace_bytes >= cb200_points_bwd_workspace(a.n_points), "points_bwd: workspace too small (%zu bytes)", workspace_bytes);
            CB_REQUIRE(a.n_points < (1ll << 32), "points_bwd: compaction indexes points with 32 bits");
            unsigned long long* count = reinterpret_cast<unsigned long long*>(workspace);
            uint32_t* idx = reinterpret_cast<uint32_t*>(count + 2);
            if (int rc = cuda_result(cb200::launch_compact_live(d_raw, a.n_points, count, idx, (cudaStream_t)stream), "compact_live")) return rc;
            a.live_idx = idx;
            a.live_count = count;
        }
        return cuda_result(cb200::launch_points_bwd_tc(a, (cudaStream_t)stream), "points_bwd_tc");
    }
    return cuda_result(cb200::launch_points_bwd(a, (cudaStream_t)stream), "points_bwd");
}

int cb200_points_bwd(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts, const float* rays_o,
                     const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, const float* d_raw,
                     const float* d_geo, const uint32_t* act_mask, float* dtable, const cb200_mlp_grad* d_mlp, float* d_pts,
                     float* d_rays_o, float* d_rays_d, int sdf_only, void* workspace, size_t workspace_bytes, void* stream) {
    return points_bwd_impl(scene, table, mlp, pts, rays_o, rays_d, z_vals, n_rays, n_samples, d_raw, d_geo, act_mask, dtable, d_mlp, d_pts,
                           d_rays_o, d_rays_d, sdf_only, workspace, workspace_bytes, nullptr, 0, stream);
}

int cb200_points_bwd_kept(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts, const float* rays_o,
                          const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, const float* d_raw,
                          const float* d_geo, const uint32_t* act_mask, const float* feat, int with_derivatives, float* dtable,
                          const cb200_mlp_grad* d_mlp, float* d_pts, float* d_rays_o, float* d_rays_d, int sdf_only, void* workspace,
                          size_t workspace_bytes, void* stream) {
    CB_REQUIRE(feat != nullptr || n_rays == 0, "points_bwd_kept: feat is NULL");
    CB_REQUIRE(g_decoder == 1, "points_bwd_kept: the kept rows need the tensor-core decoder");
    CB_REQUIRE(d_mlp != nullptr, "points_bwd_kept: the kept rows serve the weight-gradient backward (d_mlp is NULL)");
    CB_REQUIRE(with_derivatives || (d_pts == nullptr && d_rays_o == nullptr),
               "points_bwd_kept: a position gradient needs rows kept with_derivatives (or cb200_points_bwd)");
    return points_bwd_impl(scene, table, mlp, pts, rays_o, rays_d, z_vals, n_rays, n_samples, d_raw, d_geo, act_mask, dtable, d_mlp, d_pts,
                           d_rays_o, d_rays_d, sdf_only, workspace, workspace_bytes, feat,
                           with_derivatives ? 4 * CB200_GRID_FEATS : CB200_GRID_FEATS, stream);
}

static int fill_render_args(cb200::RenderArgs& a, const cb200_render_cfg* cfg, const float* raw, const float* z_vals,
                            const float* target_rgb, const float* target_d, int64_t n_rays, int64_t n_samples) {
    CB_REQUIRE(cfg != nullptr, "render cfg is NULL");
    CB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples <= 256, "render: n_samples %lld must be in [1,256]", (long long)n_samples);
    CB_REQUIRE(n_rays == 0 || (raw && z_vals), "render: NULL raw / z_vals");
    a.raw = raw; a.z_vals = z_vals; a.target_rgb = target_rgb; a.target_d = target_d;
    a.n_rays = n_rays; a.n_samples = (int)n_samples;
    a.trunc = cfg->trunc; a.trunc_sc = cfg->trunc_sc; a.depth_trunc = cfg->depth_trunc;
    a.rgb_missing_nz = (int)cfg->rgb_missing_nz; a.white_bkgd = (int)cfg->white_bkgd;
    return 0;
}

int cb200_render_fwd(const cb200_render_cfg* cfg, const float* raw, const float* z_vals, const float* target_rgb, const float* target_d,
                     int64_t n_rays, int64_t n_samples, float* rgb, float* depth, float* depth_var, float* disp, float* acc,
                     float* weights, double* loss_sums, void* stream) {
    cb200::RenderArgs a{};
    if (int rc = fill_render_args(a, cfg, raw, z_vals, target_rgb, target_d, n_rays, n_samples)) return rc;
    if (n_rays == 0) return 0;
    CB_REQUIRE(loss_sums == nullptr || (target_rgb && target_d), "render_fwd: the loss head needs target_rgb and target_d");
    a.rgb = rgb; a.depth = depth; a.depth_var = depth_var; a.disp = disp; a.acc = acc; a.weights = weights; a.loss_sums = loss_sums;
    return cuda_result(cb200::launch_render_fwd(a, (cudaStream_t)stream), "render_fwd");
}

int cb200_loss_finish(const double* loss_sums, const int64_t* counts, int64_t n_rays_total, int64_t n_samples, float* losses, float* coef,
                      void* stream) {
    CB_REQUIRE(loss_sums && counts && losses && coef, "loss_finish: NULL buffer");
    return cuda_result(cb200::launch_loss_finish(loss_sums, reinterpret_cast<const long long*>(counts), n_rays_total, n_samples, losses, coef,
                                                 (cudaStream_t)stream),
                       "loss_finish");
}

int cb200_render_bwd(const cb200_render_cfg* cfg, const float* raw, const float* z_vals, const float* target_rgb, const float* target_d,
                     int64_t n_rays, int64_t n_samples, const cb200_loss_grads* g_loss, const float* coef, const float* g_rgb, const float* g_depth,
                     float* d_raw, void* stream) {
    cb200::RenderArgs a{};
    if (int rc = fill_render_args(a, cfg, raw, z_vals, target_rgb, target_d, n_rays, n_samples)) return rc;
    if (n_rays == 0) return 0;
    CB_REQUIRE(d_raw != nullptr, "render_bwd: d_raw is NULL");
    CB_REQUIRE(g_loss == nullptr || (coef && target_rgb && target_d), "render_bwd: loss gradients need coef and the targets");
    if (g_loss) {
        a.g_loss[0] = g_loss->rgb; a.g_loss[1] = g_loss->depth; a.g_loss[2] = g_loss->sdf; a.g_loss[3] = g_loss->fs;
        a.has_loss = (g_loss->rgb || g_loss->depth || g_loss->sdf || g_loss->fs) ? 1 : 0;
    }
    a.coef = coef; a.g_rgb = g_rgb; a.g_depth = g_depth; a.d_raw = d_raw;
    return cuda_result(cb200::launch_render_bwd(a, (cudaStream_t)stream), "render_bwd");
}

int cb200_render_loss_bwd(const cb200_render_cfg* cfg, const float* raw, const float* z_vals, const float* target_rgb, const float* target_d,
                          int64_t n_rays, int64_t n_samples, const int64_t* counts, int64_t n_rays_total, const cb200_loss_grads* g_loss,
                          double* loss_sums, float* coef, float* d_raw, void* list_workspace, size_t list_workspace_bytes,
                          int list_count_is_zero, void* stream) {
    cb200::RenderArgs a{};
    if (int rc = fill_render_args(a, cfg, raw, z_vals, target_rgb, target_d, n_rays, n_samples)) return rc;
    if (n_rays == 0) return 0;
    CB_REQUIRE(d_raw && counts && g_loss && target_rgb && target_d, "render_loss_bwd: NULL buffer");
    CB_REQUIRE(n_rays_total >= n_rays, "render_loss_bwd: n_rays_total < n_rays");
    a.g_loss[0] = g_loss->rgb; a.g_loss[1] = g_loss->depth; a.g_loss[2] = g_loss->sdf; a.g_loss[3] = g_loss->fs;
    a.has_loss = 1;
    a.loss_sums = loss_sums;
    a.d_raw = d_raw;
    uint32_t* list = nullptr;
    unsigned long long* count = nullptr;
    if (list_workspace) {
        CB_REQUIRE(list_workspace_bytes >= cb200_points_bwd_workspace(n_rays * n_samples), "render_loss_bwd: list workspace too small");
        CB_REQUIRE(n_rays * n_samples < (1ll << 32), "render_loss_bwd: the list indexes points with 32 bits");
        count = reinterpret_cast<unsigned long long*>(list_workspace);
        list = reinterpret_cast<uint32_t*>(count + 2);
    }
    return cuda_result(cb200::launch_render_loss_bwd(a, reinterpret_cast<const unsigned long long*>(counts), n_rays_total, list, count, coef,
                                                     list_count_is_zero, (cudaStream_t)stream),
                       "render_loss_bwd");
}

int cb200_debug_positions(const cb200_scene* scene, const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays,
                          int64_t n_samples, float* pts, float* x, void* stream) {
    CB_REQUIRE(scene && rays_o && rays_d && z_vals && pts && x, "debug_positions: NULL buffer");
    cb200::PointsArgs a{};
    a.scene = cb200::make_scene_dev(*scene);
    a.rays_o = rays_o; a.rays_d = rays_d; a.z = z_vals;
    a.n_points = n_rays * n_samples; a.n_samples = (int)n_samples;
    if (a.n_points == 0) return 0;
    return cuda_result(cb200::launch_debug_positions(a, pts, x, (cudaStream_t)stream), "debug_positions");
}

int cb200_selftest_mma(const float* a, const float* b, int k, int n, int mode, float* c, void* stream) {
    CB_REQUIRE(a && b && c, "selftest_mma: NULL buffer");
    CB_REQUIRE(mode == 0 || mode == 1, "selftest_mma: mode must be 0 (K-major) or 1 (MN-major)");
    CB_REQUIRE(n >= 16 && n <= 128 && n % 16 == 0, "selftest_mma: N must be a multiple of 16 in [16,128]");
    CB_REQUIRE(mode == 1 || (k >= 16 && k <= 128 && k % 16 == 0), "selftest_mma: K must be a multiple of 16 in [16,128]");
    return cuda_result(cb200::launch_tc_selftest(a, b, k, n, mode, c, (cudaStream_t)stream), "selftest_mma");
}

}  // extern "C"
