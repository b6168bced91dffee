// Kernel argument blocks and host launchers shared between the kernels (*.cu) and the C ABI (api.cu).
#pragma once
#include "common.cuh"

namespace cb200 {

struct PointsArgs {
    SceneDev scene;
    const float2* table;
    const float *w1, *w2, *w3, *w4;
    const float* pts;
    const float* rays_o;
    const float* rays_d;
    const float* z;
    int64_t n_points;
    int n_samples;
    int sdf_only;
    // tensor-core decoder: optional list of the points to process (backward: non-zero d_raw; forward: lazy evaluation)
    const uint32_t* live_idx;
    const unsigned long long* live_count;
    // forward outputs
    float* raw;
    float* geo;
    uint32_t* act_mask;
    // tensor-core decoder, optional: rows kept by the forward for every point it evaluates and read back by the weight-gradient backward
    // INSTEAD of gathering the table again.  feat_stride 32: the hash-grid features (no position gradient in that mode);
    // feat_stride 128: [features | d/dx | d/dy | d/dz of them] (x = normalised position), which also serve the position gradient
    float* feat;
    int feat_stride;
    // backward
    const float* d_raw;
    const float* d_geo;
    float2* dtable;
    float *dw1, *dw2, *dw3, *dw4;
    float* d_pts;
    float* d_rays_o;
    float* d_rays_d;
};

struct SampleArgs {
    const float* target_d;
    const float* lin_uniform;
    const float* lin_guided;
    const float* lin_nodepth;
    const float* noise;
    uint64_t seed;
    const unsigned long long* seed_dev;   // optional device counter added to `seed` (graph-captured loops)
    int64_t n_rays;
    float* z_vals;
    unsigned long long* counts;
    int n_uniform, n_guided;
    int perturb;
    float trunc_sc, depth_trunc;
};

struct ImportanceArgs {
    const float* z_vals;       // [N,S] coarse samples (ascending)
    const float* weights;      // [N,S] coarse rendering weights
    const float* u;            // uniforms: [N,I] (u_stride = I) or one shared row [I] (u_stride = 0, the det=True linspace)
    int64_t u_stride;
    const float* target_d;     // [N] or NULL (no counts)
    int64_t n_rays;
    int n_samples, n_importance;
    float* z_out;              // [N,S+I] merged, ascending
    float* z_std;              // [N] or NULL
    unsigned long long* counts;
    float trunc_sc, depth_trunc;
    // optional: where the samples went in the merged row.  The coarse samples keep their positions in space, so the fine pass need not
    // evaluate them again: rank [N,S] = position of coarse sample s in z_out's row; new_list [N*I] = flat indices (ray * (S+I) + position)
    // of the I new samples of every ray, new_count (device) = N * I -- a point list for the per-point forward
    int32_t* rank;
    uint32_t* new_list;
    unsigned long long* new_count;
};

struct RenderArgs {
    const float* raw;          // [N,S,4]
    const float* z_vals;       // [N,S]
    const float* target_rgb;   // [N,3]  (loss head only)
    const float* target_d;     // [N]
    int64_t n_rays;
    int n_samples;
    float trunc, trunc_sc, depth_trunc;
    int rgb_missing_nz, white_bkgd;
    // forward outputs
    float *rgb, *depth, *depth_var, *disp, *acc, *weights;
    double* loss_sums;
    // backward inputs / outputs
    const float* g_loss[4];    // upstream gradients of rgb/depth/sdf/fs losses (each nullable)
    int has_loss;              // loss head participates in the backward
    const float* coef;         // [4] d loss_k / d sum_k
    const float* g_rgb;        // [N,3]
    const float* g_depth;      // [N]
    float* d_raw;              // [N,S,4]
};

// one tracked frame in one cooperative launch (track_frame.cu)
struct TrackFrameArgs {
    PointsArgs pts;              // scene, table, weights; the ray / sample / gradient buffers point into the workspace
    SampleArgs smp;
    RenderArgs rnd;
    cb200_track_step step;
    const float* kf_rays;        // [N,7] camera-frame direction, rgb, depth of the frame's sampled pixels
    const float* c2w_init;       // [16] optional: start the frame from this matrix (track_init_state) instead of the state in `step`
    uint32_t* live_list;         // [N*S]
    unsigned long long* live_count;
    unsigned long long* counts;  // [3] mask counts   } zero on entry, cleared again by the pose update
    double* sums;                // [4] loss sums      }
    float* d_c2w;                // [16]               }
    float* losses;               // [4] out: the last iteration's four losses
    unsigned* barrier;           // zero on entry
    unsigned* ticket;            // zero on entry
    unsigned long long* seed_dev;    // optional device noise counter: read at entry, advanced by the iterations run
    int n_iters;
};

cudaError_t launch_track_frame(const TrackFrameArgs& t, cudaStream_t st);

cudaError_t launch_assemble_rays_fwd2(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, long long n_first,
                                      const float* cur_rays, const long long* idx_cur, const long long* slot_cur, long long n_cur,
                                      const float* poses, long long n_poses, float* rays_o, float* rays_d, float* target_rgb,
                                      float* target_d, cudaStream_t st);
cudaError_t launch_assemble_rays_bwd2(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, long long n_first,
                                      const float* cur_rays, const long long* idx_cur, const long long* slot_cur, long long n_cur,
                                      const float* d_rays_o, const float* d_rays_d, long long n_poses, float* d_poses, cudaStream_t st);
cudaError_t launch_assemble_rays_fwd(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, const float* poses,
                                     long long n, long long n_poses, float* rays_o, float* rays_d, float* target_rgb, float* target_d,
                                     cudaStream_t st);
cudaError_t launch_assemble_rays_bwd(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, const float* d_rays_o,
                                     const float* d_rays_d, long long n, long long n_poses, float* d_poses, cudaStream_t st);
cudaError_t launch_pose_matrices(const float* rot, const float* trans, long long n, int n_rot, float* c2w, cudaStream_t st);
cudaError_t launch_pose_adam(float* rot, float* trans, float* m_rot, float* v_rot, float* m_trans, float* v_trans, float* d_c2w, float* c2w,
                             long long n, int n_rot, float lr_rot, float lr_trans, long long step, int zero_grad, cudaStream_t st);
cudaError_t launch_smooth_tv(const float* f, int n, int n_feat, float scale, float* d_f, double* tv_sum, cudaStream_t st);
cudaError_t launch_smooth_points(int n, float vox, const float* jitter, const float* offset, const float* lo, const float* ext, float* x,
                                 cudaStream_t st);
cudaError_t launch_track_update(const cb200_track_step& s, cudaStream_t st);
cudaError_t launch_pose_params(const float* c2w, long long n, int n_rot, float* rot, float* trans, cudaStream_t st);
cudaError_t launch_track_init(const float* c2w_init, const cb200_track_step& s, cudaStream_t st);
cudaError_t launch_adam_multi(const cb200_adam_param* params, int n, int64_t step, int zero_grad, cudaStream_t st);
cudaError_t launch_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
                             float weight_decay, const int64_t* step_dev, int64_t step, int zero_grad, cudaStream_t st);
cudaError_t launch_lazy_prefix(const float* z_vals, const float* target_d, int64_t n_rays, int n_samples, float trunc_sc, int* k_done,
                               uint32_t* list, unsigned long long* count, cudaStream_t st);
cudaError_t launch_lazy_resolve(const float* z_vals, const float* target_d, const float* raw, int64_t n_rays, int n_samples, float trunc_sc,
                                int* k_done, uint32_t* list, unsigned long long* count, cudaStream_t st);
cudaError_t launch_importance_sample(const ImportanceArgs& a, cudaStream_t st);
cudaError_t launch_merge_coarse(const int32_t* rank, int64_t n_rays, int n_coarse, int n_fine, const float* src4, float* dst4, const uint32_t* src_mask,
                                uint32_t* dst_mask, int accumulate, cudaStream_t st);
cudaError_t launch_sample_rays(const SampleArgs& a, cudaStream_t st);
cudaError_t launch_render_fwd(const RenderArgs& a, cudaStream_t st);
cudaError_t launch_render_bwd(const RenderArgs& a, cudaStream_t st);
cudaError_t launch_render_loss_bwd(const RenderArgs& r, const unsigned long long* counts, long long n_rays_total, uint32_t* list,
                                   unsigned long long* list_count, float* coef_out, int count_is_zero, cudaStream_t st);
cudaError_t launch_track_rays_sample(const SampleArgs& a, const float* kf_rays, const float* c2w, float* rays_o, float* rays_d, float* target_rgb,
                                     float* target_d, cudaStream_t st);
cudaError_t launch_track_pose_update(const float* kf_rays, const float* d_rays_o, const float* d_rays_d, long long n, const cb200_track_step& s,
                                     unsigned* ticket, unsigned long long* noise_counter, cudaStream_t st);
cudaError_t launch_loss_finish(const double* sums, const long long* counts, long long n_rays_total, long long n_samples,
                               float* losses, float* coef, cudaStream_t st);
cudaError_t launch_points_fwd(const PointsArgs& a, cudaStream_t st);
cudaError_t launch_compact_live(const float* d_raw, int64_t n_points, unsigned long long* count, uint32_t* idx, cudaStream_t st);
cudaError_t launch_points_bwd_tc(const PointsArgs& a, cudaStream_t st);
cudaError_t launch_points_fwd_tc(const PointsArgs& a, cudaStream_t st);
long long fwd_ws_min_points();
void set_fwd_ws_min_points(long long v);
cudaError_t launch_points_bwd(const PointsArgs& a, cudaStream_t st);
cudaError_t launch_debug_positions(const PointsArgs& a, float* pts_out, float* x_out, cudaStream_t st);
cudaError_t launch_hash_fwd(const SceneDev& sc, const float* x, const float* table, int64_t n, float* out, cudaStream_t st);
cudaError_t launch_hash_indices(const SceneDev& sc, const float* x, int64_t n, uint32_t* idx, cudaStream_t st);
cudaError_t launch_gather_peak(const SceneDev& sc, const uint32_t* idx, int64_t n_points, const float* table, float* dtable, int mode,
                               float* sink, cudaStream_t st);
cudaError_t launch_hash_bwd(const SceneDev& sc, const float* x, const float* table, const float* dy, int64_t n, float* dtable, float* dx,
                            cudaStream_t st);
cudaError_t launch_oneblob_fwd(const float* x, int64_t n_flat, int n_bins, float* out, cudaStream_t st);
cudaError_t launch_tc_selftest(const float* A, const float* B, int K, int N, int mode, float* C, cudaStream_t st);
cudaError_t launch_oneblob_bwd(const float* x, const float* dy, int64_t n_flat, int n_bins, float* dx, cudaStream_t st);

}  // namespace cb200

// sets cb200_last_error() of the calling thread and returns `code` (api.cu); for entry points in other translation units
int cb200_set_error(int code, const char* fmt, ...);
