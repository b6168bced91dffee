/* coslam_b200 -- C ABI of the B200-native (sm_100a) Co-SLAM render-and-optimise hot path.
 *
 * This is the drop-in boundary.  The reference has no FFI of its own: its hot path is
 * Python that calls the tinycudann torch bindings (model/encodings.py:34-46, 63-71) and ATen ops
 * (model/scene_rep.py:77-351, model/decoder.py:165-180, model/utils.py:89-161).  Each entry point below names
 * the reference lines it replaces.  The reference-side binding is a ctypes stub.
 *
 * Conventions
 *   - every pointer argument is a DEVICE pointer, except the small config structs (cb200_grid_cfg, cb200_grid,
 *     cb200_scene, cb200_render_cfg, cb200_mlp, cb200_mlp_grad) which are HOST structs passed by address;
 *     buffers are owned by the caller (PyTorch): the library never allocates, frees or synchronises;
 *   - `stream` is a cudaStream_t passed as void*;
 *   - return value: 0 = ok, negative = argument error, positive = cudaError_t of the launch;
 *     cb200_last_error() returns a thread-local message for the last non-zero return;
 *   - all tensors are dense row-major fp32 unless stated; counts are int64.
 */
#ifndef COSLAM_B200_H_
#define COSLAM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CB200_ABI_VERSION 19
#define CB200_MAX_LEVELS 16

/* decoder geometry served by the fused kernels (model/decoder.py:145-180 with every shipped YAML:
 * hidden_dim 32, geo_feat_dim 15, num_layers 2, hidden_dim_color 32, num_layers_color 2, bias-free) */
#define CB200_GRID_FEATS 32 /* 16 levels x 2 features          */
#define CB200_BLOB_FEATS 48 /* 3 dims x 16 bins                */
#define CB200_SDF_IN 80     /* [embed | embed_pos]             */
#define CB200_HIDDEN 32
#define CB200_SDF_OUT 16 /* 1 sdf + 15 geometric features   */
#define CB200_COL_IN 63  /* [embed_pos | geo_feat]          */

/* tinycudann HashGrid hyper-parameters as the reference passes them (model/encodings.py:34-43) */
typedef struct cb200_grid_cfg {
    uint32_t n_levels;
    uint32_t n_features_per_level;
    uint32_t log2_hashmap_size;
    uint32_t base_resolution;
    float per_level_scale;
} cb200_grid_cfg;

/* derived per-level geometry (host side, filled by cb200_grid_meta) */
typedef struct cb200_grid {
    uint32_t n_levels;
    uint32_t n_features;
    uint32_t offset[CB200_MAX_LEVELS + 1]; /* first entry of each level, in entries (not floats) */
    uint32_t res[CB200_MAX_LEVELS];
    uint32_t hashed[CB200_MAX_LEVELS]; /* 1: coherent-prime XOR hash, 0: dense stride index */
    float scale[CB200_MAX_LEVELS];
} cb200_grid;

/* scene = grid geometry + bounding box normalisation (model/scene_rep.py:182-186) + OneBlob bins */
typedef struct cb200_scene {
    cb200_grid grid;
    float bb_lo[3];
    float bb_ext[3];    /* bound[:,1] - bound[:,0], subtracted in fp32 by the caller */
    uint32_t normalise; /* config['grid']['tcnn_encoding']: 1 = apply (p - lo) / ext */
    uint32_t n_bins;    /* OneBlob bins per dimension; the fused kernels require 16 */
    /* optional (NULL = absent), DEVICE: cell-major copies of the DENSE levels, see cb200_cell_table_build.  cells: the 8 corner
     * values of every cell in one 64-byte block, read by the tensor-core forward instead of 8 separate table entries; dcells: a
     * gradient accumulator of the same layout, added to by cb200_points_bwd_kept instead of d_table's dense levels and folded
     * into d_table afterwards (cb200_cell_table_fold) */
    const void* cells;
    void* dcells;
} cb200_scene;

/* ray sampling + loss head parameters (model/scene_rep.py:220-255, 311-338; model/utils.py:89-161) */
typedef struct cb200_render_cfg {
    uint32_t n_uniform;       /* training.n_samples_d */
    uint32_t n_guided;        /* training.n_range_d   */
    uint32_t perturb;         /* training.perturb > 0 */
    uint32_t rgb_missing_nz;  /* bool(training.rgb_missing): weight of rays without valid depth */
    float trunc;              /* (float) training.trunc */
    float trunc_sc;           /* (float)(training.trunc * data.sc_factor), product taken in double like Python does */
    float depth_trunc;        /* cam.depth_trunc */
    uint32_t white_bkgd;      /* training.white_bkgd */
} cb200_render_cfg;

/* the four decoder matrices, PyTorch nn.Linear layout [out, in], fp32 */
typedef struct cb200_mlp {
    const float* w_sdf0;   /* [32, 80]  decoder.sdf_net.model.0.weight   */
    const float* w_sdf1;   /* [16, 32]  decoder.sdf_net.model.2.weight   */
    const float* w_col0;   /* [32, 63]  decoder.color_net.model.0.weight */
    const float* w_col1;   /* [3, 32]   decoder.color_net.model.2.weight */
} cb200_mlp;

typedef struct cb200_mlp_grad {
    float* w_sdf0;
    float* w_sdf1;
    float* w_col0;
    float* w_col1;
} cb200_mlp_grad;

int cb200_abi_version(void);
const char* cb200_last_error(void);
/* decoder implementation used by cb200_points_fwd / cb200_points_bwd: 1 = tcgen05 tensor-core decoder (default),
 * 0 = SIMT fp32 decoder (kept as the A/B reference of the tensor-core path).  Process-wide, not thread-safe. */
int cb200_set_decoder(int kind);
int cb200_get_decoder(void);
/* which tensor-core forward cb200_points_fwd / cb200_points_fwd_keep launch: min_points < 0 = the one-role kernel (every warp gathers,
 * then runs the decoder rounds), min_points >= 0 = the warp-specialised kernel (gather warps and decoder warps coupled by mbarriers,
 * csrc/points_ws.cu) for batches of at least min_points points.  Same outputs bit for bit.  Process-wide, not thread-safe; the
 * initial value comes from COSLAM_B200_FWD_WS (0 / 1) and COSLAM_B200_FWD_WS_MIN.  No counterpart in the reference (the decoder there
 * is model/decoder.py:145-180 on torch / tinycudann); an A/B switch like cb200_set_decoder. */
/* Cell-major copies of the dense hash-grid levels (tensor-core per-point kernels, optional).  The dense levels index their vertices
 * x + y*res + z*res^2 (tiny-cuda-nn grid.h, model/encodings.py:34-43), so the 8 corners of a cell lie in 4 different places of the
 * table and cost a lane pair 4 requests per level; in the copy a cell's corners are ONE 64-byte block
 * [cell = x + res*(y + res*z)][corner x][corner (y, z)][2 features], a lane pair reads it with one request (32 bytes per lane) and the
 * backward reduces into a gradient block of the same layout with two 16-byte reductions per lane.  Values are copies: the forward's
 * outputs do not change by a bit.  cb200_cell_table_bytes: size of either buffer (0 if no level is dense); _build: fills `cells` from
 * `table` (call again after every optimiser step on the table); _fold: adds the non-zero entries of `dcells` to d_table (the caller
 * zero-fills dcells in front of the backward).  No counterpart in the reference (tinycudann keeps one layout). */
size_t cb200_cell_table_bytes(const cb200_scene* scene);
int cb200_cell_table_build(const cb200_scene* scene, const float* table, void* cells, void* stream);
int cb200_cell_table_fold(const cb200_scene* scene, const void* dcells, float* d_table, void* stream);
int cb200_set_forward_pipeline(long long min_points);
long long cb200_get_forward_pipeline(void);
/* number of SMs of the current device (host query), used by callers to size persistent grids/workspaces */
int cb200_device_sm_count(void);

/* replaces tinycudann's GridEncoding constructor geometry (called via model/encodings.py:34-46);
 * `grid` is HOST memory */
int cb200_grid_meta(const cb200_grid_cfg* cfg, cb200_grid* grid);

/* ---- lower boundary: the two tinycudann encodings as stand-alone operators ------------------------------ */

/* tcnn.Encoding(HashGrid).forward  (model/scene_rep.py:137,166):  x [P,3] in [0,1] -> out [P, 2*n_levels] */
int cb200_hash_fwd(const cb200_grid* grid, const float* x, const float* table, int64_t n_points, float* out,
                   void* stream);
/* its backward: dy [P, 2L] -> dtable (+=, may be NULL) and dx [P,3] (=, may be NULL) */
int cb200_hash_bwd(const cb200_grid* grid, const float* x, const float* table, const float* dy, int64_t n_points,
                   float* dtable, float* dx, void* stream);
/* test hook for the bit-exact index gate: idx [P, L, 8] uint32 level-local entry indices */
int cb200_hash_indices(const cb200_grid* grid, const float* x, int64_t n_points, uint32_t* idx, void* stream);
/* tcnn.Encoding(OneBlob)  (model/scene_rep.py:141,167):  x [P,D] -> out [P, D*n_bins];  backward dy -> dx */
int cb200_oneblob_fwd(const float* x, int64_t n_points, uint32_t n_dims, uint32_t n_bins, float* out, void* stream);
int cb200_oneblob_bwd(const float* x, const float* dy, int64_t n_points, uint32_t n_dims, uint32_t n_bins, float* dx,
                      void* stream);

/* ---- keyframe ray assembly ----------------------------------------------------- */

/* Mapping / BA ray assembly (coslam.py:346-355 after model/keyframe.py:96-106): for ray i
 *   r = kf_rays[ray_idx[i]] (7 floats: camera-frame direction, rgb, depth);  T = poses[pose_slot[i]] (negative slot counts
 *   from the end: -1 = current frame);  rays_d[i] = sum_k r.dir[k] * T[:3, k];  rays_o[i] = T[:3, 3];
 *   target_rgb[i] = r[3:6];  target_d[i] = r[6].
 * kf_rays [M,7], poses [K,4,4] row-major, all device memory. */
int cb200_assemble_rays_fwd(const float* kf_rays, const int64_t* ray_idx, const int64_t* pose_slot, const float* poses,
                            int64_t n_rays, int64_t n_poses, float* rays_o, float* rays_d, float* target_rgb,
                            float* target_d, void* stream);
/* its backward with respect to the poses: d_poses [K,4,4] += (rows 0..2: d_rays_d (outer) dir | d_rays_o), row 3 untouched */
int cb200_assemble_rays_bwd(const float* kf_rays, const int64_t* ray_idx, const int64_t* pose_slot, const float* d_rays_o,
                            const float* d_rays_d, int64_t n_rays, int64_t n_poses, float* d_poses, void* stream);

/* ---- optimiser ----------------------------------------------------------------- */

/* One dense torch.optim.Adam step (amsgrad off, maximize off) over a flat fp32 parameter, as set up by
 * CoSLAM.create_optimizer / get_pose_param_optim (coslam.py:610-634, 276-283):
 *   g' = g + weight_decay * p;  m = m + (1-beta1)(g' - m);  v = beta2 v + (1-beta2) g'^2;
 *   p -= lr / (1 - beta1^t) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)
 * t = *step_dev when step_dev != NULL (device counter, for graph-captured loops) else `step` (1-based).
 * zero_grad != 0 clears g in the same pass (set_to_none=False semantics). */
int cb200_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                    float beta2, float eps, float weight_decay, const int64_t* step_dev, int64_t step, int zero_grad,
                    void* stream);

/* ---- tracking-loop glue (coslam.py:502-573 tracking_render) ---------------------------------------------------- */

/* matrix_from_tensor (coslam.py:526 -> optimization/utils.py:30-55): rot [n_rot] (3: axis-angle, 4: real-first quaternion,
 * un-normalised) + trans [3] -> c2w [4,4] row-major.  Device pointers. */
int cb200_track_pose(const float* rot, const float* trans, int n_rot, float* c2w, void* stream);

/* One tracking iteration's glue after the render backward, on the device (no host sync): weighted loss
 * (get_loss_from_ret, coslam.py:140-160), best / last pose bookkeeping with the patience counter and early stop
 * (coslam.py:552-566), pose-parameter gradients from dL/d c2w (backward of optimization/utils.py:5-55), one Adam step
 * on rot / trans (torch.optim.Adam defaults: betas (0.9, 0.999), eps 1e-8; coslam.py:276-283) unless the loop has
 * stopped, and the c2w of the updated parameters.  Host struct of DEVICE pointers. */
typedef struct cb200_track_step {
    float* rot;             /* [n_rot] in/out */
    float* trans;           /* [3] in/out */
    float* adam;            /* [2 n_rot + 6]: exp_avg(rot), exp_avg_sq(rot), exp_avg(trans), exp_avg_sq(trans) */
    int64_t* step;          /* Adam step counter, incremented here */
    const float* losses;    /* [4] rgb, depth, sdf, fs losses (cb200_loss_finish) */
    const float* d_c2w;     /* [4,4] dL/d c2w (cb200_assemble_rays_bwd) */
    float* c2w;             /* [4,4] in: the pose this iteration rendered with; out: the pose of the updated parameters */
    float* best_loss;       /* [1] */
    float* best_c2w;        /* [4,4] */
    float* last_loss;       /* [1] */
    float* last_c2w;        /* [4,4] */
    float* psnr;            /* [1], may be NULL */
    float* loss;            /* [1] this iteration's weighted loss, may be NULL */
    float* g_rot_out;       /* [n_rot] the rotation gradient (test hook), may be NULL */
    int64_t* patience;      /* [1] */
    uint8_t* alive;         /* [1] 0 once the patience counter exceeded wait_iters */
    float loss_w[4];        /* rgb_weight, depth_weight, sdf_weight, fs_weight */
    float lr_rot, lr_trans;
    int32_t n_rot;          /* 3 or 4 */
    int32_t first_iter;     /* != 0 for iteration 0 of a frame */
    int32_t may_stop;       /* wait_iters < iters */
    int32_t wait_iters;
    /* optional (sums != NULL): take the four losses from the squared-residual sums [4] and mask counts [3] of this iteration
     * (cb200_render_loss_bwd) instead of reading `losses`, which then RECEIVES them: the cb200_loss_finish arithmetic inside the update */
    const double* sums;
    const int64_t* counts;
    int64_t n_rays_total;
    int64_t n_samples;
} cb200_track_step;
int cb200_track_update(const cb200_track_step* step, void* stream);

/* matrix_to_tensor of the reference (optimization/utils.py:23-28 + pytorch3d, coslam.py:276-283) for n poses on the device:
 * c2w [n,4,4] -> rot [n, n_rot] (axis-angle, or the quaternion with non-negative real part), trans [n,3] */
int cb200_pose_params(const float* c2w, int64_t n_poses, int32_t n_rot, float* rot, float* trans, void* stream);

/* Two fusions of the tracking iteration (coslam.py:546-547 + model/scene_rep.py:227-255, and optimization/utils.py + coslam.py:552-573):
 * cb200_track_rays_sample = cb200_assemble_rays_fwd for ONE pose c2w [16] over kf_rays [N,7] (rays_o, rays_d, target_rgb [N,3],
 *   target_d [N] out) + cb200_sample_rays on those targets (z_vals, counts; noise / seed / seed_dev as there), one launch;
 * cb200_track_pose_update = cb200_assemble_rays_bwd for that pose (step->d_c2w [16] +=, zero on entry) + cb200_track_update in the
 *   CTA that finishes last (*ticket: uint32, zero on entry) + *noise_counter += 1 (may be NULL), one launch. */
int cb200_track_rays_sample(const cb200_render_cfg* cfg, const float* kf_rays, const float* c2w, const float* lin_uniform,
                            const float* lin_guided, const float* lin_nodepth, const float* noise, uint64_t seed,
                            const uint64_t* seed_dev, int64_t n_rays, float* rays_o, float* rays_d, float* target_rgb,
                            float* target_d, float* z_vals, int64_t* counts, void* stream);
int cb200_track_pose_update(const float* kf_rays, const float* d_rays_o, const float* d_rays_d, int64_t n_rays,
                            const cb200_track_step* step, uint32_t* ticket, uint64_t* noise_counter, void* stream);

/* Fresh per-frame tracking state from an initial camera-to-world matrix c2w_init [16] (device): the pose parameters the reference
 * derives on the host with pytorch3d (optimization/utils.py:23-28 matrix_to_axis_angle = quaternion_to_axis_angle(matrix_to_quaternion),
 * coslam.py:276-283; n_rot == 4: the quaternion with non-negative real part), zero Adam moments / step / patience, alive = 1, and
 * c2w = the matrix of those parameters (coslam.py:508-526).  Uses rot, trans, adam, step, patience, alive, c2w, n_rot of `step`. */
int cb200_track_init(const float* c2w_init, const cb200_track_step* step, void* stream);

/* One tracked frame -- all `n_iters` iterations of coslam.py:523-573 (pose -> rays -> sampling -> render -> weighted loss -> backward
 * to the pose -> Adam, with the best / last / patience bookkeeping of cb200_track_update) -- in ONE cooperative launch
 * (csrc/track_frame.cu): phases separated by grid-wide barriers, decoder weights staged once, early exit when the patience counter
 * trips.  kf_rays [N,7] = (camera-frame direction, rgb, depth) of the frame's sampled pixels (coslam.py:537-544); `step` as for
 * cb200_track_update (first_iter is ignored, d_c2w is not used: the gradient lives in the workspace; losses [4] receives the last
 * iteration's losses); seed / seed_dev as for cb200_sample_rays, *seed_dev is advanced by the iterations run.
 * c2w_init [16] (device, may be NULL): start the frame from this matrix as cb200_track_init does; NULL = the state in `step`.
 * workspace: cb200_track_frame_workspace(N, n_samples) bytes of device memory, 256-byte aligned; contents need not be preserved.
 * Served: tensor-core decoder, 16-level grids with 5 or 6 dense levels, n_samples <= 64 (5 dense) / 97..128 (6 dense). */
size_t cb200_track_frame_workspace(int64_t n_rays, int64_t n_samples);
int cb200_track_frame(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const cb200_render_cfg* cfg,
                      const float* lin_uniform, const float* lin_guided, const float* lin_nodepth, uint64_t seed,
                      uint64_t* seed_dev, const float* kf_rays, const float* c2w_init, int64_t n_rays, int32_t n_iters,
                      const cb200_track_step* step, void* workspace, size_t workspace_bytes, void* stream);

/* ---- bundle-adjustment glue (coslam.py:285-397 global_BA, :251-274 smoothness) -------------------------------------- */

/* matrix_from_tensor for n poses (coslam.py:317, :378): rot [n, n_rot], trans [n, 3] -> c2w [n, 4, 4] */
int cb200_pose_matrices(const float* rot, const float* trans, int64_t n_poses, int n_rot, float* c2w, void* stream);

/* pose_optimizer.step() of get_pose_param_optim (coslam.py:276-283, :374): torch.optim.Adam defaults on rot / trans of n
 * poses, gradients taken from the accumulated d_c2w [n,4,4] (cb200_assemble_rays_bwd) through the backward of
 * optimization/utils.py:5-55; writes the matrices of the updated parameters to c2w [n,4,4]; zero_grad != 0 clears d_c2w.
 * `step` is the 1-based optimiser step (host). */
int cb200_pose_adam(float* rot, float* trans, float* exp_avg_rot, float* exp_avg_sq_rot, float* exp_avg_trans,
                    float* exp_avg_sq_trans, float* d_c2w, float* c2w, int64_t n_poses, int n_rot, float lr_rot,
                    float lr_trans, int64_t step, int zero_grad, void* stream);

/* lattice of CoSLAM.smoothness (coslam.py:258-266), normalised: x [n^3, 3] = (((coords + jitter) * voxel + lo + offset) - lo)
 * / ext with torch's operation order; jitter / offset / bb_lo / bb_ext are HOST arrays of 3 floats.  Row p of x is the lattice
 * node (ix, iy, iz) with p = (iz * n + iy) * n + ix (x fastest: the term sums over the whole lattice, and this order makes
 * neighbouring threads touch neighbouring table entries). */
int cb200_smooth_points(int64_t n_side, float voxel, const float* jitter, const float* offset, const float* bb_lo,
                        const float* bb_ext, float* x, void* stream);

/* total variation of the lattice features (coslam.py:268-270): f [n,n,n,F] (cb200_hash_fwd on the lattice) ->
 * tv_sum (double, +=, may be NULL) = tv_x + tv_y + tv_z,  d_f [n,n,n,F] (=) = grad_scale * d tv / d f.
 * The caller passes grad_scale = smooth_weight / smooth_pts^3 and feeds d_f to cb200_hash_bwd.  F must be a multiple of 4. */
int cb200_smooth_tv(const float* f, int64_t n_side, int64_t n_feat, float grad_scale, float* d_f, double* tv_sum,
                    void* stream);

/* ---- upper boundary: the fused JointEncoding path -------------------------------------------------------- */

/* render_rays sampling (model/scene_rep.py:227-255) + the mask counts of get_masks (model/utils.py:103-115)
 * and of valid_depth_mask (model/scene_rep.py:319).
 *   target_d [N]; lin_uniform [n_uniform], lin_guided [n_guided], lin_nodepth [n_guided] are the three
 *   torch.linspace constants of :227,:239,:242 (device copies);
 *   noise [N,S] = the torch.rand of :255, or NULL to draw Philox noise in-kernel from (seed, ray, sample);
 *   seed_dev: optional device counter ADDED to seed when the kernel runs (a CUDA-graph replay then draws fresh noise);
 *   z_vals [N,S] out;  counts [3] int64, ACCUMULATED (+=): n_front, n_sdf, n_valid_depth  (may be NULL). */
int cb200_sample_rays(const cb200_render_cfg* cfg, const float* target_d, const float* lin_uniform,
                      const float* lin_guided, const float* lin_nodepth, const float* noise, uint64_t seed,
                      const uint64_t* seed_dev, int64_t n_rays, float* z_vals, int64_t* counts, void* stream);

/* Importance resampling of render_rays (model/scene_rep.py:265-272) with sample_pdf (model/utils.py:32-71):
 *   z_vals [N,S] ascending coarse samples, weights [N,S] their rendering weights (cb200_render_fwd);
 *   u: the uniforms of utils.py:53 as [N,I] (u_stride = I), or ONE row [I] shared by all rays (u_stride = 0) holding the
 *   det=True linspace of utils.py:50;   z_out [N,S+I] = sort(cat(z_vals, samples)) (scene_rep.py:271);
 *   z_std [N] = std(samples, unbiased=False) (scene_rep.py:292), may be NULL;
 *   target_d [N] + counts [3] (ACCUMULATED n_front, n_sdf, n_valid_depth of the merged samples, as in cb200_sample_rays),
 *   both may be NULL.  Requires 3 <= S, 1 <= I, S + I <= 256. */
int cb200_importance_sample(const cb200_render_cfg* cfg, const float* z_vals, const float* weights, const float* u,
                            int64_t u_stride, const float* target_d, int64_t n_rays, int64_t n_samples,
                            int64_t n_importance, float* z_out, float* z_std, int64_t* counts, void* stream);

/* The same, plus where the samples went: rank [N,S] int32 = position of coarse sample s in its merged row; new_list [N*I] uint32 =
 * flat indices (ray * (S+I) + position) of the NEW samples, *new_count = N * I (both may be NULL) -- a point list for cb200_points_fwd.
 * The coarse samples keep their positions in space, hence their network outputs: the fine pass of the autograd-free loops evaluates
 * only the new points and takes the coarse ones over with cb200_merge_coarse (accumulate = 0: dst[ray, rank[ray, s]] = src[ray, s] for
 * raw [.,4] and, when given, the activation masks [.,2]); in the backward the coarse pass's d_raw is ADDED to the merged rows'
 * (accumulate = 1), so one cb200_points_bwd over the merged samples serves both passes (model/scene_rep.py:262-295, :327-329). */
int cb200_importance_sample_merge(const cb200_render_cfg* cfg, const float* z_vals, const float* weights, const float* u,
                                  int64_t u_stride, const float* target_d, int64_t n_rays, int64_t n_samples,
                                  int64_t n_importance, float* z_out, float* z_std, int64_t* counts, int32_t* rank,
                                  uint32_t* new_list, uint64_t* new_count, void* stream);
int cb200_merge_coarse(const int32_t* rank, int64_t n_rays, int64_t n_coarse, int64_t n_fine, const float* src, float* dst,
                       const uint32_t* src_mask, uint32_t* dst_mask, int accumulate, void* stream);

/* run_network / query_color_sdf / query_sdf forward (model/scene_rep.py:126-191 + model/decoder.py:165-180):
 * positions are either explicit (pts [P,3], rays_o == NULL) or pts = rays_o + rays_d * z (scene_rep.py:258,
 * separate multiply and add) with P = n_rays * n_samples.
 *   raw   [P,4]  (rgb logits, sdf)            -- may be NULL
 *   geo   [P,15] geometric feature            -- may be NULL
 *   act_mask [P,2] uint32: ReLU sign bits of the two hidden layers, saved for the backward -- may be NULL
 *   sdf_only != 0 skips the colour network (query_sdf).
 *   point_list / point_count (device, optional, tensor-core decoder): process only the listed point indices. */
int cb200_points_fwd(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts,
                     const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays,
                     int64_t n_samples, float* raw, float* geo, uint32_t* act_mask, int sdf_only,
                     const uint32_t* point_list, const uint64_t* point_count, void* stream);

/* Lazy training forward: the step only needs the SDF of the samples up to the end of the truncation band (loss masks,
 * model/utils.py:103-109) and, for rendering, of the samples below z[first sign change] + T (sdf2weights mask,
 * scene_rep.py:89-95) -- both prefixes of a ray's ascending samples.  Everything else keeps raw = 0 (the caller zero-fills
 * raw), which the rendering mask removes exactly as in the reference, so outputs and gradients are unchanged.
 *   workspace (device): [count u64 | pad u64 | k_done int32 x N | list u32 x N*S], cb200_lazy_workspace bytes;
 *   cb200_lazy_prefix  fills `list`/`count` with the loss prefix of every ray (all samples where target_d <= 0);
 *   after cb200_points_fwd(point_list = list, point_count = count) on it, cb200_lazy_resolve rewrites `list`/`count`
 *   with the samples still missing (up to the cut-off of the first sign change, or the rest of the ray if the prefix has
 *   none); a second cb200_points_fwd on that list completes the forward.  Tensor-core decoder only. */
size_t cb200_lazy_workspace(int64_t n_rays, int64_t n_samples);
int cb200_lazy_prefix(const cb200_render_cfg* cfg, const float* z_vals, const float* target_d, int64_t n_rays,
                      int64_t n_samples, void* workspace, size_t workspace_bytes, void* stream);
int cb200_lazy_resolve(const cb200_render_cfg* cfg, const float* z_vals, const float* target_d, const float* raw,
                       int64_t n_rays, int64_t n_samples, void* workspace, size_t workspace_bytes, void* stream);

/* its backward.  d_raw [P,4] (d_geo [P,15] optional, NULL = zero).  Outputs, each optional (NULL = skip):
 *   dtable (+=)  [n_params]     d_mlp (+=) four matrices
 *   d_pts [P,3] (=)             d_rays_o, d_rays_d [N,3] (+=, ray mode only)
 * act_mask is the array written by cb200_points_fwd on the same inputs.
 * workspace: optional device scratch of at least cb200_points_bwd_workspace(P) bytes.  When given (ray mode, tensor-core
 * decoder) the points whose d_raw is exactly zero -- samples behind the truncation band carry no loss and no rendering
 * weight -- are compacted away before the per-point kernel runs (they contribute exact zeros to every output).
 * workspace != NULL with workspace_bytes == 0: the workspace already holds that list (cb200_render_loss_bwd). */
size_t cb200_points_bwd_workspace(int64_t n_points);
int cb200_points_bwd(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts,
                     const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays,
                     int64_t n_samples, const float* d_raw, const float* d_geo, const uint32_t* act_mask,
                     float* dtable, const cb200_mlp_grad* d_mlp, float* d_pts, float* d_rays_o, float* d_rays_d,
                     int sdf_only, void* workspace, size_t workspace_bytes, void* stream);

/* The pair that does not gather twice (tensor-core decoder, weight-gradient backward: d_mlp must be given).  cb200_points_bwd reads
 * the 128 table entries of every point a second time, for the features (operand of the weight-gradient products) and for the position
 * gradient (d feature / dx is a combination of the corner values).  Here the forward keeps what the backward needs, per point it
 * evaluates (rows of other points are left untouched), and the backward's level loop is index arithmetic + red.v2 only:
 *   with_derivatives == 0: feat [P,32] fp32, the tcnn.Encoding(HashGrid) output (level-major, 2 per level) -- 128 B per point of
 *       streamed HBM traffic instead of 1 KB of L2 gathers.  No position gradient (d_pts, d_rays_o, d_rays_d must be NULL): the step
 *       of model/scene_rep.py:297-351 on fixed rays (first-frame mapping, mapping on given poses).
 *   with_derivatives != 0: feat [P,128] = [features | d/dx | d/dy | d/dz of them] (x = the normalised position the encoding sees);
 *       the position gradient is d embed . those rows, so d_pts / d_rays_o / d_rays_d work as in cb200_points_bwd (512 B per point).
 * Every point with non-zero d_raw must have been evaluated by the forward that wrote `feat`.  Results equal cb200_points_fwd /
 * cb200_points_bwd up to fp32 summation order. */
int cb200_points_fwd_keep(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts,
                          const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays,
                          int64_t n_samples, float* raw, float* geo, uint32_t* act_mask, int sdf_only,
                          const uint32_t* point_list, const uint64_t* point_count, float* feat, int with_derivatives,
                          void* stream);
int cb200_points_bwd_kept(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* pts,
                          const float* rays_o, const float* rays_d, const float* z_vals, int64_t n_rays,
                          int64_t n_samples, const float* d_raw, const float* d_geo, const uint32_t* act_mask,
                          const float* feat, int with_derivatives, float* dtable, const cb200_mlp_grad* d_mlp, float* d_pts,
                          float* d_rays_o, float* d_rays_d, int sdf_only, void* workspace, size_t workspace_bytes,
                          void* stream);

/* raw2outputs + sdf2weights (model/scene_rep.py:77-124) and, when loss_sums != NULL, the loss head
 * (model/scene_rep.py:319-338, model/utils.py:138-161).
 *   raw [N,S,4], z_vals [N,S], target_rgb [N,3] / target_d [N] (only for the loss head)
 *   outputs (each may be NULL): rgb [N,3], depth [N], depth_var [N], disp [N], acc [N], weights [N,S]
 *   loss_sums [4] double, ACCUMULATED: sum of squared residuals of rgb, depth, sdf, free-space terms. */
int cb200_render_fwd(const cb200_render_cfg* cfg, const float* raw, const float* z_vals, const float* target_rgb,
                     const float* target_d, int64_t n_rays, int64_t n_samples, float* rgb, float* depth,
                     float* depth_var, float* disp, float* acc, float* weights, double* loss_sums, void* stream);

/* finishes the four losses from the (possibly all-reduced) sums and counts:
 *   losses [4] = rgb_loss, depth_loss, sdf_loss, fs_loss;  coef [4] = d loss_k / d (sum of squares)_k, saved for
 *   the backward.  totals [2] int64 = global number of rays and of samples per ray (host-known but kept on the
 *   device so that a sharded run can pass global values without a sync). */
int cb200_loss_finish(const double* loss_sums, const int64_t* counts, int64_t n_rays_total, int64_t n_samples,
                      float* losses, float* coef, void* stream);

/* upstream gradients of the four losses: DEVICE pointers to 0-d fp32 tensors, each may be NULL (= no gradient) */
typedef struct cb200_loss_grads {
    const float* rgb;
    const float* depth;
    const float* sdf;
    const float* fs;
} cb200_loss_grads;

/* backward of cb200_render_fwd (+ loss head).  g_loss: upstream gradients of the four losses (host struct of device
 * pointers, may be NULL), coef [4] from cb200_loss_finish, g_rgb [N,3] / g_depth [N] upstream gradients of the
 * rendered maps (may be NULL).  d_raw [N,S,4] out (=). */
int cb200_render_bwd(const cb200_render_cfg* cfg, const float* raw, const float* z_vals, const float* target_rgb,
                     const float* target_d, int64_t n_rays, int64_t n_samples, const cb200_loss_grads* g_loss,
                     const float* coef, const float* g_rgb, const float* g_depth, float* d_raw, void* stream);

/* Loss head forward AND backward in one pass (the autograd-free loops): per ray the rendering, its terms of the four squared-residual
 * sums (loss_sums [4] double, += ; may be NULL), and d_raw [N,S,4] (=) from the loss weights g_loss (device scalars) -- d loss / d sum
 * depends only on the batch-global mask counts `counts` [3] (cb200_sample_rays / cb200_importance_sample), so no cb200_loss_finish
 * has to run in between; coef [4] (may be NULL) receives those normalisers.  n_rays_total: rays of the whole batch (all ranks).
 * list_workspace (may be NULL): cb200_points_bwd_workspace(N * S) bytes that receive the list of the points with non-zero d_raw;
 * pass the same pointer to cb200_points_bwd with workspace_bytes = 0 and it skips its own compaction pass;
 * list_count_is_zero != 0: the caller has cleared the first 8 bytes of the workspace already (no memset is issued).
 * Replaces cb200_render_fwd(loss_sums) + cb200_loss_finish + cb200_render_bwd (+ the compaction) of model/scene_rep.py:311-338 in those loops. */
int cb200_render_loss_bwd(const cb200_render_cfg* cfg, const float* raw, const float* z_vals, const float* target_rgb,
                          const float* target_d, int64_t n_rays, int64_t n_samples, const int64_t* counts,
                          int64_t n_rays_total, const cb200_loss_grads* g_loss, double* loss_sums, float* coef, float* d_raw,
                          void* list_workspace, size_t list_workspace_bytes, int list_count_is_zero, void* stream);

/* ---- native iteration runtime of global_BA (coslam.py:336-390) ------------------------------------------------------ */

typedef struct cb200_adam_param {
    float* param;
    float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int64_t n;
    float lr, beta1, beta2, eps, weight_decay;
} cb200_adam_param;

/* cb200_adam_step for several parameter tensors in ONE launch (HOST array of descriptors; all share the 1-based `step`) */
int cb200_adam_multi(const cb200_adam_param* params, int64_t n_params, int64_t step, int zero_grad, void* stream);

/* Everything one bundle-adjustment call needs, set up once by the host; cb200_ba_iteration then issues all launches of
 * iteration `it` (ray assembly from the keyframe store and the current frame, sampling, per-point forward -- lazily when
 * lazy_ws is given --, rendering + losses, backward into the gradient buffers, pose-gradient segment sums, the smoothness
 * term, and the optimiser steps that are due) on `stream` without touching the host again.  All pointers are device memory
 * unless marked HOST. */
typedef struct cb200_ba_plan {
    const cb200_scene* scene;          /* HOST */
    const cb200_render_cfg* render;    /* HOST */
    const cb200_grid* grid;            /* HOST: geometry of the hash grid for the smoothness term */
    const float* table;
    cb200_mlp mlp;
    const float *lin_uniform, *lin_guided, *lin_nodepth;
    /* rays */
    const float* kf_rays;              /* [M,7] keyframe store */
    const int64_t* idx_global;         /* [iters, n_global] rows of kf_rays */
    const int64_t* slot_global;        /* [iters, n_global] keyframe slots */
    int64_t n_global;
    const float* cur_rays;             /* [H*W,7] current frame */
    const int64_t* idx_cur;            /* [iters, n_cur] */
    const int64_t* slot_cur;           /* [n_cur], all -1 */
    int64_t n_cur;
    float* poses_all;                  /* [n_poses,4,4] */
    int64_t n_poses;
    int64_t n_samples;
    int64_t n_rays_total;              /* rays of the GLOBAL batch when this rank holds a shard (0: N) */
    const float* noise;                /* [iters, N, S] host-drawn jitter, or NULL: device Philox with seeds[it] */
    const uint64_t* seeds;             /* HOST [iters] */
    const uint64_t* seed_dev;          /* optional device counter added to the seed */
    /* per-iteration buffers, N = n_global + n_cur */
    float *rays_o, *rays_d, *target_rgb, *target_d;      /* [N,3] x3, [N] */
    float* z_vals;                     /* [N,S] */
    float* raw;                        /* [N,S,4] */
    uint32_t* act_mask;                /* [N,S,2] */
    float* d_raw;                      /* [N,S,4] */
    void* lazy_ws;                     /* cb200_lazy_workspace bytes, or NULL: plain forward */
    size_t lazy_ws_bytes;
    void* bwd_ws;                      /* cb200_points_bwd_workspace bytes */
    size_t bwd_ws_bytes;
    double* ws;                        /* [10]: 4 int64 counts | 4 loss sums | tv sum | pad */
    float* loss_out;                   /* [8]: 4 losses | 4 d loss / d sum */
    cb200_loss_grads loss_w;           /* device scalars: the four loss weights */
    /* hierarchical pass (model/scene_rep.py:262-277, :327-329); n_importance == 0 disables it.  n_samples is then the number of
     * COARSE samples; z_vals / raw / act_mask / d_raw hold n_samples + n_importance samples per ray */
    int64_t n_importance;
    float* z0;                         /* [N,S] coarse samples */
    float* raw0;                       /* [N,S,4] */
    uint32_t* act_mask0;               /* [N,S,2] */
    float* weights0;                   /* [N,S] coarse rendering weights */
    float* d_raw0;                     /* [N,S,4] */
    const float* u;                    /* uniforms of sample_pdf: [iters,N,I] (u_stride = I) or one row [I] (u_stride = 0) */
    int64_t u_stride;
    double* sums0;                     /* [4] scratch: loss sums of the coarse pass (colour and depth are added to ws) */
    /* gradients */
    float* g_table;
    cb200_mlp_grad g_mlp;
    float* d_rays;                     /* [2,N,3], or NULL when no pose is optimised */
    float* d_poses;                    /* [n_poses,4,4] accumulated between pose steps */
    /* smoothness (coslam.py:251-274); n_side == 0 disables it */
    int64_t smooth_side, smooth_feat;
    float smooth_voxel, smooth_grad_scale;
    const float* smooth_jitter;        /* HOST [iters,3] */
    const float* smooth_offset;        /* HOST [iters,3] */
    float *smooth_x, *smooth_f, *smooth_df;
    /* optimisers */
    const cb200_adam_param* map_params;   /* HOST [n_map_params] */
    int64_t n_map_params;
    float *rot, *trans, *pose_adam;    /* [n_free, n_rot], [n_free,3], Adam state [2 n_free (n_rot + 3)] as in mapping.py */
    int64_t n_free;
    int32_t n_rot;
    float lr_rot, lr_trans;
    /* hierarchical pass, optional (rank != NULL): scratch of cb200_importance_sample_merge -- rank [N,S] int32, new_list [N*I] uint32,
     * new_count [1] -- so that the fine pass evaluates the new samples only and ONE per-point backward serves both passes */
    int32_t* rank;
    uint32_t* new_list;
    uint64_t* new_count;
    /* kept rows, optional (n_importance == 0): [N*S, 32] fp32, or [N*S, 128] when kept_derivatives != 0 (required when d_rays != NULL);
     * the iteration then runs cb200_points_fwd_keep / cb200_points_bwd_kept */
    float* kept;
    int64_t kept_derivatives;
} cb200_ba_plan;

/* phases: which part of iteration `it` to issue.  One GPU: all three in one call.  Rays sharded over several GPUs
 * (SURVEY.md section 8e): FORWARD, then the caller all-reduces ws[0..7] (mask counts and loss sums are batch-global,
 * model/utils.py:111-115), BACKWARD, then the caller all-reduces the gradient buffers that an optimiser step is about to
 * consume, OPTIMISE.
 * map_step / pose_step: 1-based optimiser step to take after this iteration's backward, 0 = none;
 * zero_map_grads != 0 clears the map gradients afterwards (map_optimizer.zero_grad(), coslam.py:370).
 * Without importance samples the mask counts are complete after the SAMPLING, and the backward needs only them (the loss sums are
 * for the reported losses): a sharded caller issues CB200_BA_RAYS (ray assembly + sampling), starts the all-reduce of ws[0..2]
 * asynchronously, issues CB200_BA_FORWARD | CB200_BA_AFTER_RAYS (the per-point forward, which hides the collective), waits, issues
 * BACKWARD (cb200_render_loss_bwd: sums, d_raw and the live-point list in one launch), all-reduces ws[4..7] and the gradients that are
 * due, OPTIMISE (which starts with cb200_loss_finish). */
#define CB200_BA_FORWARD 1
#define CB200_BA_BACKWARD 2
#define CB200_BA_OPTIMISE 4
#define CB200_BA_ALL 7
#define CB200_BA_RAYS 8        /* ray assembly + sampling only */
#define CB200_BA_AFTER_RAYS 16 /* with FORWARD: the rays of this iteration were issued by an earlier CB200_BA_RAYS call */
int cb200_ba_iteration(const cb200_ba_plan* plan, int64_t it, int phases, int64_t map_step, int zero_map_grads,
                       int64_t pose_step, void* stream);

/* HOST: `random.sample(range(n[d]), k[d])` for d = 0 .. n_draws-1 with CPython's exact algorithm (Lib/random.py sample /
 * _randbelow_with_getrandbits over the MT19937 of Modules/_randommodule.c), replacing the reference's Python-level draws of
 * model/keyframe.py:45-46, :100 and coslam.py:131-138, :339-341.  state [625]: the generator state of `random.getstate()[1]`
 * (624 words + position), updated in place; out: the samples, concatenated.  Requires n < 2^32.  No CUDA involved. */
int cb200_host_random_sample(uint32_t* state, int64_t n_draws, const int64_t* n, const int64_t* k, int64_t* out);

/* sizeof of the host structs above, for the bindings' layout self-check: 0 = cb200_ba_plan, 1 = cb200_track_step,
 * 2 = cb200_adam_param */
size_t cb200_sizeof(int which);

/* ---- allocation-free variant (NO counterpart in the reference) ------------- */

/* Largest n_rays * n_samples one cb200_points_bwd_noalloc call accepts: its scratch is a file-scope device array of
 * 256 B + 4 B per point = 4 MiB + 256 B, held by every GPU that loads the library until exit. */
#define CB200_NOALLOC_MAX_POINTS (1 << 20)
int64_t cb200_points_bwd_noalloc_max_points(void);

/* cb200_points_bwd in ray mode for callers that cannot allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc): same inputs, outputs (all caller-owned, accumulated into like cb200_points_bwd) and
 * arithmetic, but the live-point compaction scratch is the library's static device array instead of a caller workspace.
 * Calls no cudaMalloc / cudaFree / allocating routine.  Returns -1 with a message when n_rays * n_samples exceeds
 * CB200_NOALLOC_MAX_POINTS.  The array is shared by every call on a device and keeps the last call's contents (the
 * counter is cleared on `stream` at the start of each call): calls that overlap on the GPU race -- serialise them. */
int cb200_points_bwd_noalloc(const cb200_scene* scene, const float* table, const cb200_mlp* mlp, const float* rays_o,
                             const float* rays_d, const float* z_vals, int64_t n_rays, int64_t n_samples, const float* d_raw,
                             const uint32_t* act_mask, float* dtable, const cb200_mlp_grad* d_mlp, float* d_rays_o,
                             float* d_rays_d, int sdf_only, void* stream);

/* HOST routine: `n` fp32 uniforms exactly as `torch.rand(n)` draws them from torch's default CPU generator (model/scene_rep.py:255
 * `torch.rand(z_vals.shape)`), on the caller's copy of the generator state: `state` = the 5 056 bytes of torch.get_rng_state(),
 * updated in place (write it back with torch.set_rng_state and every later draw is unchanged).  at::mt19937 + the 24-bit
 * uniform_real transformation in a tight loop instead of one virtual call per element.  Returns 0, or -1 for a state it does not understand. */
int cb200_host_torch_rand(uint8_t* state, int64_t n, float* out);
/* The same stream produced on the DEVICE for large batches (scene_rep.py:255 then `.to(device)`: here the values never cross PCIe).
 * The Mersenne-Twister recurrence is sequential, so the HOST still walks it -- cb200_host_torch_rand_segments advances `state` as if n
 * floats had been drawn (the twist without the outputs, a fifth of the time) and records the 624 state words in front of every
 * `blocks_per_segment`-th regeneration into `checkpoints` ([capacity][624] uint32, host; -2 if more are needed), returning the number
 * of segments, the outputs the block current at entry still serves (`head`) and its read index.  cb200_torch_rand_device takes a
 * DEVICE copy of the checkpoints and writes the n floats with one CTA per segment.  Together: out == torch.rand(n).to(device) bit for
 * bit, generator state as after that call.  cb200_host_torch_rand_skip: the state advance alone. */
int cb200_host_torch_rand_segments(uint8_t* state, int64_t n, int64_t blocks_per_segment, uint32_t* checkpoints, int64_t capacity,
                                   int64_t* n_segments, int64_t* head, int32_t* head_next);
int cb200_torch_rand_device(const uint32_t* checkpoints, int64_t n_segments, int64_t blocks_per_segment, int64_t head, int32_t head_next,
                            int64_t n, float* out, void* stream);
int cb200_host_torch_rand_skip(uint8_t* state, int64_t n);

/* test hook for the bit-exact sample-position gate: pts [P,3] and normalised x [P,3] exactly as the fused
 * kernels compute them (model/scene_rep.py:258 and :186) */
int cb200_debug_positions(const cb200_scene* scene, const float* rays_o, const float* rays_d, const float* z_vals,
                          int64_t n_rays, int64_t n_samples, float* pts, float* x, void* stream);

/* measurement hook (tools/l2_gather_peak.py): float2 gathers (mode bit 0), red.add.v2.f32 reductions (bit 1) or both over
 * the corner list idx [P, L, 8] of cb200_hash_indices, lane pair per point like the fused kernels, no other arithmetic --
 * the measured L2 gather / reduction ceiling the roofline fractions of points_fwd / points_bwd are quoted against.
 * Optional fields of `mode` (tools/level_probe.py): bits 4-19 level mask (0 = all levels); bits 20-23 log2(copies) and bits 24-28
 * n of an experiment that sends the reductions of the first n levels to copy (SM id % copies) of those levels' rows, stored
 * behind the table gradient (dtable then holds n_params + copies * 2 * offset[n] floats) */
int cb200_debug_gather_peak(const cb200_grid* grid, const uint32_t* idx, int64_t n_points, const float* table, float* dtable,
                            int mode, float* sink, void* stream);

/* test hook for the tcgen05 decoder plumbing (one 128-row tile, bf16 hi/lo split, three MMAs per reduction step):
 *   mode 0: c[128,n] = a[128,k] * b[n,k]^T     mode 1: c[128,n] = a[128,128]^T * b[128,n]     (fp32 row-major) */
int cb200_selftest_mma(const float* a, const float* b, int k, int n, int mode, float* c, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* COSLAM_B200_H_ */
