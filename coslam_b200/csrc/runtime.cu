// Native iteration runtime: one C call issues every launch of a bundle-adjustment iteration (coslam.py:336-390), so the
// host pays ~30 cudaLaunchKernel calls instead of ~30 Python -> ctypes round trips with tensor bookkeeping in between
// (1.5 ms -> the GPU time of the kernels at room0's 2048 + rays).  It only sequences the library's own public entry
// points (include/coslam_b200.h) on the caller's stream -- plus the two-source ray assembly of assemble.cu, which gathers the
// keyframe rays and the current frame's rays of a batch in one launch; no allocation, no synchronisation.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <vector>

#include "../../include/coslam_b200.h"

namespace cb200 {
cudaError_t launch_assemble_rays_fwd2(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, long long n_first,
                                      const float* cur_rays, const long long* idx_cur, const long long* slot_cur, long long n_cur,
                                      const float* poses, long long n_poses, float* rays_o, float* rays_d, float* target_rgb,
                                      float* target_d, cudaStream_t st);
cudaError_t launch_assemble_rays_bwd2(const float* kf_rays, const long long* ray_idx, const long long* pose_slot, long long n_first,
                                      const float* cur_rays, const long long* idx_cur, const long long* slot_cur, long long n_cur,
                                      const float* d_rays_o, const float* d_rays_d, long long n_poses, float* d_poses, cudaStream_t st);
}  // namespace cb200

namespace {

int cuda_rc(cudaError_t e) { return e == cudaSuccess ? 0 : (int)e; }

}  // namespace

extern "C" size_t cb200_sizeof(int which) {       // binding self-check: 0 = cb200_ba_plan, 1 = cb200_track_step, 2 = cb200_adam_param
    switch (which) {
        case 0: return sizeof(cb200_ba_plan);
        case 1: return sizeof(cb200_track_step);
        case 2: return sizeof(cb200_adam_param);
        default: return 0;
    }
}

#define CB_TRY(call)             \
    do {                         \
        int rc_ = (call);        \
        if (rc_ != 0) return rc_; \
    } while (0)

namespace {

// the coarse pass has colour and depth terms only (scene_rep.py:327-329): its two squared-residual sums join the fine ones
__global__ void add_coarse_sums_kernel(double* sums, const double* sums0) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        sums[0] += sums0[0];
        sums[1] += sums0[1];
    }
}

// rays (coslam.py:336-355) -> JointEncoding.forward up to the loss sums (scene_rep.py:297-338)
// without importance samples: rendering, loss sums, d_raw and the live-point list in one launch (cb200_render_loss_bwd), the losses
// at the start of the optimise phase
bool fused_loss_head(const cb200_ba_plan* p) { return p->n_importance == 0 && p->bwd_ws != nullptr; }

// kept rows (cb200_points_fwd_keep / cb200_points_bwd_kept): the single-pass iteration with the fused loss head, on the tensor-core
// decoder, with rows wide enough for the ray gradients when poses are optimised
bool kept_rows(const cb200_ba_plan* p) {
    return p->kept != nullptr && fused_loss_head(p) && cb200_get_decoder() == 1 && (p->kept_derivatives != 0 || p->d_rays == nullptr);
}

// per-point forward of the merged / only sample set over an optional point list
int points_fwd(const cb200_ba_plan* p, const uint32_t* list, const uint64_t* count, void* stream) {
    const int64_t n = p->n_global + p->n_cur, S = p->n_samples + p->n_importance;
    if (kept_rows(p))
        return cb200_points_fwd_keep(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z_vals, n, S, p->raw, nullptr, p->act_mask, 0,
                                     list, count, p->kept, (int)p->kept_derivatives, stream);
    return cb200_points_fwd(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z_vals, n, S, p->raw, nullptr, p->act_mask, 0, list,
                            count, stream);
}

int ba_forward(const cb200_ba_plan* p, int64_t it, bool do_rays, bool do_points, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t g = p->n_global, n = p->n_global + p->n_cur, S0 = p->n_samples, I = p->n_importance, S = S0 + I;
    int64_t* counts = reinterpret_cast<int64_t*>(p->ws);
    if (do_rays) {
    CB_TRY(cuda_rc(cudaMemsetAsync(p->ws, 0, 10 * sizeof(double), st)));
    if (p->d_rays) CB_TRY(cuda_rc(cudaMemsetAsync(p->d_rays, 0, sizeof(float) * 6 * n, st)));
    if (!p->cur_rays || !p->poses_all || !p->idx_cur || !p->slot_cur || (g > 0 && (!p->kf_rays || !p->idx_global || !p->slot_global))) return -1;
    CB_TRY(cuda_rc(cb200::launch_assemble_rays_fwd2(
        p->kf_rays, reinterpret_cast<const long long*>(p->idx_global) + it * g, reinterpret_cast<const long long*>(p->slot_global) + it * g, g,
        p->cur_rays, reinterpret_cast<const long long*>(p->idx_cur) + it * p->n_cur, reinterpret_cast<const long long*>(p->slot_cur), p->n_cur,
        p->poses_all, p->n_poses, p->rays_o, p->rays_d, p->target_rgb, p->target_d, st)));
    }
    const float* noise = p->noise ? p->noise + it * n * S0 : nullptr;
    const uint64_t seed = p->seeds ? p->seeds[it] : 0;
    if (I == 0) {
        if (do_rays)
            CB_TRY(cb200_sample_rays(p->render, p->target_d, p->lin_uniform, p->lin_guided, p->lin_nodepth, noise, seed, p->seed_dev, n, p->z_vals,
                                     counts, stream));
        if (!do_points) return 0;
    } else if (!do_points) {
        return 0;                                 // with importance samples the counts come out of the coarse pass: nothing to split
    }
    if (I > 0) {
        // coarse samples -> weights -> I extra depths per ray, merged and sorted; the mask counts belong to the merged samples
        CB_TRY(cb200_sample_rays(p->render, p->target_d, p->lin_uniform, p->lin_guided, p->lin_nodepth, noise, seed, p->seed_dev, n, p->z0,
                                 nullptr, stream));
        CB_TRY(cb200_points_fwd(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z0, n, S0, p->raw0, nullptr, p->act_mask0, 0,
                                nullptr, nullptr, stream));
        CB_TRY(cuda_rc(cudaMemsetAsync(p->sums0, 0, 4 * sizeof(double), st)));
        CB_TRY(cb200_render_fwd(p->render, p->raw0, p->z0, p->target_rgb, p->target_d, n, S0, nullptr, nullptr, nullptr, nullptr, nullptr,
                                p->weights0, p->sums0, stream));
        if (p->rank) {
            // the coarse samples sit in the merged rows with the same positions, i.e. the same network outputs: evaluate the new
            // samples only and take the coarse rows over
            CB_TRY(cb200_importance_sample_merge(p->render, p->z0, p->weights0, p->u + (p->u_stride ? it * n * I : 0), p->u_stride, p->target_d, n,
                                                 S0, I, p->z_vals, nullptr, counts, p->rank, p->new_list, p->new_count, stream));
            CB_TRY(cb200_points_fwd(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z_vals, n, S, p->raw, nullptr, p->act_mask, 0,
                                    p->new_list, p->new_count, stream));
            CB_TRY(cb200_merge_coarse(p->rank, n, S0, S, p->raw0, p->raw, p->act_mask0, p->act_mask, 0, stream));
        } else {
            CB_TRY(cb200_importance_sample(p->render, p->z0, p->weights0, p->u + (p->u_stride ? it * n * I : 0), p->u_stride, p->target_d, n, S0, I,
                                           p->z_vals, nullptr, counts, stream));
        }
    }
    if (I > 0 && p->rank) {
        // (fine forward done above)
    } else if (p->lazy_ws) {
        CB_TRY(cuda_rc(cudaMemsetAsync(p->raw, 0, sizeof(float) * 4 * n * S, st)));
        unsigned char* base = static_cast<unsigned char*>(p->lazy_ws);
        const uint64_t* count = reinterpret_cast<const uint64_t*>(base);
        const uint32_t* list = reinterpret_cast<const uint32_t*>(base + 16 + 4 * n);      // layout of cb200_lazy_workspace
        CB_TRY(cb200_lazy_prefix(p->render, p->z_vals, p->target_d, n, S, p->lazy_ws, p->lazy_ws_bytes, stream));
        CB_TRY(points_fwd(p, list, count, stream));
        CB_TRY(cb200_lazy_resolve(p->render, p->z_vals, p->target_d, p->raw, n, S, p->lazy_ws, p->lazy_ws_bytes, stream));
        CB_TRY(points_fwd(p, list, count, stream));
    } else {
        CB_TRY(points_fwd(p, nullptr, nullptr, stream));
    }
    if (fused_loss_head(p)) return 0;             // the loss sums are taken by the backward phase's fused launch
    CB_TRY(cb200_render_fwd(p->render, p->raw, p->z_vals, p->target_rgb, p->target_d, n, S, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                            p->ws + 4, stream));
    if (I > 0) {
        add_coarse_sums_kernel<<<1, 32, 0, st>>>(p->ws + 4, p->sums0);
        CB_TRY(cuda_rc(cudaGetLastError()));
    }
    return 0;
}

// losses from the (global) sums, backward into the gradient buffers, pose-gradient sums, smoothness (coslam.py:154-158, :251-274)
int ba_backward(const cb200_ba_plan* p, int64_t it, void* stream) {
    const int64_t g = p->n_global, n = p->n_global + p->n_cur, S0 = p->n_samples, I = p->n_importance, S = S0 + I;
    int64_t* counts = reinterpret_cast<int64_t*>(p->ws);
    double* sums = p->ws + 4;
    const int64_t n_total = p->n_rays_total > 0 ? p->n_rays_total : n;
    if (fused_loss_head(p)) {
        CB_TRY(cb200_render_loss_bwd(p->render, p->raw, p->z_vals, p->target_rgb, p->target_d, n, S, counts, n_total, &p->loss_w, sums,
                                     p->loss_out + 4, p->d_raw, p->bwd_ws, p->bwd_ws_bytes, 0, stream));
        if (kept_rows(p))
            CB_TRY(cb200_points_bwd_kept(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z_vals, n, S, p->d_raw, nullptr, p->act_mask,
                                         p->kept, (int)p->kept_derivatives, p->g_table, &p->g_mlp, nullptr, p->d_rays,
                                         p->d_rays ? p->d_rays + 3 * n : nullptr, 0, p->bwd_ws, 0, stream));
        else
            CB_TRY(cb200_points_bwd(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z_vals, n, S, p->d_raw, nullptr, p->act_mask,
                                    p->g_table, &p->g_mlp, nullptr, p->d_rays, p->d_rays ? p->d_rays + 3 * n : nullptr, 0, p->bwd_ws, 0, stream));
    } else {
        CB_TRY(cb200_loss_finish(sums, counts, n_total, S, p->loss_out, p->loss_out + 4, stream));
        CB_TRY(cb200_render_bwd(p->render, p->raw, p->z_vals, p->target_rgb, p->target_d, n, S, &p->loss_w, p->loss_out + 4, nullptr, nullptr,
                                p->d_raw, stream));
        if (I > 0) {                             // the coarse maps carry the colour and depth terms only
            cb200_loss_grads lw = p->loss_w;
            lw.sdf = nullptr;
            lw.fs = nullptr;
            CB_TRY(cb200_render_bwd(p->render, p->raw0, p->z0, p->target_rgb, p->target_d, n, S0, &lw, p->loss_out + 4, nullptr, nullptr, p->d_raw0,
                                    stream));
            if (p->rank)                         // ... on points the merged rows contain: one backward for both passes
                CB_TRY(cb200_merge_coarse(p->rank, n, S0, S, p->d_raw0, p->d_raw, nullptr, nullptr, 1, stream));
        }
        CB_TRY(cb200_points_bwd(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z_vals, n, S, p->d_raw, nullptr, p->act_mask,
                                p->g_table, &p->g_mlp, nullptr, p->d_rays, p->d_rays ? p->d_rays + 3 * n : nullptr, 0, p->bwd_ws, p->bwd_ws_bytes,
                                stream));
        if (I > 0 && !p->rank)
            CB_TRY(cb200_points_bwd(p->scene, p->table, &p->mlp, nullptr, p->rays_o, p->rays_d, p->z0, n, S0, p->d_raw0, nullptr, p->act_mask0,
                                    p->g_table, &p->g_mlp, nullptr, p->d_rays, p->d_rays ? p->d_rays + 3 * n : nullptr, 0, p->bwd_ws,
                                    p->bwd_ws_bytes, stream));
    }
    if (p->d_rays) {
        if (!p->d_poses) return -1;
        CB_TRY(cuda_rc(cb200::launch_assemble_rays_bwd2(
            p->kf_rays, reinterpret_cast<const long long*>(p->idx_global) + it * g, reinterpret_cast<const long long*>(p->slot_global) + it * g, g,
            p->cur_rays, reinterpret_cast<const long long*>(p->idx_cur) + it * p->n_cur, reinterpret_cast<const long long*>(p->slot_cur), p->n_cur,
            p->d_rays, p->d_rays + 3 * n, p->n_poses, p->d_poses, (cudaStream_t)stream)));
    }
    if (p->smooth_side > 0) {
        const int64_t m = p->smooth_side * p->smooth_side * p->smooth_side;
        CB_TRY(cb200_smooth_points(p->smooth_side, p->smooth_voxel, p->smooth_jitter + 3 * it, p->smooth_offset + 3 * it, p->scene->bb_lo,
                                   p->scene->bb_ext, p->smooth_x, stream));
        CB_TRY(cb200_hash_fwd(p->grid, p->smooth_x, p->table, m, p->smooth_f, stream));
        CB_TRY(cb200_smooth_tv(p->smooth_f, p->smooth_side, p->smooth_feat, p->smooth_grad_scale, p->smooth_df, p->ws + 8, stream));
        CB_TRY(cb200_hash_bwd(p->grid, p->smooth_x, p->table, p->smooth_df, m, p->g_table, nullptr, stream));
    }
    return 0;
}

// optimiser steps that are due (coslam.py:363-390)
int ba_optimise(const cb200_ba_plan* p, int64_t map_step, int zero_map_grads, int64_t pose_step, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (fused_loss_head(p)) {                     // the reported losses, from the (by now batch-global) sums
        const int64_t n = p->n_global + p->n_cur;
        CB_TRY(cb200_loss_finish(p->ws + 4, reinterpret_cast<int64_t*>(p->ws), p->n_rays_total > 0 ? p->n_rays_total : n, p->n_samples,
                                 p->loss_out, p->loss_out + 4, stream));
    }
    if (map_step > 0) {
        CB_TRY(cb200_adam_multi(p->map_params, p->n_map_params, map_step, zero_map_grads ? 1 : 0, stream));
    } else if (zero_map_grads) {
        for (int64_t k = 0; k < p->n_map_params; ++k)
            CB_TRY(cuda_rc(cudaMemsetAsync(p->map_params[k].grad, 0, sizeof(float) * p->map_params[k].n, st)));
    }
    if (pose_step > 0 && p->n_free > 0) {
        const int64_t nf = p->n_free, nr = p->n_rot;
        float* m_rot = p->pose_adam;
        float* v_rot = m_rot + nf * nr;
        float* m_tr = v_rot + nf * nr;
        float* v_tr = m_tr + 3 * nf;
        CB_TRY(cb200_pose_adam(p->rot, p->trans, m_rot, v_rot, m_tr, v_tr, p->d_poses + 16, p->poses_all + 16, nf, (int)nr, p->lr_rot, p->lr_trans,
                               pose_step, 1, stream));
        CB_TRY(cuda_rc(cudaMemsetAsync(p->d_poses, 0, sizeof(float) * 16 * p->n_poses, st)));
    }
    return 0;
}

}  // namespace

extern "C" int cb200_ba_iteration(const cb200_ba_plan* p, int64_t it, int phases, int64_t map_step, int zero_map_grads, int64_t pose_step,
                                  void* stream) {
    if (p == nullptr || it < 0) return -1;
    if (phases & CB200_BA_RAYS) CB_TRY(ba_forward(p, it, true, false, stream));
    if (phases & CB200_BA_FORWARD) CB_TRY(ba_forward(p, it, !(phases & CB200_BA_AFTER_RAYS), true, stream));
    if (phases & CB200_BA_BACKWARD) CB_TRY(ba_backward(p, it, stream));
    if (phases & CB200_BA_OPTIMISE) CB_TRY(ba_optimise(p, map_step, zero_map_grads, pose_step, stream));
    return 0;
}
#undef CB_TRY

// ---- host: CPython's random.sample(range(n), k), bit for bit ----------------------------------------------------------
// The reference draws its ray / pixel indices with Python's `random.sample` (model/keyframe.py:45-46, :100; coslam.py:131-138,
// :339-341): ~1 us per element in the interpreter, i.e. more host time per BA iteration than the GPU needs for the whole
// iteration.  This is the same algorithm (Lib/random.py `sample` + `_randbelow_with_getrandbits`, Modules/_randommodule.c
// genrand_uint32) on the caller's copy of the generator state, so the indices -- and every later draw -- are unchanged.
namespace {

struct Mt19937 {
    uint32_t* mt;      // [624]
    uint32_t& idx;     // position, 624 = regenerate
    uint32_t next() {
        constexpr int N = 624, M = 397;
        constexpr uint32_t A = 0x9908b0dfu, UP = 0x80000000u, LO = 0x7fffffffu;
        if (idx >= (uint32_t)N) {
            int kk;
            uint32_t y;
            for (kk = 0; kk < N - M; ++kk) {
                y = (mt[kk] & UP) | (mt[kk + 1] & LO);
                mt[kk] = mt[kk + M] ^ (y >> 1) ^ ((y & 1u) ? A : 0u);
            }
            for (; kk < N - 1; ++kk) {
                y = (mt[kk] & UP) | (mt[kk + 1] & LO);
                mt[kk] = mt[kk + (M - N)] ^ (y >> 1) ^ ((y & 1u) ? A : 0u);
            }
            y = (mt[N - 1] & UP) | (mt[0] & LO);
            mt[N - 1] = mt[M - 1] ^ (y >> 1) ^ ((y & 1u) ? A : 0u);
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    uint32_t below(uint32_t n, int bits) {      // _randbelow_with_getrandbits, n < 2^32
        uint32_t r = next() >> (32 - bits);
        while (r >= n) r = next() >> (32 - bits);
        return r;
    }
};

}  // namespace

extern "C" int cb200_host_random_sample(uint32_t* state, int64_t n_draws, const int64_t* n, const int64_t* k, int64_t* out) {
    if (!state || !n || !k || !out || n_draws < 0) return -1;
    if (state[624] > 624u) return -1;
    Mt19937 g{state, state[624]};
    std::vector<uint32_t> table;
    std::vector<int64_t> pool;
    for (int64_t d = 0; d < n_draws; ++d) {
        const int64_t nn = n[d], kk = k[d];
        if (kk < 0 || kk > nn || nn >= (int64_t)1 << 32) return -2;
        if (kk == 0) continue;
        int bits = 0;
        while ((nn >> bits) != 0) ++bits;
        double setsize = 21.0;
        if (kk > 5) setsize += pow(4.0, ceil(log((double)(kk * 3)) / log(4.0)));
        if ((double)nn <= setsize) {                          // pool branch
            pool.resize((size_t)nn);
            for (int64_t i = 0; i < nn; ++i) pool[(size_t)i] = i;
            for (int64_t i = 0; i < kk; ++i) {
                const uint32_t m = (uint32_t)(nn - i);
                int b = 0;
                while ((m >> b) != 0) ++b;
                const uint32_t j = g.below(m, b);
                out[i] = pool[j];
                pool[j] = pool[(size_t)(nn - i - 1)];
            }
        } else {                                              // set branch: redraw while already selected
            size_t cap = 16;
            while (cap < (size_t)kk * 4) cap <<= 1;
            table.assign(cap, 0u);                            // stores j + 1 (j < 2^32 - 1 because j < n <= 2^32 - 1)
            for (int64_t i = 0; i < kk; ++i) {
                for (;;) {
                    const uint32_t j = g.below((uint32_t)nn, bits);
                    size_t h = ((size_t)j * 2654435761u) & (cap - 1);
                    bool seen = false;
                    while (table[h] != 0u) {
                        if (table[h] == j + 1u) { seen = true; break; }
                        h = (h + 1) & (cap - 1);
                    }
                    if (seen) continue;
                    table[h] = j + 1u;
                    out[i] = (int64_t)j;
                    break;
                }
            }
        }
        out += kk;
    }
    return 0;
}
