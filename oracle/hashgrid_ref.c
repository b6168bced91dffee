/* ORACLE (test infrastructure, not product code).
 *
 * Plain-C restatement of the integer / rounding-order-sensitive part of Co-SLAM's hot path:
 *   - tiny-cuda-nn HashGrid level metadata, corner indices and trilinear forward
 *     (tiny-cuda-nn include/tiny-cuda-nn/encodings/grid.h: grid_scale, grid_resolution, pos_fract,
 *      grid_index, coherent_prime_hash, kernel_grid; called from the reference at
 *      model/encodings.py:34-46 and model/scene_rep.py:137,166)
 *   - tiny-cuda-nn OneBlob forward (oneblob.h one_blob_subwarp_aligned, common_device.h quartic_cdf;
 *      model/encodings.py:63-71, model/scene_rep.py:141,167)
 *   - depth-guided sampling, stratified jitter, sample positions and box normalisation
 *      (model/scene_rep.py:227-258 and :186)
 * It uses a real fmaf() where tiny-cuda-nn does, and is compiled with -ffp-contract=off so that every other
 * multiply/add rounds separately like torch eager ops do.  tests/test_oracle.py cross-checks it against the
 * pure-torch shim (oracle/tcnn_shim.py) and against the reference's own scene_rep.py run in this container.
 * PARITY UNPINNED at the tinycudann boundary (no reference test or golden vector exists; see tcnn_shim.py).
 *
 * Build: make -C oracle   (-> oracle/_build/libcoslam_oracle.so)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CBO_MAX_LEVELS 32

typedef struct {
    uint32_t n_levels;
    uint32_t n_features;
    uint32_t log2_hashmap_size;
    uint32_t base_resolution;
    float per_level_scale;
    float scale[CBO_MAX_LEVELS];
    uint32_t res[CBO_MAX_LEVELS];
    uint32_t offset[CBO_MAX_LEVELS + 1];
} cbo_grid;

/* grid.h grid_scale / grid_resolution + GridEncodingTemplated ctor offset loop */
int cbo_grid_init(cbo_grid* g, uint32_t n_levels, uint32_t n_features, uint32_t log2_hashmap_size,
                  uint32_t base_resolution, float per_level_scale) {
    if (n_levels > CBO_MAX_LEVELS) return -1;
    g->n_levels = n_levels;
    g->n_features = n_features;
    g->log2_hashmap_size = log2_hashmap_size;
    g->base_resolution = base_resolution;
    g->per_level_scale = per_level_scale;
    /* exp2f / log2f := correctly rounded fp32 (fp64 evaluation rounded once) so the definition does not depend on
     * the libm; see oracle/tcnn_shim.py grid_meta for the tiny-cuda-nn device-exp2f caveat. */
    const float log2_pls = (float)log2((double)per_level_scale);
    uint32_t offset = 0;
    for (uint32_t l = 0; l < n_levels; ++l) {
        const float arg = (float)l * log2_pls;
        const float e = (float)exp2((double)arg);
        const float scale = e * (float)base_resolution - 1.0f;
        const uint32_t res = (uint32_t)ceilf(scale) + 1u;
        const uint32_t max_params = UINT32_MAX / 2;
        uint32_t n = powf((float)res, 3.0f) > (float)max_params ? max_params : res * res * res;
        n = (n + 7u) / 8u * 8u;
        const uint32_t cap = 1u << log2_hashmap_size;
        if (n > cap) n = cap;
        g->scale[l] = scale;
        g->res[l] = res;
        g->offset[l] = offset;
        offset += n;
    }
    g->offset[n_levels] = offset;
    return 0;
}

static inline uint32_t cbo_grid_index(uint32_t size, uint32_t res, const uint32_t p[3]) {
    uint32_t stride = 1, index = 0;
    for (uint32_t d = 0; d < 3 && stride <= size; ++d) {
        index += p[d] * stride;
        stride *= res;
    }
    if (size < stride) {
        index = (p[0] * 1u) ^ (p[1] * 2654435761u) ^ (p[2] * 805459861u);
    }
    return index % size;
}

/* idx_out [P, n_levels, 8] uint32 (level-local entry index), frac_out [P, n_levels, 3] float */
void cbo_hash_indices(const cbo_grid* g, const float* x, int64_t P, uint32_t* idx_out, float* frac_out) {
    for (int64_t i = 0; i < P; ++i) {
        for (uint32_t l = 0; l < g->n_levels; ++l) {
            const uint32_t size = g->offset[l + 1] - g->offset[l];
            uint32_t cell[3];
            float w[3];
            for (int d = 0; d < 3; ++d) {
                const float pos = fmaf(g->scale[l], x[i * 3 + d], 0.5f);
                const float fl = floorf(pos);
                cell[d] = (uint32_t)(int)fl;
                w[d] = pos - fl;
                frac_out[(i * g->n_levels + l) * 3 + d] = w[d];
            }
            for (uint32_t c = 0; c < 8; ++c) {
                uint32_t p[3];
                for (int d = 0; d < 3; ++d) p[d] = cell[d] + ((c >> d) & 1u);
                idx_out[(i * g->n_levels + l) * 8 + c] = cbo_grid_index(size, g->res[l], p);
            }
        }
    }
}

/* out [P, n_levels * n_features]; corner order and fma accumulation as in grid.h kernel_grid */
void cbo_hash_fwd(const cbo_grid* g, const float* x, const float* table, int64_t P, float* out) {
    const uint32_t F = g->n_features;
    for (int64_t i = 0; i < P; ++i) {
        for (uint32_t l = 0; l < g->n_levels; ++l) {
            const uint32_t size = g->offset[l + 1] - g->offset[l];
            const float* lvl = table + (size_t)g->offset[l] * F;
            uint32_t cell[3];
            float w[3];
            for (int d = 0; d < 3; ++d) {
                const float pos = fmaf(g->scale[l], x[i * 3 + d], 0.5f);
                const float fl = floorf(pos);
                cell[d] = (uint32_t)(int)fl;
                w[d] = pos - fl;
            }
            float acc[8] = {0};
            for (uint32_t c = 0; c < 8; ++c) {
                uint32_t p[3];
                float wc = 1.0f;
                for (int d = 0; d < 3; ++d) {
                    if ((c >> d) & 1u) { wc *= w[d]; p[d] = cell[d] + 1u; }
                    else { wc *= 1.0f - w[d]; p[d] = cell[d]; }
                }
                const uint32_t idx = cbo_grid_index(size, g->res[l], p);
                for (uint32_t f = 0; f < F; ++f) acc[f] = fmaf(wc, lvl[(size_t)idx * F + f], acc[f]);
            }
            for (uint32_t f = 0; f < F; ++f) out[i * g->n_levels * F + l * F + f] = acc[f];
        }
    }
}

/* common_device.h quartic_cdf */
static inline float cbo_quartic_cdf(float t, float inv_radius) {
    const float u = t * inv_radius;
    const float u2 = u * u;
    const float u4 = u2 * u2;
    return fmaxf(0.0f, fminf(1.0f, (15.0f / 16.0f) * u * (1.0f - (2.0f / 3.0f) * u2 + (1.0f / 5.0f) * u4) + 0.5f));
}

/* out [P, n_dims * n_bins] */
void cbo_oneblob_fwd(const float* x, int64_t P, uint32_t n_dims, uint32_t n_bins, float* out) {
    float left[64];
    const float B = (float)n_bins;
    for (int64_t i = 0; i < P; ++i) {
        for (uint32_t d = 0; d < n_dims; ++d) {
            const float v = x[i * n_dims + d];
            for (uint32_t b = 0; b < n_bins; ++b) {
                const float lb = (float)b / B;
                left[b] = cbo_quartic_cdf(lb - v, B) + cbo_quartic_cdf(lb - v - 1.0f, B) + cbo_quartic_cdf(lb - v + 1.0f, B);
            }
            for (uint32_t b = 0; b < n_bins; ++b) {
                float right = left[(b + 1) % n_bins];
                if (b == n_bins - 1) right += 1.0f;
                out[(i * n_dims + d) * n_bins + b] = right - left[b];
            }
        }
    }
}

/* model/scene_rep.py:227-255.  lin_guided / lin_nodepth [S_g], lin_uniform [S_u] are the torch.linspace outputs the
 * reference builds on the CPU; noise [N, S] is torch.rand(z_vals.shape) (NULL = perturb 0).  z_out [N, S]. */
void cbo_sample_z(const float* target_d, const float* noise, const float* lin_uniform, const float* lin_guided,
                  const float* lin_nodepth, int64_t N, uint32_t S_u, uint32_t S_g, float* z_out) {
    const uint32_t S = S_u + S_g;
    float* g = (float*)malloc(sizeof(float) * S_g);
    float* z = (float*)malloc(sizeof(float) * S);
    for (int64_t r = 0; r < N; ++r) {
        const float d = target_d[r];
        for (uint32_t k = 0; k < S_g; ++k) g[k] = d <= 0.0f ? lin_nodepth[k] : lin_guided[k] + d;
        /* torch.sort of cat(uniform, guided): both halves ascending -> plain 2-way merge gives the same multiset order */
        uint32_t a = 0, b = 0;
        for (uint32_t k = 0; k < S; ++k) {
            if (b >= S_g || (a < S_u && lin_uniform[a] <= g[b])) z[k] = lin_uniform[a++];
            else z[k] = g[b++];
        }
        if (noise) {
            for (uint32_t k = 0; k < S; ++k) {
                const float lower = k == 0 ? z[0] : 0.5f * (z[k] + z[k - 1]);
                const float upper = k == S - 1 ? z[S - 1] : 0.5f * (z[k + 1] + z[k]);
                z_out[r * S + k] = lower + (upper - lower) * noise[r * S + k];
            }
        } else {
            memcpy(z_out + r * S, z, sizeof(float) * S);
        }
    }
    free(g);
    free(z);
}

/* model/scene_rep.py:258 (pts = o + d * z, separate multiply and add) and :186 (x = (pts - lo) / (hi - lo)). */
void cbo_points(const float* rays_o, const float* rays_d, const float* z, const float* bb_lo, const float* bb_hi,
                int64_t N, uint32_t S, float* pts_out, float* x_out) {
    for (int64_t r = 0; r < N; ++r) {
        for (uint32_t k = 0; k < S; ++k) {
            for (int d = 0; d < 3; ++d) {
                const float m = rays_d[r * 3 + d] * z[r * S + k];
                const float p = rays_o[r * 3 + d] + m;
                if (pts_out) pts_out[(r * S + k) * 3 + d] = p;
                if (x_out) x_out[(r * S + k) * 3 + d] = (p - bb_lo[d]) / (bb_hi[d] - bb_lo[d]);
            }
        }
    }
}
