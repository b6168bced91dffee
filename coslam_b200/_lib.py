"""ctypes binding of libcoslam_b200.so (include/coslam_b200.h).

PyTorch is plumbing here: it owns device memory and streams; every call below passes raw device pointers and the
current CUDA stream handle across the C ABI.  There is NO fallback: if the library is missing or a call fails, a
RuntimeError is raised.
"""
import ctypes
import os

import torch

from . import build as _build

MAX_LEVELS = 16
ABI_VERSION = 19

c_f32p = ctypes.c_void_p
c_stream = ctypes.c_void_p


class GridCfg(ctypes.Structure):
    _fields_ = [('n_levels', ctypes.c_uint32), ('n_features_per_level', ctypes.c_uint32),
                ('log2_hashmap_size', ctypes.c_uint32), ('base_resolution', ctypes.c_uint32),
                ('per_level_scale', ctypes.c_float)]


class Grid(ctypes.Structure):
    _fields_ = [('n_levels', ctypes.c_uint32), ('n_features', ctypes.c_uint32),
                ('offset', ctypes.c_uint32 * (MAX_LEVELS + 1)), ('res', ctypes.c_uint32 * MAX_LEVELS),
                ('hashed', ctypes.c_uint32 * MAX_LEVELS), ('scale', ctypes.c_float * MAX_LEVELS)]


class Scene(ctypes.Structure):
    _fields_ = [('grid', Grid), ('bb_lo', ctypes.c_float * 3), ('bb_ext', ctypes.c_float * 3),
                ('normalise', ctypes.c_uint32), ('n_bins', ctypes.c_uint32), ('cells', ctypes.c_void_p), ('dcells', ctypes.c_void_p)]


class RenderCfg(ctypes.Structure):
    _fields_ = [('n_uniform', ctypes.c_uint32), ('n_guided', ctypes.c_uint32), ('perturb', ctypes.c_uint32),
                ('rgb_missing_nz', ctypes.c_uint32), ('trunc', ctypes.c_float), ('trunc_sc', ctypes.c_float),
                ('depth_trunc', ctypes.c_float), ('white_bkgd', ctypes.c_uint32)]


class Mlp(ctypes.Structure):
    _fields_ = [('w_sdf0', ctypes.c_void_p), ('w_sdf1', ctypes.c_void_p), ('w_col0', ctypes.c_void_p),
                ('w_col1', ctypes.c_void_p)]


class TrackStep(ctypes.Structure):
    """cb200_track_step (include/coslam_b200.h)."""
    _fields_ = [(n, ctypes.c_void_p) for n in ('rot', 'trans', 'adam', 'step', 'losses', 'd_c2w', 'c2w', 'best_loss', 'best_c2w',
                                               'last_loss', 'last_c2w', 'psnr', 'loss', 'g_rot_out', 'patience', 'alive')] + \
               [('loss_w', ctypes.c_float * 4), ('lr_rot', ctypes.c_float), ('lr_trans', ctypes.c_float), ('n_rot', ctypes.c_int32),
                ('first_iter', ctypes.c_int32), ('may_stop', ctypes.c_int32), ('wait_iters', ctypes.c_int32),
                ('sums', ctypes.c_void_p), ('counts', ctypes.c_void_p), ('n_rays_total', ctypes.c_int64), ('n_samples', ctypes.c_int64)]


class MlpGrad(ctypes.Structure):
    _fields_ = [('w_sdf0', ctypes.c_void_p), ('w_sdf1', ctypes.c_void_p), ('w_col0', ctypes.c_void_p), ('w_col1', ctypes.c_void_p)]


class LossGradPtrs(ctypes.Structure):
    _fields_ = [('rgb', ctypes.c_void_p), ('depth', ctypes.c_void_p), ('sdf', ctypes.c_void_p), ('fs', ctypes.c_void_p)]


class AdamParam(ctypes.Structure):
    """cb200_adam_param."""
    _fields_ = [('param', ctypes.c_void_p), ('grad', ctypes.c_void_p), ('exp_avg', ctypes.c_void_p), ('exp_avg_sq', ctypes.c_void_p),
                ('n', ctypes.c_int64), ('lr', ctypes.c_float), ('beta1', ctypes.c_float), ('beta2', ctypes.c_float),
                ('eps', ctypes.c_float), ('weight_decay', ctypes.c_float)]


class BaPlan(ctypes.Structure):
    """cb200_ba_plan (include/coslam_b200.h), field for field."""
    _fields_ = [('scene', ctypes.c_void_p), ('render', ctypes.c_void_p), ('grid', ctypes.c_void_p), ('table', ctypes.c_void_p),
                ('mlp', Mlp), ('lin_uniform', ctypes.c_void_p), ('lin_guided', ctypes.c_void_p), ('lin_nodepth', ctypes.c_void_p),
                ('kf_rays', ctypes.c_void_p), ('idx_global', ctypes.c_void_p), ('slot_global', ctypes.c_void_p), ('n_global', ctypes.c_int64),
                ('cur_rays', ctypes.c_void_p), ('idx_cur', ctypes.c_void_p), ('slot_cur', ctypes.c_void_p), ('n_cur', ctypes.c_int64),
                ('poses_all', ctypes.c_void_p), ('n_poses', ctypes.c_int64), ('n_samples', ctypes.c_int64), ('n_rays_total', ctypes.c_int64),
                ('noise', ctypes.c_void_p), ('seeds', ctypes.c_void_p), ('seed_dev', ctypes.c_void_p),
                ('rays_o', ctypes.c_void_p), ('rays_d', ctypes.c_void_p), ('target_rgb', ctypes.c_void_p), ('target_d', ctypes.c_void_p),
                ('z_vals', ctypes.c_void_p), ('raw', ctypes.c_void_p), ('act_mask', ctypes.c_void_p), ('d_raw', ctypes.c_void_p),
                ('lazy_ws', ctypes.c_void_p), ('lazy_ws_bytes', ctypes.c_size_t), ('bwd_ws', ctypes.c_void_p), ('bwd_ws_bytes', ctypes.c_size_t),
                ('ws', ctypes.c_void_p), ('loss_out', ctypes.c_void_p), ('loss_w', LossGradPtrs),
                ('n_importance', ctypes.c_int64), ('z0', ctypes.c_void_p), ('raw0', ctypes.c_void_p), ('act_mask0', ctypes.c_void_p),
                ('weights0', ctypes.c_void_p), ('d_raw0', ctypes.c_void_p), ('u', ctypes.c_void_p), ('u_stride', ctypes.c_int64),
                ('sums0', ctypes.c_void_p),
                ('g_table', ctypes.c_void_p), ('g_mlp', MlpGrad), ('d_rays', ctypes.c_void_p), ('d_poses', ctypes.c_void_p),
                ('smooth_side', ctypes.c_int64), ('smooth_feat', ctypes.c_int64), ('smooth_voxel', ctypes.c_float),
                ('smooth_grad_scale', ctypes.c_float), ('smooth_jitter', ctypes.c_void_p), ('smooth_offset', ctypes.c_void_p),
                ('smooth_x', ctypes.c_void_p), ('smooth_f', ctypes.c_void_p), ('smooth_df', ctypes.c_void_p),
                ('map_params', ctypes.c_void_p), ('n_map_params', ctypes.c_int64),
                ('rot', ctypes.c_void_p), ('trans', ctypes.c_void_p), ('pose_adam', ctypes.c_void_p), ('n_free', ctypes.c_int64),
                ('n_rot', ctypes.c_int32), ('lr_rot', ctypes.c_float), ('lr_trans', ctypes.c_float),
                ('rank', ctypes.c_void_p), ('new_list', ctypes.c_void_p), ('new_count', ctypes.c_void_p),
                ('kept', ctypes.c_void_p), ('kept_derivatives', ctypes.c_int64)]


EXPORTS = ['cb200_abi_version', 'cb200_last_error', 'cb200_device_sm_count', 'cb200_grid_meta', 'cb200_hash_fwd',
           'cb200_hash_bwd', 'cb200_hash_indices', 'cb200_oneblob_fwd', 'cb200_oneblob_bwd', 'cb200_sample_rays', 'cb200_importance_sample', 'cb200_importance_sample_merge', 'cb200_merge_coarse', 'cb200_points_bwd_workspace', 'cb200_lazy_workspace', 'cb200_lazy_prefix', 'cb200_lazy_resolve', 'cb200_adam_step', 'cb200_assemble_rays_fwd', 'cb200_assemble_rays_bwd', 'cb200_track_pose', 'cb200_track_update', 'cb200_track_frame', 'cb200_track_init', 'cb200_track_rays_sample', 'cb200_track_pose_update', 'cb200_pose_params', 'cb200_track_frame_workspace', 'cb200_pose_matrices', 'cb200_pose_adam', 'cb200_smooth_points', 'cb200_smooth_tv', 'cb200_ba_iteration', 'cb200_sizeof', 'cb200_host_random_sample', 'cb200_host_torch_rand', 'cb200_adam_multi',
           'cb200_points_fwd', 'cb200_points_bwd', 'cb200_points_fwd_keep', 'cb200_points_bwd_kept', 'cb200_render_fwd', 'cb200_loss_finish', 'cb200_render_bwd', 'cb200_render_loss_bwd',
           'cb200_debug_positions', 'cb200_debug_gather_peak', 'cb200_points_bwd_noalloc', 'cb200_points_bwd_noalloc_max_points', 'cb200_selftest_mma', 'cb200_set_decoder',
           'cb200_get_decoder', 'cb200_set_forward_pipeline', 'cb200_get_forward_pipeline', 'cb200_cell_table_bytes', 'cb200_cell_table_build',
           'cb200_cell_table_fold', 'cb200_torch_rand_device', 'cb200_host_torch_rand_skip', 'cb200_host_torch_rand_segments']

_lib = None


def library_path():
    return _build.LIB_PATH


def load():
    """Loads (building first if the in-tree .so is stale and nvcc is present) and returns the CDLL."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    override = os.environ.get('COSLAM_B200_LIB')          # A/B runs of two builds on one box (tools/): load this file as it is
    if override:
        path = override
    elif _build.needs_build():
        if os.path.isfile(_build.NVCC):
            _build.build()
        elif not os.path.isfile(path):
            raise RuntimeError('coslam_b200: %s is missing and nvcc is not available to build it; '
                               'there is no fallback path' % path)
    lib = ctypes.CDLL(path)
    lib.cb200_last_error.restype = ctypes.c_char_p
    lib.cb200_points_bwd_workspace.restype = ctypes.c_size_t
    lib.cb200_lazy_workspace.restype = ctypes.c_size_t
    lib.cb200_lazy_workspace.argtypes = [ctypes.c_int64, ctypes.c_int64]
    lib.cb200_points_bwd_workspace.argtypes = [ctypes.c_int64]
    lib.cb200_track_frame_workspace.restype = ctypes.c_size_t
    lib.cb200_track_frame_workspace.argtypes = [ctypes.c_int64, ctypes.c_int64]
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise RuntimeError('coslam_b200: %s does not export %s' % (path, name))
    if lib.cb200_abi_version() != ABI_VERSION:
        raise RuntimeError('coslam_b200: ABI version mismatch (library %d, binding %d); rebuild with '
                           '`python -m coslam_b200.build --force`' % (lib.cb200_abi_version(), ABI_VERSION))
    lib.cb200_sizeof.restype = ctypes.c_size_t
    lib.cb200_get_forward_pipeline.restype = ctypes.c_longlong
    lib.cb200_cell_table_bytes.restype = ctypes.c_size_t
    for which, cls in ((0, BaPlan), (1, TrackStep), (2, AdamParam)):
        if lib.cb200_sizeof(which) != ctypes.sizeof(cls):
            raise RuntimeError('coslam_b200: %s layout differs from the library (%d vs %d bytes)'
                               % (cls.__name__, ctypes.sizeof(cls), lib.cb200_sizeof(which)))
    _lib = lib
    if os.environ.get('COSLAM_B200_DECODER'):
        lib.cb200_set_decoder({'simt': 0, 'tc': 1}[os.environ['COSLAM_B200_DECODER']])
    return lib


KERNEL_LAUNCHES = 0        # every successful device call below launches exactly one kernel of this library


def set_decoder(kind):
    """'tc' (tcgen05 tensor-core decoder, default) or 'simt' (fp32 FMA decoder)."""
    rc = load().cb200_set_decoder({'simt': 0, 'tc': 1}[kind])
    if rc != 0:
        raise RuntimeError('cb200_set_decoder failed')


def set_forward_pipeline(min_points):
    """< 0 / None: one-role tensor-core forward; >= 0: warp-specialised forward for batches of at least that many points."""
    load().cb200_set_forward_pipeline(ctypes.c_longlong(-1 if min_points is None else int(min_points)))


def get_forward_pipeline():
    return int(load().cb200_get_forward_pipeline())


def check(rc, what):
    global KERNEL_LAUNCHES
    KERNEL_LAUNCHES += 1
    if rc != 0:
        msg = load().cb200_last_error().decode('utf-8', 'replace')
        raise RuntimeError('coslam_b200 %s failed (%d): %s' % (what, rc, msg))


def ptr(t):
    """Device pointer of a tensor (None -> NULL).  The tensor must be contiguous CUDA memory."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError('coslam_b200: expected a CUDA tensor (there is no CPU path)')
    if not t.is_contiguous():
        raise RuntimeError('coslam_b200: internal error, non-contiguous tensor reached the C ABI')
    return ctypes.c_void_p(t.data_ptr())


def stream():
    """The current CUDA stream of the current device as a raw handle (called ~10 times per forward + backward: the raw query is
    ~10x cheaper than building a torch.cuda.Stream object)."""
    raw = getattr(torch._C, '_cuda_getCurrentRawStream', None)
    if raw is not None:
        return ctypes.c_void_p(raw(torch.cuda.current_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def grid_meta(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale):
    """Host-side geometry; works without a GPU."""
    cfg = GridCfg(n_levels, n_features, log2_hashmap_size, base_resolution, float(per_level_scale))
    g = Grid()
    global KERNEL_LAUNCHES
    check(load().cb200_grid_meta(ctypes.byref(cfg), ctypes.byref(g)), 'grid_meta')
    KERNEL_LAUNCHES -= 1       # host-only call
    return g
