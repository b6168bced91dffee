"""Host side of the fused path: scene/render parameter blocks, thin wrappers over the C ABI and the autograd glue.

Every compute call goes through libcoslam_b200.so (coslam_b200/_lib.py); torch only allocates buffers and records
the autograd graph.  Reference lines each piece stands in for are cited at the wrappers.
"""
import ctypes
import os
import math

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr, stream

N_GEO = 15
N_GRID_FEATS = 32      # 16 levels x 2 hash-grid features (CB200_GRID_FEATS)

# tests set KEEP_LAST_MASK to read back the ReLU sign bits the last ray forward saved for its backward
KEEP_LAST_MASK = False
LAST_MASK = None
LAST_EVALUATED = None   # bool [P]: the points the (lazy) forward evaluated, None = all; kept with LAST_MASK
MASK_LOG = []          # every mask saved while KEEP_LAST_MASK is set, in call order (two entries for an importance-pass step)

# bench.py sets this to a list to collect (name, start_event, end_event) around the two per-point kernels
KERNEL_EVENTS = None
KEEP_LAST_D_RAW = False   # bench.py: keep the last d_raw of the loss head to count the points that carry a gradient
LAST_D_RAW = None


def _timed(name, fn):
    if KERNEL_EVENTS is None:
        return fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    KERNEL_EVENTS.append((name, e0, e1))


def _f32c(t):
    """fp32 + contiguous view of a CUDA tensor (copy only when needed)."""
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class LossGrads(ctypes.Structure):
    _fields_ = [('rgb', ctypes.c_void_p), ('depth', ctypes.c_void_p), ('sdf', ctypes.c_void_p), ('fs', ctypes.c_void_p)]


class SceneSpec(object):
    """Plain-python description of the scene + sampling/loss parameters the kernels need.

    Holds no device pointers and no ctypes objects in its pickled state, so the owning nn.Module survives
    share_memory(), copy.deepcopy and spawn-pickling (coslam_mp.py:52, mp_slam/tracker.py:27); the ctypes blocks and
    the per-device linspace constants are rebuilt lazily in each process.
    """

    def __init__(self, config, bound_box, grid_cfg, n_bins):
        t, cam = config['training'], config['cam']
        bb = torch.as_tensor(bound_box, dtype=torch.float32).detach().cpu()
        self.bb_lo = [float(v) for v in bb[:, 0]]
        self.bb_ext = [float(v) for v in (bb[:, 1] - bb[:, 0])]          # fp32 subtraction, model/scene_rep.py:186
        self.normalise = bool(config['grid']['tcnn_encoding'])
        self.grid_cfg = dict(grid_cfg)
        self.n_bins = int(n_bins)
        self.n_uniform = int(t['n_samples_d'])
        self.n_guided = int(t['n_range_d'])
        self.range_d = t['range_d']
        self.near, self.far = cam['near'], cam['far']
        self.perturb = float(t['perturb']) > 0.
        self.rgb_missing_nz = bool(t['rgb_missing'])
        self.trunc = float(t['trunc'])
        self.trunc_sc = float(t['trunc']) * float(config['data']['sc_factor'])   # python double product, like the reference
        self.depth_trunc = float(cam['depth_trunc'])
        self.white_bkgd = bool(t['white_bkgd'])
        self.n_importance = int(t.get('n_importance', 0))
        self._cache = {}

    def __getstate__(self):
        d = dict(self.__dict__)
        d['_cache'] = {}
        return d

    @property
    def n_samples(self):
        return self.n_uniform + self.n_guided

    def grid(self):
        g = self._cache.get('grid')
        if g is None:
            c = self.grid_cfg
            g = _lib.grid_meta(c['n_levels'], c['n_features_per_level'], c['log2_hashmap_size'], c['base_resolution'],
                               c['per_level_scale'])
            self._cache['grid'] = g
        return g

    def n_table_params(self):
        g = self.grid()
        return int(g.offset[g.n_levels]) * int(g.n_features)

    def scene(self):
        s = self._cache.get('scene')
        if s is None:
            s = _lib.Scene()
            s.grid = self.grid()
            for i in range(3):
                s.bb_lo[i] = self.bb_lo[i]
                s.bb_ext[i] = self.bb_ext[i]
            s.normalise = 1 if self.normalise else 0
            s.n_bins = self.n_bins
            self._cache['scene'] = s
        return s

    def scene_unnormalised(self):
        """Same grid, inputs already in the unit cube (query_sdf / query_color_sdf take normalised points)."""
        s = self._cache.get('scene_raw')
        if s is None:
            s = _lib.Scene()
            s.grid = self.grid()
            for i in range(3):
                s.bb_lo[i] = 0.0
                s.bb_ext[i] = 1.0
            s.normalise = 0
            s.n_bins = self.n_bins
            self._cache['scene_raw'] = s
        return s

    def render_cfg(self, perturb=None):
        key = ('rcfg', perturb)
        r = self._cache.get(key)
        if r is None:
            r = _lib.RenderCfg()
            r.n_uniform, r.n_guided = self.n_uniform, self.n_guided
            r.perturb = 1 if (self.perturb if perturb is None else perturb) else 0
            r.rgb_missing_nz = 1 if self.rgb_missing_nz else 0
            r.trunc = self.trunc
            r.trunc_sc = self.trunc_sc
            r.depth_trunc = self.depth_trunc
            r.white_bkgd = 1 if self.white_bkgd else 0
            self._cache[key] = r
        return r

    def linspaces(self, device):
        """The three torch.linspace constants of model/scene_rep.py:227, :239, :242 -- created on the CPU by
        torch itself (bit-identical to the reference) and uploaded once per device."""
        key = ('lin', str(device))
        v = self._cache.get(key)
        if v is None:
            lin_g = torch.linspace(-self.range_d, self.range_d, steps=self.n_guided)
            lin_n = torch.linspace(self.near, self.far, steps=self.n_guided)
            lin_u = torch.linspace(self.near, self.far, self.n_uniform) if self.n_uniform > 0 else torch.zeros(1)
            v = tuple(t.float().to(device) for t in (lin_u, lin_g, lin_n))
            self._cache[key] = v
        return v


# --------------------------------------------------------------------------------------------------------------------
# host noise: torch.rand's CPU stream, bit for bit, from a tight loop (csrc/host_rng.cu)
# --------------------------------------------------------------------------------------------------------------------

def torch_rand_into(out):
    """Fills the contiguous fp32 CPU tensor `out` with exactly the values `torch.rand(out.shape)` would draw from torch's default
    CPU generator and leaves the generator in exactly the state that call would (model/scene_rep.py:255 draws its jitter this way)."""
    if out.device.type != 'cpu' or out.dtype != torch.float32 or not out.is_contiguous():
        raise ValueError('torch_rand_into: a contiguous fp32 CPU tensor')
    st = torch.get_rng_state()
    rc = _lib.load().cb200_host_torch_rand(ctypes.c_void_p(st.data_ptr()), ctypes.c_int64(out.numel()), ctypes.c_void_p(out.data_ptr()))
    if rc != 0:                                    # a generator state the routine does not know: torch draws the same values itself
        torch.rand(out.shape, out=out)
        return out
    torch.set_rng_state(st)
    return out


# batches of at least this many uniforms are generated on the device (cb200_torch_rand_device: the same stream, no upload); the host
# only advances its generator state.  Below it the host loop + pinned upload is quicker than a single-CTA kernel's latency.
DEVICE_STREAM_MIN = int(os.environ.get('COSLAM_B200_DEVICE_STREAM_MIN', str(1 << 20)))


DEVICE_STREAM_SEGMENT = 64      # Mersenne-Twister blocks (of 624 floats) per device CTA
DEVICE_STREAM_CHUNK = 1 << 24   # floats per host walk + launch: the device generates chunk k while the host walks chunk k + 1
_STREAM_BUFFERS = {}            # device index -> [pinned checkpoints, device checkpoints, event of the last upload]


def device_uniform(n, device):
    """torch.rand(n).to(device), bit for bit, generated on the device (cb200_host_torch_rand_segments + cb200_torch_rand_device);
    None if the generator state is not one the routines know."""
    lib = _lib.load()
    device = torch.device(device)
    if n == 0:
        return torch.empty(0, dtype=torch.float32, device=device)
    C, words = DEVICE_STREAM_SEGMENT, 624
    cap_total = n // (words * C) + 2 * ((n + DEVICE_STREAM_CHUNK - 1) // DEVICE_STREAM_CHUNK) + 2
    key = device.index if device.index is not None else torch.cuda.current_device()
    buf = _STREAM_BUFFERS.get(key)
    if buf is None or buf[0].shape[0] < cap_total:
        buf = _STREAM_BUFFERS[key] = [torch.empty((cap_total, words), dtype=torch.int32).pin_memory(),
                                      torch.empty((cap_total, words), dtype=torch.int32, device=device), None]
    if buf[2] is not None:
        buf[2].synchronize()                       # the previous call's uploads have left the pinned block
    st = torch.get_rng_state()
    out = torch.empty(n, dtype=torch.float32, device=device)
    n_seg, head, head_next = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int32(0)
    done, used = 0, 0
    with torch.cuda.device(device):
        while done < n:
            take = min(DEVICE_STREAM_CHUNK, n - done)
            rc = lib.cb200_host_torch_rand_segments(ctypes.c_void_p(st.data_ptr()), ctypes.c_int64(take), ctypes.c_int64(C),
                                                    ctypes.c_void_p(buf[0].data_ptr() + used * words * 4), ctypes.c_int64(cap_total - used),
                                                    ctypes.byref(n_seg), ctypes.byref(head), ctypes.byref(head_next))
            if rc != 0:
                if done == 0 and rc == -1:
                    return None                    # nothing consumed yet: the caller falls back to the host loop
                raise RuntimeError('cb200_host_torch_rand_segments failed (%d) in the middle of a draw' % rc)
            k = n_seg.value
            buf[1][used:used + k].copy_(buf[0][used:used + k], non_blocking=True)
            check(lib.cb200_torch_rand_device(ctypes.c_void_p(buf[1].data_ptr() + used * words * 4), n_seg, ctypes.c_int64(C), head, head_next,
                                              ctypes.c_int64(take), ctypes.c_void_p(out.data_ptr() + done * 4), stream()), 'torch_rand_device')
            used += k
            done += take
        buf[2] = torch.cuda.Event()
        buf[2].record()
    torch.set_rng_state(st)
    return out


_PINNED = {}      # numel -> [[pinned buffer, event of its last upload] x 2, next]


def host_uniform(shape, device):
    """`torch.rand(shape).to(device)` (the reference's jitter, CPU generator + upload) through torch_rand_into and a pinned staging buffer."""
    n = 1
    for d in shape:
        n *= int(d)
    if n >= DEVICE_STREAM_MIN and torch.device(device).type == 'cuda':
        out = device_uniform(n, torch.device(device))
        if out is not None:
            return out.view(shape)
    slot = _PINNED.get(n)
    if slot is None:
        slot = _PINNED[n] = [[torch.empty(n, dtype=torch.float32).pin_memory(), None], [torch.empty(n, dtype=torch.float32).pin_memory(), None], 0]
        if len(_PINNED) > 8:                       # a handful of shapes in practice (tracking batch, mapping batch); do not hoard pinned memory
            _PINNED.pop(next(iter(_PINNED)))
    buf = slot[slot[2]]
    slot[2] ^= 1
    if buf[1] is not None:
        buf[1].synchronize()                       # the upload that last used this buffer has finished
    torch_rand_into(buf[0])
    out = buf[0].view(shape).to(device, non_blocking=True)
    buf[1] = torch.cuda.Event()
    buf[1].record()
    return out


# --------------------------------------------------------------------------------------------------------------------
# raw wrappers (no autograd)
# --------------------------------------------------------------------------------------------------------------------

def sample_rays(spec, target_d, noise=None, seed=0, counts=None, perturb=None, seed_dev=None):
    """model/scene_rep.py:227-255.  target_d [N] or [N,1]; noise [N,S] CPU-drawn uniforms or None (device Philox).
    Returns z_vals [N,S]; accumulates (n_front, n_sdf, n_valid_depth) into counts (int64 [3]) when given."""
    d = _f32c(target_d.reshape(-1))
    n = d.shape[0]
    z = torch.empty((n, spec.n_samples), dtype=torch.float32, device=d.device)
    lin_u, lin_g, lin_n = spec.linspaces(d.device)
    if noise is not None:
        noise = _f32c(noise)
        if tuple(noise.shape) != (n, spec.n_samples):
            raise ValueError('noise must have shape [N, S]')
    check(_lib.load().cb200_sample_rays(ctypes.byref(spec.render_cfg(perturb)), ptr(d), ptr(lin_u), ptr(lin_g), ptr(lin_n),
                                        ptr(noise), ctypes.c_uint64(int(seed) & (2 ** 64 - 1)), ptr(seed_dev), ctypes.c_int64(n), ptr(z),
                                        ptr(counts), stream()), 'sample_rays')
    return z


def importance_sample(spec, z_vals, weights, u, target_d=None, counts=None, want_std=True, merge=False):
    """sample_pdf + sorted merge (model/scene_rep.py:265-272, model/utils.py:32-71).  u is [N,I] or one shared row [I].
    Returns (z [N,S+I], z_std [N] | None); accumulates the mask counts of the merged samples into counts when given.
    merge=True: also returns (rank [N,S] int32, new_list [N*I] int32 storage of uint32 indices, new_count int64 [1]) -- where the coarse
    samples sit in the merged rows and the point list of the new samples (cb200_importance_sample_merge)."""
    z, w = _f32c(z_vals), _f32c(weights)
    n, s = z.shape
    u = _f32c(u)
    n_imp = u.shape[-1]
    if u.dim() == 2 and u.shape[0] != n:
        raise ValueError('u must be [N, I] or [I]')
    stride = n_imp if u.dim() == 2 else 0
    z_out = torch.empty((n, s + n_imp), dtype=torch.float32, device=z.device)
    z_std = torch.empty((n,), dtype=torch.float32, device=z.device) if want_std else None
    d = _f32c(target_d.reshape(-1)) if target_d is not None else None
    if merge:
        rank = torch.empty((n, s), dtype=torch.int32, device=z.device)
        new_list = torch.empty((n * n_imp,), dtype=torch.int32, device=z.device)
        new_count = torch.empty((1,), dtype=torch.int64, device=z.device)
        check(_lib.load().cb200_importance_sample_merge(ctypes.byref(spec.render_cfg()), ptr(z), ptr(w), ptr(u), ctypes.c_int64(stride), ptr(d),
                                                        ctypes.c_int64(n), ctypes.c_int64(s), ctypes.c_int64(n_imp), ptr(z_out), ptr(z_std),
                                                        ptr(counts), ptr(rank), ptr(new_list), ptr(new_count), stream()), 'importance_sample_merge')
        return z_out, z_std, (rank, new_list, new_count)
    check(_lib.load().cb200_importance_sample(ctypes.byref(spec.render_cfg()), ptr(z), ptr(w), ptr(u), ctypes.c_int64(stride), ptr(d),
                                              ctypes.c_int64(n), ctypes.c_int64(s), ctypes.c_int64(n_imp), ptr(z_out), ptr(z_std),
                                              ptr(counts), stream()), 'importance_sample')
    return z_out, z_std


def merge_coarse(rank, src, dst, src_mask=None, dst_mask=None, accumulate=False):
    """dst[ray, rank[ray, s]] (=, or += with accumulate) src[ray, s] over float4 rows ([N,S,4] into [N,S+I,4]); optionally the [.,2]
    uint32 activation masks as well (cb200_merge_coarse)."""
    n, s = rank.shape
    f = dst.numel() // (4 * n)
    check(_lib.load().cb200_merge_coarse(ptr(rank), ctypes.c_int64(n), ctypes.c_int64(s), ctypes.c_int64(f), ptr(src), ptr(dst), ptr(src_mask),
                                         ptr(dst_mask), 1 if accumulate else 0, stream()), 'merge_coarse')


def _mlp_block(w1, w2, w3, w4):
    m = _lib.Mlp()
    m.w_sdf0, m.w_sdf1, m.w_col0, m.w_col1 = w1.data_ptr(), w2.data_ptr(), w3.data_ptr(), w4.data_ptr()
    return m


def _check_weights(w1, w2, w3, w4):
    if tuple(w1.shape) != (32, 80) or tuple(w2.shape) != (16, 32) or tuple(w3.shape) != (32, 63) or tuple(w4.shape) != (3, 32):
        raise RuntimeError('coslam_b200: the fused decoder kernels serve hidden_dim 32, geo_feat_dim 15, 2+2 layers and '
                           'a 32+48 wide input (every shipped YAML); got weight shapes %s' %
                           [tuple(w.shape) for w in (w1, w2, w3, w4)])


# training steps keep the hash-grid features of the forward for the backward (RayPointsFunction, BundleAdjuster) ...
KEEP_FEATURES = os.environ.get('COSLAM_B200_KEEP_ROWS', '1') != '0'
# ... and, when the rays carry a gradient, the features' position derivatives as well
KEEP_DERIVATIVES = os.environ.get('COSLAM_B200_KEEP_ROWS', '2') not in ('0', '1')
# opt-in (COSLAM_B200_CELL_TABLES=1): large training steps read / reduce the dense levels through cell-major copies
# (cb200_cell_table_build / _fold, csrc/cells.cu).  Measured SLOWER on B200 (profiles/r2/NOTES.md section 15: forward +14 %, backward +1 %
# at 2^20 rays -- adjacent cells no longer share their corners' lines in L1), so it is off by default and kept as the A/B twin.
CELL_TABLES = os.environ.get('COSLAM_B200_CELL_TABLES', '0') == '1'
CELL_MIN_POINTS = 1 << 20   # below this the build / clear / fold launches (~25 us together) are not worth it
_CELL_BUFFERS = {}          # (device, bytes) -> [cells, dcells or None]: reused across steps, rebuilt from the table in front of every forward
LAZY_FORWARD = True      # training forward evaluates only the sample points the step can depend on (lazy_ray_forward)
LAZY_MIN_POINTS = 1 << 18   # below ~7 waves of tiles (tracking / BA batches) the three extra launches cost more than the skipped points save


def _kept_rows_fit(nbytes, device, fraction=0.6):
    """Kept rows are an optimisation, never a reason to run out of memory: large ones (> 8 GB) are only allocated when they take at
    most `fraction` of what the device still has (driver-free memory + the caching allocator's unused reserve)."""
    if nbytes <= (8 << 30):                      # (the driver query costs ~40 us: not worth asking for a few GB on a 180 GB device)
        return True
    free, _ = torch.cuda.mem_get_info(device)
    free += torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
    return nbytes <= fraction * free


def cell_tables(scene, table, want_grad=False):
    """Scene copy whose `cells` points at a fresh cell-major copy of `table`'s dense levels (None if the grid has no leading dense
    level).  Returns (scene_with_cells, buffers); buffers[1] is the gradient block buffer when want_grad (allocated, not cleared)."""
    lib = _lib.load()
    nbytes = int(lib.cb200_cell_table_bytes(ctypes.byref(scene)))
    if nbytes == 0:
        return None, None
    key = (table.device.index, nbytes)
    buf = _CELL_BUFFERS.get(key)
    if buf is None:
        buf = [torch.empty(nbytes // 4, dtype=torch.float32, device=table.device), None]
        _CELL_BUFFERS[key] = buf
    if want_grad and buf[1] is None:
        buf[1] = torch.empty(nbytes // 4, dtype=torch.float32, device=table.device)
    check(lib.cb200_cell_table_build(ctypes.byref(scene), ptr(table), ptr(buf[0]), stream()), 'cell_table_build')
    s2 = _lib.Scene.from_buffer_copy(scene)
    s2.cells = buf[0].data_ptr()
    return s2, buf


def _kept_width(feat):
    """with_derivatives of cb200_points_fwd_keep / cb200_points_bwd_kept from the shape of the kept rows: [P,32] features, [P,128]
    features and their three position derivatives."""
    if feat.dim() != 2 or feat.shape[1] not in (N_GRID_FEATS, 4 * N_GRID_FEATS) or not feat.is_contiguous() or feat.dtype != torch.float32:
        raise ValueError('kept rows must be a contiguous fp32 [P,32] or [P,128] tensor')
    return 1 if feat.shape[1] == 4 * N_GRID_FEATS else 0


def points_forward(scene, table, w, pts=None, rays=None, want_raw=True, want_geo=False, want_mask=True, sdf_only=False,
                   out=None, point_list=None, feat=None):
    """query_color_sdf / run_network forward through the fused kernel.  rays = (rays_o [N,3], rays_d [N,3], z [N,S]).
    out = (raw, mask) to write into existing buffers; point_list = (list ptr, count ptr) restricts the launch to them;
    feat [P,32]: also keep the hash-grid features of the evaluated points for points_backward(feat=) (cb200_points_fwd_keep)."""
    _check_weights(*w)
    if rays is not None:
        o, d, z = rays
        n, s = z.shape
        dev = z.device
    else:
        n, s = pts.shape[0], 1
        dev = pts.device
        o = d = z = None
    P = n * s
    if out is not None:
        raw, mask = out
    else:
        raw = torch.empty((P, 4), dtype=torch.float32, device=dev) if want_raw else None
        mask = torch.empty((P, 2), dtype=torch.int32, device=dev) if want_mask else None
    geo = torch.empty((P, N_GEO), dtype=torch.float32, device=dev) if want_geo else None
    mlp = _mlp_block(*w)
    lst, cnt = point_list if point_list is not None else (None, None)
    if feat is not None:
        _timed('points_fwd', lambda: check(_lib.load().cb200_points_fwd_keep(
            ctypes.byref(scene), ptr(table), ctypes.byref(mlp), ptr(pts), ptr(o), ptr(d), ptr(z), ctypes.c_int64(n),
            ctypes.c_int64(s), ptr(raw), ptr(geo), ptr(mask), 1 if sdf_only else 0, lst, cnt, ptr(feat), _kept_width(feat), stream()),
            'points_fwd_keep'))
        return raw, geo, mask
    _timed('points_fwd', lambda: check(_lib.load().cb200_points_fwd(
        ctypes.byref(scene), ptr(table), ctypes.byref(mlp), ptr(pts), ptr(o), ptr(d), ptr(z), ctypes.c_int64(n),
        ctypes.c_int64(s), ptr(raw), ptr(geo), ptr(mask), 1 if sdf_only else 0, lst, cnt, stream()), 'points_fwd'))
    return raw, geo, mask


def lazy_ray_forward(spec, table, w, rays, target_d, want_mask=True, feat=None, scene=None):
    """Training forward over the sample points the step can depend on (cb200_lazy_prefix / cb200_lazy_resolve): the loss
    prefix of every ray first, then whatever the first sign change found there still exposes.  Points never listed keep
    raw = 0; the rendering mask gives them zero weight and the loss masks exclude them, as in the reference."""
    o, d, z = rays
    n, s = z.shape
    dev = z.device
    lib = _lib.load()
    raw = torch.zeros((n * s, 4), dtype=torch.float32, device=dev)
    mask = torch.empty((n * s, 2), dtype=torch.int32, device=dev) if want_mask else None
    nbytes = int(lib.cb200_lazy_workspace(ctypes.c_int64(n), ctypes.c_int64(s)))
    wsp = torch.empty((nbytes + 7) // 8, dtype=torch.int64, device=dev)
    base = wsp.data_ptr()
    plist = (ctypes.c_void_p(base + 16 + 4 * n), ctypes.c_void_p(base))
    td = _f32c(target_d.reshape(-1))
    cfg = spec.render_cfg()
    check(lib.cb200_lazy_prefix(ctypes.byref(cfg), ptr(z), ptr(td), ctypes.c_int64(n), ctypes.c_int64(s), ptr(wsp),
                                ctypes.c_size_t(nbytes), stream()), 'lazy_prefix')
    scene = scene if scene is not None else spec.scene()
    points_forward(scene, table, w, rays=rays, out=(raw, mask), point_list=plist, feat=feat)
    check(lib.cb200_lazy_resolve(ctypes.byref(cfg), ptr(z), ptr(td), ptr(raw), ctypes.c_int64(n), ctypes.c_int64(s), ptr(wsp),
                                 ctypes.c_size_t(nbytes), stream()), 'lazy_resolve')
    points_forward(scene, table, w, rays=rays, out=(raw, mask), point_list=plist, feat=feat)
    return raw, mask


def points_backward(scene, table, w, mask, d_raw, d_geo=None, pts=None, rays=None, need_table=True, need_mlp=True,
                    need_pos=True, sdf_only=False, accumulate=None, ready_list=None, feat=None):
    """Backward of points_forward.  Returns (d_table, (dW1..dW4), d_pts | (d_rays_o, d_rays_d)).
    accumulate = (d_table, (dW1..dW4), d_rays [2,N,3]): existing buffers the kernel adds into (any entry may be None = not
    needed) instead of fresh zero-filled ones -- the autograd-free loops (mapping.py) keep their gradient buffers.
    feat: the features kept by points_forward(feat=); used (cb200_points_bwd_kept) when the decoder gradients are wanted and no
    position gradient is."""
    if rays is not None:
        o, d, z = rays
        n, s = z.shape
        dev = z.device
    else:
        n, s = pts.shape[0], 1
        dev = pts.device
        o = d = z = None
    if accumulate is not None:
        d_table, d_w, d_od = accumulate
        gblock = _mlp_block(*d_w) if d_w is not None else None
        d_pts, d_o, d_d = None, (d_od[0] if d_od is not None else None), (d_od[1] if d_od is not None else None)
        need_mlp = need_pos = False               # buffers are set up: skip the allocations below
    else:
        d_table = torch.zeros_like(table) if need_table else None
        d_w = None
        gblock = None
        d_pts = d_o = d_d = None
    if need_mlp:
        flat = torch.zeros(sum(x.numel() for x in w), dtype=torch.float32, device=dev)
        d_w, off = [], 0
        for x in w:
            d_w.append(flat[off:off + x.numel()].view_as(x))
            off += x.numel()
        gblock = _mlp_block(*d_w)
    if need_pos:
        if rays is not None:
            d_od = torch.zeros((2, n, 3), dtype=torch.float32, device=dev)
            d_o, d_d = d_od[0], d_od[1]
        else:
            d_pts = torch.empty((n, 3), dtype=torch.float32, device=dev)
    mlp = _mlp_block(*w)
    # scratch for the live-point compaction (ray mode): index list + counter, see cb200_points_bwd
    wsp, wsp_bytes = None, 0
    if ready_list is not None:                     # render_loss_backward has written the live-point list already
        wsp = ready_list
    elif rays is not None and d_geo is None:
        wsp_bytes = int(_lib.load().cb200_points_bwd_workspace(ctypes.c_int64(n * s)))
        wsp = torch.empty((wsp_bytes + 3) // 4, dtype=torch.int32, device=dev)
    if feat is not None and gblock is not None and (_kept_width(feat) or (d_pts is None and d_o is None)):
        _timed('points_bwd', lambda: check(_lib.load().cb200_points_bwd_kept(
            ctypes.byref(scene), ptr(table), ctypes.byref(mlp), ptr(pts), ptr(o), ptr(d), ptr(z), ctypes.c_int64(n),
            ctypes.c_int64(s), ptr(d_raw), ptr(d_geo), ptr(mask), ptr(feat), _kept_width(feat), ptr(d_table), ctypes.byref(gblock),
            ptr(d_pts), ptr(d_o), ptr(d_d), 1 if sdf_only else 0, ptr(wsp), ctypes.c_size_t(wsp_bytes), stream()), 'points_bwd_kept'))
    else:
        _timed('points_bwd', lambda: check(_lib.load().cb200_points_bwd(
            ctypes.byref(scene), ptr(table), ctypes.byref(mlp), ptr(pts), ptr(o), ptr(d), ptr(z), ctypes.c_int64(n),
            ctypes.c_int64(s), ptr(d_raw), ptr(d_geo), ptr(mask), ptr(d_table),
            ctypes.byref(gblock) if gblock is not None else None, ptr(d_pts), ptr(d_o), ptr(d_d), 1 if sdf_only else 0,
            ptr(wsp), ctypes.c_size_t(wsp_bytes), stream()), 'points_bwd'))
    if wsp is not None and ready_list is None and _lib.load().cb200_get_decoder() == 1:
        _lib.KERNEL_LAUNCHES += 1          # the compaction kernel in front of the per-point kernel
    return d_table, d_w, (d_pts if rays is None else (d_o, d_d))


# --------------------------------------------------------------------------------------------------------------------
# autograd
# --------------------------------------------------------------------------------------------------------------------

class RayPointsFunction(torch.autograd.Function):
    """raw[N,S,4] = decoder(encodings(normalise(rays_o + rays_d * z)))   (model/scene_rep.py:258 + run_network)."""

    @staticmethod
    def forward(ctx, rays_o, rays_d, z_vals, table, w1, w2, w3, w4, spec, target_d=None):
        o, d, z = _f32c(rays_o), _f32c(rays_d), _f32c(z_vals)
        tb, w = _f32c(table), tuple(_f32c(x) for x in (w1, w2, w3, w4))
        need_grad = any(ctx.needs_input_grad)
        ctx.set_materialize_grads(False)
        tc = _lib.load().cb200_get_decoder() == 1
        lazy = target_d is not None and LAZY_FORWARD and tc and LAZY_MIN_POINTS <= z.numel() < 2 ** 32
        # a map-only step (decoder gradients wanted, rays fixed): keep the hash-grid features for the backward, which then does not
        # gather the table a second time (128 B per point of scratch; without the memory the step runs as before)
        feat = None
        need = ctx.needs_input_grad
        if KEEP_FEATURES and tc and z.numel() > 0 and any(need[4:8]):
            # with ray gradients the rows also hold d feature / d position (512 B per point instead of 128)
            width = 4 * N_GRID_FEATS if (need[0] or need[1]) else N_GRID_FEATS
            if (width == N_GRID_FEATS or KEEP_DERIVATIVES) and _kept_rows_fit(z.numel() * width * 4, z.device):
                try:
                    feat = torch.empty((z.numel(), width), dtype=torch.float32, device=z.device)
                except torch.cuda.OutOfMemoryError:
                    feat = None
        scene, cellbuf = spec.scene(), None
        if CELL_TABLES and tc and z.numel() >= CELL_MIN_POINTS:
            # the backward reduces into gradient blocks of the same layout when it runs on kept rows and a table gradient is wanted
            with_cells, cellbuf = cell_tables(scene, tb, want_grad=feat is not None and need[3])
            scene = with_cells if with_cells is not None else scene
        if lazy:
            _check_weights(*w)
            raw, mask = lazy_ray_forward(spec, tb, w, (o, d, z), target_d, want_mask=need_grad, feat=feat, scene=scene)
        else:
            raw, _, mask = points_forward(scene, tb, w, rays=(o, d, z), want_mask=need_grad, feat=feat)
        if KEEP_LAST_MASK:
            global LAST_MASK, LAST_EVALUATED
            LAST_MASK = mask
            LAST_EVALUATED = (raw != 0).any(-1) if lazy else None
            MASK_LOG.append(mask)
        ctx.spec = spec
        ctx.kept = feat is not None
        ctx.dcells = cellbuf[1] if (cellbuf is not None and feat is not None and need[3]) else None
        ctx.save_for_backward(o, d, z, tb, *w, mask if mask is not None else torch.empty(0), *((feat,) if feat is not None else ()))
        return raw.view(z.shape[0], z.shape[1], 4)

    @staticmethod
    def backward(ctx, g_raw):
        if g_raw is None:
            return (None,) * 10
        o, d, z, tb, w1, w2, w3, w4, mask = ctx.saved_tensors[:9]
        feat = ctx.saved_tensors[9] if ctx.kept else None
        need = ctx.needs_input_grad
        scene = ctx.spec.scene()
        if ctx.dcells is not None:
            scene = _lib.Scene.from_buffer_copy(scene)
            scene.dcells = ctx.dcells.zero_().data_ptr()
        d_table, d_w, d_ray = points_backward(scene, tb, (w1, w2, w3, w4), mask, _f32c(g_raw), rays=(o, d, z),
                                              need_table=need[3], need_mlp=any(need[4:8]), need_pos=need[0] or need[1], feat=feat)
        if ctx.dcells is not None:
            check(_lib.load().cb200_cell_table_fold(ctypes.byref(scene), ptr(ctx.dcells), ptr(d_table), stream()), 'cell_table_fold')
        d_o, d_d = d_ray if d_ray is not None else (None, None)
        d_w = d_w or (None, None, None, None)
        return d_o, d_d, None, d_table, d_w[0], d_w[1], d_w[2], d_w[3], None, None


class PointsFunction(torch.autograd.Function):
    """raw[P,4] (and optionally geo[P,15]) for explicit points; `normalised` = points already in the unit cube
    (query_color_sdf / query_sdf, model/scene_rep.py:126-171) or world points (run_network, :173-191)."""

    @staticmethod
    def forward(ctx, pts, table, w1, w2, w3, w4, spec, normalised, sdf_only):
        p = _f32c(pts)
        tb, w = _f32c(table), tuple(_f32c(x) for x in (w1, w2, w3, w4))
        scene = spec.scene_unnormalised() if (normalised or not spec.normalise) else spec.scene()
        need_grad = any(ctx.needs_input_grad)
        ctx.set_materialize_grads(False)
        raw, geo, mask = points_forward(scene, tb, w, pts=p, want_geo=sdf_only, want_mask=need_grad, sdf_only=sdf_only)
        ctx.scene, ctx.sdf_only = scene, sdf_only
        ctx.save_for_backward(p, tb, *w, mask if mask is not None else torch.empty(0))
        if sdf_only:
            return raw, geo
        return raw, torch.empty(0, device=p.device)

    @staticmethod
    def backward(ctx, g_raw, g_geo):
        if g_raw is None and (g_geo is None or g_geo.numel() == 0):
            return (None,) * 9
        p, tb, w1, w2, w3, w4, mask = ctx.saved_tensors
        need = ctx.needs_input_grad
        if g_raw is None:
            g_raw = torch.zeros((p.shape[0], 4), dtype=torch.float32, device=p.device)
        d_geo = _f32c(g_geo) if (ctx.sdf_only and g_geo is not None and g_geo.numel() > 0) else None
        d_table, d_w, d_pts = points_backward(ctx.scene, tb, (w1, w2, w3, w4), mask, _f32c(g_raw), d_geo=d_geo, pts=p,
                                              need_table=need[1], need_mlp=any(need[2:6]), need_pos=need[0],
                                              sdf_only=ctx.sdf_only)
        d_w = d_w or (None, None, None, None)
        return d_pts, d_table, d_w[0], d_w[1], d_w[2], d_w[3], None, None, None


def _render_forward(spec, raw, z, target_rgb=None, target_d=None, loss_sums=None, want_aux=False, want_weights=False):
    n, s = z.shape
    dev = z.device
    rgb = torch.empty((n, 3), dtype=torch.float32, device=dev)
    depth = torch.empty((n,), dtype=torch.float32, device=dev)
    aux = torch.empty((3, n), dtype=torch.float32, device=dev) if want_aux else None
    weights = torch.empty((n, s), dtype=torch.float32, device=dev) if want_weights else None
    check(_lib.load().cb200_render_fwd(ctypes.byref(spec.render_cfg()), ptr(raw), ptr(z), ptr(target_rgb), ptr(target_d),
                                       ctypes.c_int64(n), ctypes.c_int64(s), ptr(rgb), ptr(depth),
                                       ptr(aux[0]) if want_aux else None, ptr(aux[1]) if want_aux else None,
                                       ptr(aux[2]) if want_aux else None, ptr(weights), ptr(loss_sums), stream()),
          'render_fwd')
    return rgb, depth, aux, weights


def _render_backward(spec, raw, z, target_rgb, target_d, g_losses, coef, g_rgb, g_depth):
    n, s = z.shape
    d_raw = torch.empty((n, s, 4), dtype=torch.float32, device=z.device)
    lg = None
    keep = []
    if g_losses is not None and any(g is not None for g in g_losses):
        lg = LossGrads()
        for name, g in zip(('rgb', 'depth', 'sdf', 'fs'), g_losses):
            if g is not None:
                g = _f32c(g)
                keep.append(g)
                setattr(lg, name, g.data_ptr())
    check(_lib.load().cb200_render_bwd(ctypes.byref(spec.render_cfg()), ptr(raw), ptr(z), ptr(target_rgb), ptr(target_d),
                                       ctypes.c_int64(n), ctypes.c_int64(s), ctypes.byref(lg) if lg is not None else None,
                                       ptr(coef), ptr(g_rgb), ptr(g_depth), ptr(d_raw), stream()), 'render_bwd')
    return d_raw


def render_loss_backward(spec, raw, z, target_rgb, target_d, counts, loss_w, loss_sums, n_rays_total=None, list_ws=None, coef=None,
                         list_count_is_zero=False, d_raw=None):
    """cb200_render_loss_bwd: loss-head forward (sums) and backward (d_raw) of every ray in one launch, with the list of the points
    whose d_raw is non-zero written into `list_ws` (int32 tensor of cb200_points_bwd_workspace bytes) for points_backward(ready_list=).
    counts: the batch-global int64 [3] mask counts of the sampling; loss_w: four device scalars (or None)."""
    n, s = z.shape
    if d_raw is None:
        d_raw = torch.empty((n, s, 4), dtype=torch.float32, device=z.device)
    lg = LossGrads()
    for name, g in zip(('rgb', 'depth', 'sdf', 'fs'), loss_w):
        if g is not None:
            setattr(lg, name, g.data_ptr())
    check(_lib.load().cb200_render_loss_bwd(ctypes.byref(spec.render_cfg()), ptr(raw), ptr(z), ptr(target_rgb), ptr(target_d),
                                            ctypes.c_int64(n), ctypes.c_int64(s), ptr(counts), ctypes.c_int64(n_rays_total or n),
                                            ctypes.byref(lg), ptr(loss_sums), ptr(coef), ptr(d_raw), ptr(list_ws),
                                            ctypes.c_size_t(list_ws.numel() * list_ws.element_size() if list_ws is not None else 0),
                                            1 if list_count_is_zero else 0, stream()), 'render_loss_bwd')
    return d_raw


class RenderFunction(torch.autograd.Function):
    """raw2outputs + sdf2weights (model/scene_rep.py:77-124): (raw, z) -> rgb, depth, depth_var, disp, acc, weights.
    Gradients flow back from rgb and depth; the diagnostic maps are not differentiable here."""

    @staticmethod
    def forward(ctx, raw, z_vals, spec):
        r, z = _f32c(raw), _f32c(z_vals)
        rgb, depth, aux, weights = _render_forward(spec, r, z, want_aux=True, want_weights=True)
        ctx.spec = spec
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(r, z)
        ctx.mark_non_differentiable(aux, weights)
        return rgb, depth, aux, weights

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_aux, g_weights):
        if g_rgb is None and g_depth is None:
            return None, None, None
        r, z = ctx.saved_tensors
        d_raw = _render_backward(ctx.spec, r, z, None, None, None, None,
                                 _f32c(g_rgb) if g_rgb is not None else None, _f32c(g_depth) if g_depth is not None else None)
        return d_raw, None, None


class RenderLossFunction(torch.autograd.Function):
    """Rendering + the loss head of JointEncoding.forward (model/scene_rep.py:319-338, model/utils.py:138-161).

    `ws` is the zero-initialised step workspace: 8 x 8 bytes = [n_front, n_sdf, n_valid (int64) | pad | 4 double sums];
    the sampling kernel has already accumulated the counts.  With a process group the sums and counts are all-reduced
    before the losses are finished, so every rank sees the global losses.
    """

    @staticmethod
    def forward(ctx, raw, z_vals, target_rgb, target_d, ws, spec, n_rays_total, group):
        r, z = _f32c(raw), _f32c(z_vals)
        t_rgb, t_d = _f32c(target_rgb), _f32c(target_d.reshape(-1))
        counts, sums = ws[:4].view(torch.int64), ws[4:8]
        rgb, depth, _, _ = _render_forward(spec, r, z, t_rgb, t_d, loss_sums=sums)
        if group is not None:
            import torch.distributed as dist
            dist.all_reduce(counts, group=group)
            dist.all_reduce(sums, group=group)
        out = torch.empty(8, dtype=torch.float32, device=z.device)
        losses, coef = out[:4], out[4:]
        check(_lib.load().cb200_loss_finish(ptr(sums), ptr(counts), ctypes.c_int64(int(n_rays_total)),
                                            ctypes.c_int64(z.shape[1]), ptr(losses), ptr(coef), stream()), 'loss_finish')
        ctx.spec = spec
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(r, z, t_rgb, t_d, coef)
        l = losses.unbind(0)
        return rgb, depth, l[0], l[1], l[2], l[3]

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g0, g1, g2, g3):
        if all(g is None for g in (g_rgb, g_depth, g0, g1, g2, g3)):
            return (None,) * 8
        r, z, t_rgb, t_d, coef = ctx.saved_tensors
        d_raw = _render_backward(ctx.spec, r, z, t_rgb, t_d, (g0, g1, g2, g3), coef,
                                 _f32c(g_rgb) if g_rgb is not None else None, _f32c(g_depth) if g_depth is not None else None)
        if KEEP_LAST_D_RAW:
            global LAST_D_RAW
            LAST_D_RAW = d_raw
        return d_raw, None, None, None, None, None, None, None


def new_step_workspace(device):
    """8 x float64 zeroed: int64 counts in the first 4 slots, double loss sums in the last 4."""
    return torch.zeros(8, dtype=torch.float64, device=device)
