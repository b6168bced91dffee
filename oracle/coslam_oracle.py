"""ORACLE (test infrastructure, not product code) -- CPU restatement, in plain
PyTorch fp32, of Co-SLAM's per-ray render-and-optimise hot path.

Every function cites the reference file:line it follows (paths relative to the
reference tree).  The restatement is validated in THIS container against the
reference's own unmodified `model.scene_rep.JointEncoding` (imported from
/root/reference with oracle/tcnn_shim.py standing in for tinycudann) by
oracle/gen_golden.py, which also writes the fixtures under tests/golden/.
The encodings themselves come from oracle/tcnn_shim.py, whose header explains
why parity is UNPINNED at the tinycudann boundary.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this file.  It never runs on the product path.
"""
import math
import random

import numpy as np
import torch
import torch.nn.functional as F

from . import tcnn_shim


# --------------------------------------------------------------------------------------
# configuration
# --------------------------------------------------------------------------------------

def room0_config():
    """Hyper-parameters of configs/Replica/replica.yaml + configs/Replica/room0.yaml that the hot path reads."""
    return {
        'grid': {'enc': 'HashGrid', 'tcnn_encoding': True, 'hash_size': 16, 'voxel_color': 0.08, 'voxel_sdf': 0.02,
                 'oneGrid': True},
        'pos': {'enc': 'OneBlob', 'n_bins': 16},
        'decoder': {'geo_feat_dim': 15, 'hidden_dim': 32, 'num_layers': 2, 'num_layers_color': 2,
                    'hidden_dim_color': 32, 'tcnn_network': False},
        'cam': {'H': 680, 'W': 1200, 'fx': 600.0, 'fy': 600.0, 'cx': 599.5, 'cy': 339.5, 'near': 0, 'far': 5,
                'depth_trunc': 100.},
        'training': {'rgb_weight': 5.0, 'depth_weight': 0.1, 'sdf_weight': 1000, 'fs_weight': 10,
                     'smooth_weight': 0.000001, 'smooth_pts': 32, 'smooth_vox': 0.1, 'smooth_margin': 0.05,
                     'n_samples_d': 32, 'range_d': 0.1, 'n_range_d': 11, 'n_importance': 0, 'perturb': 1,
                     'white_bkgd': False, 'trunc': 0.1, 'rot_rep': 'axis_angle', 'rgb_missing': 0.05},
        'data': {'sc_factor': 1, 'downsample': 1, 'translation': 0},
        'mapping': {'bound': [[-1.0, 7.0], [-1.3, 3.7], [-1.7, 1.4]], 'sample': 2048, 'keyframe_every': 5,
                    'min_pixels_cur': 100, 'n_pixels': 0.05, 'iters': 10},
        'tracking': {'sample': 1024, 'iter': 10},
    }


def scannet0000_config(hash_size=16):
    """configs/ScanNet/scannet.yaml + scene0000.yaml (scene0000.yaml:8 overrides hash_size 19 -> 16)."""
    c = room0_config()
    c['grid'].update({'hash_size': hash_size, 'voxel_sdf': 0.04, 'voxel_color': 0.08})
    c['cam'].update({'H': 480, 'W': 640, 'fx': 577.590698, 'fy': 578.729797, 'cx': 318.905426, 'cy': 242.683609,
                     'near': 0, 'far': 8, 'depth_trunc': 100., 'crop_edge': 10})
    c['training'].update({'smooth_weight': 0.001, 'smooth_pts': 64, 'smooth_vox': 0.1, 'n_samples_d': 96,
                          'range_d': 0.25, 'n_range_d': 21, 'trunc': 0.1, 'rot_rep': 'quat', 'rgb_missing': 0.0})
    c['mapping'].update({'bound': [[-0.1, 8.6], [-0.1, 8.9], [-0.3, 3.3]], 'min_pixels_cur': 20, 'optim_cur': False})
    return c


def tum_desk_config(n_importance=0):
    """configs/Tum/tum.yaml + fr1_desk.yaml.  This fork's tum.yaml:89 sets n_importance 128; the hierarchical
    pass is restated in `render_rays` below and enabled with n_importance=128."""
    c = room0_config()
    c['grid'].update({'hash_size': 16, 'voxel_sdf': 0.02, 'voxel_color': 0.04})
    c['cam'].update({'H': 480, 'W': 640, 'fx': 517.3, 'fy': 516.5, 'cx': 318.6, 'cy': 255.3, 'near': 0, 'far': 5,
                     'depth_trunc': 5., 'crop_edge': 8, 'crop_size': [384, 512]})
    c['training'].update({'rgb_weight': 1.0, 'depth_weight': 0.1, 'sdf_weight': 5000, 'fs_weight': 10,
                          'smooth_weight': 0.00000001, 'smooth_pts': 64, 'smooth_vox': 0.04, 'smooth_margin': 0.,
                          'n_samples_d': 64,
                          'range_d': 0.25, 'n_range_d': 21, 'n_importance': n_importance, 'trunc': 0.05,
                          'rgb_missing': 1.0})
    c['mapping'].update({'bound': [[-3.5, 3], [-3, 3], [-3, 3]], 'iters': 20})
    return c


def sdf_resolution(cfg, bound, voxel_key='voxel_sdf'):
    """model/scene_rep.py:23-35  get_resolution: int(dim_max / voxel_sdf) (and / voxel_color) evaluated with a torch fp32 scalar."""
    dim_max = (bound[:, 1] - bound[:, 0]).max()
    v = cfg['grid'][voxel_key]
    return v if v > 10 else int(dim_max / v)


def hashgrid_encoding_config(cfg, bound, n_levels=16, level_dim=2, base_resolution=16, voxel_key='voxel_sdf'):
    """model/encodings.py:31-46  (HashGrid branch of get_encoder)."""
    desired = sdf_resolution(cfg, bound, voxel_key)
    per_level_scale = np.exp2(np.log2(desired / base_resolution) / (n_levels - 1))
    return {'otype': 'HashGrid', 'n_levels': n_levels, 'n_features_per_level': level_dim,
            'log2_hashmap_size': cfg['grid']['hash_size'], 'base_resolution': base_resolution,
            'per_level_scale': per_level_scale}


# --------------------------------------------------------------------------------------
# the scene model
# --------------------------------------------------------------------------------------

class OracleScene(object):
    """Restatement of model/scene_rep.py:10-351 (JointEncoding) for the nn.Linear configurations, with the decoders of
    model/decoder.py:145-180 (ColorSDFNet_v2, `oneGrid: True`) and :110-143 (ColorSDFNet, `oneGrid: False`), :57-108 (SDFNet),
    :7-55 (ColorNet)."""

    def __init__(self, cfg, bound, seed=0):
        self.cfg = cfg
        self.bound = torch.as_tensor(bound, dtype=torch.float32)
        self.grid = tcnn_shim.Encoding(3, hashgrid_encoding_config(cfg, self.bound))
        self.blob = tcnn_shim.Encoding(3, {'otype': 'OneBlob', 'n_bins': cfg['pos']['n_bins']})
        d = cfg['decoder']
        in_sdf = self.grid.n_output_dims + self.blob.n_output_dims
        in_col = self.blob.n_output_dims + d['geo_feat_dim']
        # two-grid variant (`oneGrid: False`): a second hash grid at voxel_color (scene_rep.py:56-58, :32-33) whose feature
        # joins the colour net's input (ColorSDFNet, decoder.py:110-143: cat([embed_pos, embed_color, geo_feat]))
        self.grid_color = None
        if not cfg['grid']['oneGrid']:
            self.grid_color = tcnn_shim.Encoding(3, hashgrid_encoding_config(cfg, self.bound, voxel_key='voxel_color'))
            in_col += self.grid_color.n_output_dims
        g = torch.Generator().manual_seed(seed)

        def lin(o, i):   # nn.Linear default init: U(-1/sqrt(in), 1/sqrt(in)), bias=False (decoder.py:51, 104)
            k = 1.0 / math.sqrt(i)
            return ((torch.rand(o, i, generator=g) * 2 - 1) * k).requires_grad_(True)

        assert d['num_layers'] == 2 and d['num_layers_color'] == 2
        self.W1 = lin(d['hidden_dim'], in_sdf)                # sdf_net.model.0.weight   [32,80]
        self.W2 = lin(1 + d['geo_feat_dim'], d['hidden_dim'])  # sdf_net.model.2.weight   [16,32]
        self.W3 = lin(d['hidden_dim_color'], in_col)           # color_net.model.0.weight [32,63]
        self.W4 = lin(3, d['hidden_dim_color'])                # color_net.model.2.weight [3,32]
        self.training = True

    @property
    def table(self):
        return self.grid.params

    @property
    def table_color(self):
        return self.grid_color.params

    def to(self, device):
        """The restatement is plain torch, so it also runs on a CUDA device (used by the trajectory test to keep it short)."""
        self.bound = self.bound.to(device)
        self.grid.to(device)
        if self.grid_color is not None:
            self.grid_color.to(device)
        for k in ('W1', 'W2', 'W3', 'W4'):
            setattr(self, k, getattr(self, k).detach().to(device).requires_grad_(True))
        return self

    def parameters(self):
        return [self.grid.params, self.W1, self.W2, self.W3, self.W4] + ([self.grid_color.params] if self.grid_color is not None else [])

    def load_weights(self, table, W1, W2, W3, W4, table_color=None):
        with torch.no_grad():
            self.grid.params.copy_(torch.as_tensor(table))
            if table_color is not None:
                self.grid_color.params.copy_(torch.as_tensor(table_color))
            for dst, src in zip((self.W1, self.W2, self.W3, self.W4), (W1, W2, W3, W4)):
                dst.copy_(torch.as_tensor(src))

    # ---- decoder.py:165-180 ---------------------------------------------------------
    # `mask_override` = (m1 [P,32] bool, m3 [P,32] bool) replaces the ReLU activation pattern by a given one (used by the
    # parity tests to compare gradients under an identical activation pattern: the gradient of a ReLU network is
    # discontinuous where a pre-activation crosses zero, so two fp32-accurate forwards can legitimately disagree on a
    # handful of units whose pre-activation is ~1e-6).  A LIST of such pairs is consumed one per decode call (coarse pass,
    # importance pass).  `last_pre` keeps the pre-activations of the last call, `pre_log` of every call, for that audit.
    mask_override = None
    last_pre = None
    pre_log = None
    z_all_override = None
    last_z_all = None

    def decode(self, embed, embed_pos, embed_color=None):
        ov = self.mask_override
        if isinstance(ov, list):
            ov = ov.pop(0)
        h1 = torch.cat([embed, embed_pos], -1) @ self.W1.t()
        a1 = torch.relu(h1) if ov is None else torch.where(ov[0], h1, torch.zeros_like(h1))
        h = a1 @ self.W2.t()
        sdf, geo = h[..., :1], h[..., 1:]
        col_in = [embed_pos, geo] if embed_color is None else [embed_pos, embed_color, geo]      # decoder.py:176 / :138
        h3 = torch.cat(col_in, -1) @ self.W3.t()
        a3 = torch.relu(h3) if ov is None else torch.where(ov[1], h3, torch.zeros_like(h3))
        rgb = a3 @ self.W4.t()
        self.last_pre = (h1.detach(), h3.detach())
        if self.pre_log is not None:
            self.pre_log.append(self.last_pre)
        return torch.cat([rgb, sdf], -1)

    # ---- scene_rep.py:155-171 -------------------------------------------------------
    def query_color_sdf(self, x):
        flat = x.reshape(-1, x.shape[-1])
        if self.grid_color is not None:                            # scene_rep.py:168-170
            return self.decode(self.grid(flat), self.blob(flat), self.grid_color(flat))
        return self.decode(self.grid(flat), self.blob(flat))

    # ---- scene_rep.py:126-153 -------------------------------------------------------
    def query_sdf(self, x, return_geo=False, embed=False):
        flat = x.reshape(-1, x.shape[-1])
        e = self.grid(flat)
        if embed:
            return e.reshape(list(x.shape[:-1]) + [e.shape[-1]])
        h = torch.relu(torch.cat([e, self.blob(flat)], -1) @ self.W1.t()) @ self.W2.t()
        sdf = h[..., 0].reshape(list(x.shape[:-1]))
        if not return_geo:
            return sdf
        return sdf, h[..., 1:].reshape(list(x.shape[:-1]) + [h.shape[-1] - 1])

    def query_color(self, x):
        return torch.sigmoid(self.query_color_sdf(x)[..., :3])

    # ---- scene_rep.py:173-191 -------------------------------------------------------
    def normalise(self, pts):
        lo, hi = self.bound[:, 0], self.bound[:, 1]
        return (pts - lo) / (hi - lo)

    def run_network(self, pts):
        flat = pts.reshape(-1, pts.shape[-1])
        if self.cfg['grid']['tcnn_encoding']:
            flat = self.normalise(flat)
        return self.query_color_sdf(flat).reshape(list(pts.shape[:-1]) + [4])

    # ---- scene_rep.py:77-97 ---------------------------------------------------------
    def sdf2weights(self, sdf, z):
        tr = self.cfg['training']['trunc']
        w = torch.sigmoid(sdf / tr) * torch.sigmoid(-sdf / tr)
        crossing = (sdf[:, 1:] * sdf[:, :-1] < 0.0).to(sdf.dtype)
        first = torch.argmax(crossing, dim=1, keepdim=True)        # 0 when no sign change
        z_min = torch.gather(z, 1, first)
        keep = (z < z_min + self.cfg['data']['sc_factor'] * tr).to(sdf.dtype)
        w = w * keep
        return w / (w.sum(-1, keepdim=True) + 1e-8)

    # ---- scene_rep.py:99-124 --------------------------------------------------------
    def raw2outputs(self, raw, z):
        rgb = torch.sigmoid(raw[..., :3])
        w = self.sdf2weights(raw[..., 3], z)
        rgb_map = (w[..., None] * rgb).sum(-2)
        depth = (w * z).sum(-1)
        depth_var = (w * (z - depth[:, None]) ** 2).sum(-1)
        acc = w.sum(-1)
        disp = 1.0 / torch.maximum(1e-10 * torch.ones_like(depth), depth / acc)
        if self.cfg['training']['white_bkgd']:
            rgb_map = rgb_map + (1.0 - acc[..., None])
        return {'rgb': rgb_map, 'depth': depth, 'depth_var': depth_var, 'disp_map': disp, 'acc_map': acc,
                'weights': w}

    # ---- scene_rep.py:220-255 -------------------------------------------------------
    def sample_z(self, target_d, noise=None):
        """Depth-guided samples merged with uniform ones, then the stratified jitter.  `noise` stands for the
        CPU `torch.rand(z_vals.shape)` of scene_rep.py:255; None draws it from the default generator exactly there."""
        t, cam = self.cfg['training'], self.cfg['cam']
        n = target_d.shape[0]
        guided = torch.linspace(-t['range_d'], t['range_d'], steps=t['n_range_d']).to(target_d)
        zs = guided[None, :].repeat(n, 1) + target_d
        zs[target_d.squeeze(-1) <= 0] = torch.linspace(cam['near'], cam['far'], steps=t['n_range_d']).to(target_d)
        if t['n_samples_d'] > 0:
            uni = torch.linspace(cam['near'], cam['far'], t['n_samples_d'])[None, :].repeat(n, 1).to(target_d)
            z, _ = torch.sort(torch.cat([uni, zs], -1), -1)
        else:
            z = zs
        if t['perturb'] > 0.:
            mids = .5 * (z[..., 1:] + z[..., :-1])
            upper = torch.cat([mids, z[..., -1:]], -1)
            lower = torch.cat([z[..., :1], mids], -1)
            u = torch.rand(z.shape) if noise is None else noise
            z = lower + (upper - lower) * u.to(z)
        return z

    # ---- scene_rep.py:211-295 -------------------------------------------------------
    def render_rays(self, rays_o, rays_d, target_d, noise=None, noise_imp=None):
        z = self.sample_z(target_d, noise)
        pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
        raw = self.run_network(pts)
        out = self.raw2outputs(raw, z)
        n_imp = self.cfg['training']['n_importance']
        if n_imp > 0:                                              # scene_rep.py:265-277
            coarse = out
            mid = .5 * (z[..., 1:] + z[..., :-1])
            zi = sample_pdf(mid, coarse['weights'][..., 1:-1], n_imp,
                            det=(self.cfg['training']['perturb'] == 0.), u=noise_imp).detach()
            z, _ = torch.sort(torch.cat([z, zi], -1), -1)
            self.last_z_all = z
            if self.z_all_override is not None:                    # tests: continue from a given merged sample set
                z = self.z_all_override
            pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
            raw = self.run_network(pts)
            out = self.raw2outputs(raw, z)
            out.update({'rgb0': coarse['rgb'], 'depth0': coarse['depth'], 'disp0': coarse['disp_map'],
                        'acc0': coarse['acc_map'], 'depth_var0': coarse['depth_var'],
                        'z_std': torch.std(zi, dim=-1, unbiased=False)})
        out['z_vals'] = z
        out['raw'] = raw
        out['pts'] = pts
        return out

    # ---- scene_rep.py:297-351 + model/utils.py:89-161 -------------------------------
    def forward(self, rays_o, rays_d, target_rgb, target_d, noise=None, noise_imp=None):
        r = self.render_rays(rays_o, rays_d, target_d, noise, noise_imp)
        if not self.training:
            return r
        d = target_d.squeeze(-1)
        valid = (d > 0.) * (d < self.cfg['cam']['depth_trunc'])
        # scene_rep.py:320-321: rgb_weight is a BOOL tensor, so assigning rgb_missing casts it to bool(rgb_missing)
        rgb_w = valid.clone().unsqueeze(-1)
        rgb_w[rgb_w == 0] = self.cfg['training']['rgb_missing']
        rgb_loss = F.mse_loss(r['rgb'] * rgb_w, target_rgb * rgb_w)
        psnr = -10. * torch.log(rgb_loss + 1e-6) / torch.log(torch.Tensor([10.])).to(rgb_loss.device)   # model/utils.py:8-12
        depth_loss = F.mse_loss(r['depth'][valid], d[valid])
        if 'rgb0' in r:
            rgb_loss = rgb_loss + F.mse_loss(r['rgb0'] * rgb_w, target_rgb * rgb_w)
            depth_loss = depth_loss + F.mse_loss(r['depth0'][valid], d[valid])
        z, sdf = r['z_vals'], r['raw'][..., -1]
        T = self.cfg['training']['trunc'] * self.cfg['data']['sc_factor']
        front = (z < (target_d - T)).to(z.dtype)                  # model/utils.py:103
        back = (z > (target_d + T)).to(z.dtype)                   # :105
        dmask = (target_d > 0.0).to(z.dtype)                      # :107
        sdf_mask = (1.0 - front) * (1.0 - back) * dmask           # :109
        n_fs = torch.count_nonzero(front)
        n_sdf = torch.count_nonzero(sdf_mask)
        fs_w = 1.0 - n_fs / (n_fs + n_sdf)                        # :113-115
        sdf_w = 1.0 - n_sdf / (n_fs + n_sdf)
        fs_loss = F.mse_loss(sdf * front, torch.ones_like(sdf) * front) * fs_w           # :152-153
        sdf_loss = F.mse_loss((z + sdf * T) * sdf_mask, target_d * sdf_mask) * sdf_w     # :154-155
        out = dict(r)
        out.update({'rgb_loss': rgb_loss, 'depth_loss': depth_loss, 'sdf_loss': sdf_loss, 'fs_loss': fs_loss,
                    'psnr': psnr, 'n_front': n_fs, 'n_sdf': n_sdf, 'n_valid': valid.sum()})
        return out

    # ---- coslam.py:140-160 ----------------------------------------------------------
    def total_loss(self, ret, smooth=None):
        t = self.cfg['training']
        loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] \
            + t['sdf_weight'] * ret['sdf_loss'] + t['fs_weight'] * ret['fs_loss']
        if smooth is not None and t['smooth_weight'] > 0:
            loss = loss + t['smooth_weight'] * smooth
        return loss

    # ---- coslam.py:251-274 ----------------------------------------------------------
    def smoothness(self, offset_u=None, jitter_u=None):
        """Feature-grid total variation on a randomly placed (P-1)^3 lattice.  offset_u [3] and jitter_u [1,1,1,3]
        stand for the two CPU torch.rand draws of coslam.py:260 and :262."""
        t = self.cfg['training']
        n, vox, margin = t['smooth_pts'], t['smooth_vox'], t['smooth_margin']
        lo, hi = self.bound[:, 0], self.bound[:, 1]
        grid_size = (n - 1) * vox
        offset_max = hi - lo - grid_size - 2 * margin
        offset_u = torch.rand(3) if offset_u is None else offset_u
        offset = offset_u.to(offset_max) * offset_max + margin
        ar = torch.arange(0, n - 1, dtype=torch.long, device=lo.device)
        gx, gy, gz = torch.meshgrid(ar, ar, ar, indexing='ij')
        coords = torch.stack([gx, gy, gz], -1).float()
        jitter_u = torch.rand((1, 1, 1, 3)) if jitter_u is None else jitter_u
        pts = (coords + jitter_u.to(coords)) * vox + lo + offset
        x = (pts - lo) / (hi - lo)
        f = self.query_sdf(x, embed=True)
        tv = ((f[1:] - f[:-1]) ** 2).sum() + ((f[:, 1:] - f[:, :-1]) ** 2).sum() + ((f[:, :, 1:] - f[:, :, :-1]) ** 2).sum()
        return tv / (n ** 3)


# ---- model/utils.py:32-71 ---------------------------------------------------------------
def sample_pdf(bins, weights, n_imp, det=False, u=None):
    weights = weights + 1e-5
    pdf = weights / weights.sum(-1, keepdim=True)
    cdf = torch.cat([torch.zeros_like(pdf[..., :1]), torch.cumsum(pdf, -1)], -1)
    if det:
        u = torch.linspace(0. + 0.5 / n_imp, 1. - 0.5 / n_imp, steps=n_imp).expand(list(cdf.shape[:-1]) + [n_imp])
    elif u is None:
        u = torch.rand(list(cdf.shape[:-1]) + [n_imp])
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_b, cdf_a = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    bin_b, bin_a = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    return bin_b + (u - cdf_b) / denom * (bin_a - bin_b)


# --------------------------------------------------------------------------------------
# pose parametrisations (optimization/utils.py, optimization/pose_array.py)
# --------------------------------------------------------------------------------------

def axis_angle_to_matrix(r):
    """optimization/utils.py:5-21 -- Rodrigues; divides by theta, so r = 0 yields NaN exactly like the reference."""
    theta = torch.norm(r, dim=-1, keepdim=True)
    w = r / theta
    z = torch.zeros_like(w[..., 0:1])
    K = torch.stack([torch.cat([z, -w[..., 2:3], w[..., 1:2]], -1),
                     torch.cat([w[..., 2:3], z, -w[..., 0:1]], -1),
                     torch.cat([-w[..., 1:2], w[..., 0:1], z], -1)], dim=-2)
    eye = torch.eye(3).expand(*r.shape[:-1], 3, 3).to(r)
    return eye + torch.sin(theta).unsqueeze(-1) * K + (1. - torch.cos(theta).unsqueeze(-1)) * (K @ K)


def quaternion_to_matrix(q):
    """pytorch3d.transforms.quaternion_to_matrix (real-first), used by optimization/utils.py:43-54; identical
    arithmetic to optimization/pose_array.py:6-29 quad2rotation (two_s = 2 / |q|^2)."""
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def pose_matrix(rot, trans):
    """optimization/utils.py:30-55: [K,3|4] + [K,3] -> [K,4,4]."""
    R = axis_angle_to_matrix(rot) if rot.shape[-1] == 3 else quaternion_to_matrix(rot)
    T = torch.eye(4).to(rot)[None].repeat(rot.shape[0], 1, 1)
    T[:, :3, :3] = R
    T[:, :3, 3] = trans
    return T


def rays_from_pose(dir_cam, c2w):
    """coslam.py:192-193 / 546-547: rays_o = t (repeated), rays_d = sum_k dir_cam[k] * R[:, k]."""
    rays_o = c2w[None, :3, -1].repeat(dir_cam.shape[0], 1)
    rays_d = torch.sum(dir_cam[..., None, :] * c2w[:3, :3], -1)
    return rays_o, rays_d


def rays_from_pose_table(dir_cam, poses_all, ids):
    """coslam.py:353-355: per-ray pose gathered by keyframe slot (id -1 = last row = current frame)."""
    rays_d = torch.sum(dir_cam[..., None, None, :] * poses_all[ids, None, :3, :3], -1)
    rays_o = poses_all[ids, None, :3, -1].repeat(1, rays_d.shape[1], 1).reshape(-1, 3)
    return rays_o, rays_d.reshape(-1, 3)


def camera_rays(H, W, fx, fy, cx, cy):
    """datasets/utils.py:24-57 (OpenGL convention)."""
    i, j = torch.meshgrid(torch.arange(W, dtype=torch.float32), torch.arange(H, dtype=torch.float32), indexing='xy')
    return torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)


# ---- model/keyframe.py:96-106 and coslam.py:131-138 ------------------------------------------
def sample_global_ray_indices(num_kf, rays_per_kf, bs, rng=random):
    return rng.sample(range(num_kf * rays_per_kf), bs)


def select_samples(H, W, samples, rng=random):
    return rng.sample(range(H * W), int(samples))
