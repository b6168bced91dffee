"""Drop-in `JointEncoding` (upper boundary): same constructor, methods, attributes, return
dictionaries and state_dict keys as the reference's model/scene_rep.py:10-351, with every method routed through the
sm_100a kernels of libcoslam_b200.so.  coslam.py / coslam_mp.py use it unchanged:

    from coslam_b200.scene_rep import JointEncoding          # instead of  from model.scene_rep import JointEncoding

Sampling noise: the reference draws `torch.rand(z_vals.shape)` on the CPU generator and uploads it every call
(model/scene_rep.py:255).  `noise_mode='host'` (default, env COSLAM_B200_NOISE=host) does exactly that, so that a run
seeded like the reference consumes the same random stream; `noise_mode='device'` draws counter-based Philox noise
inside the sampling kernel (no host work, different stream -- throughput mode).
"""
import os

import numpy as np
import torch
import torch.nn as nn

from . import ops, tcnn
from .decoder import ColorSDFNet, ColorSDFNet_v2


def get_encoder(encoding, input_dim=3, degree=4, n_bins=16, n_frequencies=12, n_levels=16, level_dim=2,
                base_resolution=16, log2_hashmap_size=19, desired_resolution=512):
    """model/encodings.py:6-97 restricted to the two branches Co-SLAM configs select."""
    name = encoding.lower()
    if 'hash' in name or 'tiled' in name:
        print('Hash size', log2_hashmap_size)
        per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (n_levels - 1))
        embed = tcnn.Encoding(n_input_dims=input_dim, dtype=torch.float, encoding_config={
            'otype': 'HashGrid', 'n_levels': n_levels, 'n_features_per_level': level_dim,
            'log2_hashmap_size': log2_hashmap_size, 'base_resolution': base_resolution,
            'per_level_scale': per_level_scale})
    elif 'blob' in name:
        print('Use blob')
        embed = tcnn.Encoding(n_input_dims=input_dim, dtype=torch.float,
                              encoding_config={'otype': 'OneBlob', 'n_bins': n_bins})
    else:
        raise NotImplementedError('coslam_b200 serves the HashGrid and OneBlob encodings (every shipped config); got %r'
                                  % encoding)
    return embed, embed.n_output_dims


class JointEncoding(nn.Module):
    def __init__(self, config, bound_box):
        super().__init__()
        self.config = config
        self.bounding_box = bound_box
        self.noise_mode = os.environ.get('COSLAM_B200_NOISE', 'host')
        self.noise_seed = 0
        self.process_group = None          # set by coslam_b200.dist for ray-sharded mapping
        self._step = 0
        self._noise_counter = None
        self.get_resolution()
        self.get_encoding(config)
        self.get_decoder(config)
        self.spec = ops.SceneSpec(config, bound_box, self.embed_fn.grid_cfg, self.embedpos_fn.n_bins)

    # ---- construction (model/scene_rep.py:23-75) ------------------------------------------------------------------
    def get_resolution(self):
        dim_max = (self.bounding_box[:, 1] - self.bounding_box[:, 0]).max()
        g = self.config['grid']
        self.resolution_sdf = g['voxel_sdf'] if g['voxel_sdf'] > 10 else int(dim_max / g['voxel_sdf'])
        self.resolution_color = g['voxel_color'] if g['voxel_color'] > 10 else int(dim_max / g['voxel_color'])
        print('SDF resolution:', self.resolution_sdf)

    def get_encoding(self, config):
        self.embedpos_fn, self.input_ch_pos = get_encoder(config['pos']['enc'], n_bins=config['pos']['n_bins'])
        self.embed_fn, self.input_ch = get_encoder(config['grid']['enc'], log2_hashmap_size=config['grid']['hash_size'],
                                                   desired_resolution=self.resolution_sdf)
        if not config['grid']['oneGrid']:
            print('Color resolution:', self.resolution_color)
            self.embed_fn_color, self.input_ch_color = get_encoder(config['grid']['enc'],
                                                                   log2_hashmap_size=config['grid']['hash_size'],
                                                                   desired_resolution=self.resolution_color)

    def get_decoder(self, config):
        cls = ColorSDFNet_v2 if config['grid']['oneGrid'] else ColorSDFNet
        self.decoder = cls(config, input_ch=self.input_ch, input_ch_pos=self.input_ch_pos)
        # the reference registers the two sub-nets a second time (batchify(fn, None) returns fn itself), which is
        # where the duplicate `sdf_net.*` / `color_net.*` state_dict keys come from (model/scene_rep.py:74-75)
        self.color_net = self.decoder.color_net
        self.sdf_net = self.decoder.sdf_net

    # ---- helpers -----------------------------------------------------------------------------------------------
    def _fused(self):
        """(table, W1..W4) when the configuration is the one the fused kernels serve, else a loud error."""
        if self._modular():
            raise NotImplementedError('coslam_b200: the fused path serves oneGrid: True, tcnn_network: False (every shipped config)')
        w = self.decoder.fused_weights()
        if w is None:
            raise NotImplementedError('coslam_b200: the fused path serves the 2-layer SDF / 2-layer colour decoder')
        return (self.embed_fn.params,) + tuple(w)

    def _modular(self):
        """True for the two config-reachable variants no shipped config uses: `oneGrid: False` and
        `tcnn_network: True`.  They compose the stand-alone CUDA encodings with the torch decoder, like the reference does."""
        return (not self.config['grid']['oneGrid']) or bool(self.config['decoder']['tcnn_network'])

    def _decode_modular(self, flat):
        if self.config['grid']['oneGrid']:
            out = self.decoder(self.embed_fn(flat), self.embedpos_fn(flat))
        else:
            out = self.decoder(self.embed_fn(flat), self.embedpos_fn(flat), self.embed_fn_color(flat))
        return out.float()                                           # fp16 from tcnn.Network; the rendering kernels take fp32

    def ray_points(self, rays_o, rays_d, z_vals, target_d=None):
        """raw [N,S,4] of the sample points o + d z.  oneGrid (every shipped config): the fused per-point kernels, lazily over
        the points the step can depend on when `target_d` is given (training).  Two-grid variant (`oneGrid: False`,
        model/scene_rep.py:56-58, 168-170): the stand-alone CUDA encodings and the torch decoder, like the reference."""
        if not self._modular():
            return ops.RayPointsFunction.apply(rays_o, rays_d, z_vals, *self._fused(), self.spec, target_d)
        pts = rays_o[..., None, :] + rays_d[..., None, :] * z_vals[..., :, None]        # model/scene_rep.py:258
        return self.run_network(pts)

    @staticmethod
    def _require_cuda(t):
        if not t.is_cuda:
            raise RuntimeError('coslam_b200 has no CPU path: tensors must live on a CUDA device (got %s)' % t.device)

    def _noise(self, n_rays, n_samples, device):
        if self.noise_mode == 'host':
            return ops.host_uniform((n_rays, n_samples), device), 0    # torch.rand((N, S)).to(device): CPU generator, model/scene_rep.py:255
        if self._noise_counter is not None:                               # graph-captured loops: the step lives on the device
            return None, int(self.noise_seed) << 32
        self._step += 1
        return None, (int(self.noise_seed) << 32) + self._step

    def use_device_noise_counter(self, device):
        """Device-Philox noise whose step counter is a device tensor (incremented by a captured torch op after every
        sampling call), so a CUDA-graph replay of forward() draws fresh jitter.  Implies noise_mode = 'device'."""
        self.noise_mode = 'device'
        self._noise_counter = torch.zeros(1, dtype=torch.int64, device=device)
        return self._noise_counter

    # ---- model/scene_rep.py:77-124 ---------------------------------------------------------------------------------
    def sdf2weights(self, sdf, z_vals, args=None):
        self._require_cuda(sdf)
        raw = torch.cat([torch.zeros(sdf.shape + (3,), device=sdf.device, dtype=sdf.dtype), sdf[..., None]], -1)
        return ops.RenderFunction.apply(raw, z_vals, self.spec)[3]

    def raw2outputs(self, raw, z_vals, white_bkgd=False):
        self._require_cuda(raw)
        rgb_map, depth_map, aux, weights = ops.RenderFunction.apply(raw, z_vals, self.spec)
        if bool(white_bkgd) != self.spec.white_bkgd:
            acc = aux[2][..., None]
            rgb_map = rgb_map + (1. - acc) if white_bkgd else rgb_map - (1. - acc)
        return rgb_map, aux[1], aux[2], weights, depth_map, aux[0]

    # ---- model/scene_rep.py:126-171 --------------------------------------------------------------------------------
    def query_sdf(self, query_points, return_geo=False, embed=False):
        self._require_cuda(query_points)
        flat = torch.reshape(query_points, [-1, query_points.shape[-1]])
        lead = list(query_points.shape[:-1])
        if embed:
            e = self.embed_fn(flat)
            return torch.reshape(e, lead + [e.shape[-1]])
        if self._modular():
            h = self.sdf_net(torch.cat([self.embed_fn(flat), self.embedpos_fn(flat)], dim=-1), return_geo=True).float()
            raw, geo = h[..., :1].expand(-1, 4), h[..., 1:]
        else:
            raw, geo = ops.PointsFunction.apply(flat, *self._fused(), self.spec, True, True)
        sdf = torch.reshape(raw[..., 3], lead)
        if not return_geo:
            return sdf
        return sdf, torch.reshape(geo, lead + [geo.shape[-1]])

    def query_color(self, query_points):
        return torch.sigmoid(self.query_color_sdf(query_points)[..., :3])

    def query_color_sdf(self, query_points):
        self._require_cuda(query_points)
        flat = torch.reshape(query_points, [-1, query_points.shape[-1]])
        if self._modular():
            return self._decode_modular(flat)
        return ops.PointsFunction.apply(flat, *self._fused(), self.spec, True, False)[0]

    # ---- model/scene_rep.py:173-209 --------------------------------------------------------------------------------
    def run_network(self, inputs):
        self._require_cuda(inputs)
        flat = torch.reshape(inputs, [-1, inputs.shape[-1]])
        if self._modular():
            if self.config['grid']['tcnn_encoding']:
                flat = (flat - self.bounding_box[:, 0]) / (self.bounding_box[:, 1] - self.bounding_box[:, 0])
            out = self.query_color_sdf(flat)
        else:
            out = ops.PointsFunction.apply(flat, *self._fused(), self.spec, False, False)[0]
        return torch.reshape(out, list(inputs.shape[:-1]) + [out.shape[-1]])

    def render_surface_color(self, rays_o, normal):
        n_rays = rays_o.shape[0]
        trunc = self.config['training']['trunc']
        z_vals = torch.linspace(-trunc, trunc, steps=self.config['training']['n_range_d']).to(rays_o)
        z_vals = z_vals.repeat(n_rays, 1)
        pts = rays_o[..., :] + normal[..., None, :] * z_vals[..., :, None]
        raw = self.run_network(pts)
        return self.raw2outputs(raw, z_vals, self.config['training']['white_bkgd'])[0]

    # ---- model/scene_rep.py:211-295 --------------------------------------------------------------------------------
    def _sample(self, target_d, counts=None):
        n = target_d.shape[0]
        noise, seed = (None, 0)
        if self.spec.perturb:
            noise, seed = self._noise(n, self.spec.n_samples, target_d.device)
        z = ops.sample_rays(self.spec, target_d, noise=noise, seed=seed, counts=counts,
                            seed_dev=self._noise_counter if noise is None else None)
        if noise is None and self._noise_counter is not None:
            self._noise_counter.add_(1)
        return z

    def render_rays(self, rays_o, rays_d, target_d=None):
        self._require_cuda(rays_o)
        if target_d is None:
            # the reference reads config['training']['n_samples'] here, a key no shipped YAML defines
            # (model/scene_rep.py:246-248); keep the failure explicit
            raise KeyError("n_samples: render_rays without target_d is not configured in any Co-SLAM YAML")
        t = self.config['training']
        z_vals = self._sample(target_d)
        raw = self.ray_points(rays_o, rays_d, z_vals)
        rgb_map, depth_map, aux, weights = ops.RenderFunction.apply(raw, z_vals, self.spec)
        ret = {'rgb': rgb_map, 'depth': depth_map, 'disp_map': aux[1], 'acc_map': aux[2], 'depth_var': aux[0]}
        if t['n_importance'] > 0:
            from .hierarchical import importance_pass
            ret, z_vals, raw = importance_pass(self, rays_o, rays_d, z_vals, weights, ret)
        ret['z_vals'] = z_vals
        ret['raw'] = raw
        return ret

    # ---- model/scene_rep.py:297-351 --------------------------------------------------------------------------------
    def forward(self, rays_o, rays_d, target_rgb, target_d, global_step=0):
        if not self.training:
            return self.render_rays(rays_o, rays_d, target_d=target_d)
        self._require_cuda(rays_o)
        if self.config['training']['n_importance'] > 0:
            from .hierarchical import forward_with_importance
            return forward_with_importance(self, rays_o, rays_d, target_rgb, target_d)
        ws = ops.new_step_workspace(rays_o.device)
        z_vals = self._sample(target_d, counts=ws[:4].view(torch.int64))
        # training needs rgb / depth / losses only: sample points no loss mask and no rendering weight can reach are skipped
        raw = self.ray_points(rays_o, rays_d, z_vals, target_d)
        n_total = rays_o.shape[0]
        if self.process_group is not None:
            import torch.distributed as dist
            n_total = n_total * dist.get_world_size(self.process_group)
        rgb, depth, rgb_loss, depth_loss, sdf_loss, fs_loss = ops.RenderLossFunction.apply(
            raw, z_vals, target_rgb, target_d, ws, self.spec, n_total, self.process_group)
        return {'rgb': rgb, 'depth': depth, 'rgb_loss': rgb_loss, 'depth_loss': depth_loss, 'sdf_loss': sdf_loss,
                'fs_loss': fs_loss, 'psnr': mse2psnr(rgb_loss)}


_LOG10 = {}


def mse2psnr(x):
    """model/utils.py:8-12 (result has shape [1], like the reference's division by a 1-element tensor)."""
    c = _LOG10.get(x.device)
    if c is None:
        c = torch.log(torch.Tensor([10.])).to(x.device)
        _LOG10[x.device] = c
    return -10. * torch.log(x + 1e-6) / c
