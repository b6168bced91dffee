"""Seeded synthetic inputs of the benchmark configurations.
Host-side only (numpy / torch CPU); shared by the tests, the golden-fixture generator and bench.py."""
import math
import random

import numpy as np
import torch


def room0_config():
    """Hyper-parameters the hot path reads from configs/Replica/replica.yaml + configs/Replica/room0.yaml."""
    return {
        'grid': {'enc': 'HashGrid', 'tcnn_encoding': True, 'hash_size': 16, 'voxel_color': 0.08, 'voxel_sdf': 0.02,
                 'oneGrid': True},
        'pos': {'enc': 'OneBlob', 'n_bins': 16},
        'decoder': {'geo_feat_dim': 15, 'hidden_dim': 32, 'num_layers': 2, 'num_layers_color': 2,
                    'hidden_dim_color': 32, 'tcnn_network': False},
        'cam': {'H': 680, 'W': 1200, 'fx': 600.0, 'fy': 600.0, 'cx': 599.5, 'cy': 339.5, 'png_depth_scale': 6553.5,
                'crop_edge': 0, 'near': 0, 'far': 5, 'depth_trunc': 100.},
        'training': {'rgb_weight': 5.0, 'depth_weight': 0.1, 'sdf_weight': 1000, 'fs_weight': 10, 'eikonal_weight': 0,
                     'smooth_weight': 0.000001, 'smooth_pts': 32, 'smooth_vox': 0.1, 'smooth_margin': 0.05,
                     'n_samples_d': 32, 'range_d': 0.1, 'n_range_d': 11, 'n_importance': 0, 'perturb': 1,
                     'white_bkgd': False, 'trunc': 0.1, 'rot_rep': 'axis_angle', 'rgb_missing': 0.05},
        'data': {'sc_factor': 1, 'downsample': 1, 'translation': 0, 'num_workers': 4},
        'mapping': {'bound': [[-1.0, 7.0], [-1.3, 3.7], [-1.7, 1.4]], 'marching_cubes_bound': [[-1.0, 7.0], [-1.3, 3.7], [-1.7, 1.4]],
                    'sample': 2048, 'first_mesh': True, 'iters': 10, 'cur_frame_iters': 0, 'lr_embed': 0.01,
                    'lr_decoder': 0.01, 'lr_rot': 0.001, 'lr_trans': 0.001, 'keyframe_every': 5, 'map_every': 5,
                    'n_pixels': 0.05, 'first_iters': 200, 'optim_cur': True, 'min_pixels_cur': 100, 'map_accum_step': 1,
                    'pose_accum_step': 5, 'map_wait_step': 0, 'filter_depth': False},
        'tracking': {'iter': 10, 'sample': 1024, 'pc_samples': 40960, 'lr_rot': 0.001, 'lr_trans': 0.001,
                     'ignore_edge_W': 20, 'ignore_edge_H': 20, 'iter_point': 0, 'wait_iters': 100, 'const_speed': True,
                     'best': True},
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


def tum_desk_config(n_importance=0, perturb=1):
    """configs/Tum/tum.yaml + fr1_desk.yaml; this fork's tum.yaml:89 sets n_importance 128 (perturb 1, tum.yaml:90;
    perturb 0 selects sample_pdf's deterministic uniforms, model/utils.py:50)."""
    c = room0_config()
    c['grid'].update({'hash_size': 16, 'voxel_sdf': 0.02, 'voxel_color': 0.04})
    c['cam'].update({'H': 480, 'W': 640, 'fx': 517.3, 'fy': 516.5, 'cx': 318.6, 'cy': 255.3, 'near': 0, 'far': 5,
                     'depth_trunc': 5., 'crop_edge': 8, 'crop_size': [384, 512]})
    c['training'].update({'rgb_weight': 1.0, 'depth_weight': 0.1, 'sdf_weight': 5000, 'fs_weight': 10,
                          'smooth_weight': 0.00000001, 'smooth_pts': 64, 'smooth_vox': 0.04, 'smooth_margin': 0.,
                          'n_samples_d': 64, 'range_d': 0.25, 'n_range_d': 21, 'n_importance': n_importance,
                          'perturb': perturb, 'trunc': 0.05, 'rgb_missing': 1.0})
    c['mapping'].update({'bound': [[-3.5, 3], [-3, 3], [-3, 3]], 'iters': 20})
    return c


def room0_twogrid_config():
    """room0 with `grid.oneGrid: False`: a second hash grid at voxel_color 0.08 for the colour net (model/scene_rep.py:56-58,
    ColorSDFNet model/decoder.py:110-143).  Config-reachable, selected by no shipped YAML."""
    c = room0_config()
    c['grid']['oneGrid'] = False
    c['mapping']['lr_embed_color'] = 0.01
    return c


CONFIGS = {'room0': room0_config, 'scannet0000': scannet0000_config, 'tum_desk': tum_desk_config,
           'room0_twogrid': room0_twogrid_config}


def bound_tensor(cfg):
    return torch.tensor(cfg['mapping']['bound'], dtype=torch.float32)


def camera_ray_dirs(H, W, fx, fy, cx, cy):
    """Per-pixel camera-frame directions in the OpenGL convention (datasets/utils.py:24-57)."""
    i, j = torch.meshgrid(torch.arange(W, dtype=torch.float32), torch.arange(H, dtype=torch.float32), indexing='xy')
    return torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)


def _rodrigues(r):
    theta = float(np.linalg.norm(r))
    k = r / theta
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]], dtype=np.float64)
    return np.eye(3) + math.sin(theta) * K + (1 - math.cos(theta)) * (K @ K)


def ray_batch(cfg, n_rays, seed=0, invalid_frac=0.05, over_trunc_frac=0.0):
    """Config-1 style batch: one random camera inside the bound, `n_rays` distinct pixels, synthetic targets.
    Returns CPU fp32 tensors rays_o [N,3], rays_d [N,3], target_rgb [N,3], target_d [N,1]."""
    rng = np.random.RandomState(seed)
    pyrng = random.Random(seed)
    cam = cfg['cam']
    bound = np.asarray(cfg['mapping']['bound'], dtype=np.float64)
    centre = bound.mean(1)
    t = centre + rng.uniform(-0.5, 0.5, 3)
    axis = rng.normal(size=3)
    axis /= np.linalg.norm(axis)
    R = _rodrigues(axis * rng.uniform(0.1, 1.0))
    H, W = cam['H'], cam['W']
    dirs = camera_ray_dirs(H, W, cam['fx'], cam['fy'], cam['cx'], cam['cy']).reshape(-1, 3)
    if n_rays <= H * W:
        idx = torch.as_tensor(pyrng.sample(range(H * W), n_rays), dtype=torch.long)
    else:
        idx = torch.as_tensor(rng.randint(0, H * W, size=n_rays), dtype=torch.long)
    d_cam = dirs[idx]
    Rt = torch.as_tensor(R, dtype=torch.float32)
    rays_d = torch.sum(d_cam[..., None, :] * Rt, -1)                       # coslam.py:193
    rays_o = torch.as_tensor(t, dtype=torch.float32)[None, :].repeat(n_rays, 1)
    depth = torch.as_tensor(rng.uniform(0.5, min(4.5, cam['far'] - 0.5), size=(n_rays, 1)), dtype=torch.float32)
    u = rng.uniform(size=n_rays)
    depth[torch.as_tensor(u < invalid_frac)] = 0.0
    if over_trunc_frac > 0:
        depth[torch.as_tensor((u >= invalid_frac) & (u < invalid_frac + over_trunc_frac))] = cam['depth_trunc'] + 0.5
    rgb = torch.as_tensor(rng.uniform(size=(n_rays, 3)), dtype=torch.float32)
    return rays_o, rays_d, rgb, depth


def table_init(n_params, kind='small', seed=1337):
    """'small' = tiny-cuda-nn's U(-1e-4, 1e-4); 'trained' = U(-0.5, 0.5) so that the SDF changes sign along rays."""
    g = torch.Generator().manual_seed(seed)
    scale = 1e-4 if kind == 'small' else 0.5
    return (torch.rand(n_params, generator=g) * 2.0 - 1.0) * scale


def decoder_init(seed=0, col_in=63):
    """nn.Linear default init U(-1/sqrt(in), 1/sqrt(in)) for the four bias-free matrices [out, in]; col_in = 63 for the
    one-grid colour net (48 OneBlob + 15 geo), 95 with the colour grid's 32 features (two-grid variant)."""
    g = torch.Generator().manual_seed(seed)

    def lin(o, i):
        k = 1.0 / math.sqrt(i)
        return (torch.rand(o, i, generator=g) * 2 - 1) * k

    return lin(32, 80), lin(16, 32), lin(32, col_in), lin(3, 32)


def keyframe_ray_batch(cfg, n_rays, n_kf=64, seed=0, invalid_frac=0.05):
    """Mapping-BA style batch (coslam.py:336-355): rays drawn from `n_kf` keyframe cameras scattered in the bound.
    Returns CPU fp32 rays_o [N,3], rays_d [N,3], target_rgb [N,3], target_d [N,1] and the keyframe id of each ray."""
    rng = np.random.RandomState(seed)
    cam = cfg['cam']
    bound = np.asarray(cfg['mapping']['bound'], dtype=np.float64)
    centre, half = bound.mean(1), 0.25 * (bound[:, 1] - bound[:, 0])
    t = centre + rng.uniform(-1, 1, size=(n_kf, 3)) * half
    R = np.stack([_rodrigues(a / np.linalg.norm(a) * rng.uniform(0.1, 3.0)) for a in rng.normal(size=(n_kf, 3))])
    ids = torch.as_tensor(rng.randint(0, n_kf, size=n_rays), dtype=torch.long)
    H, W = cam['H'], cam['W']
    px = torch.as_tensor(rng.uniform(0, W, size=n_rays), dtype=torch.float32).floor()
    py = torch.as_tensor(rng.uniform(0, H, size=n_rays), dtype=torch.float32).floor()
    d_cam = torch.stack([(px - cam['cx']) / cam['fx'], -(py - cam['cy']) / cam['fy'], -torch.ones_like(px)], -1)
    Rt = torch.as_tensor(R, dtype=torch.float32)[ids]                      # [N,3,3]
    rays_d = torch.sum(d_cam[:, None, :] * Rt, -1)
    rays_o = torch.as_tensor(t, dtype=torch.float32)[ids]
    depth = torch.as_tensor(rng.uniform(0.5, min(4.5, cam['far'] - 0.5), size=(n_rays, 1)), dtype=torch.float32)
    depth[torch.as_tensor(rng.uniform(size=n_rays) < invalid_frac)] = 0.0
    rgb = torch.as_tensor(rng.uniform(size=(n_rays, 3)), dtype=torch.float32)
    return rays_o, rays_d, rgb, depth, ids
