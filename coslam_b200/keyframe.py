"""Device-resident keyframe database and fused ray assembly for mapping / bundle adjustment.

`KeyFrameDatabase` keeps the interface of model/keyframe.py:4-106 (constructor, `add_keyframe`, `sample_global_rays`,
`sample_global_keyframe`, `attach_ids`, `__len__`) but stores `rays` [num_kf, R, 7] in HBM (458 MB for room0's 401 x 40 800
rays), so a BA iteration moves `bs` int64 indices to the device instead of fancy-indexing a CPU tensor and uploading the
result (coslam.py:336-349).  The Python-`random` index draws are unchanged, hence the sampled ray indices are bit-identical
to the reference's for the same seed; `sample_global_indices_device` is the non-parity sampler for the large sweeps
(`random.sample` of 2^22 out of 16 M takes seconds on the host).

`assemble_rays` fuses the gather of the 7-float rays, the pose gather by keyframe slot and `rays_d = R . dir`
(coslam.py:346-355) into one kernel, with a backward that segment-sums dL/d rays into dL/d poses.
"""
import ctypes
import random

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr, stream


def sample_ranges(draws):
    """[`random.sample(range(n), k)` for (n, k) in draws] -- the same values in the same order and the same state of Python's
    global generator afterwards -- computed by the library's host routine `cb200_host_random_sample` (CPython's algorithm over
    a copy of the Mersenne-Twister state, written back with `random.setstate`) instead of the interpreter: the reference draws
    2048 + ray indices per BA iteration and 40 800 per keyframe this way (model/keyframe.py:45-46, :100; coslam.py:339-341),
    ~1 us per element in CPython.  Returns a list of int64 numpy arrays."""
    draws = [(int(n), int(k)) for n, k in draws]
    for n, k in draws:
        if not 0 <= k <= n:
            raise ValueError('Sample larger than population or is negative')
    if not draws:
        return []
    if any(n >= 2 ** 32 for n, _ in draws):
        return [np.asarray(random.sample(range(n), k), dtype=np.int64) for n, k in draws]
    version, mt, gauss = random.getstate()
    state = np.array(mt, dtype=np.uint32)
    ns = np.array([n for n, _ in draws], dtype=np.int64)
    ks = np.array([k for _, k in draws], dtype=np.int64)
    out = np.empty(int(ks.sum()), dtype=np.int64)
    rc = _lib.load().cb200_host_random_sample(ctypes.c_void_p(state.ctypes.data), ctypes.c_int64(len(draws)), ctypes.c_void_p(ns.ctypes.data),
                                              ctypes.c_void_p(ks.ctypes.data), ctypes.c_void_p(out.ctypes.data))
    if rc != 0:
        raise RuntimeError('cb200_host_random_sample failed (%d)' % rc)
    random.setstate((version, tuple(state.tolist()), gauss))
    return np.split(out, np.cumsum(ks)[:-1])


def sample_range(n, k):
    """`random.sample(range(n), k)` as an int64 numpy array (see sample_ranges)."""
    return sample_ranges([(n, k)])[0]


class _AssembleRays(torch.autograd.Function):
    @staticmethod
    def forward(ctx, kf_rays, ray_idx, pose_slot, poses):
        if not kf_rays.is_cuda:
            raise RuntimeError('coslam_b200.keyframe: the keyframe rays must live on a CUDA device (no CPU path)')
        kf = kf_rays.reshape(-1, 7)
        if kf.dtype != torch.float32 or not kf.is_contiguous():
            raise RuntimeError('coslam_b200.keyframe: keyframe rays must be contiguous fp32 [*, 7]')
        dev = kf.device
        idx = ray_idx.to(device=dev, dtype=torch.int64).contiguous()
        slot = pose_slot.to(device=dev, dtype=torch.int64).contiguous()
        T = poses.detach().to(torch.float32).contiguous()
        n, k = idx.shape[0], T.shape[0]
        rays_o, rays_d, rgb = (torch.empty((n, 3), dtype=torch.float32, device=dev) for _ in range(3))
        d = torch.empty((n, 1), dtype=torch.float32, device=dev)
        check(_lib.load().cb200_assemble_rays_fwd(ptr(kf), ptr(idx), ptr(slot), ptr(T), ctypes.c_int64(n), ctypes.c_int64(k),
                                                  ptr(rays_o), ptr(rays_d), ptr(rgb), ptr(d), stream()), 'assemble_rays_fwd')
        ctx.save_for_backward(kf, idx, slot)
        ctx.n_poses = k
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(rgb, d)
        return rays_o, rays_d, rgb, d

    @staticmethod
    def backward(ctx, g_o, g_d, g_rgb, g_depth):
        if g_o is None and g_d is None:
            return None, None, None, None
        kf, idx, slot = ctx.saved_tensors
        n = idx.shape[0]
        z = None
        if g_o is None or g_d is None:
            z = torch.zeros((n, 3), dtype=torch.float32, device=kf.device)
        g_o = z if g_o is None else g_o.to(torch.float32).contiguous()
        g_d = z if g_d is None else g_d.to(torch.float32).contiguous()
        d_poses = torch.zeros((ctx.n_poses, 4, 4), dtype=torch.float32, device=kf.device)
        check(_lib.load().cb200_assemble_rays_bwd(ptr(kf), ptr(idx), ptr(slot), ptr(g_o), ptr(g_d), ctypes.c_int64(n),
                                                  ctypes.c_int64(ctx.n_poses), ptr(d_poses), stream()), 'assemble_rays_bwd')
        return None, None, None, d_poses


def assemble_rays(kf_rays, ray_idx, pose_slot, poses):
    """kf_rays [..., 7] device fp32, ray_idx [N] rows of its [-1, 7] view, pose_slot [N] keyframe slots (negative counts from
    the end), poses [K,4,4] -> rays_o [N,3], rays_d [N,3], target_rgb [N,3], target_d [N,1]; differentiable in `poses`."""
    return _AssembleRays.apply(kf_rays, ray_idx, pose_slot, poses)


class KeyFrameDatabase(object):
    """model/keyframe.py:4-106 with the ray store on the device."""

    def __init__(self, config, H, W, num_kf, num_rays_to_save, device):
        self.config = config
        self.keyframes = {}
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise RuntimeError('coslam_b200.keyframe.KeyFrameDatabase keeps its rays in device memory; got %s' % device)
        self.rays = torch.zeros((num_kf, num_rays_to_save, 7), dtype=torch.float32, device=self.device)
        self.num_rays_to_save = num_rays_to_save
        self.frame_ids = None
        self.H = H
        self.W = W

    def __len__(self):
        return len(self.frame_ids)

    def get_length(self):
        return self.__len__()

    def sample_single_keyframe_rays(self, rays, option='random'):
        """model/keyframe.py:40-63: same Python-`random` draws, the gather runs where `rays` lives."""
        if option == 'random':
            idxs = sample_range(self.H * self.W, self.num_rays_to_save)
        elif option == 'filter_depth':
            valid = (rays[..., -1] > 0.0) & (rays[..., -1] <= self.config['cam']['depth_trunc'])
            rays_valid = rays[valid, :]
            idxs = sample_range(len(rays_valid), self.num_rays_to_save)
        else:
            raise NotImplementedError()
        return rays[:, torch.as_tensor(idxs, device=rays.device)]

    def attach_ids(self, frame_ids):
        frame_ids = frame_ids.cpu()
        self.frame_ids = frame_ids if self.frame_ids is None else torch.cat([self.frame_ids, frame_ids], dim=0)

    def add_keyframe(self, batch, filter_depth=False):
        rays = torch.cat([batch['direction'], batch['rgb'], batch['depth'][..., None]], dim=-1)
        rays = rays.reshape(1, -1, rays.shape[-1])
        rays = self.sample_single_keyframe_rays(rays, 'filter_depth' if filter_depth else 'random')
        if not isinstance(batch['frame_id'], torch.Tensor):
            batch['frame_id'] = torch.tensor([batch['frame_id']])
        self.attach_ids(batch['frame_id'])
        self.rays[len(self.frame_ids) - 1] = rays.to(self.device)

    # ---- model/keyframe.py:96-106 ------------------------------------------------------------------------------------
    def sample_global_indices(self, bs):
        """The reference's draw: `random.sample(range(num_kf * R), bs)` (without replacement) as an int64 tensor (host)."""
        return torch.from_numpy(sample_range(len(self) * self.num_rays_to_save, bs))

    def sample_global_indices_device(self, bs, generator=None):
        """Non-parity sampler for the large sweeps: device uniform draws (with replacement)."""
        return torch.randint(0, len(self) * self.num_rays_to_save, (bs,), device=self.device, generator=generator)

    def sample_global_rays(self, bs):
        idxs = self.sample_global_indices(bs)
        sample_rays = self.rays[:len(self)].reshape(-1, 7)[idxs.to(self.device)]
        frame_ids = self.frame_ids[idxs // self.num_rays_to_save]
        return sample_rays, frame_ids

    def sample_global_keyframe(self, window_size, n_fixed=1):
        if window_size >= len(self.frame_ids):
            return self.rays[:len(self.frame_ids)], self.frame_ids
        current_num_kf = len(self.frame_ids)
        last_frame_ids = self.frame_ids[-n_fixed:]
        idx = random.sample(range(0, len(self.frame_ids) - n_fixed), window_size)
        idx_rays = idx + list(range(current_num_kf - n_fixed, current_num_kf))
        return self.rays[idx_rays], torch.cat([self.frame_ids[idx], last_frame_ids], dim=0)

    # ---- fused BA batch (coslam.py:336-355) ---------------------------------------------------------------------------
    def assemble_ba_batch(self, bs, poses_all, keyframe_every, current_rays=None, n_current=0, device_sampler=False):
        """Global rays + `n_current` rays of the current frame (`current_rays` [H*W,7] on the device, appended after the
        keyframe store's rows as in coslam.py:338-344) -> (rays_o, rays_d, target_rgb, target_d, ids_all).
        `poses_all` [K,4,4] may require grad; slot -1 is the current frame."""
        idxs = self.sample_global_indices_device(bs) if device_sampler else self.sample_global_indices(bs)
        ids = (self.frame_ids.to(idxs.device)[idxs // self.num_rays_to_save] // keyframe_every).to(torch.int64)
        store = self.rays[:len(self)].reshape(-1, 7)
        out = assemble_rays(store, idxs, ids, poses_all)
        ids = ids.to(self.device)
        if current_rays is not None and n_current > 0:              # the store is never copied: second, small assembly
            idx_cur = torch.from_numpy(sample_range(self.H * self.W, n_current))
            cur = current_rays.reshape(-1, 7).to(self.device)
            ids_cur = -torch.ones(n_current, dtype=torch.int64, device=self.device)
            out_cur = assemble_rays(cur, idx_cur, ids_cur, poses_all)
            out = tuple(torch.cat([a, b], 0) for a, b in zip(out, out_cur))
            ids = torch.cat([ids, ids_cur])
        return out + (ids,)
