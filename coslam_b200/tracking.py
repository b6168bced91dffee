"""Per-frame camera tracking with the iteration loop captured in ONE CUDA graph.

`tracking_render` (coslam.py:502-573) runs `tracking.iter` dependent iterations of: pose parameters -> c2w -> rays ->
JointEncoding.forward -> weighted loss -> backward to the pose -> Adam step, with two `.cpu().item()` syncs per iteration for
the best-pose bookkeeping.  At 1024 rays the kernels take tens of microseconds and the loop is bound by launches and syncs.
FrameTracker keeps the same arithmetic and bookkeeping (first loss initialises the best pose; `loss < best` resets the patience
counter, otherwise it grows; stop once it exceeds `wait_iters`; return the best or the last pose) but keeps every decision on
the device, so the whole frame is one graph launch and one read-back.

coslam.py stays unchanged by design; this class is the opt-in fast path with the same inputs
(`batch['direction']` pixels, rgb, depth, initial c2w) and the same result (`est_c2w_data[frame_id]`).
"""
import ctypes

import torch

from . import _lib, pose
from ._lib import check, ptr, stream


class FrameTracker(object):
    def __init__(self, model, config, n_rays=None, device='cuda:0', use_graph=True):
        t, tr = config['tracking'], config['training']
        self.model, self.config, self.device = model, config, torch.device(device)
        self.n_rays = int(n_rays if n_rays is not None else t['sample'])
        self.iters, self.wait_iters, self.use_best = int(t['iter']), int(t['wait_iters']), bool(t['best'])
        self.lr_rot, self.lr_trans = float(t['lr_rot']), float(t['lr_trans'])
        self.quat = tr['rot_rep'] != 'axis_angle'                   # coslam.py:54-63
        self.loss_w = [float(tr[k]) for k in ('rgb_weight', 'depth_weight', 'sdf_weight', 'fs_weight')]
        self.use_graph = bool(use_graph)
        dev, n = self.device, self.n_rays
        f32 = dict(dtype=torch.float32, device=dev)
        self.dir_cam, self.target_s, self.target_d = torch.zeros((n, 3), **f32), torch.zeros((n, 3), **f32), torch.zeros((n, 1), **f32)
        self.rot = torch.zeros((1, 4 if self.quat else 3), **f32).requires_grad_(True)
        self.trans = torch.zeros((1, 3), **f32).requires_grad_(True)
        self.init_rot, self.init_trans = torch.zeros_like(self.rot), torch.zeros_like(self.trans)
        self.state = [torch.zeros_like(x) for x in (self.rot, self.rot, self.trans, self.trans)]   # m, v per parameter
        self.step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.best_loss, self.last_loss = torch.zeros((), **f32), torch.zeros((), **f32)
        self.best_c2w, self.last_c2w = torch.zeros((1, 4, 4), **f32), torch.zeros((1, 4, 4), **f32)
        self.patience = torch.zeros((), dtype=torch.int64, device=dev)
        self.alive = torch.ones((), dtype=torch.bool, device=dev)
        self.psnr = torch.zeros((1,), **f32)
        self._graph = None
        if self.use_graph:
            model.use_device_noise_counter(dev)                     # CPU torch.rand cannot be replayed by a graph

    # ---- one frame, eager or under capture --------------------------------------------------------------------------
    def _adam(self, p, g, m, v, lr):
        check(_lib.load().cb200_adam_step(ptr(p), ptr(g), ptr(m), ptr(v), ctypes.c_int64(p.numel()), ctypes.c_float(lr),
                                          ctypes.c_float(0.9), ctypes.c_float(0.999), ctypes.c_float(1e-8), ctypes.c_float(0.0),
                                          ptr(self.step), ctypes.c_int64(0), 0, stream()), 'adam_step')

    def _reset(self):
        with torch.no_grad():
            self.rot.copy_(self.init_rot)
            self.trans.copy_(self.init_trans)
            for s in self.state:
                s.zero_()
            self.step.zero_()
            self.patience.zero_()
            self.alive.fill_(True)

    def _iteration(self, k):
        c2w = pose.pose_matrix(self.rot, self.trans)                 # matrix_from_tensor, coslam.py:526
        rays_o, rays_d = pose.rays_from_pose(self.dir_cam, c2w)      # coslam.py:546-547
        ret = self.model.forward(rays_o, rays_d, self.target_s, self.target_d)
        w = self.loss_w                                              # get_loss_from_ret, coslam.py:140-160
        loss = w[0] * ret['rgb_loss'] + w[1] * ret['depth_loss'] + w[2] * ret['sdf_loss'] + w[3] * ret['fs_loss']
        may_stop = self.wait_iters < self.iters                      # otherwise the patience counter can never trip
        with torch.no_grad():
            lv, cd = loss.detach(), c2w.detach()
            if k == 0:                                               # coslam.py:552-554, then :559-563 sees loss == best
                self.best_loss.copy_(lv)
                self.best_c2w.copy_(cd)
                self.patience.fill_(1)
            else:
                better = (lv < self.best_loss) & self.alive
                self.best_loss.copy_(torch.where(better, lv, self.best_loss))
                self.best_c2w.copy_(torch.where(better, cd, self.best_c2w))
                if may_stop:
                    grown = torch.where(better, torch.zeros_like(self.patience), self.patience + 1)
                    self.patience.copy_(torch.where(self.alive, grown, self.patience))
            keep = self.alive.clone() if may_stop else None
            self.last_c2w.copy_(torch.where(keep, cd, self.last_c2w) if may_stop else cd)
            self.last_loss.copy_(torch.where(keep, lv, self.last_loss) if may_stop else lv)
            self.psnr.copy_(ret['psnr'].detach().reshape(1))
            if may_stop:
                self.alive.copy_(self.alive & (self.patience <= self.wait_iters))      # coslam.py:565-566: break before the update
        g_rot, g_trans = torch.autograd.grad(loss, (self.rot, self.trans))
        with torch.no_grad():
            old = (self.rot.detach().clone(), self.trans.detach().clone()) if may_stop else None
            self.step.add_(1)
            self._adam(self.rot, g_rot.contiguous(), self.state[0], self.state[1], self.lr_rot)
            self._adam(self.trans, g_trans.contiguous(), self.state[2], self.state[3], self.lr_trans)
            if may_stop:
                self.rot.copy_(torch.where(self.alive, self.rot, old[0]))
                self.trans.copy_(torch.where(self.alive, self.trans, old[1]))

    def _frame(self):
        self._reset()
        for k in range(self.iters):
            self._iteration(k)

    # ---- public -----------------------------------------------------------------------------------------------------
    def track(self, dir_cam, target_s, target_d, c2w_init):
        """dir_cam [N,3] camera-frame ray directions of the sampled pixels, target_s [N,3], target_d [N,1] (any device),
        c2w_init [4,4].  Returns the estimated c2w [4,4] (best or last, `tracking.best`) and a dict of device scalars."""
        c2w0 = torch.as_tensor(c2w_init, dtype=torch.float32).detach().cpu().reshape(1, 4, 4)
        rot0 = pose.matrix_to_quaternion(c2w0[:, :3, :3]) if self.quat else pose.matrix_to_axis_angle(c2w0[:, :3, :3])
        self.init_rot.copy_(rot0, non_blocking=True)
        self.init_trans.copy_(c2w0[:, :3, 3], non_blocking=True)
        self.dir_cam.copy_(dir_cam.reshape(self.n_rays, 3), non_blocking=True)
        self.target_s.copy_(target_s.reshape(self.n_rays, 3), non_blocking=True)
        self.target_d.copy_(target_d.reshape(self.n_rays, 1), non_blocking=True)
        frozen = [p for p in self.model.parameters() if p.requires_grad]
        for p in frozen:                                             # tracking needs dL/d rays only (no table scatter, no dW)
            p.requires_grad_(False)
        try:
            if not self.use_graph:
                self._frame()
            else:
                if self._graph is None:
                    self._capture()
                self._graph.replay()
        finally:
            for p in frozen:
                p.requires_grad_(True)
        est = (self.best_c2w if self.use_best else self.last_c2w)[0]
        return est, {'best_loss': self.best_loss, 'last_loss': self.last_loss, 'psnr': self.psnr, 'last_c2w': self.last_c2w[0],
                     'best_c2w': self.best_c2w[0]}

    def _capture(self):
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(2):                                       # warm-up: lazy constants, allocator, attribute calls
                self._frame()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._frame()
        self._graph = g
