"""Per-frame camera tracking with the iteration loop captured in ONE CUDA graph.

`tracking_render` (coslam.py:502-573) runs `tracking.iter` dependent iterations of: pose parameters -> c2w -> rays ->
JointEncoding.forward -> weighted loss -> backward to the pose -> Adam step, with two `.cpu().item()` syncs per iteration for
the best-pose bookkeeping.  At 1024 rays the kernels take tens of microseconds and the loop is bound by launches and syncs.
FrameTracker keeps the same arithmetic and bookkeeping (first loss initialises the best pose; `loss < best` resets the patience
counter, otherwise it grows; stop once it exceeds `wait_iters`; return the best or the last pose) but keeps every decision on
the device, so the whole frame is one graph launch and one read-back.

coslam.py stays unchanged by design; this class is the opt-in fast path with the same inputs
(`batch['direction']` pixels, rgb, depth, initial c2w) and the same result (`est_c2w_data[frame_id]`).

Two implementations of one iteration:
  * `fused=True` (default for the shipped single-grid configs): no autograd graph at all.  Without importance samples an iteration is
    SIX launches (`_iteration_lean`): one fill | `cb200_track_rays_sample` (rays from the device c2w + depth-guided samples) |
    per-point forward | `cb200_render_loss_bwd` (rendering, loss sums, d_raw and the live-point list: the loss head's backward needs the
    mask counts only) | per-point backward (dL/d rays only) | `cb200_track_pose_update` (pose-gradient segment sum and, in its last
    CTA, the weighted loss, the best / last bookkeeping, the backward of the pose parametrisation, both Adam steps, the next c2w and
    the noise counter).  With importance samples (this fork's TUM config) the coarse pass, `cb200_importance_sample_merge` and the
    separate loss-head kernels run; the fine pass evaluates only the new samples and one backward serves both passes.
  * `persistent=True` (opt-in; needs `fused`, `use_graph` and no importance pass): the whole frame -- every iteration of the list
    above -- is ONE cooperative launch (`cb200_track_frame`, csrc/track_frame.cu): CTAs stay resident, phases are separated by
    grid-wide barriers, the decoder weights are staged once per frame and the loop leaves early when the patience counter trips.
    Bit-identical to the graph-captured loop at perturb 0; measured 0.86 ms against 0.82 ms of device time per room0 frame
    (profiles/r2/NOTES.md: the two per-point phases are the same tile loops and take 52 of the 87 us of an iteration either way,
    six grid barriers at ~2 us cost more than thirteen in-graph kernel boundaries at ~1 us), so the graph stays the default.
  * `fused=False`: the same arithmetic through torch autograd (`pose.py`, `JointEncoding.forward`), ~100 graph nodes per
    iteration; kept as the A/B twin and for the two-grid / importance-sampling variants.
"""
import ctypes

import torch

from . import _lib, ops, pose
from ._lib import check, ptr, stream


class FrameTracker(object):
    def __init__(self, model, config, n_rays=None, device='cuda:0', use_graph=True, fused=None, persistent=None):
        t, tr = config['tracking'], config['training']
        can_fuse = bool(config['grid']['oneGrid']) and not config['decoder']['tcnn_network']
        if fused and not can_fuse:
            raise NotImplementedError('FrameTracker(fused=True) serves oneGrid: True, tcnn_network: False (every shipped config)')
        self.n_imp = int(tr['n_importance'])
        self.fused = can_fuse if fused is None else bool(fused)
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
        if self.fused:
            nr = self.rot.shape[1]
            self.rot.requires_grad_(False)
            self.trans.requires_grad_(False)
            # [rot | trans | exp_avg(rot) | exp_avg_sq(rot) | exp_avg(trans) | exp_avg_sq(trans)]: one copy resets a frame
            self.pstate, self.pinit = torch.zeros(3 * nr + 9, **f32), torch.zeros(3 * nr + 9, **f32)
            self.rot, self.trans, self.adam = self.pstate[:nr].view(1, nr), self.pstate[nr:nr + 3].view(1, 3), self.pstate[nr + 3:]
            self.init_rot, self.init_trans = self.pinit[:nr].view(1, nr), self.pinit[nr:nr + 3].view(1, 3)
            self.istate = torch.zeros(2, dtype=torch.int64, device=dev)            # Adam step, patience
            self.step, self.patience = self.istate[0:1], self.istate[1:2]
            self.alive = torch.ones(1, dtype=torch.uint8, device=dev)
            self.best_loss, self.last_loss, self.loss = torch.zeros(1, **f32), torch.zeros(1, **f32), torch.zeros(1, **f32)
            self.c2w = torch.zeros((1, 4, 4), **f32)
            self.kf7 = torch.zeros((n, 7), **f32)                                   # (camera-frame direction, rgb, depth) per ray
            self.ray_idx = torch.arange(n, dtype=torch.int64, device=dev)
            self.pose_slot = torch.zeros(n, dtype=torch.int64, device=dev)
            self.rays_o, self.rays_d = torch.zeros((n, 3), **f32), torch.zeros((n, 3), **f32)
            # everything an iteration starts from zero, cleared by ONE fill: dL/d rays_o, dL/d rays_d [2, n, 3] (rounded up to 256 bytes) |
            # a 256-byte header: step workspace (8 doubles: 4 counts, 4 loss sums) | dL/d c2w (16 floats) | coarse-pass loss sums
            # (4 doubles) | ticket of cb200_track_pose_update | ... | count of the live-point list in its last 16 bytes.  The list
            # itself follows the header (cb200_points_bwd_workspace layout: count, pad, indices) and is not cleared.
            S_all = model.spec.n_samples + self.n_imp
            R = (24 * n + 255) // 256 * 256
            self.zbytes = torch.zeros(R + 256 + 4 * n * S_all, dtype=torch.uint8, device=dev)
            self.zclear = self.zbytes[:R + 256]
            hdr = self.zbytes[R:R + 256]
            self.zbuf = hdr[:128].view(torch.float64)
            self.d_c2w = self.zbuf[8:].view(torch.float32)
            self.sums0 = hdr[128:160].view(torch.float64)
            self.ticket = hdr[160:164].view(torch.int32)
            self.d_rays = self.zbytes[:24 * n].view(torch.float32).view(2, n, 3)
            self._list_ws = self.zbytes[R + 240:].view(torch.int32)
            self.z = torch.zeros((n, model.spec.n_samples), **f32)
            self.raw, self.d_raw = torch.zeros((n * model.spec.n_samples, 4), **f32), torch.zeros((n, model.spec.n_samples, 4), **f32)
            self.mask = torch.zeros((n * model.spec.n_samples, 2), dtype=torch.int32, device=dev)
            self.loss_out = torch.zeros(8, **f32)                                   # 4 losses | 4 d loss / d sum
            self.w_dev = torch.tensor(self.loss_w, **f32).unbind(0)
            self.g_rot = torch.zeros(nr, **f32)
            self.c2w_init = torch.zeros(16, **f32)
            self._kf7_host = torch.zeros((n, 7), dtype=torch.float32).pin_memory()
            self._steps = [self._track_step(first) for first in (True, False)]
            self._steps_own = [self._track_step(first, own_losses=True) for first in (True, False)]
        can_persist = self.fused and self.use_graph and self.n_imp == 0
        if persistent and not can_persist:
            raise NotImplementedError('FrameTracker(persistent=True) needs fused=True, use_graph=True (device noise) and n_importance 0')
        self.persistent = bool(persistent)
        self._wsp = None
        self._kf7_sent = None

    def _track_step(self, first, own_losses=False):
        s = _lib.TrackStep()
        if own_losses:                                               # losses from the iteration's sums / counts inside the update kernel
            s.sums, s.counts = self.zbuf[4:8].data_ptr(), self.zbuf[:4].data_ptr()
            s.n_rays_total, s.n_samples = self.n_rays, self.model.spec.n_samples
        for name, t in (('rot', self.rot), ('trans', self.trans), ('adam', self.adam), ('step', self.step), ('losses', self.loss_out),
                        ('d_c2w', self.d_c2w), ('c2w', self.c2w), ('best_loss', self.best_loss), ('best_c2w', self.best_c2w),
                        ('last_loss', self.last_loss), ('last_c2w', self.last_c2w), ('psnr', self.psnr), ('loss', self.loss),
                        ('g_rot_out', self.g_rot), ('patience', self.patience), ('alive', self.alive)):
            setattr(s, name, t.data_ptr())
        s.loss_w = (ctypes.c_float * 4)(*self.loss_w)
        s.lr_rot, s.lr_trans, s.n_rot = self.lr_rot, self.lr_trans, self.rot.shape[1]
        s.first_iter, s.may_stop, s.wait_iters = int(first), int(self.wait_iters < self.iters), self.wait_iters
        return s

    # ---- one frame without autograd -----------------------------------------------------------------------------------
    def _reset_fused(self):
        """Pose parameters of the initial matrix + a fresh Adam / patience state, on the device (cb200_track_init)."""
        check(_lib.load().cb200_track_init(ptr(self.c2w_init), ctypes.byref(self._steps[0]), stream()), 'track_init')

    def _iteration_lean(self, k):
        """Six launches: fill | rays from the pose + sampling | per-point forward | rendering, loss sums, d_raw and live-point list |
        per-point backward over the list | pose-gradient sum + pose update (losses, bookkeeping, Adam, next c2w, noise counter)."""
        lib, m, n = _lib.load(), self.model, self.n_rays
        spec = m.spec
        fw = m._fused()
        tb, w = fw[0].detach(), tuple(x.detach() for x in fw[1:])
        self.zclear.zero_()
        ws = self.zbuf[:8]
        counts = ws[:4].view(torch.int64)
        td = self.target_d.view(-1)
        noise, seed = m._noise(n, spec.n_samples, self.device) if spec.perturb else (None, 0)      # host mode: torch.rand's CPU stream, uploaded
        counter = m._noise_counter if noise is None else None
        lin_u, lin_g, lin_n = spec.linspaces(self.device)
        check(lib.cb200_track_rays_sample(ctypes.byref(spec.render_cfg()), ptr(self.kf7), ptr(self.c2w), ptr(lin_u), ptr(lin_g), ptr(lin_n),
                                          ptr(noise), ctypes.c_uint64(int(seed) & (2 ** 64 - 1)), ptr(counter), ctypes.c_int64(n), ptr(self.rays_o),
                                          ptr(self.rays_d), ptr(self.target_s), ptr(td), ptr(self.z), ptr(counts), stream()),
              'track_rays_sample')
        rays = (self.rays_o, self.rays_d, self.z)
        ops.points_forward(spec.scene(), tb, w, rays=rays, out=(self.raw, self.mask))
        ops.render_loss_backward(spec, self.raw, self.z, self.target_s, td, counts, self.w_dev, ws[4:8], list_ws=self._list_ws,
                                 list_count_is_zero=True, d_raw=self.d_raw)
        ops.points_backward(spec.scene(), tb, w, self.mask, self.d_raw, rays=rays, accumulate=(None, None, self.d_rays),
                            ready_list=self._list_ws)
        check(lib.cb200_track_pose_update(ptr(self.kf7), ptr(self.d_rays[0]), ptr(self.d_rays[1]), ctypes.c_int64(n),
                                          ctypes.byref(self._steps_own[0 if k == 0 else 1]), ptr(self.ticket), ptr(counter), stream()),
              'track_pose_update')

    def _iteration_fused(self, k):
        lib, m, n = _lib.load(), self.model, self.n_rays
        spec = m.spec
        if self.n_imp == 0 and lib.cb200_get_decoder() == 1:
            return self._iteration_lean(k)
        fw = m._fused()
        tb, w = fw[0].detach(), tuple(x.detach() for x in fw[1:])
        self.zclear.zero_()
        ws = self.zbuf[:8]
        td = self.target_d.view(-1)
        check(lib.cb200_assemble_rays_fwd(ptr(self.kf7), ptr(self.ray_idx), ptr(self.pose_slot), ptr(self.c2w), ctypes.c_int64(n),
                                          ctypes.c_int64(1), ptr(self.rays_o), ptr(self.rays_d), ptr(self.target_s), ptr(td),
                                          stream()), 'assemble_rays_fwd')
        counts = ws[:4].view(torch.int64)
        coarse = None
        if self.n_imp > 0:
            # hierarchical pass (scene_rep.py:262-277, :327-329): coarse samples -> rendering weights -> n_importance extra depths
            # per ray (cb200_importance_sample, counts of the MERGED samples) -> all samples again.  The coarse colour / depth
            # terms share the normalisers of the fine ones, so their squared-residual sums are simply added to the fine sums.
            from .hierarchical import draw_uniforms
            z0 = m._sample(td)
            rays0 = (self.rays_o, self.rays_d, z0)
            raw0, _, mask0 = ops.points_forward(spec.scene(), tb, w, rays=rays0)
            sums0 = self.sums0
            _, _, _, weights0 = ops._render_forward(spec, raw0, z0, self.target_s, td, loss_sums=sums0, want_weights=True)
            u = draw_uniforms(n, self.n_imp, z0.device, det=(self.config['training']['perturb'] == 0.))
            z, _, (rank, new_list, new_count) = ops.importance_sample(spec, z0, weights0, u, target_d=td, counts=counts, want_std=False, merge=True)
            coarse = (rays0, raw0, mask0, sums0, rank)
        else:
            z = m._sample(td, counts=counts)
        rays = (self.rays_o, self.rays_d, z)
        if coarse is not None and lib.cb200_get_decoder() == 1:
            # the coarse samples are part of the merged rows with the same positions, i.e. the same network outputs: evaluate the NEW
            # samples only and take the coarse rows over
            P_f = n * z.shape[1]
            raw = torch.empty((P_f, 4), dtype=torch.float32, device=self.device)
            mask = torch.empty((P_f, 2), dtype=torch.int32, device=self.device)
            ops.points_forward(spec.scene(), tb, w, rays=rays, out=(raw, mask), point_list=(ptr(new_list), ptr(new_count)))
            ops.merge_coarse(rank, raw0, raw, mask0, mask)
        else:
            raw, _, mask = ops.points_forward(spec.scene(), tb, w, rays=rays)
        into = (None, None, self.d_rays)                             # both passes add into the same ray-gradient buffer
        if coarse is None and lib.cb200_get_decoder() == 1:
            # rendering, loss sums, d_raw and the live-point list of every ray in ONE launch: d loss / d sum needs the mask counts only,
            # which the sampling produced; the losses themselves are taken from the sums inside cb200_track_update
            d_raw = ops.render_loss_backward(spec, raw, z, self.target_s, td, counts, self.w_dev, ws[4:8], list_ws=self._list_ws,
                                             list_count_is_zero=True)
            ops.points_backward(spec.scene(), tb, w, mask, d_raw, rays=rays, accumulate=into, ready_list=self._list_ws)
            step = self._steps_own[0 if k == 0 else 1]
        else:
            ops._render_forward(spec, raw, z, self.target_s, td, loss_sums=ws[4:8])
            if coarse is not None:
                ws[4:6].add_(coarse[3][:2])                          # colour and depth only: the coarse pass has no SDF / free-space terms
            check(lib.cb200_loss_finish(ptr(ws[4:8]), ptr(counts), ctypes.c_int64(n), ctypes.c_int64(z.shape[1]),
                                        ptr(self.loss_out[:4]), ptr(self.loss_out[4:]), stream()), 'loss_finish')
            d_raw = ops._render_backward(spec, raw, z, self.target_s, td, self.w_dev, self.loss_out[4:], None, None)
            if coarse is not None:
                # the coarse maps' gradient (colour and depth terms only) lands on the same points: add it to the merged rows' d_raw and
                # run ONE per-point backward
                rays0, raw0, mask0, _, rank = coarse
                d_raw0 = ops._render_backward(spec, raw0, rays0[2], self.target_s, td, (self.w_dev[0], self.w_dev[1], None, None),
                                              self.loss_out[4:], None, None)
                ops.merge_coarse(rank, d_raw0, d_raw, accumulate=True)
            ops.points_backward(spec.scene(), tb, w, mask, d_raw, rays=rays, accumulate=into)
            step = self._steps[0 if k == 0 else 1]
        d_o, d_d = self.d_rays[0], self.d_rays[1]
        check(lib.cb200_assemble_rays_bwd(ptr(self.kf7), ptr(self.ray_idx), ptr(self.pose_slot), ptr(d_o), ptr(d_d), ctypes.c_int64(n),
                                          ctypes.c_int64(1), ptr(self.d_c2w), stream()), 'assemble_rays_bwd')
        check(lib.cb200_track_update(ctypes.byref(step), stream()), 'track_update')

    def _frame_persistent(self):
        """The whole frame in one cooperative launch (csrc/track_frame.cu)."""
        lib, m, n = _lib.load(), self.model, self.n_rays
        spec = m.spec
        fw = m._fused()
        tb, w = fw[0].detach(), tuple(x.detach() for x in fw[1:])
        if self._wsp is None:
            nbytes = int(lib.cb200_track_frame_workspace(ctypes.c_int64(n), ctypes.c_int64(spec.n_samples)))
            self._wsp = torch.empty(nbytes + 256, dtype=torch.uint8, device=self.device)
            self._wsp_off = (-self._wsp.data_ptr()) % 256
            self._wsp_bytes = nbytes
        lin_u, lin_g, lin_n = spec.linspaces(self.device)
        mlp = ops._mlp_block(*w)
        check(lib.cb200_track_frame(ctypes.byref(spec.scene()), ptr(tb), ctypes.byref(mlp), ctypes.byref(spec.render_cfg()), ptr(lin_u),
                                    ptr(lin_g), ptr(lin_n), ctypes.c_uint64((int(m.noise_seed) << 32) & (2 ** 64 - 1)),
                                    ptr(m._noise_counter), ptr(self.kf7), ptr(self.c2w_init), ctypes.c_int64(n), ctypes.c_int32(self.iters),
                                    ctypes.byref(self._steps[0]), ctypes.c_void_p(self._wsp.data_ptr() + self._wsp_off),
                                    ctypes.c_size_t(self._wsp_bytes), stream()), 'track_frame')
        _lib.KERNEL_LAUNCHES += 1

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
        if self.fused:
            with torch.no_grad():
                self._reset_fused()
                for k in range(self.iters):
                    self._iteration_fused(k)
            return
        self._reset()
        for k in range(self.iters):
            self._iteration(k)

    # ---- public -----------------------------------------------------------------------------------------------------
    def track(self, dir_cam, target_s, target_d, c2w_init):
        """dir_cam [N,3] camera-frame ray directions of the sampled pixels, target_s [N,3], target_d [N,1] (any device),
        c2w_init [4,4].  Returns the estimated c2w [4,4] (best or last, `tracking.best`) and a dict of device scalars."""
        n = self.n_rays
        if self.fused:
            # no host arithmetic and no synchronisation: the initial matrix goes to the device as it is (cb200_track_init derives
            # the pose parameters there), the rays as ONE [N,7] upload (direction, rgb, depth rows -- what the kernels read)
            c2w_src = torch.as_tensor(c2w_init, dtype=torch.float32).detach().reshape(16)
            self.c2w_init.copy_(c2w_src, non_blocking=True)
            parts = [dir_cam.reshape(n, 3), target_s.reshape(n, 3), target_d.reshape(n, 1)]
            if all(p.device.type == 'cpu' for p in parts):
                if self._kf7_sent is not None:
                    self._kf7_sent.synchronize()                     # the previous frame's upload has left the staging buffer
                torch.cat(parts, -1, out=self._kf7_host)
                self.kf7.copy_(self._kf7_host, non_blocking=True)
                self._kf7_sent = torch.cuda.Event()
                self._kf7_sent.record()
            else:
                self.kf7.copy_(torch.cat([p.to(self.device, non_blocking=True).float() for p in parts], -1), non_blocking=True)
            if self.persistent:
                try:
                    with torch.no_grad():
                        self._frame_persistent()
                except RuntimeError as e:                            # a grid / sample count the kernel is not instantiated for
                    if 'no instantiation' not in str(e):
                        raise
                    self.persistent = False
            if not self.persistent:
                if not self.use_graph:
                    self._frame()
                else:
                    # the graph holds raw parameter addresses: model.to(), share_memory_(), load_state_dict(assign=True) or a
                    # re-pointed parameter would make a replay read stale memory, so a moved parameter means a new capture
                    if self._graph is None or self._graph_ptrs != self._param_ptrs():
                        self._capture()
                    self._graph.replay()
        else:
            c2w0 = torch.as_tensor(c2w_init, dtype=torch.float32).detach().cpu().reshape(1, 4, 4)
            rot0 = pose.matrix_to_quaternion(c2w0[:, :3, :3]) if self.quat else pose.matrix_to_axis_angle(c2w0[:, :3, :3])
            self.init_rot.copy_(rot0, non_blocking=True)
            self.init_trans.copy_(c2w0[:, :3, 3], non_blocking=True)
            self.dir_cam.copy_(dir_cam.reshape(n, 3), non_blocking=True)
            self.target_s.copy_(target_s.reshape(n, 3), non_blocking=True)
            self.target_d.copy_(target_d.reshape(n, 1), non_blocking=True)
            frozen = [p for p in self.model.parameters() if p.requires_grad]
            for p in frozen:                                         # tracking needs dL/d rays only (no table scatter, no dW)
                p.requires_grad_(False)
            try:
                if not self.use_graph:
                    self._frame()
                else:
                    if self._graph is None or self._graph_ptrs != self._param_ptrs():
                        self._capture()
                    self._graph.replay()
            finally:
                for p in frozen:
                    p.requires_grad_(True)
        est = (self.best_c2w if self.use_best else self.last_c2w)[0]
        return est, {'best_loss': self.best_loss.reshape(()), 'last_loss': self.last_loss.reshape(()), 'psnr': self.psnr,
                     'last_c2w': self.last_c2w[0], 'best_c2w': self.best_c2w[0]}

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
        self._graph_ptrs = self._param_ptrs()

    def _param_ptrs(self):
        """Addresses the captured launches have baked in: the model's parameters and the noise counter."""
        ptrs = tuple(p.data_ptr() for p in self.model.parameters())
        c = self.model._noise_counter
        return ptrs + ((c.data_ptr(),) if c is not None else ())
