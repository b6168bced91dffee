"""Global bundle adjustment without an autograd graph.

`CoSLAM.global_BA` (coslam.py:285-397) runs `mapping.iters` iterations of: sample rays from the keyframe database and the
current frame -> pose gather and ray rotation -> JointEncoding.forward -> weighted loss (+ smoothness) -> backward into the
hash table, the decoder and the keyframe poses -> optimiser steps every `map_accum_step` / `pose_accum_step` iterations.
Through torch autograd one such iteration is ~250 small kernels and is bound by the host (6.9 ms on a B200 at room0's 2048 +
rays); `BundleAdjuster.run` issues the same arithmetic as ~25 library launches per iteration with no host synchronisation:

    host (once per call, the reference's draws in the reference's order): `random.sample` ray indices of every iteration
        (model/keyframe.py:96-106, coslam.py:339-341), the CPU `torch.rand` of the stratified jitter (scene_rep.py:255, host
        noise mode only) and of the smoothness lattice (coslam.py:260-262) -> one upload
    per iteration: cb200_assemble_rays_fwd x2 | cb200_sample_rays | cb200_points_fwd (lazy above 2^16 points) |
        cb200_render_fwd | cb200_loss_finish | cb200_render_bwd | cb200_points_bwd (table, decoder, ray gradients) |
        cb200_assemble_rays_bwd x2 | smoothness: cb200_smooth_points, cb200_hash_fwd, cb200_smooth_tv, cb200_hash_bwd |
        when due: cb200_adam_step per map parameter (coslam_b200.optim.Adam), cb200_pose_adam

Gradients accumulate in the parameters' `.grad` buffers (the map optimiser is the caller's, as in coslam.py:610-624) and in
one [K,4,4] pose-gradient buffer: between two pose steps the reference back-propagates through the same pose matrices
(`retain_graph=True`), so chaining the summed dL/d c2w through the pose parametrisation once is the same gradient.
coslam.py stays unchanged by design; this class is the opt-in fast path (tools/slam_synth.py uses it).
"""
import ctypes
import random

import torch
import torch.distributed as dist

from . import _lib, ops, pose
from .dist import shard_bounds
from ._lib import check, ptr, stream
from .keyframe import sample_ranges


class _StagedDraws(object):
    """The `random.sample` index draws of a BA call (model/keyframe.py:100, coslam.py:339-341), made iteration by iteration in
    front of each iteration's launches instead of all at once in front of the first: the GPU starts after one iteration's draws
    (~0.05 ms) instead of the call's (~0.9 ms for 10 iterations), and the host stays ahead of it (an iteration's draws + launches
    take less host time than its kernels take).  Same values, same order, same state of Python's generator afterwards as
    keyframe.sample_ranges on the whole list: the Mersenne-Twister state lives in one numpy array for the duration of the call.
    Layout per iteration, identical in the pinned block and on the device: [ray indices (bs) | current-frame pixels (n_cur) |
    keyframe slots (bs)], one upload per iteration."""

    def __init__(self, adj, iters, bs, n_cur):
        import numpy as np
        kf = adj.kfdb
        n_store, n_pix = len(kf) * kf.num_rays_to_save, kf.H * kf.W
        if not (0 <= bs <= n_store and 0 <= n_cur <= n_pix):
            raise ValueError('Sample larger than population or is negative')
        self.bs, self.n_cur, self.T = bs, n_cur, 2 * bs + n_cur
        # two pinned blocks used in turn: a call only waits for the uploads of the call before the previous one, so back-to-back
        # calls do not make the host wait for the GPU to reach the previous call's last iteration
        key = (iters, bs, n_cur)
        st = adj._pin_staged
        if st is None or st['key'] != key:
            st = adj._pin_staged = {'key': key, 'bufs': [torch.empty(iters * self.T, dtype=torch.int64).pin_memory() for _ in range(2)],
                                    'sent': [None, None], 'turn': 0}
        self.slot = st['turn']
        st['turn'] ^= 1
        if st['sent'][self.slot] is not None:
            st['sent'][self.slot].synchronize()
        self.adj = adj
        pin = st['bufs'][self.slot]
        self.pin_rows = pin.view(iters, self.T)
        self.pin_np = pin.numpy().reshape(iters, self.T)
        self.pin_base = pin.data_ptr()
        self.smooth = None                                               # (jitter, offset ctypes arrays, offset_max, margin) when staged too
        self.dev = torch.empty(iters * self.T, dtype=torch.int64, device=adj.device)
        self.dev_rows = self.dev.view(iters, self.T)
        self.version, mt, self.gauss = random.getstate()
        self.state = np.array(mt, dtype=np.uint32)
        self.ns = np.array([n_store, n_pix], dtype=np.int64)
        self.ks = np.array([bs, n_cur], dtype=np.int64)
        self.args = (ctypes.c_void_p(self.state.ctypes.data), ctypes.c_int64(2), ctypes.c_void_p(self.ns.ctypes.data),
                     ctypes.c_void_p(self.ks.ctypes.data))
        self.fn = _lib.load().cb200_host_random_sample
        self.frame_ids = kf.frame_ids.numpy()
        self.R, self.every = kf.num_rays_to_save, adj.config['mapping']['keyframe_every']

    def stage(self, it, plan):
        """draws of iteration `it` -> pinned row -> device row; points the plan's index arrays at that row"""
        row = self.pin_np[it]
        if self.fn(*self.args, ctypes.c_void_p(self.pin_base + 8 * it * self.T)) != 0:
            raise RuntimeError('cb200_host_random_sample failed')
        bs, n_cur, T = self.bs, self.n_cur, self.T
        row[bs + n_cur:] = self.frame_ids[row[:bs] // self.R] // self.every                   # keyframe.py:103, coslam.py:345
        self.dev_rows[it].copy_(self.pin_rows[it], non_blocking=True)
        base = self.dev.data_ptr() + 8 * it * T
        # the runtime addresses iteration `it` as array + it * length: bias the three bases accordingly
        plan.idx_global = base - 8 * it * bs
        plan.idx_cur = base + 8 * bs - 8 * it * n_cur
        plan.slot_global = base + 8 * (bs + n_cur) - 8 * it * bs
        if self.smooth is not None:                                     # coslam.py:260, :262 -- the call's only draws from torch's CPU generator
            jit, off, offset_max, margin = self.smooth
            o = (torch.rand(3).to(offset_max) * offset_max + margin).tolist()
            j = torch.rand((1, 1, 1, 3)).reshape(3).tolist()
            off[3 * it], off[3 * it + 1], off[3 * it + 2] = o
            jit[3 * it], jit[3 * it + 1], jit[3 * it + 2] = j

    def finish(self):
        random.setstate((self.version, tuple(self.state.tolist()), self.gauss))
        ev = torch.cuda.Event()
        ev.record()
        self.adj._pin_staged['sent'][self.slot] = ev


class BundleAdjuster(object):
    def __init__(self, model, config, kfdb, map_optimizer, device='cuda:0', group=None, device_sampler=False):
        if not config['grid']['oneGrid'] or config['decoder']['tcnn_network']:
            raise NotImplementedError('BundleAdjuster serves oneGrid: True, tcnn_network: False (every shipped config); '
                                      'use the autograd path otherwise')
        self.n_imp = int(config['training']['n_importance'])
        self.model, self.config, self.kfdb, self.opt, self.device = model, config, kfdb, map_optimizer, torch.device(device)
        t = config['training']
        self.loss_w = [float(t[k]) for k in ('rgb_weight', 'depth_weight', 'sdf_weight', 'fs_weight')]
        self.w_dev = torch.tensor(self.loss_w, dtype=torch.float32, device=self.device).unbind(0)
        self.n_rot = 3 if t['rot_rep'] == 'axis_angle' else 4
        self.smooth_w = float(t['smooth_weight'])
        self.last_losses = None                   # device [4] of the last iteration (rgb, depth, sdf, fs)
        self.last_tv = None                       # device double [1]: raw TV sum of the last iteration
        self.last_d_poses = None                  # [K+1,4,4] pose gradient accumulated since the last pose step
        self.last_kept_width = 0                  # floats per point the last call's forward kept for its backward (0, 32 or 128)
        self._kept_buf = None                     # those rows (large batches only), reused across calls
        # Opt-in here: the rows pay for themselves when the forward evaluates little more than the points that end up with a gradient
        # (a depth-consistent batch: +3 % on the autograd step); over a store whose depths the map does not explain yet the lazy forward
        # evaluates every sample, writes rows for all of them and the call comes out even (tools/ba_sweep.py, profiles/r2/NOTES.md 13)
        self.keep_rows = False
        self._noise_calls = 0
        # non-parity index draws on the device (uniform, with replacement) for the large sweeps of BASELINE.json config 4:
        # `random.sample` of 2^22 out of 16 M rays is host-bound even in C
        self.device_sampler = bool(device_sampler)
        # rays sharded over the ranks of `group`: parameters, poses and the keyframe store are replicated,
        # every rank works on a contiguous shard of each iteration's globally drawn ray lists (rank 0's draws, broadcast)
        self.group = group
        self.world = dist.get_world_size(group) if group is not None else 1
        self.rank = dist.get_rank(group) if group is not None else 0
        self._flat = None
        self._pin, self._pin_sent = None, None     # pinned staging block of a call's index draws, event of its upload
        self.staged_draws = True                   # one GPU, host sampler: draw iteration by iteration in front of each launch (_StagedDraws)
        self._pin_staged = None
        self.phase_times = None                   # set to [] to collect CUDA events per sharded iteration: forward | AR-1 | backward | AR-2 | optimise
        bb = model.bounding_box.detach().float().cpu()
        self._lo = (ctypes.c_float * 3)(*bb[:, 0].tolist())
        self._ext = (ctypes.c_float * 3)(*(bb[:, 1] - bb[:, 0]).tolist())
        self._bb = bb

    # ---- host-side draws of one BA call, in the reference's order ------------------------------------------------------
    def _draw(self, iters, bs, n_cur, n_rays, n_samples, smooth_on=True, staged=False):
        kf, every = self.kfdb, self.config['mapping']['keyframe_every']
        t = self.config['training']
        host_noise = self.model.spec.perturb and self.model.noise_mode == 'host'
        noise, smooth = [], []
        if bs == 0:
            # first-frame mapping (coslam.py:183-189): select_samples draws i in range(H W) and reads pixel (i % H, i // H)
            draws = sample_ranges([(kf.H * kf.W, n_cur)] * iters)
            idx_c = [torch.from_numpy((d % kf.H) * kf.W + d // kf.H) for d in draws]
            if host_noise:
                noise = [ops.torch_rand_into(torch.empty((n_rays, n_samples))) for _ in range(iters)]
            dev = self.device
            empty = torch.zeros((iters, 0), dtype=torch.int64, device=dev)
            return empty, empty, torch.stack(idx_c).to(dev, non_blocking=True), \
                (torch.stack(noise).to(dev, non_blocking=True) if host_noise else None), []
        # Python-`random` and torch's CPU generator are independent streams: the index draws of all iterations go first, in the
        # reference's order (keyframe.py:100 then coslam.py:339-341 per iteration), as ONE call of the host sampler
        n_store = len(kf) * kf.num_rays_to_save
        if self.device_sampler:
            dev = self.device
            idx_g = torch.randint(0, n_store, (iters, bs), device=dev)
            idx_c = torch.randint(0, kf.H * kf.W, (iters, n_cur), device=dev)
            slot_g = (kf.frame_ids.to(dev)[idx_g // kf.num_rays_to_save] // every).to(torch.int64)
            for _ in range(iters):
                if host_noise:
                    noise.append(ops.torch_rand_into(torch.empty((n_rays, n_samples))))
                if self.smooth_w > 0 and smooth_on:
                    vol = self._bb[:, 1] - self._bb[:, 0]
                    offset_max = vol - (t['smooth_pts'] - 1) * t['smooth_vox'] - 2 * t['smooth_margin']
                    smooth.append((torch.rand(3).to(offset_max) * offset_max + t['smooth_margin'], torch.rand((1, 1, 1, 3)).reshape(3)))
            return idx_g, slot_g, idx_c, (torch.stack(noise).to(dev, non_blocking=True) if host_noise else None), smooth
        if staged:
            idx_g = slot_g = idx_c = None            # made per iteration by _StagedDraws (the torch-generator draws below are a separate stream)
            if not host_noise:
                return idx_g, slot_g, idx_c, None, None      # smoothness draws per iteration as well (no other user of that generator)
            for _ in range(iters):                           # host jitter and smoothness draws alternate on torch's CPU generator
                noise.append(ops.torch_rand_into(torch.empty((n_rays, n_samples))))                         # scene_rep.py:255
                if self.smooth_w > 0 and smooth_on:
                    vol = self._bb[:, 1] - self._bb[:, 0]
                    offset_max = vol - (t['smooth_pts'] - 1) * t['smooth_vox'] - 2 * t['smooth_margin']
                    smooth.append((torch.rand(3).to(offset_max) * offset_max + t['smooth_margin'], torch.rand((1, 1, 1, 3)).reshape(3)))
            return idx_g, slot_g, idx_c, torch.stack(noise).to(self.device, non_blocking=True), smooth
        draws = sample_ranges([(n_store, bs), (kf.H * kf.W, n_cur)] * iters)
        # one pinned block [ray indices [iters, bs] | keyframe slots [iters, bs] | current-frame pixels [iters, n_cur]] filled with numpy
        # over all iterations at once and sent in ONE copy (was ~60 small torch ops and three pageable uploads per call)
        key = (iters, bs, n_cur)
        if self._pin is None or self._pin[0] != key:
            self._pin = (key, torch.empty(iters * (2 * bs + n_cur), dtype=torch.int64).pin_memory())
        elif self._pin_sent is not None:
            self._pin_sent.synchronize()                             # the previous call's upload has left the staging block
        pin = self._pin[1].numpy()
        pg, ps, pc = pin[:iters * bs].reshape(iters, bs), pin[iters * bs:2 * iters * bs].reshape(iters, bs), pin[2 * iters * bs:].reshape(iters, n_cur)
        for it_ in range(iters):
            pg[it_] = draws[2 * it_]
            pc[it_] = draws[2 * it_ + 1]
        ps[:] = kf.frame_ids.numpy()[pg // kf.num_rays_to_save] // every                                 # keyframe.py:103, coslam.py:345
        all_dev = self._pin[1].to(self.device, non_blocking=True)
        self._pin_sent = torch.cuda.Event()
        self._pin_sent.record()
        idx_g, slot_g = all_dev[:iters * bs].view(iters, bs), all_dev[iters * bs:2 * iters * bs].view(iters, bs)
        idx_c = all_dev[2 * iters * bs:].view(iters, n_cur)
        for _ in range(iters):
            if host_noise:
                noise.append(ops.torch_rand_into(torch.empty((n_rays, n_samples))))                     # scene_rep.py:255
            if self.smooth_w > 0 and smooth_on:
                vol = self._bb[:, 1] - self._bb[:, 0]
                n = t['smooth_pts']
                offset_max = vol - (n - 1) * t['smooth_vox'] - 2 * t['smooth_margin']
                offset = torch.rand(3).to(offset_max) * offset_max + t['smooth_margin']              # coslam.py:260
                smooth.append((offset, torch.rand((1, 1, 1, 3)).reshape(3)))                         # coslam.py:262
        dev = self.device
        return idx_g, slot_g, idx_c, (torch.stack(noise).to(dev, non_blocking=True) if host_noise else None), smooth

    def _map_grads(self):
        """(table.grad, [dW1..dW4]) as zero-initialised persistent buffers of the caller's parameters."""
        fw = self.model._fused()
        if self.world > 1:                        # one flat bucket so that AR-2 is a single all-reduce; the .grad tensors are views
            n = sum(p.numel() for p in fw)
            if self._flat is None or self._flat.numel() != n or any(p.grad is None or p.grad.data_ptr() != self._flat.data_ptr() + 4 * o
                                                                    for p, o in zip(fw, self._offsets(fw))):
                self._flat = torch.zeros(n, dtype=torch.float32, device=fw[0].device)
                for p, o in zip(fw, self._offsets(fw)):
                    p.grad = self._flat[o:o + p.numel()].view_as(p)
        for p in fw:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
        return fw[0].grad, [p.grad for p in fw[1:]]

    @staticmethod
    def _offsets(params):
        out, o = [], 0
        for p in params:
            out.append(o)
            o += p.numel()
        return out

    def _adam_params(self):
        """The map optimiser's parameters as a cb200_adam_param array (state created like coslam_b200.optim.Adam.step does);
        every parameter must share one step count, as they do when only this class and Adam.step drive the optimiser."""
        items, states = [], []
        for group in self.opt.param_groups:
            b1, b2 = group['betas']
            for p in group['params']:
                if p.grad is None:
                    continue
                st = self.opt.state[p]
                if len(st) == 0:
                    st['step'] = torch.zeros((), dtype=torch.float32)
                    st['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                a = _lib.AdamParam(p.data_ptr(), p.grad.data_ptr(), st['exp_avg'].data_ptr(), st['exp_avg_sq'].data_ptr(), p.numel(),
                                   group['lr'], b1, b2, group['eps'], group['weight_decay'])
                items.append(a)
                states.append(st)
        if len({int(s['step']) for s in states}) > 1:
            raise RuntimeError('BundleAdjuster: the map parameters have different optimiser step counts')
        return (_lib.AdamParam * len(items))(*items), states

    def _zero_map_grads(self):
        for p in self.model._fused():
            if p.grad is not None:
                p.grad.zero_()

    # ---- coslam.py:285-397 -----------------------------------------------------------------------------------------------
    @torch.no_grad()
    def first_frame(self, c2w, current_rays, n_iters):
        """first_frame_mapping (coslam.py:162-207): `n_iters` map-only iterations on `mapping.sample` pixels of the first frame
        under its fixed pose (no smoothness term, an optimiser step every iteration).  current_rays [H*W,7] on the device."""
        empty = torch.zeros((0, 4, 4), dtype=torch.float32, device=self.device)
        self.run(empty, c2w, current_rays, _first_iters=int(n_iters))

    @torch.no_grad()
    def run(self, kf_poses, cur_pose, current_rays, _first_iters=None):
        """kf_poses [K,4,4]: estimated poses of the keyframes 0, every, 2 every, ... (coslam.py:299-300); cur_pose [4,4]: the
        current frame's estimate; current_rays [H*W,7] on the device (direction, rgb, depth; coslam.py:333-334).
        Returns (kf_poses_new [K,4,4], cur_pose_new [4,4]) -- unchanged tensors where the reference leaves them unchanged."""
        lib, model, dev = _lib.load(), self.model, self.device
        m = dict(self.config['mapping'])
        spec = model.spec
        kf = self.kfdb
        first = _first_iters is not None
        if first:
            m.update({'iters': _first_iters, 'map_accum_step': 1, 'map_wait_step': 0})
            iters, bs, n_cur = _first_iters, 0, int(m['sample'])
        else:
            iters, bs = int(m['iters']), int(m['sample'])
            n_cur = max(bs // len(kf), int(m['min_pixels_cur']))
        smooth_w = 0.0 if first else self.smooth_w
        n, S = bs + n_cur, spec.n_samples
        K = kf_poses.shape[0]
        nr = self.n_rot
        f32 = dict(dtype=torch.float32, device=dev)
        # ---- pose parameters (coslam.py:305-323) -------------------------------------------------------------------------
        optimise = (not first) and len(kf) >= 2
        optim_cur = bool(m['optim_cur'])
        poses_all = torch.cat([kf_poses.to(dev).float(), cur_pose.to(dev).float()[None]]).contiguous()     # slot -1 = current frame
        n_free = 0
        if optimise:
            free = poses_all[1:] if optim_cur else poses_all[1:K]
            n_free = free.shape[0]
        if n_free > 0:
            # matrix_to_tensor (coslam.py:279) on the device: no host round trip of the poses, no synchronisation
            rot, trans = torch.empty((n_free, nr), **f32), torch.empty((n_free, 3), **f32)
            check(lib.cb200_pose_params(ptr(free.contiguous()), ctypes.c_int64(n_free), nr, ptr(rot), ptr(trans), stream()), 'pose_params')
            state = torch.zeros(2 * n_free * (nr + 3), **f32)
            m_rot, v_rot = state[:n_free * nr], state[n_free * nr:2 * n_free * nr]
            m_tr, v_tr = state[2 * n_free * nr:2 * n_free * nr + 3 * n_free], state[2 * n_free * nr + 3 * n_free:]
            free_c2w = poses_all[1:1 + n_free]                                           # view: the kernels write the matrices in place
            check(lib.cb200_pose_matrices(ptr(rot), ptr(trans), ctypes.c_int64(n_free), nr, ptr(free_c2w), stream()), 'pose_matrices')
        d_poses = torch.zeros((K + 1, 4, 4), **f32)
        pose_step = 0
        # ---- buffers and the plan of the native iteration runtime (csrc/runtime.cu) ---------------------------------------
        staged = self.staged_draws and self.world == 1 and not self.device_sampler and bs > 0 and self.phase_times is None
        idx_g, slot_g, idx_c, noise, smooth = self._draw(iters, bs, n_cur, n, S, smooth_on=not first, staged=staged)
        stager = _StagedDraws(self, iters, bs, n_cur) if staged else None
        n_total = n
        if self.world > 1:
            for t_ in (idx_g, slot_g, idx_c) + ((noise,) if noise is not None else ()):
                if t_.numel() > 0:
                    dist.broadcast(t_, src=dist.get_global_rank(self.group, 0), group=self.group)
            if smooth:
                sm = torch.stack([torch.cat(p_) for p_ in smooth]).to(dev)
                dist.broadcast(sm, src=dist.get_global_rank(self.group, 0), group=self.group)
                smooth = [(r_[:3], r_[3:]) for r_ in sm.cpu()]
            (g0, g1), (c0, c1) = shard_bounds(bs, self.world, self.rank), shard_bounds(n_cur, self.world, self.rank)
            if noise is not None:
                noise = torch.cat([noise[:, g0:g1], noise[:, bs + c0:bs + c1]], 1).contiguous()
            idx_g, slot_g, idx_c = idx_g[:, g0:g1].contiguous(), slot_g[:, g0:g1].contiguous(), idx_c[:, c0:c1].contiguous()
            bs, n_cur = g1 - g0, c1 - c0
            n = bs + n_cur
            smooth_w = smooth_w / self.world       # every rank adds the full regulariser gradient before AR-2 sums them
        slot_c = torch.full((n_cur,), -1, dtype=torch.int64, device=dev)
        store = kf.rays[:len(kf)].reshape(-1, 7) if not first else kf.rays[:1].reshape(-1, 7)
        cur = current_rays.reshape(-1, 7)
        if cur.device != dev or cur.dtype != torch.float32 or not cur.is_contiguous():
            cur = cur.to(dev, torch.float32).contiguous()
        rays_o, rays_d, t_rgb = (torch.empty((n, 3), **f32) for _ in range(3))
        t_d = torch.empty((n,), **f32)
        I = self.n_imp
        Sf = S + I                                                       # samples per ray of the (fine) pass the losses are taken on
        z = torch.empty((n, Sf), **f32)
        raw, d_raw = torch.empty((n, Sf, 4), **f32), torch.empty((n, Sf, 4), **f32)
        mask = torch.empty((n, Sf, 2), dtype=torch.int32, device=dev)
        d_rays = torch.empty((2, n, 3), **f32) if n_free > 0 else None
        ws = torch.empty(10, dtype=torch.float64, device=dev)           # 4 counts | 4 loss sums | tv sum | pad
        loss_out = torch.empty(8, **f32)
        fw = model._fused()
        self._zero_map_grads()                                           # map_optimizer.zero_grad(), coslam.py:326
        g_table, g_w = self._map_grads()
        scene, render = spec.scene(), spec.render_cfg()
        lin_u, lin_g, lin_n = spec.linspaces(dev)
        lazy = ops.LAZY_FORWARD and lib.cb200_get_decoder() == 1 and ops.LAZY_MIN_POINTS <= n * Sf < 2 ** 32
        keep = [idx_g, slot_g, idx_c, slot_c, noise, cur, stager]       # referenced by raw pointer below
        P = _lib.BaPlan()
        dp = lambda t: t.data_ptr() if t is not None else None
        P.scene, P.render = ctypes.addressof(scene), ctypes.addressof(render)
        P.table = dp(fw[0])
        P.mlp = ops._mlp_block(*fw[1:])
        P.lin_uniform, P.lin_guided, P.lin_nodepth = dp(lin_u), dp(lin_g), dp(lin_n)
        P.kf_rays, P.idx_global, P.slot_global, P.n_global = dp(store), dp(idx_g), dp(slot_g), bs
        P.cur_rays, P.idx_cur, P.slot_cur, P.n_cur = dp(cur), dp(idx_c), dp(slot_c), n_cur
        P.poses_all, P.n_poses, P.n_samples, P.n_rays_total = dp(poses_all), K + 1, S, n_total
        seeds = (ctypes.c_uint64 * iters)()
        if noise is not None:
            P.noise = dp(noise)
        elif spec.perturb:
            for it in range(iters):                                      # device Philox: one fresh seed per iteration
                seeds[it] = (model._noise(n, S, dev)[1] + (self.rank << 24)) & (2 ** 64 - 1)
                if model._noise_counter is not None:                     # graph-tracker mode: `_noise` hands out one constant seed and
                    self._noise_calls += 1                               # the device counter only moves with the tracker's replays
                    seeds[it] = (seeds[it] + (1 << 31) + self._noise_calls) & (2 ** 64 - 1)
            if model._noise_counter is not None:
                P.seed_dev = dp(model._noise_counter)
        P.seeds = ctypes.cast(seeds, ctypes.c_void_p).value
        P.rays_o, P.rays_d, P.target_rgb, P.target_d = dp(rays_o), dp(rays_d), dp(t_rgb), dp(t_d)
        P.z_vals, P.raw, P.act_mask, P.d_raw = dp(z), dp(raw), dp(mask), dp(d_raw)
        if lazy:
            nbytes = int(lib.cb200_lazy_workspace(ctypes.c_int64(n), ctypes.c_int64(Sf)))
            lazy_ws = torch.empty((nbytes + 7) // 8, dtype=torch.int64, device=dev)
            P.lazy_ws, P.lazy_ws_bytes = dp(lazy_ws), nbytes
        if lazy and I == 0 and self.keep_rows and ops.KEEP_FEATURES and lib.cb200_get_decoder() == 1:
            # large batches: the forward keeps rows for the backward (features, + their position derivatives when poses are optimised)
            width = 4 * ops.N_GRID_FEATS if d_rays is not None else ops.N_GRID_FEATS
            if width == ops.N_GRID_FEATS or ops.KEEP_DERIVATIVES:
                # one buffer per adjuster, grown when a call needs more: allocated per call next to the other scratch, a block of this
                # size does not find its way back through the caching allocator (94 GB at 2^22 rays: a cudaMalloc + cudaFree per call)
                need = n * Sf * width
                if self._kept_buf is None or self._kept_buf.numel() < need:
                    self._kept_buf = None
                    if ops._kept_rows_fit(need * 4, dev):
                        try:
                            self._kept_buf = torch.empty(need, **f32)
                        except torch.cuda.OutOfMemoryError:
                            self._kept_buf = None
                if self._kept_buf is not None:
                    P.kept, P.kept_derivatives = dp(self._kept_buf), int(width != ops.N_GRID_FEATS)
        bbytes = int(lib.cb200_points_bwd_workspace(ctypes.c_int64(n * Sf)))
        bwd_ws = torch.empty((bbytes + 3) // 4, dtype=torch.int32, device=dev)
        P.bwd_ws, P.bwd_ws_bytes = dp(bwd_ws), bbytes
        if I > 0:                                                        # hierarchical pass, model/scene_rep.py:262-277
            from .hierarchical import draw_uniforms
            det = self.config['training']['perturb'] == 0.
            u = draw_uniforms(n, I, dev, det) if det else torch.stack([draw_uniforms(n, I, dev, det) for _ in range(iters)])
            z0, w0 = torch.empty((n, S), **f32), torch.empty((n, S), **f32)
            raw0, d_raw0 = torch.empty((n, S, 4), **f32), torch.empty((n, S, 4), **f32)
            mask0 = torch.empty((n, S, 2), dtype=torch.int32, device=dev)
            sums0 = torch.empty(4, dtype=torch.float64, device=dev)
            P.n_importance, P.z0, P.raw0, P.act_mask0, P.weights0, P.d_raw0 = I, dp(z0), dp(raw0), dp(mask0), dp(w0), dp(d_raw0)
            P.u, P.u_stride, P.sums0 = dp(u), (0 if det else I), dp(sums0)
            keep += [u, z0, w0, raw0, d_raw0, mask0, sums0]
            if lib.cb200_get_decoder() == 1 and n * Sf < 2 ** 32:
                # the fine pass evaluates the I new samples per ray only (the coarse ones are taken over) and one backward serves both passes
                rank = torch.empty((n, S), dtype=torch.int32, device=dev)
                new_list = torch.empty((n * I,), dtype=torch.int32, device=dev)
                new_count = torch.empty((1,), dtype=torch.int64, device=dev)
                P.rank, P.new_list, P.new_count = dp(rank), dp(new_list), dp(new_count)
                keep += [rank, new_list, new_count]
        P.ws, P.loss_out = dp(ws), dp(loss_out)
        P.loss_w.rgb, P.loss_w.depth, P.loss_w.sdf, P.loss_w.fs = (t_.data_ptr() for t_ in self.w_dev)
        P.g_table = dp(g_table)
        P.g_mlp.w_sdf0, P.g_mlp.w_sdf1, P.g_mlp.w_col0, P.g_mlp.w_col1 = (g.data_ptr() for g in g_w)
        P.d_rays, P.d_poses = dp(d_rays), dp(d_poses)
        if smooth_w > 0:                                                 # coslam.py:154-158, :251-274
            t = self.config['training']
            n_side = int(t['smooth_pts']) - 1
            grid = model.embed_fn.grid()
            n_feat = model.embed_fn.n_output_dims
            sx = torch.empty((n_side ** 3, 3), **f32)
            sf = torch.empty((n_side ** 3, n_feat), **f32)
            sdf_ = torch.empty_like(sf)
            if smooth is None:                                           # staged: filled iteration by iteration
                jit, off = (ctypes.c_float * (3 * iters))(), (ctypes.c_float * (3 * iters))()
                vol = self._bb[:, 1] - self._bb[:, 0]
                stager.smooth = (jit, off, vol - (t['smooth_pts'] - 1) * t['smooth_vox'] - 2 * t['smooth_margin'], t['smooth_margin'])
            else:
                jit = (ctypes.c_float * (3 * iters))(*[v for _, j in smooth for v in j.tolist()])
                off = (ctypes.c_float * (3 * iters))(*[v for o, _ in smooth for v in o.tolist()])
            P.grid = ctypes.addressof(grid)
            P.smooth_side, P.smooth_feat, P.smooth_voxel = n_side, n_feat, float(t['smooth_vox'])
            P.smooth_grad_scale = smooth_w / float(int(t['smooth_pts']) ** 3)
            P.smooth_jitter, P.smooth_offset = ctypes.cast(jit, ctypes.c_void_p).value, ctypes.cast(off, ctypes.c_void_p).value
            P.smooth_x, P.smooth_f, P.smooth_df = dp(sx), dp(sf), dp(sdf_)
            keep += [sx, sf, sdf_, jit, off, grid]
        adam, states = self._adam_params()
        P.map_params, P.n_map_params = ctypes.cast(adam, ctypes.c_void_p).value, len(adam)
        if n_free > 0:
            P.rot, P.trans, P.pose_adam, P.n_free, P.n_rot = dp(rot), dp(trans), dp(state), n_free, nr
            P.lr_rot, P.lr_trans = float(m['lr_rot']), float(m['lr_trans'])

        st = stream()
        for it in range(iters):
            map_step, zero, p_step = 0, 0, 0
            if (it + 1) % int(m['map_accum_step']) == 0:                 # coslam.py:363-370
                zero = 1
                if (it + 1) > int(m['map_wait_step']):
                    for s_ in states:
                        s_['step'] += 1
                    map_step = int(states[0]['step'])
            if n_free > 0 and (it + 1) % int(m['pose_accum_step']) == 0:   # coslam.py:372-390
                pose_step += 1
                p_step = pose_step
            if stager is not None:
                stager.stage(it, P)
            if self.world == 1 and self.phase_times is None:
                check(lib.cb200_ba_iteration(ctypes.byref(P), ctypes.c_int64(it), 7, ctypes.c_int64(map_step), zero, ctypes.c_int64(p_step), st),
                      'ba_iteration')
                continue
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)] if self.phase_times is not None else None
            mark = (lambda i: ev[i].record()) if ev else (lambda i: None)
            call = lambda phases, ms=0, z_=0, ps=0: check(lib.cb200_ba_iteration(ctypes.byref(P), ctypes.c_int64(it), phases, ctypes.c_int64(ms), z_,
                                                                                 ctypes.c_int64(ps), st), 'ba_iteration')
            shard = self.world > 1
            mark(0)
            if I == 0:
                # the mask counts are complete after the sampling and are all the backward needs of the other ranks: their all-reduce
                # (AR-1) runs behind the per-point forward; the loss sums only feed the reported losses and are reduced behind the backward
                call(8)                                                        # CB200_BA_RAYS
                h1 = dist.all_reduce(ws[:4].view(torch.int64), group=self.group, async_op=True) if shard else None
                call(1 | 16)                                                   # FORWARD | AFTER_RAYS
                mark(1)
                if h1 is not None:
                    h1.wait()
                mark(2)
                call(2)
                h2 = dist.all_reduce(ws[4:8], group=self.group, async_op=True) if shard else None
            else:
                call(1)
                mark(1)
                if shard:
                    dist.all_reduce(ws[:4].view(torch.int64), group=self.group)    # AR-1: batch-global mask counts ...
                    dist.all_reduce(ws[4:8], group=self.group)                     #       ... and loss sums
                mark(2)
                call(2)
                h2 = None
            mark(3)
            if map_step > 0 and shard:                                         # AR-2, only in front of a step that consumes the sums
                dist.all_reduce(self._flat, group=self.group)
            if p_step > 0 and shard:
                dist.all_reduce(d_poses, group=self.group)
            if h2 is not None:
                h2.wait()
            mark(4)
            call(4, map_step, zero, p_step)
            mark(5)
            if ev:
                self.phase_times.append(ev)
        if stager is not None:
            stager.finish()
        _lib.KERNEL_LAUNCHES += iters * (17 + (3 if lazy else 0) + (4 if smooth_w > 0 else 0)) - iters
        del keep
        self.last_losses, self.last_tv, self.last_d_poses = loss_out[:4], ws[8:9], d_poses
        self.last_kept_width = (4 * ops.N_GRID_FEATS if P.kept_derivatives else ops.N_GRID_FEATS) if P.kept else 0
        # ---- coslam.py:392-397 --------------------------------------------------------------------------------------------
        if n_free == 0:
            return kf_poses, cur_pose
        check(lib.cb200_pose_matrices(ptr(rot), ptr(trans), ctypes.c_int64(n_free), nr, ptr(free_c2w), stream()), 'pose_matrices')
        new_kf = poses_all[:K].clone()
        new_cur = poses_all[K].clone() if optim_cur else cur_pose
        return new_kf, new_cur
