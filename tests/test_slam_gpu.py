"""Trajectory parity (north_star: "tracked ATE over a sequence must stay within 0.2 cm of the reference"): the SLAM loop of
tools/slam_synth.py (coslam.py:640-700 restated over the coslam_b200 modules) runs twice on the same synthetic RGB-D frames
with the same seeds -- once through the CUDA path (host-noise mode: it then consumes the same `random.sample` and CPU
`torch.rand` streams as the reference loop), once through the oracle (reference model code restated + tinycudann shim in
plain torch ops, torch.optim.Adam, keyframe store as in model/keyframe.py; executed by torch on the same device to keep the
test short -- none of our kernels on that arm) -- and the two estimated trajectories are compared.
"""
import os
import random
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tools'))

pytestmark = pytest.mark.gpu


def _oracle_slam_class():
    import slam_synth
    from coslam_b200 import pose
    from oracle import coslam_oracle

    class _Params(object):
        def __init__(self, ps):
            self.ps = ps

        def parameters(self):
            return self.ps

    class RefKeyframes(object):
        """model/keyframe.py:4-106 restated: torch fancy indexing, Python-`random` draws."""

        def __init__(self, cfg, H, W, n_kf, rays_per_kf, dev):
            self.H, self.W, self.R = H, W, rays_per_kf
            self.rays = torch.zeros((n_kf, rays_per_kf, 7), device=dev)
            self.frame_ids = None

        def __len__(self):
            return len(self.frame_ids)

        def add_keyframe(self, batch, filter_depth=False):
            rays = torch.cat([batch['direction'], batch['rgb'], batch['depth'][..., None]], dim=-1).reshape(1, -1, 7)
            idxs = random.sample(range(0, self.H * self.W), self.R)                       # keyframe.py:45-46
            fid = torch.tensor([batch['frame_id']])
            self.frame_ids = fid if self.frame_ids is None else torch.cat([self.frame_ids, fid])
            self.rays[len(self.frame_ids) - 1] = rays[:, idxs]

        def sample_global_rays(self, bs):                                                 # keyframe.py:96-106
            n = len(self.frame_ids)
            idxs = torch.tensor(random.sample(range(n * self.R), bs))
            return self.rays[:n].reshape(-1, 7)[idxs.to(self.rays.device)], self.frame_ids[idxs // self.R]

    class OracleSLAM(slam_synth.SynthSLAM):
        def make_model(self):
            m = coslam_oracle.OracleScene(self.cfg, self.bound.cpu()).to(self.dev)
            m.decoder = _Params([m.W1, m.W2, m.W3, m.W4])
            m.embed_fn = _Params([m.grid.params])
            return m

        def make_map_optimizer(self):                                                     # coslam.py:610-624
            mp = self.cfg['mapping']
            return torch.optim.Adam([{'params': self.model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                                     {'params': self.model.embed_fn.parameters(), 'eps': 1e-15, 'lr': mp['lr_embed']}],
                                    betas=(0.9, 0.99))

        def make_kfdb(self, n_kf, rays_per_kf):
            return RefKeyframes(self.cfg, self.H, self.W, n_kf, rays_per_kf, self.dev)

        def make_tracker(self):
            return None

        def make_bundle_adjuster(self):
            return None

        def sample_pixels(self, n, k):                                                    # coslam.py:131-138
            return torch.tensor(random.sample(range(n), k), device=self.dev)

        def track_pose(self, d_cam, rgb, depth, c2w_init):                                # coslam.py:526-573
            t = self.cfg['tracking']
            c2w0 = c2w_init[None]
            trans = torch.nn.Parameter(c2w0[:, :3, 3].clone())
            rot = torch.nn.Parameter(pose.matrix_to_axis_angle(c2w0[:, :3, :3].cpu()).to(self.dev))
            opt = torch.optim.Adam([{'params': rot, 'lr': t['lr_rot']}, {'params': trans, 'lr': t['lr_trans']}])
            best, best_c2w, thresh = None, None, 0
            for _ in range(t['iter']):
                opt.zero_grad()
                c2w = pose.pose_matrix(rot, trans)
                rays_o = c2w[..., :3, -1].repeat(t['sample'], 1)
                rays_d = torch.sum(d_cam[..., None, :] * c2w[:, :3, :3], -1)
                loss = self.loss(self.model.forward(rays_o, rays_d, rgb, depth))
                if best is None:
                    best, best_c2w = loss.item(), c2w.detach()
                if loss.item() < best:
                    best, best_c2w, thresh = loss.item(), c2w.detach(), 0
                else:
                    thresh += 1
                if thresh > t['wait_iters']:
                    break
                loss.backward()
                opt.step()
            return (best_c2w if t['best'] else c2w.detach())[0]

        def ba_batch(self, poses_all, current, n_cur):                                    # coslam.py:336-355
            mp = self.cfg['mapping']
            rays, ids = self.kfdb.sample_global_rays(mp['sample'])
            idx_cur = random.sample(range(0, self.H * self.W), n_cur)
            rays = torch.cat([rays, current[idx_cur]], 0)
            ids_all = torch.cat([ids // mp['keyframe_every'], -torch.ones(len(idx_cur))]).to(torch.int64).to(self.dev)
            rays_d_cam, rgb, depth = rays[..., :3], rays[..., 3:6], rays[..., 6:7]
            rays_d = torch.sum(rays_d_cam[..., None, None, :] * poses_all[ids_all, None, :3, :3], -1).reshape(-1, 3)
            rays_o = poses_all[ids_all, None, :3, -1].repeat(1, 1, 1).reshape(-1, 3)
            return rays_o, rays_d, rgb, depth

    return slam_synth, OracleSLAM


def test_trajectory_matches_the_oracle():
    slam_synth, OracleSLAM = _oracle_slam_class()
    from coslam_b200 import synth
    n_frames, first_iters, scale = 21, 100, 0.25
    dev = 'cuda:0'

    def config():
        return slam_synth.scaled_config(synth.room0_config(), scale)

    frames = slam_synth.make_frames(config(), n_frames, dev)
    gt = torch.stack([f['c2w'] for f in frames]).cpu()

    ours = slam_synth.SynthSLAM(config(), dev, n_frames, noise='host')
    oracle = OracleSLAM(config(), dev, n_frames)
    oracle.model.load_weights(*[p.detach() for p in ours.model._fused()])

    slam_synth.seed_everything(0)
    est_ours = ours.run([dict(f) for f in frames], first_iters).cpu()
    slam_synth.seed_everything(0)
    est_oracle = oracle.run([dict(f) for f in frames], first_iters).cpu()

    ate_ours, ate_oracle = slam_synth.ate_rmse(est_ours, gt), slam_synth.ate_rmse(est_oracle, gt)
    between = slam_synth.ate_rmse(est_ours, est_oracle)
    drift = slam_synth.ate_rmse(gt[:1].expand_as(gt), gt)
    print('ATE ours %.3f cm, oracle %.3f cm, between the two %.3f cm (camera moves %.1f cm RMSE from frame 0)'
          % (100 * ate_ours, 100 * ate_oracle, 100 * between, 100 * drift))
    assert ate_oracle < 0.25 * drift                 # the loop does track (a constant pose would score `drift`)
    assert abs(ate_ours - ate_oracle) < 0.002        # north_star: ATE within 0.2 cm of the reference
    # The loop itself is chaotic at the few-mm level: Adam's first steps move a pose component by lr * sign(g), so a
    # rounding-level change of a near-zero gradient moves it by 1e-3.  Two runs of the SAME CUDA path with the same seeds
    # (atomics order differs) end 0.4-0.6 cm apart, as far as either is from the oracle run (tests/traj_probe.py), while
    # the error against ground truth is stable to < 0.1 cm.  Hence a loose bound here and the tight one on the ATE.
    assert between < 0.01
