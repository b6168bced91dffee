"""BASELINE.json config 2 in miniature: tracking + mapping on synthetic RGB-D frames of an analytic scene, through the
coslam_b200 modules (JointEncoding, FrameTracker, KeyFrameDatabase + assemble_rays, optim.Adam) arranged like the loop of
coslam.py:640-700 (first-frame mapping, per-frame tracking, global BA every `map_every` frames, keyframe every
`keyframe_every` frames).  The frames are sphere-traced here with torch ops on the GPU (no graphics API): an axis-aligned
room matching room0's bound with three spheres, procedural colour, a smooth camera trajectory (<= 2 cm, <= 1 deg per frame).

    python tools/slam_synth.py [--frames 60] [--scale 0.5] [--first-iters 200]

Prints one JSON line: ATE RMSE of the estimated trajectory against the generator's poses, tracking ms/frame, mapping
ms/iteration, frames/s.  Not a parity test (the reference cannot run here: no tinycudann) -- a functional end-to-end check.
"""
import argparse
import json
import math
import os
import random
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                                   # noqa: E402

from coslam_b200 import optim as cb_optim                       # noqa: E402
from coslam_b200 import pose, synth                             # noqa: E402
from coslam_b200.keyframe import KeyFrameDatabase, assemble_rays   # noqa: E402
from coslam_b200.scene_rep import JointEncoding                 # noqa: E402
from coslam_b200.tracking import FrameTracker                   # noqa: E402


# ---- the synthetic scene -----------------------------------------------------------------------------------------------
ROOM_LO = torch.tensor([-0.6, -0.9, -1.4])
ROOM_HI = torch.tensor([6.6, 3.3, 1.1])
SPHERES = torch.tensor([[2.0, 1.2, -0.9, 0.5], [4.2, 0.3, -0.8, 0.6], [3.2, 2.4, -0.2, 0.45]])


SHIFT = torch.zeros(3)                        # the same room translated / scaled into another config's bound (see main)
SCALE = [1.0]


def scene_distance(p):
    """Distance to the nearest surface for points inside the room (walls from the inside, spheres from the outside)."""
    p = (p - SHIFT.to(p)) / SCALE[0]
    lo, hi, sph = ROOM_LO.to(p), ROOM_HI.to(p), SPHERES.to(p)
    d_wall = torch.minimum(p - lo, hi - p).min(-1)[0]
    d_sph = (torch.norm(p[..., None, :] - sph[:, :3], dim=-1) - sph[:, 3]).min(-1)[0]
    return torch.minimum(d_wall, d_sph) * SCALE[0]


def scene_colour(p):
    p = (p - SHIFT.to(p)) / SCALE[0]
    k = torch.tensor([1.7, 2.3, 2.9], device=p.device)
    return 0.5 + 0.5 * torch.stack([torch.sin(k[0] * p[..., 0] + 0.5 * p[..., 1]), torch.sin(k[1] * p[..., 1] + 0.7 * p[..., 2]),
                                    torch.sin(k[2] * p[..., 2] + 0.3 * p[..., 0])], -1) * 0.8


def render_frame(c2w, dirs_cam, steps=96):
    """Sphere tracing along the camera rays; depth = distance along -z of the camera (like a depth image), colour at the hit."""
    rays_d = torch.sum(dirs_cam[..., None, :] * c2w[:3, :3], -1)
    norm = torch.norm(rays_d, dim=-1, keepdim=True)
    unit = rays_d / norm
    o = c2w[:3, 3]
    t = torch.zeros(rays_d.shape[:-1], device=rays_d.device)
    for _ in range(steps):
        t = t + scene_distance(o + unit * t[..., None])
    hit = o + unit * t[..., None]
    depth = t / norm[..., 0]                     # ray parameter with the un-normalised direction (z component is -1)
    return scene_colour(hit), depth


def trajectory(n_frames):
    """Smooth arc inside the room looking toward its middle; c2w in the OpenGL convention (camera looks along -z)."""
    poses = []
    centre = torch.tensor([3.0, 1.2, -0.3])
    for i in range(n_frames):
        a = 0.012 * i
        eye = torch.tensor([1.2 + 0.9 * math.sin(a), 0.6 + 0.012 * i, 0.1 + 0.1 * math.sin(0.5 * a)])
        look = centre + torch.tensor([0.3 * math.sin(0.7 * a), 0.0, 0.0])
        fwd = (look - eye) / torch.norm(look - eye)
        up = torch.tensor([0.0, 0.0, 1.0])
        right = torch.linalg.cross(fwd, up)
        right = right / torch.norm(right)
        up2 = torch.linalg.cross(right, fwd)
        T = torch.eye(4)
        T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = right, up2, -fwd, eye * SCALE[0] + SHIFT
        poses.append(T)
    return torch.stack(poses)


# ---- the SLAM loop (coslam.py:640-700 and the methods it calls) --------------------------------------------------------
class SynthSLAM(object):
    """`noise`: 'device' (Philox in the sampling kernel, tracking as one CUDA graph) or 'host' (the reference's CPU torch.rand
    stream, eager tracking: consumes the same `random` / `torch.rand` draws as the reference loop for the same seed).
    The make_* hooks exist so that tests/test_slam_gpu.py can run the identical loop over the CPU oracle."""

    def __init__(self, cfg, dev, n_frames, noise='device'):
        self.cfg, self.dev, self.noise = cfg, dev, noise
        cam = cfg['cam']
        self.H, self.W = cam['H'], cam['W']
        self.dirs = synth.camera_ray_dirs(self.H, self.W, cam['fx'], cam['fy'], cam['cx'], cam['cy']).to(dev)   # [H,W,3]
        self.bound = synth.bound_tensor(cfg).to(dev)
        m, t = cfg['mapping'], cfg['training']
        self.model = self.make_model()
        self.map_optimizer = self.make_map_optimizer()
        self.loss_w = (t['rgb_weight'], t['depth_weight'], t['sdf_weight'], t['fs_weight'])
        n_kf = n_frames // m['keyframe_every'] + 2
        self.kfdb = self.make_kfdb(n_kf, int(self.H * self.W * m['n_pixels']))
        self.tracker = self.make_tracker()
        self.adjuster = self.make_bundle_adjuster()
        self.est, self.est_rel = {}, {}
        self.t_track, self.t_map, self.n_map_iters = [], [], 0

    # ---- hooks ---------------------------------------------------------------------------------------------------------
    def make_model(self):
        model = JointEncoding(self.cfg, self.bound).to(self.dev)
        model.noise_mode = self.noise
        model.train()
        return model

    def make_map_optimizer(self):                                   # coslam.py:610-624
        m = self.cfg['mapping']
        return cb_optim.Adam([{'params': self.model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': m['lr_decoder']},
                              {'params': self.model.embed_fn.parameters(), 'eps': 1e-15, 'lr': m['lr_embed']}], betas=(0.9, 0.99))

    def make_kfdb(self, n_kf, rays_per_kf):
        return KeyFrameDatabase(self.cfg, self.H, self.W, n_kf, rays_per_kf, self.dev)

    def make_tracker(self):
        return FrameTracker(self.model, self.cfg, device=self.dev, use_graph=self.noise == 'device')

    def make_bundle_adjuster(self):
        """The autograd-free BA of coslam_b200/mapping.py (None: the torch-autograd loop below, which the oracle arm uses)."""
        if os.environ.get('COSLAM_B200_AUTOGRAD_BA'):
            return None
        from coslam_b200.mapping import BundleAdjuster
        return BundleAdjuster(self.model, self.cfg, self.kfdb, self.map_optimizer, self.dev)

    def sample_pixels(self, n, k):
        """select_samples (coslam.py:131-138): `random.sample(range(n), k)`, drawn by the library's host sampler (bit-identical)."""
        from coslam_b200.keyframe import sample_range
        return torch.from_numpy(sample_range(n, k)).to(self.dev)

    def track_pose(self, d_cam, rgb, depth, c2w_init):
        return self.tracker.track(d_cam, rgb, depth, c2w_init)[0]

    def ba_batch(self, poses_all, current, n_cur):
        m = self.cfg['mapping']
        return self.kfdb.assemble_ba_batch(m['sample'], poses_all, m['keyframe_every'], current, n_cur)[:4]

    def sync(self):
        if torch.device(self.dev).type == 'cuda':
            torch.cuda.synchronize()

    def loss(self, ret, smooth=False):                              # get_loss_from_ret, coslam.py:140-160
        w = self.loss_w
        loss = w[0] * ret['rgb_loss'] + w[1] * ret['depth_loss'] + w[2] * ret['sdf_loss'] + w[3] * ret['fs_loss']
        sw = self.cfg['training']['smooth_weight']
        if smooth and sw > 0:
            loss = loss + sw * self.smoothness()
        return loss

    def smoothness(self):                                           # coslam.py:251-274
        t = self.cfg['training']
        n, vox, margin = t['smooth_pts'], t['smooth_vox'], t['smooth_margin']
        lo, hi = self.bound[:, 0], self.bound[:, 1]
        offset = torch.rand(3).to(self.dev) * (hi - lo - (n - 1) * vox - 2 * margin) + margin
        ar = torch.arange(0, n - 1, dtype=torch.long, device=self.dev)
        coords = torch.stack(torch.meshgrid(ar, ar, ar, indexing='ij'), -1).float()
        pts = (coords + torch.rand((1, 1, 1, 3)).to(self.dev)) * vox + lo + offset
        f = self.model.query_sdf((pts - lo) / (hi - lo), embed=True)
        tv = ((f[1:] - f[:-1]) ** 2).sum() + ((f[:, 1:] - f[:, :-1]) ** 2).sum() + ((f[:, :, 1:] - f[:, :, :-1]) ** 2).sum()
        return tv / (n ** 3)

    def pixels(self, frame, idx, h_off=0, w_off=0, H=None):
        H = self.H if H is None else H
        ih, iw = idx % H + h_off, idx // H + w_off                  # coslam.py:187
        return frame['direction'][0, ih, iw], frame['rgb'][0, ih, iw], frame['depth'][0, ih, iw][:, None]

    def first_frame_mapping(self, frame, c2w, n_iters):             # coslam.py:162-207
        self.est[0] = c2w
        n = self.cfg['mapping']['sample']
        if self.adjuster is not None:
            current = torch.cat([frame['direction'], frame['rgb'], frame['depth'][..., None]], -1).reshape(-1, 7)
            self.adjuster.first_frame(c2w, current, n_iters)
            n_iters = 0
        for _ in range(n_iters):
            self.map_optimizer.zero_grad()
            idx = self.sample_pixels(self.H * self.W, n)
            d_cam, rgb, depth = self.pixels(frame, idx)
            rays_o, rays_d = pose.rays_from_pose(d_cam, c2w)
            self.loss(self.model.forward(rays_o, rays_d, rgb, depth)).backward()
            self.map_optimizer.step()
        self.kfdb.add_keyframe(frame, filter_depth=self.cfg['mapping']['filter_depth'])

    def predict_pose(self, i):                                      # coslam.py:399-412
        if i == 1 or not self.cfg['tracking']['const_speed']:
            return self.est[i - 1]
        return (self.est[i - 1] @ self.est[i - 2].inverse()) @ self.est[i - 1]

    def track(self, frame, i):                                      # coslam.py:502-573
        t = self.cfg['tracking']
        iH, iW = t['ignore_edge_H'], t['ignore_edge_W']
        idx = self.sample_pixels((self.H - 2 * iH) * (self.W - 2 * iW), t['sample'])
        d_cam, rgb, depth = self.pixels(frame, idx, iH, iW, self.H - 2 * iH)
        t0 = time.perf_counter()
        self.est[i] = self.track_pose(d_cam, rgb, depth, self.predict_pose(i)).detach().clone()
        self.sync()
        self.t_track.append(time.perf_counter() - t0)
        every = self.cfg['mapping']['keyframe_every']
        if i % every != 0:
            self.est_rel[i] = self.est[i] @ self.est[i // every * every].inverse()

    def global_ba(self, frame, cur):                                # coslam.py:285-397
        m = self.cfg['mapping']
        every = m['keyframe_every']
        kf_ids = list(range(0, cur, every))
        poses = torch.stack([self.est[k] for k in kf_ids])
        if self.adjuster is not None:
            current = torch.cat([frame['direction'], frame['rgb'], frame['depth'][..., None]], -1).reshape(-1, 7)
            self.sync()
            t0 = time.perf_counter()
            new_kf, new_cur = self.adjuster.run(poses, self.est[cur], current)
            self.sync()
            self.t_map.append(time.perf_counter() - t0)
            self.n_map_iters += m['iters']
            for j, k in enumerate(kf_ids):
                self.est[k] = new_kf[j].clone()
            self.est[cur] = new_cur.clone()
            return
        fixed = poses[:1]
        rot = trans = pose_opt = None
        if len(self.kfdb) >= 2:
            free = torch.cat([poses[1:], self.est[cur][None]])
            to_param = pose.matrix_to_axis_angle if self.cfg['training']['rot_rep'] == 'axis_angle' else pose.matrix_to_quaternion
            rot = torch.nn.Parameter(to_param(free[:, :3, :3].cpu()).to(self.dev))
            trans = torch.nn.Parameter(free[:, :3, 3].clone())
            pose_opt = torch.optim.Adam([{'params': rot, 'lr': m['lr_rot']}, {'params': trans, 'lr': m['lr_trans']}])
        self.map_optimizer.zero_grad()
        current = torch.cat([frame['direction'], frame['rgb'], frame['depth'][..., None]], -1).reshape(-1, 7)
        n_cur = max(m['sample'] // len(self.kfdb), m['min_pixels_cur'])
        self.sync()
        t0 = time.perf_counter()
        for it in range(m['iters']):
            poses_all = torch.cat([fixed, pose.pose_matrix(rot, trans)]) if pose_opt is not None else torch.cat([poses, self.est[cur][None]])
            rays_o, rays_d, rgb, depth = self.ba_batch(poses_all, current, n_cur)
            self.loss(self.model.forward(rays_o, rays_d, rgb, depth), smooth=True).backward()
            if (it + 1) % m['map_accum_step'] == 0:
                if (it + 1) > m['map_wait_step']:
                    self.map_optimizer.step()
                self.map_optimizer.zero_grad()
            if pose_opt is not None and (it + 1) % m['pose_accum_step'] == 0:
                pose_opt.step()
                pose_opt.zero_grad()
        self.sync()
        self.t_map.append(time.perf_counter() - t0)
        self.n_map_iters += m['iters']
        if pose_opt is not None:
            with torch.no_grad():
                new = pose.pose_matrix(rot, trans)
            for j, k in enumerate(kf_ids[1:]):
                self.est[k] = new[j].clone()
            self.est[cur] = new[-1].clone()

    def final_trajectory(self, n):                                  # convert_relative_pose, coslam.py:574-590
        every = self.cfg['mapping']['keyframe_every']
        return torch.stack([self.est[i] if i % every == 0 else self.est_rel[i] @ self.est[i // every * every] for i in range(n)])


    # ---- coslam.py:652-700 ----------------------------------------------------------------------------------------------
    def run(self, frames, first_iters):
        m = self.cfg['mapping']
        for i, frame in enumerate(frames):
            if i == 0:
                self.first_frame_mapping(frame, frame['c2w'], first_iters)
                continue
            if i == 3:                                                  # steady state: after the first-frame mapping and the graph capture
                self.sync()
                self.t_steady = time.perf_counter()
            self.track(frame, i)
            if i % m['map_every'] == 0:
                self.global_ba(frame, i)
            if i % m['keyframe_every'] == 0:
                self.kfdb.add_keyframe(frame, filter_depth=m['filter_depth'])
        return self.final_trajectory(len(frames))


def make_frames(cfg, n_frames, dev):
    """Frames as the reference's dataset objects serve them (datasets/dataset.py:250-265): direction [1,H,W,3], rgb
    [1,H,W,3], depth [1,H,W], frame_id, c2w."""
    cam = cfg['cam']
    dirs = synth.camera_ray_dirs(cam['H'], cam['W'], cam['fx'], cam['fy'], cam['cx'], cam['cy']).to(dev)
    frames = []
    for i, c2w in enumerate(trajectory(n_frames).to(dev)):
        rgb, depth = render_frame(c2w, dirs)
        frames.append({'frame_id': i, 'c2w': c2w, 'direction': dirs[None], 'rgb': rgb[None].contiguous(), 'depth': depth[None].contiguous()})
    return frames


def scaled_config(cfg, scale):
    """The camera of `cfg` at `scale` of its resolution (intrinsics scaled alike; edges ignored by the tracker too)."""
    cam = cfg['cam']
    cam['H'], cam['W'] = int(cam['H'] * scale), int(cam['W'] * scale)
    for k in ('fx', 'fy', 'cx', 'cy'):
        cam[k] = cam[k] * scale
    for k in ('ignore_edge_H', 'ignore_edge_W'):
        cfg['tracking'][k] = int(cfg['tracking'][k] * scale)
    return cfg


def ate_rmse(est, gt):
    """Translation RMSE without alignment (both trajectories start from the same first pose), metres."""
    return float(torch.sqrt(((est[:, :3, 3] - gt[:, :3, 3]) ** 2).sum(-1).mean()))


def seed_everything(seed=0):                                        # coslam.py:43-49
    random.seed(seed)
    torch.manual_seed(seed)
    import numpy as np
    np.random.seed(seed)


def workload_config(workload, scale=1.0):
    """Config of a synthetic workload; places the room inside the config's bound (module-level SHIFT / SCALE)."""
    if workload == 'scannet0000':
        # BASELINE.json config 3 in miniature: scene0000's hyper-parameters (640x480, S = 117, far 8, quaternion poses,
        # optim_cur False, smoothness lattice 63^3) on the same room, moved into scene0000's bound [[-0.1,8.6],[-0.1,8.9],[-0.3,3.3]]
        SHIFT.copy_(torch.tensor([0.9, 1.3, 1.4]))
    if workload == 'tum_desk':
        # BASELINE.json config 5's SHAPE in one process: this fork's fr1_desk hyper-parameters (368x496 frames after resize + crop,
        # S = 64 + 21 = 85 plus 128 importance samples, trunc 0.05, 20 BA iterations) on the room at 0.8 scale inside TUM's bound
        SCALE[0] = 0.8
        SHIFT.copy_(torch.tensor([-2.9, -1.0, 0.0]))
        cfg = synth.tum_desk_config(n_importance=128)
        cfg['cam'].update({'H': 368, 'W': 496, 'fx': 413.84, 'fy': 413.2, 'cx': 246.88, 'cy': 196.24})
        cfg = scaled_config(cfg, scale)
    else:
        cfg = scaled_config(synth.CONFIGS[workload](), scale)
    return cfg


def run_workload(workload='room0', frames=60, scale=1.0, first_iters=None, noise='device', dev='cuda:0'):
    """One synthetic SLAM run (BASELINE.json configs 2 / 3 / 5 in miniature); returns the result dictionary `main` prints.
    Re-entrant: the module-level room placement is reset first (bench.py runs the three workloads in one process)."""
    SHIFT.zero_()
    SCALE[0] = 1.0
    seed_everything(0)
    cfg = workload_config(workload, scale)
    first_iters = cfg['mapping']['first_iters'] if first_iters is None else first_iters
    frame_list = make_frames(cfg, frames, dev)
    gt = torch.stack([f['c2w'] for f in frame_list])
    slam = SynthSLAM(cfg, dev, frames, noise=noise)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    est = slam.run(frame_list, first_iters)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    wall = t1 - t0
    steady = (frames - 3) / (t1 - slam.t_steady) if getattr(slam, 't_steady', None) and frames > 3 else None
    warm = slam.t_track[2:] if len(slam.t_track) > 4 else slam.t_track      # the first frame captures the graph
    return {
        'workload': '%s synthetic %dx%d RGB-D, %d frames, tracking %d rays x %d its, BA %d rays x %d its every %d frames' % (
            workload, cfg['cam']['W'], cfg['cam']['H'], frames, cfg['tracking']['sample'], cfg['tracking']['iter'],
            cfg['mapping']['sample'], cfg['mapping']['iters'], cfg['mapping']['map_every']),
        'noise': noise, 'ate_rmse_cm': 100 * ate_rmse(est, gt), 'const_pose_ate_cm': 100 * ate_rmse(gt[:1].expand_as(gt), gt),
        'tracking_ms_per_frame': 1e3 * sum(warm) / max(len(warm), 1),
        'mapping_ms_per_iter': 1e3 * sum(slam.t_map) / max(slam.n_map_iters, 1),
        'frames_per_s': (frames - 1) / wall, 'steady_frames_per_s': steady, 'first_frame_iters': first_iters}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=60)
    ap.add_argument('--scale', type=float, default=1.0)
    ap.add_argument('--first-iters', type=int, default=None)
    ap.add_argument('--noise', default='device', choices=('device', 'host'))
    ap.add_argument('--workload', default='room0', choices=('room0', 'scannet0000', 'tum_desk'))
    a = ap.parse_args()
    print(json.dumps(run_workload(a.workload, a.frames, a.scale, a.first_iters, a.noise)))


if __name__ == '__main__':
    main()
