"""One bundle-adjustment call (room0 / scene0000 shapes, synthetic keyframe store) for profiling.
    python tools/ba_probe.py [room0|scannet0000|tum_desk|tum_desk_imp] [time|ncu|profile] [n_keyframes]
time: ms per iteration of BundleAdjuster.run for 10 and 100 iterations per call (separates the per-call host work from the
per-iteration cost); ncu: one call, for `ncu --metrics gpu__time_duration.sum`."""
import os
import random
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import optim as cb_optim
from coslam_b200 import pose, synth
from coslam_b200.keyframe import KeyFrameDatabase
from coslam_b200.mapping import BundleAdjuster
from tests import util

DEV = 'cuda:0'
name = sys.argv[1] if len(sys.argv) > 1 else 'room0'
mode = sys.argv[2] if len(sys.argv) > 2 else 'time'
n_kf = int(sys.argv[3]) if len(sys.argv) > 3 else 8
cfg = synth.tum_desk_config(n_importance=128) if name == 'tum_desk_imp' else synth.CONFIGS[name]()
H, W = cfg['cam']['H'], cfg['cam']['W']
R = int(H * W * cfg['mapping']['n_pixels'])
every = cfg['mapping']['keyframe_every']
weights = synth.decoder_init(seed=5)
from coslam_b200.scene_rep import JointEncoding
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), 'trained', seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
model.noise_mode = 'device'
model.train()
bound = synth.bound_tensor(cfg)
centre, ext = bound.mean(1), bound[:, 1] - bound[:, 0]
g = torch.Generator().manual_seed(0)
db = KeyFrameDatabase(cfg, H, W, n_kf + 1, R, DEV)
dirs = synth.camera_ray_dirs(H, W, cfg['cam']['fx'], cfg['cam']['fy'], cfg['cam']['cx'], cfg['cam']['cy'])
poses = []
random.seed(0)
for k in range(n_kf + 1):
    r = torch.randn(1, 3, generator=g)
    r = r / r.norm() * (0.2 + 0.02 * k)
    poses.append(pose.pose_matrix(r, (centre + 0.1 * ext * (torch.rand(3, generator=g) - 0.5))[None])[0])
    batch = {'direction': dirs[None].to(DEV), 'rgb': torch.rand(1, H, W, 3, generator=g).to(DEV),
             'depth': (torch.rand(1, H, W, generator=g) * 3.0 + 0.5).to(DEV), 'frame_id': torch.tensor([k * every])}
    if k < n_kf:
        db.add_keyframe(batch)
current = torch.cat([batch['direction'], batch['rgb'], batch['depth'][..., None]], -1).reshape(-1, 7)
kf_poses, cur_pose = torch.stack(poses[:n_kf]).to(DEV), poses[n_kf].to(DEV)
m = cfg['mapping']
opt = cb_optim.Adam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': m['lr_decoder']},
                     {'params': model.embed_fn.parameters(), 'eps': 1e-15, 'lr': m['lr_embed']}], betas=(0.9, 0.99))
ba = BundleAdjuster(model, cfg, db, opt, DEV)
ba.run(kf_poses, cur_pose, current)
torch.cuda.synchronize()
if mode == 'ncu':
    sys.exit(0)
if mode == 'profile':          # host side of a call: cProfile over 300 calls with a device sync after each (as bench.py's mapping line)
    import cProfile
    import pstats
    for _ in range(20):
        ba.run(kf_poses, cur_pose, current)
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(300):
        ba.run(kf_poses, cur_pose, current)
        torch.cuda.synchronize()
    print('%.3f ms per call of %d iterations without the profiler' % ((time.perf_counter() - t0) / 300 * 1e3, m['iters']))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(300):
        ba.run(kf_poses, cur_pose, current)
        torch.cuda.synchronize()
    pr.disable()
    pstats.Stats(pr).sort_stats('cumulative').print_stats(28)
    sys.exit(0)
for iters in (10, 100):
    m['iters'] = iters
    ba.run(kf_poses, cur_pose, current)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        ba.run(kf_poses, cur_pose, current)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    print('%s, %d keyframes, %d rays: %d iterations per call: %.3f ms per call, %.3f ms per iteration'
          % (name, n_kf, m['sample'] + max(m['sample'] // n_kf, m['min_pixels_cur']), iters, dt * 1e3, dt * 1e3 / iters))
