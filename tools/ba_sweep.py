"""BASELINE.json config 4: mapping-BA ray-batch sweep over a synthetic keyframe database (room0: 401 keyframes x 40 800 rays =
16.4 M rays, 458 MB on the device), N = 2^16 .. 2^22 rays per iteration, through coslam_b200.mapping.BundleAdjuster: ray
assembly by keyframe slot, forward, losses, backward into table / decoder / 400 free poses, smoothness term, Adam on the map
every iteration and on the poses every 5th.  Index draws on the device (non-parity sampler).

    python tools/ba_sweep.py [--log2 16 18 20 22] [--iters 5] [--keyframes 401]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P tools/ba_sweep.py ...

With G ranks the N rays of every iteration are sharded (N / G per rank, strong scaling), counts / loss sums and the gradient
bucket + pose gradients all-reduced over NCCL.  Prints one JSON line per N on rank 0: ms per iteration (device time, max over
ranks) and rays/s."""
import argparse
import json
import os
import random
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from coslam_b200 import optim as cb_optim
from coslam_b200 import pose, synth
from coslam_b200.keyframe import KeyFrameDatabase
from coslam_b200.mapping import BundleAdjuster
from coslam_b200.scene_rep import JointEncoding


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--log2', type=int, nargs='+', default=[16, 18, 20, 22])
    ap.add_argument('--iters', type=int, default=5)
    ap.add_argument('--keyframes', type=int, default=401)
    ap.add_argument('--phases', action='store_true', help='sharded runs: also print the device time of forward | AR-1 | backward | AR-2 | optimise per iteration')
    a = ap.parse_args()
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = 'cuda:%d' % local
    group = None
    if world > 1:
        dist.init_process_group('nccl')
        group = dist.group.WORLD
    cfg = synth.room0_config()
    m, cam = cfg['mapping'], cfg['cam']
    H, W = cam['H'], cam['W']
    R = int(H * W * m['n_pixels'])
    torch.manual_seed(0)
    random.seed(0)
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(dev)).to(dev)
    with torch.no_grad():
        model.embed_fn.params.copy_(synth.table_init(model.embed_fn.params.numel(), 'trained', seed=77).to(dev))
        for dst, src in zip(model.decoder.fused_weights(), synth.decoder_init(seed=5)):
            dst.copy_(src.to(dev))
    model.noise_mode = 'device'
    model.train()
    n_kf = a.keyframes
    db = KeyFrameDatabase(cfg, H, W, n_kf, R, dev)
    g = torch.Generator(device=dev).manual_seed(1)
    dirs = synth.camera_ray_dirs(H, W, cam['fx'], cam['fy'], cam['cx'], cam['cy']).reshape(-1, 3).to(dev)
    for k in range(n_kf):                                    # the store filled directly (the insertion path is tested elsewhere)
        pick = torch.randint(0, H * W, (R,), device=dev, generator=g)
        db.rays[k, :, :3] = dirs[pick]
        db.rays[k, :, 3:6] = torch.rand((R, 3), device=dev, generator=g)
        db.rays[k, :, 6] = torch.rand((R,), device=dev, generator=g) * 3.0 + 0.5
    db.frame_ids = torch.arange(n_kf) * m['keyframe_every']
    bound = synth.bound_tensor(cfg)
    centre, ext = bound.mean(1), bound[:, 1] - bound[:, 0]
    gh = torch.Generator().manual_seed(2)
    poses = []
    for k in range(n_kf + 1):
        r = torch.randn(1, 3, generator=gh)
        poses.append(pose.pose_matrix(r / r.norm() * (0.2 + 0.001 * k), (centre + 0.1 * ext * (torch.rand(3, generator=gh) - 0.5))[None])[0])
    kf_poses, cur_pose = torch.stack(poses[:n_kf]).to(dev), poses[n_kf].to(dev)
    current = torch.cat([dirs, torch.rand((H * W, 3), device=dev, generator=g), torch.rand((H * W, 1), device=dev, generator=g) * 3 + 0.5], -1)
    opt = cb_optim.Adam([{'params': model.decoder.parameters(), 'weight_decay': 1e-6, 'lr': m['lr_decoder']},
                         {'params': model.embed_fn.parameters(), 'eps': 1e-15, 'lr': m['lr_embed']}], betas=(0.9, 0.99))
    ba = BundleAdjuster(model, cfg, db, opt, dev, group=group, device_sampler=True)
    m['iters'] = a.iters
    for lg in a.log2:
        m['sample'] = 1 << lg
        m['min_pixels_cur'] = max(100, (1 << lg) // 64)     # keeps the current-frame share of the batch as at 2048 rays
        ba.run(kf_poses, cur_pose, current)                  # warm-up (allocator, attribute calls)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        reps = 3
        for _ in range(reps):
            ba.run(kf_poses, cur_pose, current)
        e1.record()
        torch.cuda.synchronize()
        if a.phases:
            ba.phase_times = []
            ba.run(kf_poses, cur_pose, current)
            torch.cuda.synchronize()
            names = ('forward', 'AR-1', 'backward', 'AR-2 (+ poses)', 'optimise')
            tot = [sum(ev[i].elapsed_time(ev[i + 1]) for ev in ba.phase_times) / len(ba.phase_times) for i in range(5)]
            if rank == 0:
                print('   rank 0 phases, ms per iteration: ' + ', '.join('%s %.3f' % p_ for p_ in zip(names, tot)), flush=True)
            ba.phase_times = None
        ms = torch.tensor([e0.elapsed_time(e1) / (reps * a.iters)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        n_rays = m['sample'] + max(m['sample'] // n_kf, m['min_pixels_cur'])
        if rank == 0:
            print(json.dumps({'workload': 'room0 mapping BA over %d keyframes x %d rays' % (n_kf, R), 'n_gpus': world, 'rays_per_iteration': n_rays,
                              'ms_per_iteration': float(ms), 'rays_per_s': n_rays / (float(ms) * 1e-3), 'iters_per_call': a.iters,
                              'includes': 'index draws + upload per call, ray assembly, fwd, losses, bwd (table, decoder, %d poses), smoothness 31^3, '
                                          'Adam (map: every iteration, poses: every %d)' % (n_kf, m['pose_accum_step'])}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
