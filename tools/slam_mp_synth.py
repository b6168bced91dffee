"""BASELINE.json config 5 in miniature: the tracker / mapper split of coslam_mp.py (mp_slam/tracker.py:137-159,
mp_slam/mapper.py:201-232) on synthetic RGB-D frames -- tracker on GPU 0, mapper on the remaining GPUs with the rays of its
bundle adjustment sharded over them -- as one `torchrun` job, one process per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P \
        tools/slam_mp_synth.py [--workload tum_desk|room0|scannet0000] [--frames 60]          (G >= 2)

The reference couples the two processes through shared-memory tensors polled with `time.sleep` (the tracker may run up to
1.5 x map_every frames ahead of the mapper, tracker.py:150-152, and copies the shared map before every frame, :154).  Here the
coupling is two NCCL broadcasts per mapping interval over NVLink, with the same one-interval lag:

    tracker, after tracking frame m_k = k * map_every:   A_k  estimated poses            rank 0 -> all
                                                          B_{k-1}  map + keyframe poses   rank 1 -> all   (result of BA(k-1))
    mapper:   A_k | B_{k-1} | BA(k) on frame m_k | keyframe insertion

so BA(k) runs while the tracker works on frames m_k + 1 .. m_{k+1}, and the tracker keeps using the map of BA(k-1) until then.
Every rank renders the synthetic frames itself (no image transport is modelled; the reference's processes read the dataset
independently as well).  Prints one JSON line on rank 0: end-to-end frames/s, ATE, time the tracker waited for the mapper.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                                   # noqa: E402
import torch.distributed as dist                               # noqa: E402

import slam_synth                                              # noqa: E402
from coslam_b200 import synth                                   # noqa: E402


class Mapper(slam_synth.SynthSLAM):
    def __init__(self, cfg, dev, n_frames, group):
        self.group = group
        super().__init__(cfg, dev, n_frames, noise='device')

    def make_tracker(self):
        return None

    def make_bundle_adjuster(self):
        from coslam_b200.mapping import BundleAdjuster
        return BundleAdjuster(self.model, self.cfg, self.kfdb, self.map_optimizer, self.dev, group=self.group)


def flat_params(model):
    return torch.cat([p.detach().reshape(-1) for p in model._fused()])


def load_flat(model, flat):
    off = 0
    with torch.no_grad():
        for p in model._fused():
            p.copy_(flat[off:off + p.numel()].view_as(p))
            off += p.numel()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=60)
    ap.add_argument('--workload', default='tum_desk', choices=('room0', 'scannet0000', 'tum_desk'))
    ap.add_argument('--first-iters', type=int, default=None)
    a = ap.parse_args()
    world, rank, local = int(os.environ['WORLD_SIZE']), int(os.environ['RANK']), int(os.environ.get('LOCAL_RANK', '0'))
    assert world >= 2, 'one tracker rank + at least one mapper rank'
    torch.cuda.set_device(local)
    dev = 'cuda:%d' % local
    dist.init_process_group('nccl')
    map_group = dist.new_group(list(range(1, world)))
    slam_synth.seed_everything(0)
    cfg = slam_synth.workload_config(a.workload, 1.0)
    m = cfg['mapping']
    every, F = m['map_every'], a.frames
    assert m['keyframe_every'] == every, 'the miniature assumes a keyframe at every mapping point (true for every shipped config)'
    first_iters = m['first_iters'] if a.first_iters is None else a.first_iters
    frames = slam_synth.make_frames(cfg, F, dev)
    gt = torch.stack([f['c2w'] for f in frames])
    est = torch.zeros((F, 4, 4), device=dev)                     # the shared pose table of coslam_mp.py:94-95, one replica per rank
    est[0] = frames[0]['c2w']
    K = (F - 1) // every                                         # mapping points m_1 .. m_K
    is_tracker = rank == 0
    slam = slam_synth.SynthSLAM(cfg, dev, F, noise='device') if is_tracker else Mapper(cfg, dev, F, map_group if world > 2 else None)
    if is_tracker:
        slam.adjuster = None
    flat = flat_params(slam.model)
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    waited = 0.0

    def sync_b():                                                # map + poses of the mapper's last BA: rank 1 -> everyone
        if not is_tracker:
            for j, c2w in slam.est.items():                      # the mapper's own (refined) keyframe poses, not the tracker's copy
                if j % every == 0:
                    est[j] = c2w
        if rank == 1:
            flat.copy_(flat_params(slam.model))
        dist.broadcast(flat, src=1)
        dist.broadcast(est, src=1)
        if is_tracker:
            load_flat(slam.model, flat)

    if not is_tracker:                                           # first-frame mapping (mapper.py:205-208), rays sharded over the mappers
        slam.first_frame_mapping(frames[0], frames[0]['c2w'], first_iters)
    sync_b()                                                     # B_0
    if is_tracker:
        slam.est[0] = est[0].clone()
        t_first = time.perf_counter()
        t_steady = None
        for i in range(1, F):
            if i == 3:                                           # steady state: after the tracker's graph capture
                torch.cuda.synchronize()
                t_steady = time.perf_counter()
            slam.track(frames[i], i)
            est[i] = slam.est[i]
            if i % every == 0:
                dist.broadcast(est, src=0)                       # A_k
                if i >= 2 * every:
                    torch.cuda.synchronize()
                    tw = time.perf_counter()
                    sync_b()                                     # B_{k-1}: keyframe poses and frame m_{k-1} as refined by BA(k-1)
                    torch.cuda.synchronize()
                    waited += time.perf_counter() - tw
                    for j in range(0, i - every + 1, every):
                        slam.est[j] = est[j].clone()
        if K >= 1:
            sync_b()                                             # result of BA(K)
            for j in range(0, K * every + 1, every):
                slam.est[j] = est[j].clone()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        traj = slam.final_trajectory(F)
        print(json.dumps({
            'workload': '%s synthetic %dx%d RGB-D, %d frames; tracker on GPU 0, mapper on %d GPU(s); tracking %d rays x %d its, BA %d rays x %d its '
                        'every %d frames (%d importance samples)' % (a.workload, cfg['cam']['W'], cfg['cam']['H'], F, world - 1,
                                                                     cfg['tracking']['sample'], cfg['tracking']['iter'], m['sample'], m['iters'],
                                                                     every, cfg['training']['n_importance']),
            'n_gpus': world, 'frames_per_s': (F - 1) / (t1 - t0), 'frames_per_s_after_first_frame_mapping': (F - 1) / (t1 - t_first),
            'steady_frames_per_s': (F - 3) / (t1 - t_steady) if t_steady else None,
            'tracker_waited_for_mapper_s': waited, 'ate_rmse_cm': 100 * slam_synth.ate_rmse(traj, gt),
            'const_pose_ate_cm': 100 * slam_synth.ate_rmse(gt[:1].expand_as(gt), gt),
            'tracking_ms_per_frame': 1e3 * sum(slam.t_track[2:]) / max(len(slam.t_track) - 2, 1)}), flush=True)
    else:
        for k in range(1, K + 1):
            cur = k * every
            dist.broadcast(est, src=0)                           # A_k: the tracker's poses up to frame m_k
            slam.est[cur] = est[cur].clone()
            for j in range(0, cur, every):                       # keyframes keep the mapper's own (refined) poses
                if j not in slam.est:
                    slam.est[j] = est[j].clone()
            if k >= 2:
                sync_b()                                         # B_{k-1}
            slam.global_ba(frames[cur], cur)                     # mapper.py:219
            slam.kfdb.add_keyframe(frames[cur], filter_depth=m['filter_depth'])      # mapper.py:222-223
        if K >= 1:
            sync_b()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
