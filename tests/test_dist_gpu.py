"""Ray-sharded mapping step on 2 GPUs (NCCL) against the single-GPU step on the same global batch:
losses are the GLOBAL ones on every rank, summed gradients agree within the 1e-3 gate (sharding changes summation order).
Needs two visible GPUs; skipped otherwise (run with `gpurun --gpus 2`)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from coslam_b200 import synth

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')]


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _build(cfg, dev):
    from coslam_b200.scene_rep import JointEncoding
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(dev)).to(dev)
    table = synth.table_init(model.embed_fn.params.numel(), 'trained', seed=77)
    with torch.no_grad():
        model.embed_fn.params.copy_(table.to(dev))
        for dst, src in zip(model.decoder.fused_weights(), synth.decoder_init(seed=5)):
            dst.copy_(src.to(dev))
    model.train()
    return model


def _worker(rank, world, port, n_rays, out, backend='nccl', one_gpu=False):
    """one_gpu: every rank on cuda:0 over gloo (the host logic is backend-agnostic), so a 1-GPU box can run the shard parity."""
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(0 if one_gpu else rank)
    dist.init_process_group(backend, rank=rank, world_size=world)
    try:
        from coslam_b200.dist import ShardedMapper, equal_shard
        dev = 'cuda:0' if one_gpu else 'cuda:%d' % rank
        cfg = synth.room0_config()
        cfg['training']['perturb'] = 0.0
        model = _build(cfg, dev)
        batch = synth.keyframe_ray_batch(cfg, n_rays, n_kf=8, seed=3)
        lo, hi = equal_shard(n_rays, world, rank)
        o = batch[0][lo:hi].to(dev).requires_grad_(True)
        d = batch[1][lo:hi].to(dev).requires_grad_(True)
        t = cfg['training']
        mapper = ShardedMapper(model, (t['rgb_weight'], t['depth_weight'], t['sdf_weight'], t['fs_weight']))
        loss, ret = mapper.step(o, d, batch[2][lo:hi].to(dev), batch[3][lo:hi].to(dev))
        out[rank] = {'loss': float(loss.detach()), 'losses': [float(ret[k].detach()) for k in ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss')],
                     'd_table': model.embed_fn.params.grad.cpu(), 'dW1': model.decoder.fused_weights()[0].grad.cpu(),
                     'd_rays_o': o.grad.cpu()}
    finally:
        dist.destroy_process_group()


def test_two_gpu_sharded_step_matches_single_gpu():
    check_sharded_step('nccl', False)


def check_sharded_step(backend, one_gpu, world=2):
    n_rays = 4096
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n_rays, out, backend, one_gpu), nprocs=world, join=True)

    dev = 'cuda:0'
    cfg = synth.room0_config()
    cfg['training']['perturb'] = 0.0
    model = _build(cfg, dev)
    n_use = (n_rays // world) * world                    # equal_shard drops the remainder, like DistributedSampler(drop_last=True)
    batch = [t[:n_use] for t in synth.keyframe_ray_batch(cfg, n_rays, n_kf=8, seed=3)[:4]]
    o = batch[0].to(dev).requires_grad_(True)
    d = batch[1].to(dev).requires_grad_(True)
    ret = model.forward(o, d, batch[2].to(dev), batch[3].to(dev))
    t = cfg['training']
    loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] + t['sdf_weight'] * ret['sdf_loss'] \
        + t['fs_weight'] * ret['fs_loss']
    loss.backward()

    def rel(a, b):
        return float((a.double() - b.double()).abs().max() / b.double().abs().max())
    for r in range(world):
        assert abs(out[r]['loss'] - float(loss.detach())) <= 1e-5 * abs(float(loss.detach()))
        for a, k in zip(out[r]['losses'], ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss')):
            assert abs(a - float(ret[k].detach())) <= 1e-5 * abs(float(ret[k].detach())) + 1e-9, k
        assert rel(out[r]['d_table'], model.embed_fn.params.grad.cpu()) <= 1e-3
        assert rel(out[r]['dW1'], model.decoder.fused_weights()[0].grad.cpu()) <= 1e-3
    for r in range(1, world):
        assert torch.equal(out[0]['d_table'], out[r]['d_table'])      # replicas stay identical after AR-2
    got = torch.cat([out[r]['d_rays_o'] for r in range(world)], 0)
    assert rel(got, o.grad.cpu()) <= 1e-3


def _ba_worker(rank, world, port, out, backend='nccl', one_gpu=False):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(0 if one_gpu else rank)
    dist.init_process_group(backend, rank=rank, world_size=world)
    try:
        import random
        from coslam_b200.mapping import BundleAdjuster
        from tests.test_mapping_gpu import _map_optimizer, _setup
        dev = 'cuda:0' if one_gpu else 'cuda:%d' % rank
        cfg, model, db, kf_poses, cur_pose, current = _setup('room0', dev=dev)
        cfg['mapping'].update({'iters': 6, 'map_accum_step': 2, 'map_wait_step': 2, 'pose_accum_step': 3})
        random.seed(5 + 100 * rank); torch.manual_seed(5 + 100 * rank)        # rank 0's draws are the ones that count (broadcast)
        ba = BundleAdjuster(model, cfg, db, _map_optimizer(cfg, model), dev, group=dist.group.WORLD)
        new_kf, new_cur = ba.run(kf_poses, cur_pose, current)
        torch.cuda.synchronize()
        out[rank] = {'kf': new_kf.cpu(), 'cur': new_cur.cpu(), 'params': [p.detach().cpu() for p in model._fused()],
                     'losses': ba.last_losses.cpu()}
    finally:
        dist.destroy_process_group()


def test_two_gpu_sharded_bundle_adjustment_matches_single_gpu():
    """BundleAdjuster with the rays of every iteration sharded over 2 ranks (AR-1 on the counts / loss sums, AR-2 on the flat
    gradient bucket and the pose gradients in front of the optimiser steps) against the single-GPU call on the same draws:
    replicas identical, poses and parameter updates equal up to the summation order."""
    check_sharded_bundle_adjustment('nccl', False)


def check_sharded_bundle_adjustment(backend, one_gpu, world=2):
    import random
    from coslam_b200.mapping import BundleAdjuster
    from tests.test_mapping_gpu import _map_optimizer, _setup
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_ba_worker, args=(world, _free_port(), out, backend, one_gpu), nprocs=world, join=True)
    cfg, model, db, kf_poses, cur_pose, current = _setup('room0', dev='cuda:0')
    cfg['mapping'].update({'iters': 6, 'map_accum_step': 2, 'map_wait_step': 2, 'pose_accum_step': 3})
    before = [p.detach().clone().cpu() for p in model._fused()]
    random.seed(5); torch.manual_seed(5)
    ba = BundleAdjuster(model, cfg, db, _map_optimizer(cfg, model), 'cuda:0')
    ref_kf, ref_cur = ba.run(kf_poses, cur_pose, current)
    torch.cuda.synchronize()
    a, b = out[0], out[1]
    assert torch.equal(a['kf'], b['kf']) and torch.equal(a['cur'], b['cur'])
    for x, y in zip(a['params'], b['params']):
        assert torch.equal(x, y)                                        # replicas stay bit-identical without a broadcast
    assert float((a['kf'] - ref_kf.cpu()).abs().max()) <= 3e-4 and float((a['cur'] - ref_cur.cpu()).abs().max()) <= 3e-4
    assert float((a['losses'] - ba.last_losses.cpu()).abs().max()) <= 1e-3 * float(ba.last_losses.abs().max())
    for b0, p, q in zip(before, a['params'], [p.detach().cpu() for p in model._fused()]):
        moved = float((q - b0).abs().mean())
        assert moved > 0 and float((p - q).abs().mean()) <= 0.05 * moved + 1e-9


def test_tracker_mapper_split_tracks_the_synthetic_trajectory():
    """BASELINE config 5 in miniature (tools/slam_mp_synth.py; coslam_mp.py's tracker / mapper processes): tracker on GPU 0,
    mapper on GPU 1, poses and map exchanged by NCCL broadcasts once per mapping interval.  The trajectory must come out as
    well as the one-process loop's (ATE below 1 cm where holding the first pose gives > 10 cm)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', str(_free_port()), os.path.join(root, 'tools', 'slam_mp_synth.py'), '--workload', 'room0', '--frames', '21',
           '--first-iters', '100']
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith('{')][-1]
    res = json.loads(line)
    assert res['n_gpus'] == 2 and res['const_pose_ate_cm'] > 10.0
    assert res['ate_rmse_cm'] < 1.0, res
