"""Times points_bwd (compaction + per-point kernel) and points_fwd on the arguments of a REAL mapping step: the bench
workload's rays, the loss head's d_raw (about 60 % of the points live), the saved ReLU masks.

    python tools/bwd_probe.py [--rays 262144] [--workload room0] [--reps 10]

Prints one line per variant (full, without the table gradient, without the decoder weight gradients, positions only)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from coslam_b200 import ops, synth                 # noqa: E402
from coslam_b200.scene_rep import JointEncoding    # noqa: E402

DEV = 'cuda:0'


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--rays', type=int, default=1 << 18)
    ap.add_argument('--workload', default='room0')
    ap.add_argument('--reps', type=int, default=10)
    ap.add_argument('--check', action='store_true', help='print checksums of the gradients (to compare builds)')
    args = ap.parse_args()
    cfg = synth.CONFIGS[args.workload]()
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
    model.noise_mode, model.noise_seed = 'device', 1
    with torch.no_grad():
        model.embed_fn.params.copy_(synth.table_init(model.embed_fn.params.numel(), 'trained').to(DEV))
        for dst, src in zip(model.decoder.fused_weights(), synth.decoder_init(seed=0)):
            dst.copy_(src.to(DEV))
    model.train()
    batch = [t.to(DEV) for t in synth.keyframe_ray_batch(cfg, args.rays, seed=100)[:4]]
    t = cfg['training']
    captured = {}
    real_bwd, real_fwd = ops.points_backward, ops.lazy_ray_forward

    def spy_bwd(*a, **kw):
        captured['bwd'] = (a, kw)
        return real_bwd(*a, **kw)

    def spy_fwd(*a, **kw):
        captured['fwd'] = (a, kw)
        return real_fwd(*a, **kw)
    ops.points_backward, ops.lazy_ray_forward = spy_bwd, spy_fwd
    o, d = batch[0].clone().requires_grad_(True), batch[1].clone().requires_grad_(True)
    ret = model.forward(o, d, batch[2], batch[3])
    loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] + t['sdf_weight'] * ret['sdf_loss'] + t['fs_weight'] * ret['fs_loss']
    loss.backward()
    ops.points_backward, ops.lazy_ray_forward = real_bwd, real_fwd
    a, kw = captured['bwd']
    d_raw = a[4]
    live = float((d_raw.reshape(-1, 4) != 0).any(1).float().mean())
    P = d_raw.numel() // 4
    print('workload %s rays %d points %d live %.4f' % (args.workload, args.rays, P, live), flush=True)
    kw = dict(kw, feat=None)                       # the step above may have kept rows: the lines below time the gathering backward
    if 'fwd' in captured:
        fa, fkw = captured['fwd']
        fkw = dict(fkw, feat=None)
        print('fwd (lazy, two launches)   %.3f ms' % timed(lambda: real_fwd(*fa, **fkw), args.reps), flush=True)
    for name, over in (('bwd full', {}), ('bwd no table grad', dict(need_table=False)), ('bwd no mlp grad', dict(need_mlp=False)),
                       ('bwd no pos grad', dict(need_pos=False)), ('bwd pos only', dict(need_table=False, need_mlp=False))):
        k = dict(kw)
        k.update(over)
        ms = timed(lambda: real_bwd(*a, **k), args.reps)
        print('%-26s %.3f ms   %.1f M live points/s   %.0f GB/s of 2 048 B/pt' % (name, ms, P * live / ms / 1e3, P * live * 2048 / ms / 1e6), flush=True)
    if 'fwd' in captured:
        # kept features (cb200_points_fwd_keep / cb200_points_bwd_kept): the forward streams 128 B per evaluated point out, the
        # backward reads them back instead of gathering the table a second time (no position gradient)
        feat = torch.empty((P, 32), dtype=torch.float32, device=DEV)
        fa, fkw = captured['fwd']
        print('fwd keeping the features   %.3f ms' % timed(lambda: real_fwd(*fa, **dict(fkw, feat=feat)), args.reps), flush=True)
        k = dict(kw)
        k.update(need_pos=False, feat=feat)
        ms = timed(lambda: real_bwd(*a, **k), args.reps)
        print('%-26s %.3f ms   %.1f M live points/s' % ('bwd from kept features', ms, P * live / ms / 1e3), flush=True)
        ref = real_bwd(*a, **dict(kw, need_pos=False))
        got = real_bwd(*a, **k)
        torch.cuda.synchronize()
        for n_, x, y in [('d_table', ref[0], got[0])] + [('dW%d' % (i + 1), ref[1][i], got[1][i]) for i in range(4)]:
            print('kept vs gathered %-8s max |diff| %.3e of max %.3e' % (n_, float((x - y).abs().max()), float(x.abs().max())), flush=True)
    if 'fwd' in captured:
        # the same with the derivatives kept as well (512 B per point): the full backward, ray gradients included, without a gather
        feat4 = torch.empty((P, 128), dtype=torch.float32, device=DEV)
        print('fwd keeping feat + d/dx    %.3f ms' % timed(lambda: real_fwd(*fa, **dict(fkw, feat=feat4)), args.reps), flush=True)
        k = dict(kw)
        k.update(feat=feat4)
        ms = timed(lambda: real_bwd(*a, **k), args.reps)
        print('%-26s %.3f ms   %.1f M live points/s' % ('bwd full from kept rows', ms, P * live / ms / 1e3), flush=True)
        ref = real_bwd(*a, **kw)
        got = real_bwd(*a, **k)
        torch.cuda.synchronize()
        for n_, x, y in [('d_table', ref[0], got[0])] + [('dW%d' % (i + 1), ref[1][i], got[1][i]) for i in range(4)] + \
                [('d_rays_o', ref[2][0], got[2][0]), ('d_rays_d', ref[2][1], got[2][1])]:
            print('kept vs gathered %-8s max |diff| %.3e of max %.3e' % (n_, float((x - y).abs().max()), float(x.abs().max())), flush=True)
    if args.check:
        d_table, d_w, (d_o, d_d) = real_bwd(*a, **kw)
        torch.cuda.synchronize()
        for n_, x in (('d_table', d_table), ('dW1', d_w[0]), ('dW2', d_w[1]), ('dW3', d_w[2]), ('dW4', d_w[3]), ('d_rays_o', d_o), ('d_rays_d', d_d)):
            print('check %-9s sum %.9e  abs %.9e  max %.9e' % (n_, float(x.double().sum()), float(x.double().abs().sum()), float(x.abs().max())))


if __name__ == '__main__':
    main()
