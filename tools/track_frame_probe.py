"""One-launch tracking loop (cb200_track_frame) against the graph-captured and the eager loop: same pose, then ms / frame.

    python tools/track_frame_probe.py [--workload room0] [--rays 1024]"""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from coslam_b200 import pose, synth                 # noqa: E402
from coslam_b200.tracking import FrameTracker       # noqa: E402
from tests import util                              # noqa: E402

DEV = 'cuda:0'


def case(cfg, n_rays, seed=3):
    g = torch.Generator().manual_seed(seed)
    f = torch.randn(n_rays, 3, generator=g)
    f[:, 2] = -f[:, 2].abs() - 1.0
    dir_cam = f / f.norm(dim=-1, keepdim=True)
    bound = synth.bound_tensor(cfg)
    centre = bound.mean(-1)
    quat = cfg['training']['rot_rep'] != 'axis_angle'
    rot0 = torch.tensor([[0.9, 0.1, -0.2, 0.15]]) if quat else torch.tensor([[0.1, -0.2, 0.15]])
    c2w = pose.pose_matrix(rot0, centre[None] + torch.tensor([[0.2, -0.1, 0.1]]))[0]
    rgb = torch.rand(n_rays, 3, generator=g)
    depth = torch.rand(n_rays, 1, generator=g) * 3 + 0.5
    depth[::17] = 0.0
    return dir_cam, rgb, depth, c2w


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='room0')
    ap.add_argument('--rays', type=int, default=1024)
    ap.add_argument('--frames', type=int, default=50)
    args = ap.parse_args()
    for perturb in (0.0, 1.0):
        cfg = synth.CONFIGS[args.workload]()
        cfg['training']['perturb'] = perturb
        cfg['tracking']['sample'] = args.rays
        weights = synth.decoder_init(seed=5)
        from coslam_b200.scene_rep import JointEncoding
        probe = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
        table = synth.table_init(probe.embed_fn.params.numel(), 'trained', seed=77)
        model = util.make_cuda_model(cfg, table, weights, DEV)
        dir_cam, rgb, depth, c2w0 = case(cfg, args.rays)
        res = {}
        for name, kw in (('graph', dict(persistent=False)), ('persistent', dict(persistent=True))):
            tr = FrameTracker(model, cfg, device=DEV, use_graph=True, **kw)
            model._noise_counter.zero_()
            est, info = tr.track(dir_cam, rgb, depth, c2w0)
            torch.cuda.synchronize()
            res[name] = (est.clone(), float(info['best_loss']), float(info['last_loss']), info['last_c2w'].clone(), int(model._noise_counter))
            d, r, dp = dir_cam.to(DEV), rgb.to(DEV), depth.to(DEV)
            for _ in range(5):
                tr.track(d, r, dp, c2w0)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.frames):
                est, _ = tr.track(d, r, dp, c2w0)
                est = est.cpu()
            dt = (time.perf_counter() - t0) / args.frames
            print('perturb %.0f %-10s best %.6e last %.6e noise counter %d   %.3f ms/frame (device inputs, pose read back)' %
                  (perturb, name, res[name][1], res[name][2], res[name][4], dt * 1e3), flush=True)
            assert tr.persistent == (name == 'persistent')
            # where a frame's time goes: host time of the launch path alone (no sync), device time between two events
            fn = tr._frame_persistent if tr.persistent else tr._graph.replay
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(20):
                with torch.no_grad():
                    fn()
            host = (time.perf_counter() - t0) / 20
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                with torch.no_grad():
                    fn()
            e1.record()
            torch.cuda.synchronize()
            print('   launch path: host %.3f ms per frame (not synchronised), device %.3f ms per frame (back to back)' % (host * 1e3, e0.elapsed_time(e1) / 20))
            if tr.persistent and os.environ.get('CB200_EXTRA_NVCC_FLAGS', '').find('CB200_TRACK_CLOCKS') >= 0:
                clk = tr._wsp[tr._wsp_off + 256:tr._wsp_off + 256 + 96].view(torch.int64).cpu().tolist()
                names = ['P0 rays+sample', 'P1 fwd', 'P2 render', 'P3 render bwd+list', 'P4 bwd', 'P5 pose']
                print('   phases of CTA 0 (us per iteration): ' + ', '.join('%s %.1f' % (n_, c / 1e3 / tr.iters) for n_, c in zip(names, clk[:6])))
                print('   barrier waits (us per iteration):   ' + ', '.join('%.1f' % (c / 1e3 / tr.iters) for c in clk[6:12]))
        g, p = res['graph'], res['persistent']
        print('perturb %.0f: |est| diff %.3e  |last c2w| diff %.3e  best rel %.3e  moved %.3e' %
              (perturb, float((g[0] - p[0]).abs().max()), float((g[3] - p[3]).abs().max()), abs(g[1] - p[1]) / abs(g[1]),
               float((g[3].cpu() - c2w0).abs().max())), flush=True)


if __name__ == '__main__':
    main()
