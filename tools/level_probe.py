"""Per-level cost of the table gathers / reductions at the fused kernels' pattern, and what replicated gradient rows for the
dense (coarse) levels buy (follow-up of tools/l2_gather_peak.py and tools/red_probe.cu).

    python tools/level_probe.py [--rays 49152]
"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tools'))

from coslam_b200 import _lib, ops, synth          # noqa: E402
from coslam_b200.scene_rep import JointEncoding   # noqa: E402
import l2_gather_peak as lg                       # noqa: E402

DEV = 'cuda:0'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--rays', type=int, default=49152)
    ap.add_argument('--workload', default='room0')
    args = ap.parse_args()
    cfg = synth.room0_config() if args.workload == 'room0' else synth.scannet0000_config(hash_size=19)
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
    table = model.embed_fn.params.detach()
    grid = model.spec.grid()
    offs = [int(grid.offset[i]) for i in range(17)]
    n_dense = sum(1 for i in range(16) if not grid.hashed[i])
    print('level sizes', [offs[i + 1] - offs[i] for i in range(16)], 'dense', n_dense, flush=True)
    max_repl = 64
    dtable = torch.zeros(table.numel() + 2 * max_repl * offs[n_dense], device=DEV)
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, args.rays, seed=0)
    z = ops.sample_rays(model.spec, depth.to(DEV), perturb=False)
    n, s = z.shape
    pts = torch.empty((n * s, 3), device=DEV)
    x = torch.empty((n * s, 3), device=DEV)
    o, d = rays_o.to(DEV).contiguous(), rays_d.to(DEV).contiguous()
    _lib.check(_lib.load().cb200_debug_positions(ctypes.byref(model.spec.scene()), _lib.ptr(o), _lib.ptr(d), _lib.ptr(z),
                                                 ctypes.c_int64(n), ctypes.c_int64(s), _lib.ptr(pts), _lib.ptr(x), _lib.stream()),
               'debug_positions')
    idx = lg.corner_list(model, x.contiguous())
    P = idx.shape[0]
    slots = P * 2 * 4          # lane-ops per level
    sm_hz = 148 * 1.965e9

    def t(mode, mask=0, log2r=0, nrepl=0):
        return lg.time_mode(model, idx, table, dtable, mode | (mask << 4) | (log2r << 20) | (nrepl << 24), iters=10, warmup=2)

    base = t(2, 1 << 15) * 0          # warm the library
    print('per level (ms, SM-cycles per lane-op): reduce | gather')
    tot_r = tot_g = 0.0
    for l in range(16):
        r, g = t(2, 1 << l), t(1, 1 << l)
        tot_r += r
        tot_g += g
        print('  level %2d  reduce %.4f ms  %.3f   gather %.4f ms  %.3f' % (l, r, r * 1e-3 * sm_hz / slots, g, g * 1e-3 * sm_hz / slots), flush=True)
    print('sum of levels: reduce %.3f gather %.3f' % (tot_r, tot_g))
    print('all levels: reduce %.3f gather %.3f both %.3f' % (t(2), t(1), t(3)))
    dense_mask = (1 << n_dense) - 1
    print('dense levels only: reduce %.3f   hashed only: reduce %.3f' % (t(2, dense_mask), t(2, 0xffff & ~dense_mask)))
    for nrepl in (3, n_dense):
        for log2r in (2, 4, 6):
            print('replicas %2d of levels < %d: reduce %.3f (dense only %.3f)  both %.3f' % (1 << log2r, nrepl, t(2, 0, log2r, nrepl), t(2, dense_mask, log2r, nrepl),
                                                                                            t(3, 0, log2r, nrepl)), flush=True)


if __name__ == '__main__':
    main()
