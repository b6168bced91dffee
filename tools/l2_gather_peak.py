"""Measured L2 gather / reduction ceiling at the fused kernels' access pattern.

    python tools/l2_gather_peak.py [--rays 49152] [--out gpurun_out/l2_gather_peak.json]

For each table (room0: 2^16 entries per hashed level, 6.7 MB; the ScanNet base file's 2^19: 34.7 MB) the corner list of
real sample points (rays of the bench workload, all samples) is written once by cb200_hash_indices; cb200_debug_gather_peak
then replays it -- lane pair per point, 64 LDG.64 and / or 64 RED.v2.f32 per thread, no other arithmetic -- and the time per
launch gives GB/s of table bytes (8 B per corner, 1 024 B per point and pass).  `uniform` repeats it for points drawn
uniformly in the box (no ray coherence: the worst case)."""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from coslam_b200 import _lib, ops, synth          # noqa: E402
from coslam_b200.scene_rep import JointEncoding   # noqa: E402

DEV = 'cuda:0'


def corner_list(model, x):
    n = x.shape[0]
    idx = torch.empty((n, 16, 8), dtype=torch.int32, device=DEV)
    _lib.check(_lib.load().cb200_hash_indices(ctypes.byref(model.spec.grid()), _lib.ptr(x), ctypes.c_int64(n), _lib.ptr(idx),
                                              _lib.stream()), 'hash_indices')
    return idx


def time_mode(model, idx, table, dtable, mode, iters=20, warmup=5):
    lib = _lib.load()
    sink = torch.zeros(1, device=DEV)
    n = idx.shape[0]

    def run():
        _lib.check(lib.cb200_debug_gather_peak(ctypes.byref(model.spec.grid()), _lib.ptr(idx), ctypes.c_int64(n), _lib.ptr(table),
                                               _lib.ptr(dtable), mode, _lib.ptr(sink), _lib.stream()), 'gather_peak')
    for _ in range(warmup):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--rays', type=int, default=49152)
    ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'l2_gather_peak.json'))
    args = ap.parse_args()
    out = {'gpu': torch.cuda.get_device_name(0), 'rays': args.rays, 'cases': []}
    for name, cfg in (('room0_2^16', synth.room0_config()), ('scannet0000_2^19', synth.scannet0000_config(hash_size=19))):
        model = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
        table = model.embed_fn.params.detach()
        dtable = torch.zeros_like(table)
        rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, args.rays, seed=0)
        z = ops.sample_rays(model.spec, depth.to(DEV), perturb=False)
        n, s = z.shape
        pts = torch.empty((n * s, 3), device=DEV)
        x = torch.empty((n * s, 3), device=DEV)
        o, d = rays_o.to(DEV).contiguous(), rays_d.to(DEV).contiguous()
        _lib.check(_lib.load().cb200_debug_positions(ctypes.byref(model.spec.scene()), _lib.ptr(o), _lib.ptr(d), _lib.ptr(z),
                                                     ctypes.c_int64(n), ctypes.c_int64(s), _lib.ptr(pts), _lib.ptr(x), _lib.stream()),
                   'debug_positions')
        for pattern, xs in (('rays', x), ('uniform', torch.rand_like(x))):
            idx = corner_list(model, xs.contiguous())
            P = idx.shape[0]
            case = {'table': name, 'table_mb': table.numel() * 4 / 1e6, 'pattern': pattern, 'points': P}
            for mode, key, passes in ((1, 'gather', 1), (2, 'reduce', 1), (3, 'gather+reduce', 2)):
                ms = time_mode(model, idx, table, dtable, mode)
                case[key] = {'ms': ms, 'table_GBps': P * 1024.0 * passes / (ms * 1e-3) / 1e9,
                             'with_index_stream_GBps': P * (1024.0 * passes + 512.0) / (ms * 1e-3) / 1e9,
                             'points_per_us': P / (ms * 1e3)}
            out['cases'].append(case)
            print(json.dumps(case), flush=True)
            del idx
        del model, table, dtable
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, 'w') as f:
        json.dump(out, f, indent=1)


if __name__ == '__main__':
    main()
