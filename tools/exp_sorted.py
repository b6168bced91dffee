"""Experiment: does a spatially sorted point list cut the per-point kernels' time?  The hash leaves x un-multiplied, so
points in the same (y, z) cell with neighbouring x cells read the same 128-byte lines on every hashed level.
    python tools/exp_sorted.py [n_rays]"""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import synth, ops
from tests import util

DEV = 'cuda:0'
cfg = synth.room0_config()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n, seed=11)
weights = synth.decoder_init(seed=5)
from coslam_b200.scene_rep import JointEncoding
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), 'small', seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
z = ops.sample_rays(model.spec, depth.to(DEV), noise=None, seed=1)
o, d = rays_o.to(DEV), rays_d.to(DEV)
tb = model.embed_fn.params.detach()
w = tuple(x.detach() for x in model.decoder.fused_weights())
scene = model.spec.scene()
P = n * z.shape[1]


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


pts = (o[:, None, :] + d[:, None, :] * z[..., None]).reshape(-1, 3)
bound = synth.bound_tensor(cfg).to(DEV)
xn = ((pts - bound[:, 0]) / (bound[:, 1] - bound[:, 0]))
raw = torch.zeros((P, 4), device=DEV)
mask = torch.empty((P, 2), dtype=torch.int32, device=DEV)
count = torch.tensor([P], dtype=torch.int64, device=DEV)


def run(lst):
    ops.points_forward(scene, tb, w, rays=(o, d, z), out=(raw, mask),
                       point_list=(ctypes.c_void_p(lst.data_ptr()), ctypes.c_void_p(count.data_ptr())))


ident = torch.arange(P, dtype=torch.int32, device=DEV)
print('points', P)
print('no list            %.3f ms' % timed(lambda: ops.points_forward(scene, tb, w, rays=(o, d, z), out=(raw, mask))))
print('identity list      %.3f ms' % timed(lambda: run(ident)))
print('random permutation %.3f ms' % timed(lambda: run(torch.randperm(P, device=DEV).to(torch.int32)) if False else run(perm)) if (perm := torch.randperm(P, device=DEV).to(torch.int32)) is not None else '')
for res in (400, 200, 100, 50):
    c = torch.floor(xn * (res - 1) + 0.5).clamp(0, 1023).to(torch.int64)
    for name, key in (('zyx', (c[:, 2] << 20) | (c[:, 1] << 10) | c[:, 0]),):
        t0 = timed(lambda: torch.sort(key), reps=2)
        lst = torch.sort(key)[1].to(torch.int32)
        print('sorted res %3d %s  %.3f ms   (torch.sort of int64 keys %.3f ms)' % (res, name, timed(lambda: run(lst)), t0))
ref = raw.clone()
ops.points_forward(scene, tb, w, rays=(o, d, z), out=(raw, mask))
print('max diff vs unsorted', float((ref - raw).abs().max()))
