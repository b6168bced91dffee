"""Times the per-point kernels alone (CUDA events, 2^18 rays x 43 samples) with outputs switched off one at a time, to
attribute the backward's time to gathers / table scatter / weight gradients.   python tools/time_kernels.py [n_rays]"""
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


raw, _, mask = ops.points_forward(scene, tb, w, rays=(o, d, z))
g = torch.randn_like(raw)
print('points %d' % (n * z.shape[1]))
print('fwd                     %.3f ms' % timed(lambda: ops.points_forward(scene, tb, w, rays=(o, d, z))))
print('fwd sdf_only            %.3f ms' % timed(lambda: ops.points_forward(scene, tb, w, rays=(o, d, z), sdf_only=True)))
for name, kw in (('bwd full', {}), ('bwd no table grad', dict(need_table=False)), ('bwd no mlp grad', dict(need_mlp=False)),
                 ('bwd no pos grad', dict(need_pos=False)), ('bwd pos only (tracking)', dict(need_table=False, need_mlp=False))):
    print('%-23s %.3f ms' % (name, timed(lambda: ops.points_backward(scene, tb, w, mask, g, rays=(o, d, z), **kw))))

# stand-alone HashGrid encoding (tiny-cuda-nn style thread per (point, level)) on the same points, for reference
pts = (o[:, None, :] + d[:, None, :] * z[..., None]).reshape(-1, 3)
bound = synth.bound_tensor(cfg).to(DEV)
xn = ((pts - bound[:, 0]) / (bound[:, 1] - bound[:, 0])).contiguous()
enc = model.embed_fn
with torch.no_grad():
    print('hash_fwd stand-alone    %.3f ms' % timed(lambda: enc(xn)))
out = enc(xn)
gy = torch.randn_like(out)
print('hash_bwd stand-alone    %.3f ms (table gradient only)' % timed(lambda: torch.autograd.grad(enc(xn), enc.params, gy)[0]) )
