import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import synth, _lib, ops
from tests import util
DEV='cuda:0'
cfg = synth.room0_config()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
kind = sys.argv[2] if len(sys.argv) > 2 else 'small'
rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n, seed=11)
weights = synth.decoder_init(seed=5)
from coslam_b200.scene_rep import JointEncoding
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), kind, seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
torch.manual_seed(3); noise = torch.rand(n, 43)
z = ops.sample_rays(model.spec, depth.to(DEV), noise=noise.to(DEV))
o, d = rays_o.to(DEV), rays_d.to(DEV)
tb = model.embed_fn.params.detach(); w = tuple(x.detach() for x in model.decoder.fused_weights())
res = {}
for k in ('simt', 'tc'):
    _lib.set_decoder(k)
    raw, _, mask = ops.points_forward(model.spec.scene(), tb, w, rays=(o, d, z))
    res[k] = (raw.clone(), mask.clone())
torch.cuda.synchronize()
r0, m0_ = res['simt']; r1, m1_ = res['tc']
print('raw max abs diff', float((r0 - r1).abs().max()), 'scale', float(r0.abs().max()))
x = (m0_ ^ m1_)
print('mask words differing', int((x != 0).sum()), 'of', x.numel())
bad = (x != 0).any(1).nonzero().reshape(-1)
print('first bad points', bad[:20].tolist(), 'tiles', (bad[:20] // 128).tolist(), 'rows', (bad[:20] % 128).tolist())
g = torch.randn_like(r0)
outs = {}
for k in ('simt', 'tc'):
    _lib.set_decoder('simt')
    dt, dw, (do, dd) = ops.points_backward(model.spec.scene(), tb, w, res[k][1], g, rays=(o, d, z))
    outs[k] = (do.clone(), dd.clone(), dt.clone())
print('d_o rel diff from mask difference only', float((outs['simt'][0] - outs['tc'][0]).abs().max() / outs['simt'][0].abs().max()))
