"""Per-phase cycle counts of the tensor-core forward (clock64 inside the kernel, printed by two threads of two CTAs):
    CB200_EXTRA_NVCC_FLAGS="-DCB200_PHASE_CLOCKS" python tools/phase_probe.py
(the flag changes the build stamp, so the library is rebuilt with the instrumentation; without it nothing is printed).
-DCB200_EXP_COALESCED additionally folds every gather of a warp into one or two lines (same arithmetic)."""
import os, sys
sys.path.insert(0, '.')
import torch
from coslam_b200 import synth, ops
from tests import util
from coslam_b200.scene_rep import JointEncoding
DEV='cuda:0'
cfg = synth.room0_config()
n = 1 << 18
rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n, seed=11)
weights = synth.decoder_init(seed=5)
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), 'small', seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
z = ops.sample_rays(model.spec, depth.to(DEV), noise=None, seed=1)
o, d = rays_o.to(DEV), rays_d.to(DEV)
tb = model.embed_fn.params.detach()
w = tuple(x.detach() for x in model.decoder.fused_weights())
for _ in range(2):
    ops.points_forward(model.spec.scene(), tb, w, rays=(o, d, z))
    torch.cuda.synchronize()
