"""Per-point kernels at small sizes: fixed cost of a launch and wave quantisation (tiles of 128 points, 296 resident CTAs).
   python tools/small_probe.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import synth, ops
from tests import util

DEV = 'cuda:0'
cfg = synth.room0_config()
weights = synth.decoder_init(seed=5)
from coslam_b200.scene_rep import JointEncoding
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), 'small', seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
tb = model.embed_fn.params.detach()
w = tuple(x.detach() for x in model.decoder.fused_weights())
scene = model.spec.scene()


def timed(fn, reps=50):
    g = torch.cuda.CUDAGraph()
    fn(); fn()
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for tiles in (1, 148, 296, 344, 592, 774, 888):
    n = tiles * 128
    pts = torch.rand(n, 3, device=DEV)
    raw, _, mask = ops.points_forward(scene, tb, w, pts=pts)
    g = torch.randn_like(raw)
    f = timed(lambda: ops.points_forward(scene, tb, w, pts=pts))
    b = timed(lambda: ops.points_backward(scene, tb, w, mask, g, pts=pts))
    b2 = timed(lambda: ops.points_backward(scene, tb, w, mask, g, pts=pts, need_table=False, need_mlp=False))
    print('%4d tiles (%6d points): fwd %6.1f us   bwd full %6.1f us   bwd pos-only %6.1f us' % (tiles, n, f, b, b2))
