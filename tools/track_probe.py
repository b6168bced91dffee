"""One tracked frame (1024 rays x 10 iterations, room0) for profiling: `ncu --metrics gpu__time_duration.sum python tools/track_probe.py`
lists the kernels of a frame; without ncu it prints ms/frame of the fused / autograd loops, eager and as a CUDA graph."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import synth
from coslam_b200.tracking import FrameTracker
from tests import util

DEV = 'cuda:0'
name = sys.argv[1] if len(sys.argv) > 1 else 'room0'
mode = sys.argv[2] if len(sys.argv) > 2 else 'time'
cfg = synth.CONFIGS[name]()
n = cfg['tracking']['sample']
weights = synth.decoder_init(seed=5)
from coslam_b200.scene_rep import JointEncoding
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), 'trained', seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
from tests.test_parity_gpu import _tracking_case
dir_cam, rgb, depth, c2w0 = _tracking_case(cfg, n)
if mode == 'ncu':                       # one eager fused frame (a graph replay shows up as one opaque launch under ncu)
    tr = FrameTracker(model, cfg, device=DEV, use_graph=False)
    tr.track(dir_cam, rgb, depth, c2w0)
    torch.cuda.synchronize()
    sys.exit(0)
for fused in (True, False):
    for graph in (False, True):
        tr = FrameTracker(model, cfg, device=DEV, use_graph=graph, fused=fused)
        for _ in range(3):
            tr.track(dir_cam, rgb, depth, c2w0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 20
        for _ in range(reps):
            est, info = tr.track(dir_cam, rgb, depth, c2w0)
            est.cpu()
        torch.cuda.synchronize()
        print('%s fused=%d graph=%d: %.3f ms/frame (best loss %.6f)' % (name, fused, graph, (time.perf_counter() - t0) / reps * 1e3,
                                                                        float(info['best_loss'])))
