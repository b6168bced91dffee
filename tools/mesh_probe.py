"""Lattice SDF query of extract_mesh at room0's voxel_eval (0.05) and voxel_final (0.02): one device-side pass
(coslam_b200.meshing) vs the reference's 65 536-point chunk loop with a host copy per chunk.   python tools/mesh_probe.py"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import meshing, synth
from tests import util

DEV = 'cuda:0'
cfg = synth.room0_config()
weights = synth.decoder_init(seed=5)
from coslam_b200.scene_rep import JointEncoding
m0 = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
table = synth.table_init(m0.embed_fn.params.numel(), 'trained', seed=77)
model = util.make_cuda_model(cfg, table, weights, DEV)
model.eval()
bound = synth.bound_tensor(cfg).to(DEV)
for voxel in (0.05, 0.02):
    for _ in range(2):
        sdf, _ = meshing.query_sdf_lattice(model, cfg, bound, voxel_size=voxel)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sdf, _ = meshing.query_sdf_lattice(model, cfg, bound, voxel_size=voxel)
    host = sdf.cpu()
    t_ours = time.perf_counter() - t0
    # the reference's loop (utils.py:96-131): CPU lattice, 65 536-point chunks, host copy per chunk
    t0 = time.perf_counter()
    tx, ty, tz = meshing.get_voxels(bound[0, 1], bound[0, 0], bound[1, 1], bound[1, 0], bound[2, 1], bound[2, 0], voxel)
    flat = torch.stack(torch.meshgrid(tx, ty, tz, indexing='ij'), -1).to(torch.float32).reshape(-1, 3)
    bb = bound.cpu()
    flat = (flat - bb[:, 0]) / (bb[:, 1] - bb[:, 0])
    with torch.no_grad():
        raw = [model.query_sdf(flat[i:i + 65536, None, :].to(DEV)).cpu().data.numpy() for i in range(0, flat.shape[0], 65536)]
    t_loop = time.perf_counter() - t0
    print('voxel %.2f: %d x %d x %d = %.2f M points: one pass %.1f ms (%.2f G points/s incl. lattice build and the copy of the volume to the host), '
          'chunk loop through the same kernels %.1f ms' % (voxel, sdf.shape[0], sdf.shape[1], sdf.shape[2], sdf.numel() / 1e6, t_ours * 1e3,
                                                            sdf.numel() / t_ours / 1e9, t_loop * 1e3))
