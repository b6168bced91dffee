"""SURVEY.md 8(f) rank 3 -- the inference path of mesh extraction (coslam_b200/meshing.py) against the REFERENCE's own
utils.py:29-54 (getVoxels) and :80-135 (extract_mesh's lattice, normalisation and 65 536-point chunk loop).

CPU: `get_voxels` and the lattice points `query_sdf_lattice` hands to the model, against the committed checksums the unmodified
reference produced (tests/golden/mesh_lattice.json, oracle/gen_golden.py) and -- where a reference tree is present -- against the
reference run live.  GPU: the reference's extract_mesh (from baseline/_ref, marching cubes stubbed to capture its input) drives
the CUDA model's query_sdf; the captured SDF volume must equal query_sdf_lattice's bit for bit."""
import json
import os

import numpy as np
import pytest
import torch

from coslam_b200 import meshing, synth
from oracle import ref_loader
from tests import util

GOLDEN = json.load(open(os.path.join(util.GOLDEN_DIR, 'mesh_lattice.json')))


class _RecordingModel(object):
    """Stands for JointEncoding.query_sdf: records the points it is asked for."""

    def __init__(self):
        self.seen = []

    def query_sdf(self, pts):
        self.seen.append(pts.reshape(-1, 3).clone())
        return torch.zeros(pts.shape[:-1], device=pts.device)


def _our_lattice(cfg, voxel, device='cpu', chunk=1 << 22):
    bound = synth.bound_tensor(cfg).to(device)
    rec = _RecordingModel()
    sdf, (tx, ty, tz) = meshing.query_sdf_lattice(rec, cfg, bound, marching_cube_bound=bound, voxel_size=voxel, chunk=chunk)
    return sdf, tx, ty, tz, torch.cat(rec.seen, 0)


@pytest.mark.parametrize('name', sorted(GOLDEN))
def test_lattice_matches_reference_golden(name):
    g = GOLDEN[name]
    cfg = synth.CONFIGS[g['cfg_name']]()
    sdf, tx, ty, tz, flat = _our_lattice(cfg, g['voxel_size'], chunk=100000)      # ragged chunks on purpose
    assert list(sdf.shape) == g['shape'] == [len(tx), len(ty), len(tz)]
    assert util.checksum(tx) == g['chk_tx'] and util.checksum(ty) == g['chk_ty'] and util.checksum(tz) == g['chk_tz']
    assert util.checksum(flat) == g['chk_points']
    assert [float(v) for v in flat[0]] == g['first_point'] and [float(v) for v in flat[-1]] == g['last_point']


@pytest.mark.skipif(not ref_loader.available(), reason='no reference tree (build container / baseline/_ref)')
def test_get_voxels_matches_reference_live():
    from oracle import ref_env
    ref_utils = ref_env.import_reference('utils')
    rng = np.random.RandomState(0)
    for _ in range(50):
        lo = rng.uniform(-5, 0, 3)
        hi = lo + rng.uniform(0.5, 9, 3)
        voxel = float(rng.choice([0.02, 0.04, 0.05, 0.1, 0.013]))
        a = ref_utils.getVoxels(hi[0], lo[0], hi[1], lo[1], hi[2], lo[2], voxel_size=voxel)
        b = meshing.get_voxels(hi[0], lo[0], hi[1], lo[1], hi[2], lo[2], voxel_size=voxel)
        assert all(torch.equal(x, y) for x, y in zip(a, b))
        a = ref_utils.getVoxels(torch.tensor(hi[0]), torch.tensor(lo[0]), hi[1], lo[1], hi[2], lo[2], resolution=37)
        b = meshing.get_voxels(torch.tensor(hi[0]), torch.tensor(lo[0]), hi[1], lo[1], hi[2], lo[2], resolution=37)
        assert all(torch.equal(x, y) for x, y in zip(a, b))


@pytest.mark.gpu
@pytest.mark.skipif(not ref_loader.available(), reason='baseline/_ref not installed (python -c "import __graft_entry__ as g; g.build()" in the build container)')
@pytest.mark.parametrize('name', ['room0_coarse', 'scannet0000_eval'])
def test_reference_extract_mesh_drives_the_cuda_model(name, tmp_path):
    """utils.py:80-135 unmodified, query_fn = the drop-in model's query_sdf on the GPU, color_func = its query_color."""
    from oracle import ref_env
    ref_utils = ref_env.import_reference('utils')
    g = GOLDEN[name]
    cfg = synth.CONFIGS[g['cfg_name']]()
    dev = 'cuda:0'
    from coslam_b200.scene_rep import JointEncoding
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(dev)).to(dev)
    with torch.no_grad():
        model.embed_fn.params.copy_(synth.table_init(model.embed_fn.params.numel(), 'trained', seed=12).to(dev))
    model.eval()
    bound = synth.bound_tensor(cfg).to(dev)
    del ref_env.LAST_VOLUME[:]
    ref_utils.extract_mesh(model.query_sdf, dict(cfg, mesh={'render_color': False}), bound, marching_cube_bound=bound,
                           color_func=model.query_color, voxel_size=g['voxel_size'], mesh_savepath=str(tmp_path / 'mesh.ply'))
    volume = ref_env.LAST_VOLUME[-1]
    assert list(volume.shape) == g['shape']
    sdf, _ = meshing.query_sdf_lattice(model, cfg, bound, marching_cube_bound=bound, voxel_size=g['voxel_size'])
    assert np.array_equal(sdf.cpu().numpy(), volume)            # same points, row-independent kernels: bit-identical
    # on the device the lattice is the reference's too (IEEE fp32 subtract / divide)
    _, _, _, _, flat = _our_lattice(cfg, g['voxel_size'], device=dev)
    assert util.checksum(flat) == g['chk_points']
