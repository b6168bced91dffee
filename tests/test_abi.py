"""CPU tests: the C-ABI library builds/loads, exports every symbol include/coslam_b200.h declares, its host-side
geometry matches the oracle, and the Python boundary mirrors the reference's module structure (state_dict keys)."""
import os
import re

import numpy as np
import pytest
import torch

from coslam_b200 import _lib, synth
from oracle import tcnn_shim
from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'coslam_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(cb200_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), 'libcoslam_b200.so does not export %s' % n
    assert sorted(_lib.EXPORTS) == names
    assert lib.cb200_abi_version() == _lib.ABI_VERSION


@pytest.mark.parametrize('desired,log2T', [(400, 16), (225, 16), (225, 19), (325, 16), (512, 19), (2048, 19), (64, 14)])
def test_grid_meta_matches_oracle(desired, log2T):
    pls = np.exp2(np.log2(desired / 16) / 15)
    g = _lib.grid_meta(16, 2, log2T, 16, pls)
    meta = tcnn_shim.grid_meta(16, log2T, 16, pls)
    assert [g.res[i] for i in range(16)] == meta['res']
    assert [g.offset[i] for i in range(17)] == meta['offset']
    assert np.array_equal(np.array([g.scale[i] for i in range(16)], dtype=np.float32), np.asarray(meta['scale'], dtype=np.float32))
    for l in range(16):
        dense = meta['res'][l] ** 3 <= meta['size'][l]
        assert g.hashed[l] == (0 if dense else 1)


def test_argument_errors_are_reported():
    lib = _lib.load()
    import ctypes
    cfg = _lib.GridCfg(17, 2, 16, 16, 1.5)
    g = _lib.Grid()
    rc = lib.cb200_grid_meta(ctypes.byref(cfg), ctypes.byref(g))
    assert rc < 0 and b'n_levels' in lib.cb200_last_error()


def test_module_structure_and_state_dict_keys():
    from coslam_b200.scene_rep import JointEncoding
    cfg = synth.room0_config()
    m = JointEncoding(cfg, synth.bound_tensor(cfg))
    g = util.load_golden('room0_small')
    assert sorted(m.state_dict().keys()) == [str(k) for k in g['state_dict_keys']]     # same keys as the reference
    assert m.embed_fn.params.numel() == 1666576 and m.embed_fn.n_output_dims == 32
    assert m.embedpos_fn.n_output_dims == 48 and m.embedpos_fn.params.numel() == 0
    assert sum(p.numel() for p in m.decoder.parameters()) == 5184
    assert m.resolution_sdf == 400
    # optimiser groups of coslam.py:615-622 resolve
    assert len(list(m.decoder.parameters())) == 4 and len(list(m.embed_fn.parameters())) == 1
    # survives the mp plumbing of coslam_mp.py:52 / mp_slam/tracker.py:27
    import copy
    import pickle
    m.share_memory()
    m2 = copy.deepcopy(m)
    m3 = pickle.loads(pickle.dumps(m))
    assert torch.equal(m2.embed_fn.params, m.embed_fn.params) and torch.equal(m3.embed_fn.params, m.embed_fn.params)
    assert m3.spec.n_table_params() == 1666576


def test_no_cpu_fallback():
    from coslam_b200.scene_rep import JointEncoding
    cfg = synth.room0_config()
    m = JointEncoding(cfg, synth.bound_tensor(cfg))
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, 8)
    with pytest.raises(RuntimeError, match='no CPU path'):
        m.forward(rays_o, rays_d, rgb, depth)
    with pytest.raises(RuntimeError, match='no CPU path'):
        m.embed_fn(torch.rand(4, 3))
    from coslam_b200 import tcnn
    with pytest.raises(NotImplementedError):
        tcnn.Network(3, 3, {'otype': 'FullyFusedMLP', 'activation': 'Sigmoid', 'n_neurons': 32, 'n_hidden_layers': 1})
    net = tcnn.Network(3, 3, {'otype': 'FullyFusedMLP', 'activation': 'ReLU', 'output_activation': 'None', 'n_neurons': 32,
                              'n_hidden_layers': 1})
    assert net.params.numel() == 32 * 16 + 16 * 32
    with pytest.raises(RuntimeError, match='no CPU path'):
        net(torch.rand(4, 3))
