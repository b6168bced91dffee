"""Shared helpers of the parity tests: rebuild a golden case's inputs from its seeds, build the oracle and the
CUDA model with identical weights, and the tolerance helper for the 1e-3 relative gate of BASELINE.md section 5."""
import ast
import hashlib
import os

import numpy as np
import torch

from coslam_b200 import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
GOLDEN_CASES = ['room0_small', 'room0_trained', 'scannet0000_trained', 'tum_desk_trained']
# hierarchical pass through the reference's own sample_pdf / second pass (row a11): 128 and 8 importance samples, with the
# deterministic uniforms of perturb 0 (`_det`) and with the CPU generator's draws of perturb 1
IMPORTANCE_CASES = ['tum_desk_imp128_det', 'tum_desk_imp128', 'tum_desk_imp8_det', 'tum_desk_imp8']
# BASELINE config 1's size (4096 rays, ~176 k points); per-point tensors stored as checksums + every 16th ray
LARGE_CASES = ['room0_4096']
# `oneGrid: False` (second hash grid for colour, ColorSDFNet): SURVEY 8(f) rank 4, through the unmodified reference
TWOGRID_CASES = ['room0_twogrid']

REL_TOL = 1e-3          # north_star: rendered depth/colour, SDF and gradients within 1e-3 relative (fp32 tables)


def checksum(t):
    return hashlib.sha256(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()).hexdigest()[:16]


def load_golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + '.npz'), allow_pickle=False)


def case_inputs(g):
    """Regenerates the seeded inputs of a golden case and verifies them against the stored checksums."""
    cfg = synth.CONFIGS[str(g['cfg_name'])](**ast.literal_eval(str(g['cfg_kw'])))
    n = int(g['n_rays'])
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n, seed=int(g['seed']), **ast.literal_eval(str(g['batch_kw'])))
    two_grid = not cfg['grid']['oneGrid']
    weights = synth.decoder_init(seed=int(g['seed']), col_in=95 if two_grid else 63)
    from oracle import tcnn_shim
    from oracle.coslam_oracle import hashgrid_encoding_config
    enc = hashgrid_encoding_config(cfg, synth.bound_tensor(cfg))
    meta = tcnn_shim.grid_meta(enc['n_levels'], enc['log2_hashmap_size'], enc['base_resolution'], enc['per_level_scale'])
    table = synth.table_init(meta['offset'][-1] * 2, str(g['table_kind']))
    assert checksum(rays_o) == str(g['chk_rays_o']) and checksum(rays_d) == str(g['chk_rays_d'])
    assert checksum(rgb) == str(g['chk_rgb']) and checksum(depth) == str(g['chk_depth'])
    assert checksum(table) == str(g['chk_table'])
    assert checksum(torch.cat([w.reshape(-1) for w in weights])) == str(g['chk_w'])
    return cfg, (rays_o, rays_d, rgb, depth), table, weights, case_noise(g, cfg)[0]


def case_noise(g, cfg):
    """(noise [N,S] or None, u_imp [N,I] or None): the CPU generator's draws of the reference's forward in its own order
    (scene_rep.py:255, then model/utils.py:53).  Stored for the small cases, regenerated from the seed otherwise."""
    t = cfg['training']
    if not t['perturb'] > 0.:
        return None, None
    n, S, I = int(g['n_rays']), t['n_samples_d'] + t['n_range_d'], t['n_importance']
    torch.manual_seed(int(g['noise_seed']))
    noise = torch.rand((n, S))
    u = torch.rand((n, I)) if I > 0 else None
    if g['noise'].shape[0]:
        assert np.array_equal(noise.numpy(), g['noise'])
    if u is not None:
        assert np.array_equal(u.numpy(), g['u_imp'])
    return noise, u


def color_table(g, cfg):
    """The colour grid's table of a two-grid golden case (seeded, checksum verified)."""
    from oracle import tcnn_shim
    from oracle.coslam_oracle import hashgrid_encoding_config
    enc = hashgrid_encoding_config(cfg, synth.bound_tensor(cfg), voxel_key='voxel_color')
    meta = tcnn_shim.grid_meta(enc['n_levels'], enc['log2_hashmap_size'], enc['base_resolution'], enc['per_level_scale'])
    assert np.array_equal(np.asarray(meta['offset'], dtype=np.uint32), g['color_grid_offset'])
    t = synth.table_init(meta['offset'][-1] * 2, str(g['table_kind']), seed=int(g['color_table_seed']))
    assert checksum(t) == str(g['chk_table_color'])
    return t


def make_oracle(cfg, table, weights, table_color=None):
    from oracle.coslam_oracle import OracleScene
    orc = OracleScene(cfg, synth.bound_tensor(cfg))
    orc.load_weights(table, *weights, table_color=table_color)
    return orc


def make_cuda_model(cfg, table, weights, device='cuda:0', noise_mode='host'):
    from coslam_b200.scene_rep import JointEncoding
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(device)).to(device)
    model.noise_mode = noise_mode
    with torch.no_grad():
        model.embed_fn.params.copy_(table.to(device))
        for dst, src in zip(model.decoder.fused_weights(), weights):
            dst.copy_(src.to(device))
    model.train()
    return model


def rel_err(a, b):
    """max |a - b| relative to the scale (max |b|) of the reference tensor."""
    a = torch.as_tensor(a).detach().double().cpu()
    b = torch.as_tensor(b).detach().double().cpu()
    scale = max(b.abs().max().item(), 1e-30)
    return (a - b).abs().max().item() / scale


def assert_close(a, b, what, tol=REL_TOL):
    e = rel_err(a, b)
    assert e <= tol, '%s: relative error %.3e exceeds %.1e' % (what, e, tol)
    return e


def unpack_masks(mask, evaluated=None, oracle_pre=None):
    """[P,2] int32 ReLU sign words (hidden layer of the SDF net, hidden layer of the colour net) -> two bool [P,32].
    Rows the lazy forward never evaluated (`evaluated` [P] bool) hold no sign bits: they take the oracle's own pattern."""
    m = mask.detach().cpu().to(torch.int64) & 0xFFFFFFFF
    bits = torch.arange(32, dtype=torch.int64)
    out = [((m[:, 0:1] >> bits) & 1).bool(), ((m[:, 1:2] >> bits) & 1).bool()]
    if evaluated is not None:
        ev = evaluated.detach().cpu()[:, None]
        out = [torch.where(ev, o, h > 0) for o, h in zip(out, oracle_pre)]
    return tuple(out)


def audit_activation_flips(cuda_masks, oracle_pre, max_frac=2e-5, band=1e-4):
    """The CUDA forward and the oracle may disagree on the sign of a pre-activation only where it is ~0 (within `band` of
    the layer's largest magnitude), and only on a vanishing fraction of units."""
    n_flip = 0
    for m, h in zip(cuda_masks, oracle_pre):
        diff = m != (h > 0)
        n_flip += int(diff.sum())
        if diff.any():
            worst = float(h[diff].abs().max() / h.abs().max())
            assert worst <= band, 'activation sign differs at a pre-activation of relative size %.2e' % worst
    total = sum(m.numel() for m in cuda_masks)
    assert n_flip <= max(2, max_frac * total), '%d of %d activation signs differ' % (n_flip, total)
    return n_flip
