"""ORACLE tooling (test infrastructure) -- golden-vector generator.

Runs the reference's OWN, UNMODIFIED model/scene_rep.py:JointEncoding (imported from /root/reference, with
oracle/tcnn_shim.py standing in for the un-vendored tinycudann) on the seeded synthetic batches of
coslam_b200/synth.py, checks that the restatement oracle/coslam_oracle.py reproduces it, and writes the fixtures under
tests/golden/.  Works only where /root/reference exists (the build container); the fixtures travel with the repo.

    python -m oracle.gen_golden [case names ...]
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from coslam_b200 import synth                      # noqa: E402
from oracle import coslam_oracle, ref_loader, tcnn_shim   # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, 'tests', 'golden')
NOISE_SEED = 4321

CASES = [
    # name, config factory, kwargs, n_rays, table kind, batch seed, extra batch kwargs
    ('room0_small', 'room0', {}, 128, 'small', 0, {}),
    ('room0_trained', 'room0', {}, 128, 'trained', 1, {}),
    ('scannet0000_trained', 'scannet0000', {}, 64, 'trained', 2, {}),
    ('tum_desk_trained', 'tum_desk', {'n_importance': 0}, 64, 'trained', 3, {'over_trunc_frac': 0.05}),
    # hierarchical pass (this fork's configs/Tum/tum.yaml:89 ships n_importance 128): the reference's
    # own sample_pdf (model/utils.py:32-71) and second pass (scene_rep.py:262-295, 327-332), with the deterministic
    # uniforms of perturb 0 and with the CPU generator's draws of perturb 1
    ('tum_desk_imp128_det', 'tum_desk', {'n_importance': 128, 'perturb': 0}, 64, 'trained', 5, {'over_trunc_frac': 0.05}),
    ('tum_desk_imp128', 'tum_desk', {'n_importance': 128, 'perturb': 1}, 64, 'trained', 6, {'over_trunc_frac': 0.05}),
    ('tum_desk_imp8_det', 'tum_desk', {'n_importance': 8, 'perturb': 0}, 96, 'trained', 7, {'over_trunc_frac': 0.05}),
    ('tum_desk_imp8', 'tum_desk', {'n_importance': 8, 'perturb': 1}, 96, 'trained', 8, {'over_trunc_frac': 0.05}),
    # BASELINE config 1's size: the direct (no activation-pattern injection) 1e-3 gradient gate over ~176 k points.
    # Per-point tensors are stored as checksums / strided samples to keep the fixture small.
    ('room0_4096', 'room0', {}, 4096, 'trained', 4, {}),
    # two-grid variant: `oneGrid: False`, second hash grid at voxel_color, ColorSDFNet (decoder.py:110-143)
    ('room0_twogrid', 'room0_twogrid', {}, 128, 'trained', 9, {}),
]
LARGE = {'room0_4096'}


def checksum(t):
    return hashlib.sha256(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()).hexdigest()[:16]


COLOR_TABLE_SEED = 1338


def build_reference_model(cfg, table, weights, table_color=None):
    JointEncoding = ref_loader.load_reference_model_class()
    model = JointEncoding(cfg, synth.bound_tensor(cfg))
    with torch.no_grad():
        model.embed_fn.params.copy_(table)
        if table_color is not None:
            model.embed_fn_color.params.copy_(table_color)
        model.decoder.sdf_net.model[0].weight.copy_(weights[0])
        model.decoder.sdf_net.model[2].weight.copy_(weights[1])
        model.decoder.color_net.model[0].weight.copy_(weights[2])
        model.decoder.color_net.model[2].weight.copy_(weights[3])
    return model


def table_grad_summary(g, meta, n_pick=4096, seed=0):
    g = g.detach()
    F = 2
    level_sum = np.array([float(g[meta['offset'][l] * F:meta['offset'][l + 1] * F].double().sum()) for l in range(len(meta['res']))])
    level_abs = np.array([float(g[meta['offset'][l] * F:meta['offset'][l + 1] * F].double().abs().sum()) for l in range(len(meta['res']))])
    nz = torch.nonzero(g).reshape(-1)
    rng = np.random.RandomState(seed)
    pick = nz[torch.as_tensor(rng.choice(nz.numel(), size=min(n_pick, nz.numel()), replace=False))]
    pick, _ = torch.sort(pick)
    return {'tg_level_sum': level_sum, 'tg_level_abs': level_abs, 'tg_norm': float(g.double().norm()),
            'tg_nnz': int(nz.numel()), 'tg_pick_idx': pick.numpy().astype(np.int64), 'tg_pick_val': g[pick].numpy()}


def run_case(name, cfg_name, cfg_kw, n_rays, table_kind, seed, batch_kw):
    cfg = synth.CONFIGS[cfg_name](**cfg_kw)
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n_rays, seed=seed, **batch_kw)
    two_grid = not cfg['grid']['oneGrid']
    weights = synth.decoder_init(seed=seed, col_in=95 if two_grid else 63)
    # table size comes from the reference model itself
    JointEncoding = ref_loader.load_reference_model_class()
    model = JointEncoding(cfg, synth.bound_tensor(cfg))
    n_params = model.embed_fn.params.numel()
    table = synth.table_init(n_params, table_kind)
    table_color = synth.table_init(model.embed_fn_color.params.numel(), table_kind, seed=COLOR_TABLE_SEED) if two_grid else None
    model = build_reference_model(cfg, table, weights, table_color)
    model.train()

    S = cfg['training']['n_samples_d'] + cfg['training']['n_range_d']
    n_imp = cfg['training']['n_importance']
    perturb = cfg['training']['perturb'] > 0.
    # the CPU generator's draws in the reference's order: scene_rep.py:255 (jitter), then model/utils.py:53 (sample_pdf)
    torch.manual_seed(NOISE_SEED)
    noise = torch.rand((n_rays, S)) if perturb else None
    u_imp = torch.rand((n_rays, n_imp)) if (n_imp > 0 and perturb) else None

    # ---- the reference, training mode: forward + backward of the weighted loss (coslam.py:140-160) -------------
    ro = rays_o.clone().requires_grad_(True)
    rd = rays_d.clone().requires_grad_(True)
    torch.manual_seed(NOISE_SEED)
    ret = model.forward(ro, rd, rgb, depth)
    t = cfg['training']
    loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] + t['sdf_weight'] * ret['sdf_loss'] \
        + t['fs_weight'] * ret['fs_loss']
    loss.backward()
    # ---- the reference, eval mode: the full render dictionary ---------------------------------------------------
    model.eval()
    with torch.no_grad():
        torch.manual_seed(NOISE_SEED)
        ev = model.forward(rays_o, rays_d, rgb, depth)
    model.train()

    # ---- cross-check the restatement against the reference ----------------------------------------------------
    orc = coslam_oracle.OracleScene(cfg, synth.bound_tensor(cfg))
    orc.load_weights(table, *weights, table_color=table_color)
    o2 = rays_o.clone().requires_grad_(True)
    d2 = rays_d.clone().requires_grad_(True)
    r2 = orc.forward(o2, d2, rgb, depth, noise=noise, noise_imp=u_imp)
    orc.total_loss(r2).backward()

    def close(a, b, what, rtol=1e-5, atol=1e-7):
        a, b = a.detach().double(), b.detach().double()
        err = (a - b).abs().max().item() if a.numel() else 0.0
        scale = b.abs().max().item() if b.numel() else 1.0
        assert err <= atol + rtol * scale, '%s: restatement deviates from the reference: %g (scale %g)' % (what, err, scale)

    assert torch.equal(ev['z_vals'], r2['z_vals']), 'z_vals not bit-identical'
    close(r2['raw'], ev['raw'], 'raw')
    close(r2['rgb'], ret['rgb'], 'rgb')
    close(r2['depth'], ret['depth'], 'depth')
    for k in ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss'):
        close(r2[k], ret[k], k)
    close(o2.grad, ro.grad, 'd rays_o', rtol=1e-4)
    close(d2.grad, rd.grad, 'd rays_d', rtol=1e-4)
    close(orc.table.grad, model.embed_fn.params.grad, 'd table', rtol=1e-4)
    close(orc.W1.grad, model.decoder.sdf_net.model[0].weight.grad, 'dW1', rtol=1e-4)
    close(orc.W2.grad, model.decoder.sdf_net.model[2].weight.grad, 'dW2', rtol=1e-4)
    close(orc.W3.grad, model.decoder.color_net.model[0].weight.grad, 'dW3', rtol=1e-4)
    close(orc.W4.grad, model.decoder.color_net.model[2].weight.grad, 'dW4', rtol=1e-4)
    if two_grid:
        close(orc.table_color.grad, model.embed_fn_color.params.grad, 'd colour table', rtol=1e-4)

    # ---- bit-exact gates: positions and hash indices from the reference's arithmetic --------------------------
    z = ev['z_vals']
    pts = rays_o[..., None, :] + rays_d[..., None, :] * z[..., :, None]
    bb = synth.bound_tensor(cfg)
    x = (pts.reshape(-1, 3) - bb[:, 0]) / (bb[:, 1] - bb[:, 0])
    meta = model.embed_fn.impl.meta
    n_idx = min(2048, x.shape[0])
    idx = torch.stack([tcnn_shim.hash_corner_indices(x[:n_idx], meta, l)[0] for l in range(16)], 1)   # [n,16,8]

    if n_imp > 0:
        # the merged samples, the coarse maps and the spread of the drawn depths (scene_rep.py:268-293), eval dictionary
        assert ev['z_vals'].shape == (n_rays, S + n_imp)
        for k in ('rgb0', 'depth0', 'depth_var0', 'acc0', 'z_std'):
            close(r2[k], ev[k], k)
    large = name in LARGE
    stride = 16 if large else 1

    out = {
        'cfg_name': cfg_name, 'cfg_kw': repr(cfg_kw), 'n_rays': n_rays, 'table_kind': table_kind, 'seed': seed,
        'batch_kw': repr(batch_kw), 'noise_seed': NOISE_SEED,
        'chk_rays_o': checksum(rays_o), 'chk_rays_d': checksum(rays_d), 'chk_rgb': checksum(rgb), 'chk_depth': checksum(depth),
        'chk_table': checksum(table), 'chk_w': checksum(torch.cat([w.reshape(-1) for w in weights])),
        # `noise` [N,S] and `u_imp` [N,I] are torch.manual_seed(noise_seed); torch.rand((N,S)); torch.rand((N,I)) on the CPU
        # generator (stored for the small cases; the large case regenerates them)
        'noise': (noise.numpy() if (perturb and not large) else np.zeros((0, S), np.float32)),
        'u_imp': (u_imp.numpy() if u_imp is not None else np.zeros((0, max(n_imp, 1)), np.float32)),
        'chk_z_vals': checksum(z), 'chk_pts': checksum(pts.reshape(-1, 3)), 'chk_x': checksum(x),
        'sample_stride': stride,
        'z_vals': z[::stride].numpy(), 'pts': pts[::stride].reshape(-1, 3).numpy(),
        'x': x.reshape(n_rays, -1, 3)[::stride].reshape(-1, 3).numpy(),
        'hash_idx': idx.numpy().astype(np.uint32),
        'grid_scale': np.asarray(meta['scale'], dtype=np.float32), 'grid_res': np.asarray(meta['res'], dtype=np.uint32),
        'grid_offset': np.asarray(meta['offset'], dtype=np.uint32),
        'raw': ev['raw'][::stride].numpy(), 'rgb': ev['rgb'].numpy(), 'depth': ev['depth'].numpy(), 'depth_var': ev['depth_var'].numpy(),
        'disp_map': ev['disp_map'].numpy(), 'acc_map': ev['acc_map'].numpy(),
        'rgb_loss': ret['rgb_loss'].item(), 'depth_loss': ret['depth_loss'].item(), 'sdf_loss': ret['sdf_loss'].item(),
        'fs_loss': ret['fs_loss'].item(), 'psnr': ret['psnr'].detach().numpy(), 'loss': loss.item(),
        'd_rays_o': ro.grad.numpy(), 'd_rays_d': rd.grad.numpy(),
        'dW1': model.decoder.sdf_net.model[0].weight.grad.numpy(), 'dW2': model.decoder.sdf_net.model[2].weight.grad.numpy(),
        'dW3': model.decoder.color_net.model[0].weight.grad.numpy(), 'dW4': model.decoder.color_net.model[2].weight.grad.numpy(),
        'state_dict_keys': np.array(sorted(model.state_dict().keys())),
    }
    if n_imp > 0:
        out.update({k: ev[k].numpy() for k in ('rgb0', 'depth0', 'depth_var0', 'acc0', 'disp0', 'z_std')})
        # the coarse samples (before the merge), for the resampler's own test
        model.eval()
        cfg0 = dict(cfg, training=dict(cfg['training'], n_importance=0))
        m0 = build_reference_model(cfg0, table, weights)
        m0.eval()
        with torch.no_grad():
            torch.manual_seed(NOISE_SEED)
            ev0 = m0.forward(rays_o, rays_d, rgb, depth)
        w0 = m0.raw2outputs(ev0['raw'], ev0['z_vals'], False)[3]
        out.update({'z_coarse': ev0['z_vals'].numpy(), 'w_coarse': w0.detach().numpy()})
        close(ev0['rgb'], ev['rgb0'], 'coarse rgb of the one-pass model')
    out.update(table_grad_summary(model.embed_fn.params.grad, meta))
    if two_grid:
        cm = model.embed_fn_color.impl.meta
        out.update({'c' + k: v for k, v in table_grad_summary(model.embed_fn_color.params.grad, cm).items()})
        out.update({'chk_table_color': checksum(table_color), 'color_table_seed': COLOR_TABLE_SEED,
                    'color_grid_offset': np.asarray(cm['offset'], dtype=np.uint32), 'color_grid_res': np.asarray(cm['res'], dtype=np.uint32)})
    path = os.path.join(GOLDEN_DIR, name + '.npz')
    np.savez_compressed(path, **out)
    print('%-22s N=%d S=%d loss=%.6f  -> %s (%.0f KB)' % (name, n_rays, S, loss.item(), os.path.relpath(path, ROOT),
                                                          os.path.getsize(path) / 1024))


MESH_CASES = [
    # name, config, voxel size (mesh.voxel_eval 0.05 of every shipped YAML; 0.1 as a second, coarser lattice)
    ('room0_eval', 'room0', 0.05), ('scannet0000_eval', 'scannet0000', 0.05), ('tum_desk_eval', 'tum_desk', 0.05),
    ('room0_coarse', 'room0', 0.1),
]


def reference_mesh_lattice(cfg, voxel_size):
    """Runs the reference's OWN extract_mesh (utils.py:80-135, unmodified) with a recording query function and the stubbed
    marching cubes: returns (tx, ty, tz, the normalised query points in the order of its 65 536-point chunk loop)."""
    from oracle import ref_env
    ref_utils = ref_env.import_reference('utils')
    bound = synth.bound_tensor(cfg)
    seen = []

    def query_fn(pts):                                   # [chunk, 1, 3] -> [chunk, 1]
        seen.append(pts.reshape(-1, 3).clone())
        return torch.zeros(pts.shape[:-1])
    mesh_cfg = dict(cfg, mesh={'render_color': False})
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        ref_utils.extract_mesh(query_fn, mesh_cfg, bound, marching_cube_bound=bound, voxel_size=voxel_size,
                               mesh_savepath=os.path.join(tmp, 'm', 'mesh.ply'))
    x_min, y_min, z_min = bound[:, 0]
    x_max, y_max, z_max = bound[:, 1]
    tx, ty, tz = ref_utils.getVoxels(x_max, x_min, y_max, y_min, z_max, z_min, voxel_size, None)
    return tx, ty, tz, torch.cat(seen, 0)


def write_mesh_golden():
    import json
    out = {}
    for name, cfg_name, voxel in MESH_CASES:
        cfg = synth.CONFIGS[cfg_name]()
        tx, ty, tz, flat = reference_mesh_lattice(cfg, voxel)
        out[name] = {'cfg_name': cfg_name, 'voxel_size': voxel, 'shape': [len(tx), len(ty), len(tz)],
                     'chk_tx': checksum(tx), 'chk_ty': checksum(ty), 'chk_tz': checksum(tz), 'chk_points': checksum(flat),
                     'first_point': [float(v) for v in flat[0]], 'last_point': [float(v) for v in flat[-1]]}
        print('%-22s lattice %s, %d points' % (name, out[name]['shape'], flat.shape[0]))
    with open(os.path.join(GOLDEN_DIR, 'mesh_lattice.json'), 'w') as f:
        json.dump(out, f, indent=1)


def main():
    if not ref_loader.available():
        raise SystemExit('reference tree not present; fixtures can only be regenerated in the build container')
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    only = set(sys.argv[1:])
    for case in CASES:
        if not only or case[0] in only:
            run_case(*case)
    if not only or 'mesh_lattice' in only:
        write_mesh_golden()


if __name__ == '__main__':
    main()
