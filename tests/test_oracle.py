"""CPU tests (-m "not gpu"): pin the oracle.

 1. oracle/coslam_oracle.py (+ tcnn_shim) against the golden vectors that oracle/gen_golden.py produced by running the
    reference's own, unmodified model/scene_rep.py in the build container -- including the hierarchical pass
    (model/utils.py:32-71 sample_pdf, scene_rep.py:262-295, 327-332) at n_importance 128 / 8 and a 4096-ray room0 batch;
 2. oracle/hashgrid_ref.c (plain C, real fmaf) against the torch shim: bit-exact indices and sample positions;
 3. the tiny-cuda-nn level-size formula against the parameter counts written in the reference's YAML comments
    (configs/ScanNet/scene0000.yaml:8 "1.59M", scene0106.yaml:8 "0.48", scene0059.yaml:7 "0.13") -- the only in-tree
    corroboration of the un-vendored dependency (parity is otherwise UNPINNED at that boundary, see tcnn_shim.py).
"""
import numpy as np
import pytest
import torch

from coslam_b200 import synth
from oracle import c_oracle, coslam_oracle, ref_loader, tcnn_shim
from tests import util


@pytest.mark.parametrize('name', util.GOLDEN_CASES + util.IMPORTANCE_CASES + util.LARGE_CASES + util.TWOGRID_CASES)
def test_oracle_reproduces_reference_golden(name):
    g = util.load_golden(name)
    cfg, (rays_o, rays_d, rgb, depth), table, weights, noise = util.case_inputs(g)
    u_imp = util.case_noise(g, cfg)[1]
    orc = util.make_oracle(cfg, table, weights, util.color_table(g, cfg) if name in util.TWOGRID_CASES else None)
    o = rays_o.clone().requires_grad_(True)
    d = rays_d.clone().requires_grad_(True)
    r = orc.forward(o, d, rgb, depth, noise=noise, noise_imp=u_imp)
    orc.total_loss(r).backward()
    st = int(g['sample_stride']) if 'sample_stride' in g.files else 1
    assert np.array_equal(r['z_vals'][::st].numpy(), g['z_vals'])             # bit-exact sample depths (merged, row a11)
    assert np.array_equal(r['pts'][::st].reshape(-1, 3).detach().numpy(), g['pts'])    # bit-exact sample positions
    if 'chk_z_vals' in g.files:
        assert util.checksum(r['z_vals']) == str(g['chk_z_vals']) and util.checksum(r['pts'].reshape(-1, 3)) == str(g['chk_pts'])
    if cfg['training']['n_importance'] > 0:                                    # coarse maps and the spread of the drawn depths
        for k in ('rgb0', 'depth0', 'depth_var0', 'acc0', 'z_std'):
            util.assert_close(r[k], g[k], k, 1e-5)
        mid = .5 * (torch.from_numpy(g['z_coarse'])[..., 1:] + torch.from_numpy(g['z_coarse'])[..., :-1])
        zi = coslam_oracle.sample_pdf(mid, torch.from_numpy(g['w_coarse'])[..., 1:-1], cfg['training']['n_importance'],
                                      det=not cfg['training']['perturb'] > 0., u=u_imp)
        merged = torch.sort(torch.cat([torch.from_numpy(g['z_coarse']), zi], -1), -1)[0]
        assert np.array_equal(merged.numpy(), g['z_vals'])                     # sample_pdf alone, from the stored coarse weights
    util.assert_close(r['raw'][::st], g['raw'], 'raw', 1e-5)
    util.assert_close(r['rgb'], g['rgb'], 'rgb', 1e-5)
    util.assert_close(r['depth'], g['depth'], 'depth', 1e-5)
    for k in ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss'):
        util.assert_close(r[k], float(g[k]), k, 1e-5)
    util.assert_close(o.grad, g['d_rays_o'], 'd_rays_o', 1e-4)
    util.assert_close(d.grad, g['d_rays_d'], 'd_rays_d', 1e-4)
    for k, w in zip(('dW1', 'dW2', 'dW3', 'dW4'), (orc.W1, orc.W2, orc.W3, orc.W4)):
        util.assert_close(w.grad, g[k], k, 1e-4)
    tg = orc.table.grad
    util.assert_close(tg[torch.from_numpy(g['tg_pick_idx'])], g['tg_pick_val'], 'table grad (sampled)', 1e-4)
    assert int(torch.count_nonzero(tg)) == int(g['tg_nnz'])
    if name in util.TWOGRID_CASES:                                             # decoder.py:110-143, scene_rep.py:56-58, 168-170
        cg = orc.table_color.grad
        util.assert_close(cg[torch.from_numpy(g['ctg_pick_idx'])], g['ctg_pick_val'], 'colour table grad (sampled)', 1e-4)
        assert int(torch.count_nonzero(cg)) == int(g['ctg_nnz'])


@pytest.mark.parametrize('name', util.GOLDEN_CASES)
def test_c_oracle_indices_and_positions_bit_exact(name):
    g = util.load_golden(name)
    cfg, (rays_o, rays_d, rgb, depth), table, weights, noise = util.case_inputs(g)
    bound = synth.bound_tensor(cfg)
    enc = coslam_oracle.hashgrid_encoding_config(cfg, bound)
    grid = c_oracle.make_grid(enc['n_levels'], enc['n_features_per_level'], enc['log2_hashmap_size'], enc['base_resolution'],
                              enc['per_level_scale'])
    meta = c_oracle.grid_meta(grid)
    assert np.array_equal(np.asarray(meta['scale'], dtype=np.float32), g['grid_scale'])
    assert np.array_equal(np.asarray(meta['res'], dtype=np.uint32), g['grid_res'])
    assert np.array_equal(np.asarray(meta['offset'], dtype=np.uint32), g['grid_offset'])
    t = cfg['training']
    lin_u = torch.linspace(cfg['cam']['near'], cfg['cam']['far'], t['n_samples_d']).numpy()
    lin_g = torch.linspace(-t['range_d'], t['range_d'], steps=t['n_range_d']).numpy()
    lin_n = torch.linspace(cfg['cam']['near'], cfg['cam']['far'], steps=t['n_range_d']).numpy()
    z = c_oracle.sample_z(depth.numpy(), noise.numpy(), lin_u, lin_g, lin_n)
    assert np.array_equal(z, g['z_vals'])
    pts, x = c_oracle.points(rays_o.numpy(), rays_d.numpy(), z, bound[:, 0].numpy(), (bound[:, 1]).numpy())
    assert np.array_equal(pts.reshape(-1, 3), g['pts'])
    assert np.array_equal(x.reshape(-1, 3), g['x'])
    n = g['hash_idx'].shape[0]
    idx, _ = c_oracle.hash_indices(grid, g['x'][:n])
    assert np.array_equal(idx, g['hash_idx'])


def test_c_oracle_encodings_match_shim():
    torch.manual_seed(0)
    x = torch.rand(512, 3)
    x[:8] = torch.tensor([0.0, 1.0, 0.5]).repeat(8, 1) + torch.linspace(-0.2, 0.2, 8)[:, None]   # edge / outside points
    cfg = synth.room0_config()
    enc = coslam_oracle.hashgrid_encoding_config(cfg, synth.bound_tensor(cfg))
    shim = tcnn_shim.Encoding(3, enc)
    with torch.no_grad():
        shim.params.copy_(synth.table_init(shim.params.numel(), 'trained'))
    grid = c_oracle.make_grid(16, 2, enc['log2_hashmap_size'], 16, enc['per_level_scale'])
    out_c = c_oracle.hash_fwd(grid, x.numpy(), shim.params.detach().numpy())
    util.assert_close(out_c, shim(x), 'hash fwd C vs shim', 1e-6)
    blob = tcnn_shim.Encoding(3, {'otype': 'OneBlob', 'n_bins': 16})
    util.assert_close(c_oracle.oneblob_fwd(x.numpy(), 16), blob(x), 'oneblob C vs shim', 1e-6)
    assert abs(float(blob(x).sum(-1).mean()) - 3.0) < 1e-4                     # each dimension's bins sum to one


@pytest.mark.parametrize('desired,log2T,expect', [(225, 16, 1593680), (225, 19, 8675888)])
def test_level_sizes_match_yaml_comments(desired, log2T, expect):
    # scene0000: bound extent 8.9 m (x2 see scene0000.yaml) / voxel_sdf 0.04 -> int(222.5)?  the reference computes
    # int(dim_max / voxel) with a torch fp32 scalar; SURVEY.md section 8 records 225 for this scene.
    pls = np.exp2(np.log2(desired / 16) / 15)
    meta = tcnn_shim.grid_meta(16, log2T, 16, pls)
    assert meta['offset'][-1] * 2 == expect


def test_room0_level_table():
    cfg = synth.room0_config()
    enc = coslam_oracle.hashgrid_encoding_config(cfg, synth.bound_tensor(cfg))
    meta = tcnn_shim.grid_meta(16, 16, 16, enc['per_level_scale'])
    assert meta['res'] == [16, 20, 25, 31, 38, 47, 58, 72, 90, 111, 137, 170, 211, 261, 323, 401]
    assert meta['offset'][-1] == 833288


@pytest.mark.skipif(not ref_loader.available(), reason='reference tree only exists in the build container')
def test_restatement_matches_unmodified_reference_live():
    """Runs the reference's own JointEncoding here (shim as tinycudann) on a fresh seed that is NOT in the fixtures."""
    cfg = synth.room0_config()
    JointEncoding = ref_loader.load_reference_model_class()
    ref = JointEncoding(cfg, synth.bound_tensor(cfg))
    table = synth.table_init(ref.embed_fn.params.numel(), 'trained', seed=99)
    weights = synth.decoder_init(seed=99)
    with torch.no_grad():
        ref.embed_fn.params.copy_(table)
        for dst, src in zip((ref.decoder.sdf_net.model[0].weight, ref.decoder.sdf_net.model[2].weight,
                             ref.decoder.color_net.model[0].weight, ref.decoder.color_net.model[2].weight), weights):
            dst.copy_(src)
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, 96, seed=99)
    torch.manual_seed(7)
    noise = torch.rand(96, 43)
    torch.manual_seed(7)
    ret = ref.forward(rays_o, rays_d, rgb, depth)
    orc = util.make_oracle(cfg, table, weights)
    r = orc.forward(rays_o, rays_d, rgb, depth, noise=noise)
    for k in ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss'):
        util.assert_close(r[k], ret[k], k, 1e-5)
    util.assert_close(r['rgb'], ret['rgb'], 'rgb', 1e-5)


def test_host_torch_rand_is_the_cpu_generators_stream():
    """cb200_host_torch_rand (csrc/host_rng.cu) against torch.rand on torch's default CPU generator -- the stream the reference draws
    its stratified jitter from (model/scene_rep.py:255): same values for any split into calls, same generator state afterwards (so
    every later draw, uniform or normal, is unchanged), across block boundaries of the 624-word Mersenne-Twister state."""
    import torch
    from coslam_b200 import ops
    for seed in (0, 1, 12345):
        for sizes in ([1, 5, 623, 624, 625, 1000, 2304 * 43, 3], [1024 * 43] * 3, [624, 624, 1], [0, 7]):
            torch.manual_seed(seed)
            ref = [torch.rand(k) for k in sizes]
            tail_ref, state_ref = torch.randn(9), torch.get_rng_state()
            torch.manual_seed(seed)
            got = [ops.torch_rand_into(torch.empty(k)) for k in sizes]
            tail, state = torch.randn(9), torch.get_rng_state()
            assert all(torch.equal(a, b) for a, b in zip(ref, got)), (seed, sizes)
            assert torch.equal(tail, tail_ref) and torch.equal(state, state_ref)
    torch.manual_seed(3)                                   # interleaved with normal draws (their cache lives in the same state block)
    a, b, c = torch.randn(5), torch.rand(4, 25), torch.randn(3)
    torch.manual_seed(3)
    a2, b2, c2 = torch.randn(5), ops.torch_rand_into(torch.empty(4, 25)), torch.randn(3)
    assert torch.equal(a, a2) and torch.equal(b, b2) and torch.equal(c, c2)


def test_host_torch_rand_skip_advances_the_generator_like_a_draw():
    """cb200_host_torch_rand_skip (the host half of the device-side jitter stream, csrc/host_rng.cu): the generator state after
    skipping n floats is the state after torch.rand(n), from fresh and mid-block states and across block boundaries."""
    import ctypes
    import torch
    from coslam_b200 import _lib
    lib = _lib.load()
    for seed in (0, 7):
        for warm, n in ((0, 1), (0, 624), (5, 619), (5, 620), (100, 3 * 624 + 17), (623, 1), (300, 1 << 20), (17, 0)):
            torch.manual_seed(seed)
            torch.rand(warm)
            st = torch.get_rng_state()
            torch.rand(n)
            want, tail = torch.get_rng_state(), torch.rand(5)
            assert lib.cb200_host_torch_rand_skip(ctypes.c_void_p(st.data_ptr()), ctypes.c_int64(n)) == 0
            assert torch.equal(st, want), (seed, warm, n)
            torch.set_rng_state(st)
            assert torch.equal(torch.rand(5), tail)
            # the checkpointing walk (cb200_host_torch_rand_segments) ends in the same state; its first checkpoint is the state at entry
            torch.manual_seed(seed)
            torch.rand(warm)
            st2 = torch.get_rng_state()
            words_at_entry = st2[24:24 + 8 * 624].clone().view(torch.int64).to(torch.int32)
            ck = torch.zeros((n // (624 * 4) + 3, 624), dtype=torch.int32)
            n_seg, head, head_next = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int32(0)
            assert lib.cb200_host_torch_rand_segments(ctypes.c_void_p(st2.data_ptr()), ctypes.c_int64(n), ctypes.c_int64(4),
                                                      ctypes.c_void_p(ck.data_ptr()), ctypes.c_int64(ck.shape[0]), ctypes.byref(n_seg),
                                                      ctypes.byref(head), ctypes.byref(head_next)) == 0
            assert torch.equal(st2, want) and torch.equal(ck[0], words_at_entry)
            fresh = warm == 0
            assert head.value == min(n, 0 if fresh else 624 - warm % 624 if warm % 624 else 0)
            assert 1 <= n_seg.value <= ck.shape[0]
