"""The warp-specialised tensor-core forward (csrc/points_ws.cu, cb200_set_forward_pipeline) against the one-role forward
(csrc/points_tc.cuh: fwd_tile_loop): same operand tiles, same MMA order, so every output must be IDENTICAL -- raw, ReLU masks,
geo features, kept rows -- on full batches, ragged tile counts, point lists (the lazy forward), explicit points and the SDF-only
query.  The one-role forward is what the reference-generated goldens pin (tests/test_parity_gpu.py)."""
import pytest
import torch

from coslam_b200 import _lib, ops, synth

from . import util

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def JointEncodingFor(cfg):
    from coslam_b200.scene_rep import JointEncoding
    return JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)


def _setup(cfg_name, n_rays, seed=3):
    cfg = synth.CONFIGS[cfg_name]()
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n_rays, seed=seed)
    weights = [w.clone() for w in synth.decoder_init(seed=5)]
    probe = JointEncodingFor(cfg)
    table = synth.table_init(probe.embed_fn.params.numel(), 'trained', seed=77)
    model = util.make_cuda_model(cfg, table, weights, DEV)
    return cfg, model, rays_o.to(DEV), rays_d.to(DEV), depth.to(DEV)


def _both(fn):
    out = {}
    before = _lib.get_forward_pipeline()
    try:
        for name, min_points in (('one_role', -1), ('pipelined', 0)):
            _lib.set_forward_pipeline(min_points)
            assert _lib.get_forward_pipeline() == min_points
            out[name] = fn()
            torch.cuda.synchronize()
    finally:
        _lib.set_forward_pipeline(before)
    return out['one_role'], out['pipelined']


def _same(a, b, what):
    assert a.shape == b.shape, what
    assert torch.equal(a, b), '%s differs: max abs %g' % (what, float((a.float() - b.float()).abs().max()))


@pytest.mark.parametrize('cfg_name,n_rays', [('room0', 1), ('room0', 3), ('room0', 777), ('room0', 20000), ('scannet0000', 1500),
                                             ('tum_desk', 2000)])
def test_ray_forward_is_identical(cfg_name, n_rays):
    cfg, model, o, d, depth = _setup(cfg_name, n_rays)
    spec = model.spec
    z = ops.sample_rays(spec, depth, seed=11, perturb=1)
    w = model.decoder.fused_weights()

    def run():
        return ops.points_forward(spec.scene(), model.embed_fn.params, w, rays=(o, d, z), want_geo=True)

    (raw_a, geo_a, mask_a), (raw_b, geo_b, mask_b) = _both(run)
    assert bool(torch.isfinite(raw_a).all()) and float(raw_a.abs().max()) > 0
    _same(raw_a, raw_b, 'raw')
    _same(geo_a, geo_b, 'geo')
    _same(mask_a, mask_b, 'relu masks')


@pytest.mark.parametrize('width', [32, 128])
def test_kept_rows_are_identical(width):
    cfg, model, o, d, depth = _setup('room0', 5000)
    spec = model.spec
    z = ops.sample_rays(spec, depth, seed=11, perturb=1)
    w = model.decoder.fused_weights()

    def run():
        feat = torch.zeros((z.numel(), width), dtype=torch.float32, device=DEV)
        raw, _, mask = ops.points_forward(spec.scene(), model.embed_fn.params, w, rays=(o, d, z), feat=feat)
        return raw, mask, feat

    a, b = _both(run)
    for x, y, what in zip(a, b, ('raw', 'relu masks', 'kept rows')):
        _same(x, y, what)
    assert float(a[2].abs().max()) > 0


def test_lazy_list_forward_is_identical():
    cfg, model, o, d, depth = _setup('room0', 6000)
    spec = model.spec
    z = ops.sample_rays(spec, depth, seed=11, perturb=1)
    w = model.decoder.fused_weights()

    def run():
        return ops.lazy_ray_forward(spec, model.embed_fn.params, w, (o, d, z), depth)

    (raw_a, mask_a), (raw_b, mask_b) = _both(run)
    listed = (raw_a != 0).any(dim=1)
    assert 0.2 < float(listed.float().mean()) < 0.95
    _same(raw_a, raw_b, 'raw')
    _same(mask_a[listed], mask_b[listed], 'relu masks of the listed points')       # unlisted rows of the mask are never written


@pytest.mark.parametrize('sdf_only', [False, True])
def test_explicit_points_forward_is_identical(sdf_only):
    cfg, model, o, d, depth = _setup('room0', 16)
    spec = model.spec
    w = model.decoder.fused_weights()
    g = torch.Generator().manual_seed(5)
    pts = torch.rand((70001, 3), generator=g).to(DEV)                 # normalised positions incl. a ragged last tile

    def run():
        return ops.points_forward(spec.scene_unnormalised(), model.embed_fn.params, w, pts=pts, want_geo=True, sdf_only=sdf_only)

    (raw_a, geo_a, mask_a), (raw_b, geo_b, mask_b) = _both(run)
    _same(raw_a, raw_b, 'raw')
    _same(geo_a, geo_b, 'geo')
    _same(mask_a[:, 0], mask_b[:, 0], 'relu mask of the SDF net')
    if not sdf_only:
        _same(mask_a[:, 1], mask_b[:, 1], 'relu mask of the colour net')


def test_training_step_through_the_pipelined_forward_matches():
    """The whole autograd step (lazy forward with kept rows, rendering, losses, backward) with either forward: identical losses,
    gradients equal up to the order of their atomics."""
    cfg, model, o, d, depth = _setup('room0', 4096)
    rgb = torch.rand((4096, 3), generator=torch.Generator().manual_seed(1)).to(DEV)

    def run():
        for p in model.parameters():
            p.grad = None
        oo = o.clone().requires_grad_(True)
        dd = d.clone().requires_grad_(True)
        torch.manual_seed(3)
        ret = model.forward(oo, dd, rgb, depth)
        t = cfg['training']
        loss = (t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] + t['sdf_weight'] * ret['sdf_loss'] +
                t['fs_weight'] * ret['fs_loss'])
        loss.backward()
        return [ret[k].detach().clone() for k in ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss')], \
            [oo.grad.clone(), dd.grad.clone(), model.embed_fn.params.grad.clone()] + [w.grad.clone() for w in model.decoder.fused_weights()]

    (loss_a, grad_a), (loss_b, grad_b) = _both(run)
    for x, y in zip(loss_a, loss_b):
        assert torch.equal(x, y)
    for i, (x, y) in enumerate(zip(grad_a, grad_b)):
        util.assert_close(x, y, 'gradient %d' % i, tol=2e-5)
