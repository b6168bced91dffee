"""Cell-major copies of the dense hash-grid levels (csrc/cells.cu: cb200_cell_table_build / _fold; include/coslam_b200.h).
The copies hold the table's own values, so a forward that reads them must reproduce the forward on the table BIT FOR BIT (also for
points outside the unit cube, where the dense index wraps), and a backward that reduces into gradient blocks of that layout must,
after the fold, give the table gradient of the plain backward up to the order of the atomics."""
import ctypes

import pytest
import torch

from coslam_b200 import _lib, ops, synth
from coslam_b200._lib import ptr, stream

from . import util

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _model(cfg_name):
    from coslam_b200.scene_rep import JointEncoding
    cfg = synth.CONFIGS[cfg_name]()
    probe = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
    table = synth.table_init(probe.embed_fn.params.numel(), 'trained', seed=77)
    weights = [w.clone() for w in synth.decoder_init(seed=5)]
    return cfg, util.make_cuda_model(cfg, table, weights, DEV)


@pytest.mark.parametrize('cfg_name', ['room0', 'scannet0000'])
def test_cell_table_holds_the_corner_values(cfg_name):
    cfg, model = _model(cfg_name)
    spec = model.spec
    scene, buf = ops.cell_tables(spec.scene(), model.embed_fn.params.detach())
    assert scene is not None and scene.cells == buf[0].data_ptr()
    g = spec.grid()
    tab = model.embed_fn.params.detach().view(-1, 2)
    cells = buf[0].view(-1, 2, 4, 2)                                  # [cell][corner x][corner (y, z)][feature]
    first = 0
    n_dense = 0
    for l in range(g.n_levels):
        if g.hashed[l]:
            break
        n_dense += 1
        res, off, size = int(g.res[l]), int(g.offset[l]), int(g.offset[l + 1] - g.offset[l])
        c = torch.arange(res ** 3, device=DEV)
        cx, cy, cz = c % res, (c // res) % res, c // (res * res)
        for xb in (0, 1):
            for j in range(4):
                idx = (cx + xb) + (cy + (j & 1)) * res + (cz + (j >> 1)) * res * res
                idx = torch.where(idx >= size, idx % size, idx)
                assert torch.equal(cells[first:first + res ** 3, xb, j], tab[off + idx]), (l, xb, j)
        first += res ** 3
    assert n_dense >= 5 and first * 64 == buf[0].numel() * 4


@pytest.mark.parametrize('pipelined', [False, True])
def test_forward_on_cell_tables_is_identical(pipelined):
    cfg, model = _model('room0')
    spec = model.spec
    w = model.decoder.fused_weights()
    g = torch.Generator().manual_seed(9)
    pts = torch.rand((50000, 3), generator=g) * 1.2 - 0.1              # a tenth of them outside the unit cube on some axis
    pts[0] = torch.tensor([0.0, 0.0, 0.0])
    pts[1] = torch.tensor([1.0, 1.0, 1.0])
    pts[2] = torch.tensor([1.0, 0.5, 0.0])
    pts = pts.to(DEV)
    table = model.embed_fn.params.detach()
    plain = spec.scene_unnormalised()
    with_cells, _ = ops.cell_tables(plain, table)
    before = _lib.get_forward_pipeline()
    try:
        _lib.set_forward_pipeline(0 if pipelined else -1)
        out_a = ops.points_forward(plain, table, w, pts=pts, want_geo=True)
        feat_a = torch.zeros((pts.shape[0], 128), device=DEV)
        ops.points_forward(plain, table, w, pts=pts, feat=feat_a)
        out_b = ops.points_forward(with_cells, table, w, pts=pts, want_geo=True)
        feat_b = torch.zeros((pts.shape[0], 128), device=DEV)
        ops.points_forward(with_cells, table, w, pts=pts, feat=feat_b)
    finally:
        _lib.set_forward_pipeline(before)
    for x, y, what in zip(out_a, out_b, ('raw', 'geo', 'relu masks')):
        assert torch.equal(x, y), what
    assert torch.equal(feat_a, feat_b)
    assert float(out_a[0].abs().max()) > 0


def test_training_step_on_cell_tables_matches():
    """The autograd step with the copies forced on for a small batch: losses identical (the forward does not change by a bit), every
    gradient equal up to the order of its atomics; the table gradient of the dense levels arrives through the fold."""
    cfg, model = _model('room0')
    n = 4096
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n, seed=3)
    o, d, rgb, depth = rays_o.to(DEV), rays_d.to(DEV), rgb.to(DEV), depth.to(DEV)
    t = cfg['training']
    res = {}
    saved = (ops.CELL_TABLES, ops.CELL_MIN_POINTS)
    try:
        for on in (False, True):
            ops.CELL_TABLES, ops.CELL_MIN_POINTS = on, 1
            for p in model.parameters():
                p.grad = None
            oo, dd = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
            torch.manual_seed(3)
            before = _lib.KERNEL_LAUNCHES
            ret = model.forward(oo, dd, rgb, depth)
            loss = (t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss'] + t['sdf_weight'] * ret['sdf_loss'] +
                    t['fs_weight'] * ret['fs_loss'])
            loss.backward()
            res[on] = ([ret[k].detach().clone() for k in ('rgb_loss', 'depth_loss', 'sdf_loss', 'fs_loss')],
                       [oo.grad.clone(), dd.grad.clone(), model.embed_fn.params.grad.clone()] +
                       [x.grad.clone() for x in model.decoder.fused_weights()], _lib.KERNEL_LAUNCHES - before)
    finally:
        ops.CELL_TABLES, ops.CELL_MIN_POINTS = saved
    for x, y in zip(res[False][0], res[True][0]):
        assert torch.equal(x, y)
    for i, (x, y) in enumerate(zip(res[False][1], res[True][1])):
        util.assert_close(x, y, 'gradient %d' % i, tol=2e-5)
    assert res[True][2] == res[False][2] + 2                           # build + fold
    # the dense levels' gradient is there (it came through the gradient blocks)
    g = model.spec.grid()
    dense_end = 2 * int(g.offset[5])
    assert float(res[True][1][2][:dense_end].abs().max()) > 0


def test_fold_adds_gradient_blocks_to_the_table_gradient():
    cfg, model = _model('room0')
    spec = model.spec
    scene = spec.scene()
    lib = _lib.load()
    nbytes = int(lib.cb200_cell_table_bytes(ctypes.byref(scene)))
    dcells = torch.zeros(nbytes // 4, device=DEV)
    blocks = dcells.view(-1, 2, 4, 2)
    g = spec.grid()
    res0, res1 = int(g.res[0]), int(g.res[1])
    # cell (1, 2, 3) of level 0, corner (x 1, y 0, z 1) -> vertex (2, 2, 4); cell (0, 0, 0) of level 1, corner (0, 1, 0) -> vertex (0, 1, 0)
    blocks[1 + res0 * (2 + res0 * 3), 1, 2] = torch.tensor([1.5, -2.0], device=DEV)
    blocks[res0 ** 3, 0, 1] = torch.tensor([0.25, 4.0], device=DEV)
    d_table = torch.ones(model.embed_fn.params.numel(), device=DEV)
    _lib.check(lib.cb200_cell_table_fold(ctypes.byref(scene), ptr(dcells), ptr(d_table), stream()), 'fold')
    want = torch.ones_like(d_table).view(-1, 2)
    want[int(g.offset[0]) + 2 + 2 * res0 + 4 * res0 * res0] += torch.tensor([1.5, -2.0], device=DEV)
    want[int(g.offset[1]) + 0 + 1 * res1] += torch.tensor([0.25, 4.0], device=DEV)
    assert torch.equal(d_table.view(-1, 2), want)
