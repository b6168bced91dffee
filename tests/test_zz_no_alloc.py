"""cb200_points_bwd_noalloc (include/coslam_b200.h; no counterpart in the reference):
the fused per-point backward with its compaction scratch in the library's static device array instead of a
caller workspace.  Called through the C ABI and compared with cb200_points_bwd on the same inputs, twice in a row (the second
call with fewer live points, so stale scratch contents would show).  Collected last on purpose
(the file name sorts last)."""
import ctypes

import pytest
import torch

from coslam_b200 import _lib, ops, synth
from tests import util

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _noalloc_backward(spec, table, w, mask, d_raw, o, d, z):
    """Every buffer is allocated BEFORE the call; the call itself gets raw pointers only."""
    lib = _lib.load()
    n, s = z.shape
    d_table = torch.zeros_like(table)
    flat = torch.zeros(sum(x.numel() for x in w), dtype=torch.float32, device=DEV)
    d_w, off = [], 0
    for x in w:
        d_w.append(flat[off:off + x.numel()].view_as(x))
        off += x.numel()
    d_od = torch.zeros((2, n, 3), dtype=torch.float32, device=DEV)
    mlp, gblock, scene = ops._mlp_block(*w), ops._mlp_block(*d_w), spec.scene()
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.cb200_points_bwd_noalloc(ctypes.byref(scene), _lib.ptr(table), ctypes.byref(mlp), _lib.ptr(o), _lib.ptr(d), _lib.ptr(z),
                                      ctypes.c_int64(n), ctypes.c_int64(s), _lib.ptr(d_raw), _lib.ptr(mask), _lib.ptr(d_table),
                                      ctypes.byref(gblock), _lib.ptr(d_od[0]), _lib.ptr(d_od[1]), 0, _lib.stream())
    assert rc == 0, lib.cb200_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return d_table, d_w, d_od


def test_points_bwd_noalloc_matches_points_bwd_twice():
    lib = _lib.load()
    lib.cb200_points_bwd_noalloc_max_points.restype = ctypes.c_int64
    max_points = int(lib.cb200_points_bwd_noalloc_max_points())
    assert max_points == 1 << 20
    cfg = synth.room0_config()
    n = 2304                                               # the mapping-BA batch of every shipped config
    rays_o, rays_d, rgb, depth = synth.ray_batch(cfg, n, seed=31)
    weights = synth.decoder_init(seed=3)
    from coslam_b200.scene_rep import JointEncoding
    probe = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
    model = util.make_cuda_model(cfg, synth.table_init(probe.embed_fn.params.numel(), 'trained', seed=8), weights, DEV)
    spec = model.spec
    z = ops.sample_rays(spec, depth.to(DEV), perturb=False)
    o, d = rays_o.to(DEV).contiguous(), rays_d.to(DEV).contiguous()
    table = model.embed_fn.params.detach().contiguous()
    w = tuple(x.detach().contiguous() for x in model.decoder.fused_weights())
    _, _, mask = ops.points_forward(spec.scene(), table, w, rays=(o, d, z))
    P = n * spec.n_samples
    assert P <= max_points
    gen = torch.Generator(device=DEV).manual_seed(17)
    for live_frac in (0.6, 0.15):                          # second call: fewer live points than the scratch still holds
        d_raw = torch.randn((P, 4), device=DEV, generator=gen)
        d_raw *= (torch.rand((P, 1), device=DEV, generator=gen) < live_frac)
        ref_table, ref_w, (ref_o, ref_d) = ops.points_backward(spec.scene(), table, w, mask, d_raw, rays=(o, d, z))
        got_table, got_w, got_od = _noalloc_backward(spec, table, w, mask, d_raw, o, d, z)
        util.assert_close(got_table, ref_table, 'd_table')
        for k, a, b in zip(('dW1', 'dW2', 'dW3', 'dW4'), got_w, ref_w):
            util.assert_close(a, b, k)
        util.assert_close(got_od[0], ref_o, 'd_rays_o')
        util.assert_close(got_od[1], ref_d, 'd_rays_d')
        assert int(torch.count_nonzero(got_table)) == int(torch.count_nonzero(ref_table))


def test_points_bwd_noalloc_rejects_oversized_and_non_ray_calls():
    lib = _lib.load()
    cfg = synth.room0_config()
    from coslam_b200.scene_rep import JointEncoding
    model = JointEncoding(cfg, synth.bound_tensor(cfg).to(DEV)).to(DEV)
    spec = model.spec
    w = tuple(x.detach().contiguous() for x in model.decoder.fused_weights())
    mlp, scene = ops._mlp_block(*w), spec.scene()
    table = model.embed_fn.params.detach()
    one = torch.zeros(16, device=DEV)
    n_big = (1 << 20) // spec.n_samples + 1                # one ray more than the static scratch holds
    rc = lib.cb200_points_bwd_noalloc(ctypes.byref(scene), _lib.ptr(table), ctypes.byref(mlp), _lib.ptr(one), _lib.ptr(one), _lib.ptr(one),
                                      ctypes.c_int64(n_big), ctypes.c_int64(spec.n_samples), _lib.ptr(one), _lib.ptr(one), None, None,
                                      None, None, 0, _lib.stream())
    assert rc == -1 and b'exceed the static scratch' in lib.cb200_last_error()
    rc = lib.cb200_points_bwd_noalloc(ctypes.byref(scene), _lib.ptr(table), ctypes.byref(mlp), None, None, None,
                                      ctypes.c_int64(4), ctypes.c_int64(spec.n_samples), _lib.ptr(one), _lib.ptr(one), None, None,
                                      None, None, 0, _lib.stream())
    assert rc == -1 and b'ray mode' in lib.cb200_last_error()
    torch.cuda.synchronize()
