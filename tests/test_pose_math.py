"""the pose parametrisations, pose -> ray transform and keyframe ray sampling of the
oracle against the reference's own files (imported unmodified; CPU only, skipped where /root/reference is absent)."""
import importlib
import random
import sys
import types

import pytest
import torch

from oracle import coslam_oracle as co
from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason='reference tree not present')


def _ref_module(name):
    """Imports a reference module; pytorch3d (absent here) is stubbed with the oracle's restatement of the one function
    the quaternion path calls, so only the reference's OWN arithmetic is under test."""
    if 'pytorch3d' not in sys.modules:
        p3d, tr = types.ModuleType('pytorch3d'), types.ModuleType('pytorch3d.transforms')
        tr.quaternion_to_matrix = co.quaternion_to_matrix
        tr.matrix_to_quaternion = tr.rotation_6d_to_matrix = tr.quaternion_to_axis_angle = None
        p3d.transforms = tr
        sys.modules['pytorch3d'], sys.modules['pytorch3d.transforms'] = p3d, tr
    if ref_loader.REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    return importlib.import_module(name)


def test_quad2rotation_matches_oracle_quaternion_to_matrix():
    pa = _ref_module('optimization.pose_array')              # optimization/pose_array.py:6-29
    q = torch.randn(257, 4, generator=torch.Generator().manual_seed(0)) * 3      # un-normalised on purpose
    q.requires_grad_(True)
    R_ref = pa.quad2rotation(q)
    R = co.quaternion_to_matrix(q)
    assert torch.equal(R, R_ref)
    g = torch.randn(257, 3, 3, generator=torch.Generator().manual_seed(1))
    (g_ref,) = torch.autograd.grad((R_ref * g).sum(), q, retain_graph=True)
    (g_orc,) = torch.autograd.grad((R * g).sum(), q)
    assert torch.allclose(g_orc, g_ref, rtol=1e-5, atol=1e-6)
    qn = q.detach() / q.detach().norm(dim=-1, keepdim=True)
    Rn = co.quaternion_to_matrix(qn)
    assert torch.allclose(Rn @ Rn.transpose(-1, -2), torch.eye(3).expand(257, 3, 3), atol=1e-5)


def test_pose_matrix_matches_reference_transforms():
    ou = _ref_module('optimization.utils')                   # optimization/utils.py:5-55
    g = torch.Generator().manual_seed(2)
    rot, trans = torch.randn(64, 3, generator=g), torch.randn(64, 3, generator=g)
    assert torch.equal(co.axis_angle_to_matrix(rot), ou.axis_angle_to_matrix(rot))
    assert torch.equal(co.pose_matrix(rot, trans), ou.at_to_transform_matrix(rot, trans))
    quat = torch.randn(64, 4, generator=g)
    assert torch.equal(co.pose_matrix(quat, trans), ou.qt_to_transform_matrix(quat, trans))
    # the identity rotation is singular in the reference (utils.py:4): theta = 0 divides by zero
    assert torch.isnan(ou.axis_angle_to_matrix(torch.zeros(1, 3))).any()
    assert torch.isnan(co.axis_angle_to_matrix(torch.zeros(1, 3))).any()


def test_rays_from_pose_and_backward_to_pose_parameters():
    """coslam.py:546-547 (tracking) and :353-355 (BA): rays from c2w, gradients back to (rot, trans)."""
    ou = _ref_module('optimization.utils')
    g = torch.Generator().manual_seed(3)
    dir_cam = torch.randn(100, 3, generator=g)
    rot = torch.randn(1, 3, generator=g).requires_grad_(True)
    trans = torch.randn(1, 3, generator=g).requires_grad_(True)
    c2w = ou.at_to_transform_matrix(rot, trans)[0]
    rays_o_ref = c2w[None, :3, -1].repeat(100, 1)
    rays_d_ref = torch.sum(dir_cam[..., None, :] * c2w[:3, :3], -1)
    rays_o, rays_d = co.rays_from_pose(dir_cam, co.pose_matrix(rot, trans)[0])
    assert torch.equal(rays_o, rays_o_ref) and torch.equal(rays_d, rays_d_ref)
    go, gd = torch.randn(100, 3, generator=g), torch.randn(100, 3, generator=g)
    ref = torch.autograd.grad((rays_o_ref * go).sum() + (rays_d_ref * gd).sum(), (rot, trans), retain_graph=True)
    got = torch.autograd.grad((rays_o * go).sum() + (rays_d * gd).sum(), (rot, trans))
    for a, b in zip(got, ref):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    # BA: per-ray pose gathered by keyframe slot, -1 = current frame (last row)
    poses = co.pose_matrix(torch.randn(5, 3, generator=g), torch.randn(5, 3, generator=g))
    ids = torch.tensor([0, 3, -1, 3, 1, -1])
    d6 = dir_cam[:6]
    ro, rd = co.rays_from_pose_table(d6, poses, ids)
    for i in range(6):
        assert torch.allclose(rd[i], poses[ids[i], :3, :3] @ d6[i], atol=1e-6)
        assert torch.equal(ro[i], poses[ids[i], :3, 3])


def test_keyframe_ray_indices_bit_exact():
    """model/keyframe.py:96-106 and coslam.py:131-138: Python `random.sample` without replacement, seeded."""
    kf = _ref_module('model.keyframe')
    db = kf.KeyFrameDatabase.__new__(kf.KeyFrameDatabase)
    n_kf, R = 7, 500
    db.num_rays_to_save = R
    db.rays = torch.arange(n_kf * R * 7, dtype=torch.float32).reshape(n_kf, R, 7)
    db.frame_ids = torch.arange(n_kf) * 5
    random.seed(123)
    rays_ref, ids_ref = db.sample_global_rays(2048)
    random.seed(123)
    idx = torch.tensor(co.sample_global_ray_indices(n_kf, R, 2048))
    assert torch.equal(db.rays[:n_kf].reshape(-1, 7)[idx], rays_ref)
    assert torch.equal(db.frame_ids[idx // R], ids_ref)
    assert len(set(idx.tolist())) == 2048


def test_per_frame_alignment_table():
    fa = _ref_module('optimization.frame_align')             # optimization/frame_align.py:5-18
    m = fa.PerFrameAlignment(10)
    ids = torch.tensor([0, 9, 3])
    assert torch.equal(m(ids), torch.tensor([[1., 1., 0., 0.]]).repeat(3, 1))


def test_package_pose_helpers_match_reference_and_roundtrip():
    """coslam_b200/pose.py (used by the graph-captured tracker) against optimization/utils.py and a rotation round trip."""
    from coslam_b200 import pose
    ou = _ref_module('optimization.utils')
    g = torch.Generator().manual_seed(7)
    rot, trans = torch.randn(32, 3, generator=g), torch.randn(32, 3, generator=g)
    assert torch.equal(pose.pose_matrix(rot, trans), ou.at_to_transform_matrix(rot, trans))
    quat = torch.randn(32, 4, generator=g)
    assert torch.equal(pose.pose_matrix(quat, trans), ou.qt_to_transform_matrix(quat, trans))
    R = pose.axis_angle_to_matrix(rot)
    aa = pose.matrix_to_axis_angle(R)                          # principal value: compare through the rotation
    assert torch.allclose(pose.axis_angle_to_matrix(aa), R, atol=2e-5)
    q = pose.matrix_to_quaternion(R)
    assert torch.allclose(q.norm(dim=-1), torch.ones(32), atol=1e-5)
    assert torch.allclose(pose.quaternion_to_matrix(q), R, atol=2e-5)
    dir_cam = torch.randn(50, 3, generator=g)
    c2w = pose.pose_matrix(rot[:1], trans[:1])
    ro, rd = pose.rays_from_pose(dir_cam, c2w)
    ro_ref = c2w[..., :3, -1].repeat(50, 1)
    rd_ref = torch.sum(dir_cam[..., None, :] * c2w[:, :3, :3], -1)
    assert torch.equal(ro, ro_ref) and torch.equal(rd, rd_ref)


def test_sample_range_is_random_sample_bit_for_bit():
    """coslam_b200.keyframe.sample_range: the vectorised `random.sample(range(n), k)` (model/keyframe.py:45-46, :100;
    coslam.py:131-138, :339-341) -- same values, same order, same generator state afterwards -- over the population / sample
    sizes of the shipped configs, powers of two (acceptance 1/2), tiny populations (pool branch) and near-exhaustive draws."""
    import random
    from coslam_b200.keyframe import sample_range
    cases = [(816000, 2048), (816000, 40800), (816000, 100), (40800 * 7, 2048), (40800 * 401, 2048), (1 << 16, 1000), ((1 << 16) + 1, 1000),
             (285200, 14260), (5000, 4990), (100, 7), (30, 20), (16405, 2048), (16406, 2048), (1, 1), (9, 0), (768, 300)]
    for seed, (n, k) in enumerate(cases):
        random.seed(100 + seed)
        for _ in range(seed % 3):
            random.random()                                   # start from a state in the middle of a block of 624
        state = random.getstate()
        want = random.sample(range(n), k)
        after = [random.random(), random.getrandbits(17), random.sample(range(1000), 5)]
        random.setstate(state)
        got = sample_range(n, k)
        assert got.dtype.kind == 'i' and got.tolist() == want, (n, k)
        assert [random.random(), random.getrandbits(17), random.sample(range(1000), 5)] == after, (n, k)
    # consecutive calls chain exactly like consecutive random.sample calls
    random.seed(7)
    want = [random.sample(range(326400), 2048) for _ in range(5)]
    random.seed(7)
    assert [sample_range(326400, 2048).tolist() for _ in range(5)] == want
