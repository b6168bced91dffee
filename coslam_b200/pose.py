"""Pose parametrisations of the tracking / BA loops, as torch ops on small tensors (host-side glue, differentiable).

Mirrors optimization/utils.py:5-55 (axis-angle and quaternion to 4x4) and the two pytorch3d conversions the reference
uses to turn an initial rotation matrix into a parameter (optimization/utils.py:23-28, coslam.py:276-283):
matrix_to_quaternion and quaternion_to_axis_angle, restated from their published formulas (pytorch3d is not installed here).
"""
import torch


def axis_angle_to_matrix(r):
    """optimization/utils.py:5-21 -- Rodrigues; singular at r = 0 exactly like the reference (TODO note at :4)."""
    theta = torch.norm(r, dim=-1, keepdim=True)
    w = r / theta
    z = torch.zeros_like(w[..., 0:1])
    K = torch.stack([torch.cat([z, -w[..., 2:3], w[..., 1:2]], -1),
                     torch.cat([w[..., 2:3], z, -w[..., 0:1]], -1),
                     torch.cat([-w[..., 1:2], w[..., 0:1], z], -1)], dim=-2)
    eye = torch.eye(3, device=r.device, dtype=r.dtype).expand(*r.shape[:-1], 3, 3)
    return eye + torch.sin(theta).unsqueeze(-1) * K + (1. - torch.cos(theta).unsqueeze(-1)) * (K @ K)


def quaternion_to_matrix(q):
    """Real-first quaternion, un-normalised (two_s = 2 / |q|^2): pytorch3d.transforms.quaternion_to_matrix and
    optimization/pose_array.py:6-29 (quad2rotation)."""
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def _sqrt_positive_part(x):
    return torch.where(x > 0, torch.sqrt(torch.clamp(x, min=0)), torch.zeros_like(x))


def matrix_to_quaternion(m):
    """pytorch3d.transforms.matrix_to_quaternion: the four candidate quaternions, the one with the largest denominator."""
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(m.reshape(m.shape[:-2] + (9,)), -1)
    q_abs = _sqrt_positive_part(torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                                             1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], -1))
    cand = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], -1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], -1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], -1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], -1)], -2)
    cand = cand / (2.0 * q_abs[..., None].clamp(min=0.1))
    best = torch.nn.functional.one_hot(q_abs.argmax(-1), num_classes=4) > 0.5
    out = cand[best, :].reshape(m.shape[:-2] + (4,))
    return torch.where(out[..., 0:1] < 0, -out, out)          # standardise: non-negative real part


def quaternion_to_axis_angle(q):
    """pytorch3d.transforms.quaternion_to_axis_angle (series 0.5 - angle^2 / 48 below 1e-6)."""
    norms = torch.norm(q[..., 1:], p=2, dim=-1, keepdim=True)
    half = torch.atan2(norms, q[..., :1])
    angles = 2 * half
    small = angles.abs() < 1e-6
    s = torch.empty_like(angles)
    s[~small] = torch.sin(half[~small]) / angles[~small]
    s[small] = 0.5 - (angles[small] * angles[small]) / 48
    return q[..., 1:] / s


def matrix_to_axis_angle(rot):
    """optimization/utils.py:23-28."""
    return quaternion_to_axis_angle(matrix_to_quaternion(rot))


def pose_matrix(rot, trans):
    """at_to_transform_matrix / qt_to_transform_matrix (optimization/utils.py:30-55): [K,3|4] + [K,3] -> [K,4,4]."""
    R = axis_angle_to_matrix(rot) if rot.shape[-1] == 3 else quaternion_to_matrix(rot)
    T = torch.eye(4, device=rot.device, dtype=rot.dtype)[None].repeat(rot.shape[0], 1, 1)
    T[:, :3, :3] = R
    T[:, :3, 3] = trans
    return T


def rays_from_pose(dir_cam, c2w):
    """coslam.py:192-193 / :546-547: rays_o = t repeated, rays_d = sum_k dir_cam[k] R[:, k];  c2w [1,4,4] or [4,4]."""
    c2w = c2w.reshape(-1, 4, 4)
    rays_o = c2w[..., :3, -1].repeat(dir_cam.shape[0], 1)
    rays_d = torch.sum(dir_cam[..., None, :] * c2w[:, :3, :3], -1)
    return rays_o, rays_d
