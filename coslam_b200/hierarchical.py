"""Importance (hierarchical) pass of render_rays -- model/scene_rep.py:262-295 with sample_pdf (model/utils.py:32-71).

Off in every shipped YAML (n_importance: 0) but part of the JointEncoding path: the coarse
samples are rendered, `n_importance` extra depths are drawn per ray from the coarse weights, and the merged, sorted samples
go through the fused point kernels a second time.  The loss head then adds the coarse colour / depth terms
(model/scene_rep.py:327-329).  Everything per-point and per-ray runs in the library's kernels; what is left here is glue.
"""
import torch

from . import ops


def draw_uniforms(n_rays, n_importance, device, det):
    """The `u` of sample_pdf: a linspace row when perturb == 0 (utils.py:50), else device uniforms (utils.py:53)."""
    if det:
        return torch.linspace(0. + 0.5 / n_importance, 1. - 0.5 / n_importance, steps=n_importance, device=device)
    return torch.rand((n_rays, n_importance), device=device)


def _resample(model, z_vals, weights, target_d=None, counts=None):
    t = model.config['training']
    u = draw_uniforms(z_vals.shape[0], t['n_importance'], z_vals.device, det=(t['perturb'] == 0.))
    return ops.importance_sample(model.spec, z_vals.detach(), weights.detach(), u, target_d=target_d, counts=counts)


def importance_pass(model, rays_o, rays_d, z_vals, weights, coarse):
    """Second pass for render_rays; returns (ret, z_vals, raw) with the coarse maps kept under the '*0' keys."""
    z_all, z_std = _resample(model, z_vals, weights)
    raw = model.ray_points(rays_o, rays_d, z_all)
    rgb_map, depth_map, aux, _ = ops.RenderFunction.apply(raw, z_all, model.spec)
    ret = {'rgb': rgb_map, 'depth': depth_map, 'disp_map': aux[1], 'acc_map': aux[2], 'depth_var': aux[0],
           'rgb0': coarse['rgb'], 'disp0': coarse['disp_map'], 'acc0': coarse['acc_map'], 'depth0': coarse['depth'],
           'depth_var0': coarse['depth_var'], 'z_std': z_std}
    return ret, z_all, raw


def _global_sum(local, group):
    """Value = sum over ranks, gradient = this rank's (the gradient all-reduce sums the ranks later)."""
    if group is None:
        return local
    import torch.distributed as dist
    tot = local.detach().clone()
    dist.all_reduce(tot, group=group)
    return local + (tot - local.detach())


def forward_with_importance(model, rays_o, rays_d, target_rgb, target_d):
    """Training forward of JointEncoding with n_importance > 0 (model/scene_rep.py:297-351)."""
    spec, group = model.spec, model.process_group
    ws = ops.new_step_workspace(rays_o.device)
    counts = ws[:4].view(torch.int64)
    z0 = model._sample(target_d)                                   # the mask counts belong to the FINAL samples
    raw0 = model.ray_points(rays_o, rays_d, z0)
    rgb0, depth0, _, weights0 = ops.RenderFunction.apply(raw0, z0, spec)
    z_all, _ = _resample(model, z0, weights0, target_d=target_d, counts=counts)
    raw = model.ray_points(rays_o, rays_d, z_all)
    n_total = rays_o.shape[0]
    if group is not None:
        import torch.distributed as dist
        n_total = n_total * dist.get_world_size(group)
    rgb, depth, rgb_loss, depth_loss, sdf_loss, fs_loss = ops.RenderLossFunction.apply(
        raw, z_all, target_rgb, target_d, ws, spec, n_total, group)
    psnr = _mse2psnr(rgb_loss)
    # coarse terms (scene_rep.py:327-329); `counts` now holds the (global) number of valid-depth rays
    d = target_d.reshape(-1)
    valid = (d > 0.) & (d < spec.depth_trunc)
    w = valid if not spec.rgb_missing_nz else torch.ones_like(valid)   # bool-typed rgb_weight, scene_rep.py:320-321
    w = w.to(rgb0.dtype)[:, None]
    rgb_sum = _global_sum((((rgb0 - target_rgb) * w) ** 2).sum(), group)
    dep_sum = _global_sum((((depth0 - d) * valid.to(depth0.dtype)) ** 2).sum(), group)
    rgb_loss = rgb_loss + rgb_sum / float(3 * n_total)
    depth_loss = depth_loss + dep_sum / counts[2].to(torch.float32)
    return {'rgb': rgb, 'depth': depth, 'rgb_loss': rgb_loss, 'depth_loss': depth_loss, 'sdf_loss': sdf_loss,
            'fs_loss': fs_loss, 'psnr': psnr}


def _mse2psnr(x):
    from .scene_rep import mse2psnr
    return mse2psnr(x)
