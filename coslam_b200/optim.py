"""Dense Adam for the map and pose parameters -- a drop-in for the `torch.optim.Adam(param_groups, betas=(0.9, 0.99))` of
CoSLAM.create_optimizer / get_pose_param_optim (coslam.py:610-634, 276-283), one fused pass per parameter
(cb200_adam_step).  Same per-group options (lr, betas, eps, weight_decay); amsgrad / maximize are not used by Co-SLAM and
are rejected.  State keys match torch's (`step`, `exp_avg`, `exp_avg_sq`), so a checkpointed optimizer state loads either way.
"""
import ctypes

import torch

from . import _lib
from ._lib import check, ptr, stream


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False, maximize=False,
                 fused_zero_grad=False):
        if amsgrad or maximize:
            raise NotImplementedError('coslam_b200.optim.Adam: amsgrad / maximize are not part of the Co-SLAM path')
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.fused_zero_grad = bool(fused_zero_grad)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        for group in self.param_groups:
            b1, b2 = group['betas']
            for p in group['params']:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError('coslam_b200.optim.Adam: parameters must live on a CUDA device (no CPU fallback)')
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError('coslam_b200.optim.Adam: parameters must be contiguous fp32')
                g = p.grad
                if g.is_sparse:
                    raise RuntimeError('coslam_b200.optim.Adam: dense gradients only')
                if not g.is_contiguous():
                    g = g.contiguous()
                    p.grad = g
                st = self.state[p]
                if len(st) == 0:
                    st['step'] = torch.zeros((), dtype=torch.float32)          # host scalar, like torch's non-capturable Adam
                    st['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st['step'] += 1
                check(lib.cb200_adam_step(ptr(p), ptr(g), ptr(st['exp_avg']), ptr(st['exp_avg_sq']), ctypes.c_int64(p.numel()),
                                          ctypes.c_float(group['lr']), ctypes.c_float(b1), ctypes.c_float(b2),
                                          ctypes.c_float(group['eps']), ctypes.c_float(group['weight_decay']), None,
                                          ctypes.c_int64(int(st['step'])), 1 if self.fused_zero_grad else 0, stream()), 'adam_step')
        return loss
