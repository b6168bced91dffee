"""tinycudann-compatible lower boundary: `Encoding` for the two encodings Co-SLAM selects (HashGrid at
model/encodings.py:31-46, OneBlob at :61-71), backed by the sm_100a kernels of libcoslam_b200.so.

Same constructor arguments, `.params`, `.n_output_dims`, `forward(x[B,D]) -> [B,n_out]`, differentiable in `params`
and `x`.  `Network` exists so that `tcnn_network: True` fails with a clear message instead of an AttributeError.
CUDA-only: a CPU tensor raises (there is no fallback).
"""
import ctypes

import torch
import torch.nn as nn

from . import _lib
from ._lib import check, ptr, stream


def _f32c(t):
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


class _HashGridFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, params, module):
        xc, pc = _f32c(x), _f32c(params)
        g = module.grid()
        n = xc.shape[0]
        out = torch.empty((n, module.n_output_dims), dtype=torch.float32, device=xc.device)
        check(_lib.load().cb200_hash_fwd(ctypes.byref(g), ptr(xc), ptr(pc), ctypes.c_int64(n), ptr(out), stream()), 'hash_fwd')
        ctx.module = module
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(xc, pc)
        return out

    @staticmethod
    def backward(ctx, dy):
        if dy is None:
            return None, None, None
        xc, pc = ctx.saved_tensors
        dy = _f32c(dy)
        n = xc.shape[0]
        need_x, need_p = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        dx = torch.empty_like(xc) if need_x else None
        dp = torch.zeros_like(pc) if need_p else None
        check(_lib.load().cb200_hash_bwd(ctypes.byref(ctx.module.grid()), ptr(xc), ptr(pc), ptr(dy), ctypes.c_int64(n), ptr(dp),
                                         ptr(dx), stream()), 'hash_bwd')
        return dx, dp, None


class _OneBlobFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, n_bins):
        xc = _f32c(x)
        n, d = xc.shape
        out = torch.empty((n, d * n_bins), dtype=torch.float32, device=xc.device)
        check(_lib.load().cb200_oneblob_fwd(ptr(xc), ctypes.c_int64(n), d, n_bins, ptr(out), stream()), 'oneblob_fwd')
        ctx.n_bins = n_bins
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(xc)
        return out

    @staticmethod
    def backward(ctx, dy):
        if dy is None or not ctx.needs_input_grad[0]:
            return None, None
        xc, = ctx.saved_tensors
        n, d = xc.shape
        dx = torch.empty_like(xc)
        check(_lib.load().cb200_oneblob_bwd(ptr(xc), ptr(_f32c(dy)), ctypes.c_int64(n), d, ctx.n_bins, ptr(dx), stream()),
              'oneblob_bwd')
        return dx, None


class Encoding(nn.Module):
    """Drop-in for tinycudann.Encoding(n_input_dims, encoding_config, dtype=torch.float)."""

    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        super().__init__()
        if dtype not in (None, torch.float, torch.float32):
            raise NotImplementedError('coslam_b200.tcnn.Encoding computes in fp32 (the reference passes dtype=torch.float)')
        self.n_input_dims = int(n_input_dims)
        self.encoding_config = dict(encoding_config)
        self.seed = seed
        self.dtype = torch.float32
        otype = str(encoding_config['otype']).lower()
        self._grid = None
        if otype in ('hashgrid', 'grid'):
            if otype == 'grid' and str(encoding_config.get('type', 'Hash')).lower() != 'hash':
                raise NotImplementedError('only the Hash grid type is on the Co-SLAM hot path')
            if self.n_input_dims != 3:
                raise NotImplementedError('HashGrid kernels are written for 3-D inputs')
            self.kind = 'hash'
            self.grid_cfg = {'n_levels': int(encoding_config.get('n_levels', 16)),
                             'n_features_per_level': int(encoding_config.get('n_features_per_level', 2)),
                             'log2_hashmap_size': int(encoding_config.get('log2_hashmap_size', 19)),
                             'base_resolution': int(encoding_config.get('base_resolution', 16)),
                             'per_level_scale': float(encoding_config.get('per_level_scale', 2.0))}
            g = self.grid()
            self.n_output_dims = int(g.n_levels) * int(g.n_features)
            n_params = int(g.offset[g.n_levels]) * int(g.n_features)
            gen = torch.Generator().manual_seed(seed)
            # tiny-cuda-nn initialises grid parameters U(-1e-4, 1e-4)
            self.params = nn.Parameter((torch.rand(n_params, generator=gen) * 2.0 - 1.0) * 1e-4)
        elif otype == 'oneblob':
            self.kind = 'blob'
            self.n_bins = int(encoding_config.get('n_bins', 16))
            self.n_output_dims = self.n_input_dims * self.n_bins
            self.params = nn.Parameter(torch.zeros(0))
        else:
            raise NotImplementedError('encoding %r is not on the Co-SLAM hot path (model/encodings.py selects HashGrid '
                                      'and OneBlob in every shipped config)' % encoding_config['otype'])

    def grid(self):
        if self._grid is None:
            c = self.grid_cfg
            self._grid = _lib.grid_meta(c['n_levels'], c['n_features_per_level'], c['log2_hashmap_size'],
                                        c['base_resolution'], c['per_level_scale'])
        return self._grid

    def __getstate__(self):
        d = dict(self.__dict__)
        d['_grid'] = None               # ctypes block is rebuilt per process
        return d

    def __deepcopy__(self, memo):
        import copy
        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = None if k == '_grid' else copy.deepcopy(v, memo)
        return new

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError('coslam_b200.tcnn.Encoding has no CPU path (input is on %s)' % x.device)
        x2 = x.reshape(-1, self.n_input_dims)
        if self.kind == 'hash':
            return _HashGridFunction.apply(x2, self.params, self)
        return _OneBlobFunction.apply(x2, self.n_bins)


class _HalfMlpFunction(torch.autograd.Function):
    """bias-free ReLU MLP in tcnn's FullyFusedMLP arithmetic: fp16 operands and activations, fp32 accumulation inside each
    product, fp16 output; backward in fp16 under tcnn's loss scale of 128 (bindings/torch/tinycudann/modules.py), parameter
    gradients returned in fp32.  The products are library half GEMMs (cuBLAS through torch): this variant is config-reachable
    but unused by every shipped config, so it is served, not fused."""
    LOSS_SCALE = 128.0

    @staticmethod
    def forward(ctx, x, params, shapes):
        ws, off = [], 0
        for o, i in shapes:
            ws.append(params[off:off + o * i].view(o, i).to(torch.float16))
            off += o * i
        h = x.to(torch.float16)
        pad = shapes[0][1] - h.shape[-1]
        if pad:
            h = torch.nn.functional.pad(h, (0, pad))
        acts = [h]
        for k, w in enumerate(ws):
            h = h @ w.t()
            if k != len(ws) - 1:
                h = torch.relu_(h)
            acts.append(h)
        ctx.ws, ctx.acts, ctx.shapes, ctx.n_in, ctx.x_dtype = ws, acts[:-1], shapes, x.shape[-1], x.dtype
        return h

    @staticmethod
    def backward(ctx, dy):
        scale = _HalfMlpFunction.LOSS_SCALE
        g = (dy.float() * scale).to(torch.float16)
        d_params = []
        for k in range(len(ctx.ws) - 1, -1, -1):
            a = ctx.acts[k]
            d_params.append((g.t() @ a).float().reshape(-1) / scale)
            g = g @ ctx.ws[k]
            if k:
                g = g * (a > 0)
        d_x = (g[..., :ctx.n_in].float() / scale).to(ctx.x_dtype) if ctx.needs_input_grad[0] else None
        return d_x, torch.cat(d_params[::-1]), None


class Network(nn.Module):
    """tinycudann.Network with otype FullyFusedMLP (model/decoder.py:26-37, 79-90; reachable only with
    `decoder.tcnn_network: True`, which every shipped config sets False, configs/Replica/replica.yaml:60).  One flat fp32
    `params` (the master copy the optimiser steps), layers in order, each [out, in] row-major with the input width and the
    output width padded to multiples of 16 as FullyFusedMLP requires; Xavier-uniform init over the padded shapes (upstream
    behaviour, recalled -- parity at this boundary is unpinned).  Output is fp16, sliced to n_output_dims."""

    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        super().__init__()
        cfg = dict(network_config)
        if cfg.get('otype', 'FullyFusedMLP') not in ('FullyFusedMLP', 'CutlassMLP'):
            raise NotImplementedError('coslam_b200.tcnn.Network: otype %r' % cfg.get('otype'))
        if cfg.get('activation', 'ReLU') != 'ReLU' or cfg.get('output_activation', 'None') != 'None':
            raise NotImplementedError('coslam_b200.tcnn.Network serves ReLU hidden / linear output (model/decoder.py:30-31)')
        pad16 = lambda n: (n + 15) // 16 * 16
        width, hidden = int(cfg['n_neurons']), int(cfg['n_hidden_layers'])
        dims = [pad16(n_input_dims)] + [width] * hidden + [pad16(n_output_dims)]
        self.n_input_dims, self.n_output_dims = n_input_dims, n_output_dims
        self.shapes = tuple((dims[k + 1], dims[k]) for k in range(len(dims) - 1))
        g = torch.Generator().manual_seed(seed)
        chunks = []
        for o, i in self.shapes:
            bound = (6.0 / (o + i)) ** 0.5
            chunks.append((torch.rand(o * i, generator=g) * 2 - 1) * bound)
        self.params = nn.Parameter(torch.cat(chunks))

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError('coslam_b200.tcnn.Network has no CPU path (input is on %s)' % x.device)
        lead = x.shape[:-1]
        out = _HalfMlpFunction.apply(x.reshape(-1, self.n_input_dims), self.params, self.shapes)
        return out[:, :self.n_output_dims].reshape(*lead, self.n_output_dims)
