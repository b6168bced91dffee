"""ORACLE (test infrastructure, not product code) -- pure-torch stand-in for the
`tinycudann` torch bindings on the Co-SLAM hot path.

tiny-cuda-nn is NOT vendored in the reference: the
reference imports it at model/encodings.py:3 and model/decoder.py:4 and installs
it from an unpinned git URL (requirements.txt:29; README.md:68 suggests commit
91ee479d275d322a65726435040fc20b56b9c991).  This file restates the published
algorithm of its `HashGrid` and `OneBlob` encodings (tiny-cuda-nn
include/tiny-cuda-nn/encodings/grid.h, oneblob.h, common_device.h) so that the
reference's own model/scene_rep.py can be imported UNMODIFIED on CPU with this
module injected as `sys.modules['tinycudann']`.

PARITY UNPINNED at this boundary: the reference ships no test or golden vector
for the tinycudann call sites (model/encodings.py:34-46, 63-71), and no real
tinycudann build exists in this image.  The only in-tree corroboration is the
parameter-count comments in configs/ScanNet/scene0000.yaml:8 (1.59M), which
`grid_meta` below reproduces exactly (tests/test_oracle.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this file.
"""
import math
import struct

import numpy as np
import torch
import torch.nn as nn

_PRIME1 = 2654435761
_PRIME2 = 805459861
_U32 = 0xFFFFFFFF


def _f32(x):
    """Round a python float to the nearest IEEE binary32, returned as a python float."""
    return struct.unpack('f', struct.pack('f', float(x)))[0]


def grid_meta(n_levels, log2_hashmap_size, base_resolution, per_level_scale, n_pos_dims=3):
    """Per-level scale, resolution, entry count and offsets.

    Follows tiny-cuda-nn grid.h `grid_scale` / `grid_resolution` and the offset-table loop of the
    GridEncodingTemplated constructor:
        scale_l = exp2f(l * log2f(per_level_scale)) * base - 1   (all fp32)
        res_l   = (uint32)ceilf(scale_l) + 1
        n_l     = min(round_up(res_l^3, 8), 2^T)
    """
    f = np.float32
    # exp2f / log2f are taken as the correctly rounded fp32 results (fp64 evaluation rounded once), which makes the
    # definition independent of the libm at hand; tiny-cuda-nn evaluates exp2f on the device (ex2.approx, <= 2 ulp),
    # so a real build may differ from this by 1-2 ulp of scale_l -- the ambiguity SURVEY.md section 9.7 records.
    log2_pls = f(np.log2(np.float64(f(per_level_scale))))
    scales, ress, sizes, offsets = [], [], [], [0]
    for l in range(n_levels):
        e = f(np.exp2(np.float64(f(f(l) * log2_pls))))
        scale = f(f(e * f(base_resolution)) - f(1.0))
        res = int(np.ceil(scale)) + 1
        max_params = (2 ** 32 - 1) // 2
        dense = res ** n_pos_dims
        n = max_params if float(dense) > float(max_params) else dense
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        scales.append(float(scale))
        ress.append(res)
        sizes.append(n)
        offsets.append(offsets[-1] + n)
    return {'scale': scales, 'res': ress, 'size': sizes, 'offset': offsets}


def _mul_u32(a, k):
    """(a * k) mod 2^32 for int64 tensors `a` in [0, 2^32) and python int k < 2^32, without int64 overflow."""
    lo = (a & 0xFFFF) * k
    hi = (((a >> 16) * k) & 0xFFFF) << 16
    return (lo + hi) & _U32


def hash_corner_indices(x, meta, level):
    """Corner indices [P, 8] (int64, values < size_l) and interpolation fractions [P, 3] for one level.

    grid.h `pos_fract` (pos = fmaf(scale, x, 0.5); floor; (uint32)(int) cast) and `grid_index`
    (dense stride walk while stride <= hashmap_size, else coherent-prime XOR hash; final `% size`).
    The fused multiply-add is reproduced by an exact fp64 product followed by one fp64 add and a
    rounding to fp32 (differs from a true fmaf only on a double-rounding tie, probability ~2^-29 and
    only for scale*x < 2^-6; oracle/hashgrid_ref.c uses a real fmaf and is cross-checked in tests).
    """
    scale = meta['scale'][level]
    res = meta['res'][level]
    size = meta['size'][level]
    pos = (x.detach().double() * scale + 0.5).float()          # [P,3] fp32
    cell_f = torch.floor(pos)
    frac = pos - cell_f                                         # fp32 subtract, exact
    cell = cell_f.to(torch.int64) & _U32                        # (uint32)(int) two's-complement wrap

    idx_list = []
    for c in range(8):
        p = [(cell[:, d] + ((c >> d) & 1)) & _U32 for d in range(3)]
        stride = 1
        index = torch.zeros_like(p[0])
        d = 0
        while d < 3 and stride <= size:
            index = (index + _mul_u32(p[d], stride & _U32)) & _U32
            stride = (stride * res) & _U32
            d += 1
        if size < stride:
            index = p[0] ^ _mul_u32(p[1], _PRIME1) ^ _mul_u32(p[2], _PRIME2)
        idx_list.append(index % size)
    return torch.stack(idx_list, dim=1), frac


class _HashGrid(object):
    def __init__(self, cfg, seed=1337):
        self.n_levels = int(cfg.get('n_levels', 16))
        self.n_features = int(cfg.get('n_features_per_level', 2))
        self.log2_hashmap_size = int(cfg.get('log2_hashmap_size', 19))
        self.base_resolution = int(cfg.get('base_resolution', 16))
        self.per_level_scale = float(cfg.get('per_level_scale', 2.0))
        self.meta = grid_meta(self.n_levels, self.log2_hashmap_size, self.base_resolution, self.per_level_scale)
        self.n_output_dims = self.n_levels * self.n_features
        n_params = self.meta['offset'][-1] * self.n_features
        g = torch.Generator().manual_seed(seed)
        # tcnn initialises grid parameters U(-1e-4, 1e-4) (bindings/torch modules.py + grid.h initialize_params)
        self.init_params = (torch.rand(n_params, generator=g) * 2.0 - 1.0) * 1e-4

    def __call__(self, x, params):
        F = self.n_features
        table = params.view(-1, F)
        outs = []
        for l in range(self.n_levels):
            idx, frac = hash_corner_indices(x, self.meta, l)
            scale = self.meta['scale'][l]
            # value = exact fp32 fraction; gradient d frac / d x = scale  (pos_derivative == 1 for linear interpolation)
            sx = x * scale
            w = frac + (sx - sx.detach())
            vals = table[self.meta['offset'][l] + idx]           # [P,8,F]
            acc = 0
            for c in range(8):
                wc = 1.0
                for d in range(3):
                    wd = w[:, d] if (c >> d) & 1 else 1.0 - w[:, d]
                    wc = wc * wd
                acc = acc + wc[:, None] * vals[:, c]
            outs.append(acc)
        return torch.cat(outs, dim=-1)


def _quartic_cdf(t, inv_radius):
    # common_device.h quartic_cdf: clamp(15/16 u (1 - 2/3 u^2 + 1/5 u^4) + 1/2, 0, 1), u = t * inv_radius
    u = t * inv_radius
    u2 = u * u
    u4 = u2 * u2
    return torch.clamp((15.0 / 16.0) * u * (1.0 - (2.0 / 3.0) * u2 + (1.0 / 5.0) * u4) + 0.5, 0.0, 1.0)


class _OneBlob(object):
    def __init__(self, n_input_dims, cfg):
        self.n_bins = int(cfg.get('n_bins', 16))
        self.n_input_dims = n_input_dims
        self.n_output_dims = n_input_dims * self.n_bins
        self.init_params = torch.zeros(0)

    def __call__(self, x, params=None):
        # oneblob.h one_blob_subwarp_aligned: left-edge CDF (3 periodic images) per bin; right edge = left edge of
        # the next bin (lane shuffle), the last bin's right edge = left edge of bin 0 plus 1.
        B = self.n_bins
        edges = torch.arange(B, dtype=x.dtype, device=x.device) / B           # [B]
        t = edges[None, None, :] - x[:, :, None]                                # [P,D,B]
        left = _quartic_cdf(t, float(B)) + _quartic_cdf(t - 1.0, float(B)) + _quartic_cdf(t + 1.0, float(B))
        right = torch.roll(left, shifts=-1, dims=-1)
        wrap = torch.zeros(B, dtype=x.dtype, device=x.device)
        wrap[-1] = 1.0
        out = (right + wrap) - left
        return out.reshape(x.shape[0], -1)


class Encoding(nn.Module):
    """Signature of tinycudann.Encoding as used at model/encodings.py:34-46 and 63-71."""

    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None):
        super().__init__()
        otype = encoding_config['otype'].lower()
        if otype in ('hashgrid', 'grid'):
            if otype == 'grid' and encoding_config.get('type', 'Hash') != 'Hash':
                raise NotImplementedError('oracle shim restates the Hash grid type only')
            assert n_input_dims == 3
            self.impl = _HashGrid(encoding_config, seed=seed)
        elif otype == 'oneblob':
            self.impl = _OneBlob(n_input_dims, encoding_config)
        else:
            raise NotImplementedError('oracle shim: encoding %r is not on the hot path' % encoding_config['otype'])
        self.n_input_dims = n_input_dims
        self.n_output_dims = self.impl.n_output_dims
        self.params = nn.Parameter(self.impl.init_params.clone())
        self.dtype = torch.float32

    def forward(self, x):
        return self.impl(x.float(), self.params)


class Network(nn.Module):
    """tinycudann.Network is reachable only with decoder.tcnn_network: True (model/decoder.py:26,79);
    every shipped config sets it False (configs/Replica/replica.yaml:60)."""

    def __init__(self, *a, **k):
        super().__init__()
        raise NotImplementedError('tcnn.Network (FullyFusedMLP) is outside the restated hot path')


def install():
    """Register this module as `tinycudann` so the reference's imports resolve."""
    import sys
    sys.modules['tinycudann'] = sys.modules[__name__]
