"""ORACLE (test infrastructure) -- ctypes binding of oracle/hashgrid_ref.c (built by `make -C oracle`)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, '_build', 'libcoslam_oracle.so')
_MAX_LEVELS = 32


class Grid(ctypes.Structure):
    _fields_ = [('n_levels', ctypes.c_uint32), ('n_features', ctypes.c_uint32),
                ('log2_hashmap_size', ctypes.c_uint32), ('base_resolution', ctypes.c_uint32),
                ('per_level_scale', ctypes.c_float),
                ('scale', ctypes.c_float * _MAX_LEVELS), ('res', ctypes.c_uint32 * _MAX_LEVELS),
                ('offset', ctypes.c_uint32 * (_MAX_LEVELS + 1))]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.isfile(_SO):
            subprocess.check_call(['make', '-C', _HERE, '-s'])
        _lib = ctypes.CDLL(_SO)
    return _lib


def _fp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def make_grid(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale):
    g = Grid()
    rc = lib().cbo_grid_init(ctypes.byref(g), n_levels, n_features, log2_hashmap_size, base_resolution,
                             ctypes.c_float(float(np.float32(per_level_scale))))
    assert rc == 0
    return g


def grid_meta(g):
    L = g.n_levels
    return {'scale': [g.scale[i] for i in range(L)], 'res': [g.res[i] for i in range(L)],
            'offset': [g.offset[i] for i in range(L + 1)]}


def hash_indices(g, x):
    x = _f32(x)
    P = x.shape[0]
    idx = np.empty((P, g.n_levels, 8), dtype=np.uint32)
    frac = np.empty((P, g.n_levels, 3), dtype=np.float32)
    lib().cbo_hash_indices(ctypes.byref(g), _fp(x), ctypes.c_int64(P), _fp(idx), _fp(frac))
    return idx, frac


def hash_fwd(g, x, table):
    x, table = _f32(x), _f32(table)
    P = x.shape[0]
    out = np.empty((P, g.n_levels * g.n_features), dtype=np.float32)
    lib().cbo_hash_fwd(ctypes.byref(g), _fp(x), _fp(table), ctypes.c_int64(P), _fp(out))
    return out


def oneblob_fwd(x, n_bins):
    x = _f32(x)
    P, D = x.shape
    out = np.empty((P, D * n_bins), dtype=np.float32)
    lib().cbo_oneblob_fwd(_fp(x), ctypes.c_int64(P), D, n_bins, _fp(out))
    return out


def sample_z(target_d, noise, lin_uniform, lin_guided, lin_nodepth):
    d = _f32(target_d).reshape(-1)
    lu, lg, ln = _f32(lin_uniform), _f32(lin_guided), _f32(lin_nodepth)
    N, S = d.shape[0], lu.shape[0] + lg.shape[0]
    z = np.empty((N, S), dtype=np.float32)
    nz = None if noise is None else _f32(noise)
    lib().cbo_sample_z(_fp(d), None if nz is None else _fp(nz), _fp(lu), _fp(lg), _fp(ln), ctypes.c_int64(N),
                       lu.shape[0], lg.shape[0], _fp(z))
    return z


def points(rays_o, rays_d, z, bb_lo, bb_hi):
    o, d, z = _f32(rays_o), _f32(rays_d), _f32(z)
    lo, hi = _f32(bb_lo), _f32(bb_hi)
    N, S = z.shape
    pts = np.empty((N, S, 3), dtype=np.float32)
    x = np.empty((N, S, 3), dtype=np.float32)
    lib().cbo_points(_fp(o), _fp(d), _fp(z), _fp(lo), _fp(hi), ctypes.c_int64(N), S, _fp(pts), _fp(x))
    return pts, x
