"""Where the feature-smoothness term's time goes (coslam.py:251-274): the stand-alone HashGrid backward over the regular
lattice in the reference's order (z fastest), with x fastest, and over random points.
    python tools/smooth_probe.py [tum_desk|room0|scannet0000]
B200: tum_desk 63^3: 359 us z fastest, 164 us x fastest, 315 us random in the lattice's box, 203 us random in the bound;
room0 31^3: 33.0 / 18.8 / 30.9 / 26.8 us."""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from coslam_b200 import _lib, synth
from coslam_b200._lib import check, ptr, stream
from coslam_b200.scene_rep import JointEncoding

DEV = 'cuda:0'
name = sys.argv[1] if len(sys.argv) > 1 else 'tum_desk'
cfg = synth.CONFIGS[name]()
bound = synth.bound_tensor(cfg).to(DEV)
model = JointEncoding(cfg, bound).to(DEV)
enc = model.embed_fn
t = cfg['training']
n = int(t['smooth_pts']) - 1
vox = float(t['smooth_vox'])
ext = bound[:, 1] - bound[:, 0]
idx = torch.arange(n, device=DEV, dtype=torch.float32)
lat = torch.stack(torch.meshgrid(idx, idx, idx, indexing='ij'), -1).reshape(-1, 3) * vox        # z fastest, as the reference builds it
x_lat = ((lat + 0.3) / ext).contiguous()
x_xfast = x_lat.view(n, n, n, 3).permute(2, 1, 0, 3).reshape(-1, 3).contiguous()
x_rand = torch.rand_like(x_lat)
x_box = (torch.rand_like(x_lat) * (n * vox) / ext).contiguous()                                  # random inside the lattice's box
P = x_lat.shape[0]
dy = torch.randn(P, 32, device=DEV)
f = torch.empty(P, 32, device=DEV)
dp = torch.zeros_like(enc.params)
tv = torch.zeros(1, dtype=torch.float64, device=DEV)
lib = _lib.load()
g = enc.grid()


def timed(fn, label):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print('%-40s %8.1f us' % (label, e0.elapsed_time(e1) * 1e3 / 20))


def bwd(x):
    return lambda: check(lib.cb200_hash_bwd(ctypes.byref(g), ptr(x), ptr(enc.params), ptr(dy), ctypes.c_int64(P), ptr(dp), None, stream()), 'b')


def fwd(x):
    return lambda: check(lib.cb200_hash_fwd(ctypes.byref(g), ptr(x), ptr(enc.params), ctypes.c_int64(P), ptr(f), stream()), 'f')


print('%s: lattice %d^3 = %d points, spacing %.3f m, table %d entries' % (name, n, P, vox, enc.params.numel() // 2))
timed(bwd(x_lat), 'hash_bwd lattice (z fastest)')
timed(bwd(x_xfast), 'hash_bwd lattice (x fastest)')
timed(bwd(x_box), 'hash_bwd random in the lattice box')
timed(bwd(x_rand), 'hash_bwd random in the whole bound')
timed(fwd(x_lat), 'hash_fwd lattice (z fastest)')
timed(fwd(x_xfast), 'hash_fwd lattice (x fastest)')
timed(lambda: check(lib.cb200_smooth_tv(ptr(f), ctypes.c_int64(n), ctypes.c_int64(32), ctypes.c_float(1e-6), ptr(dy), ptr(tv), stream()), 'tv'),
      'smooth_tv')
