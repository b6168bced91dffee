"""GPU unit test of the tcgen05 plumbing (descriptors, TMEM allocation, bf16 hi/lo split, K-major and MN-major operand
views) against a float64 matmul.  The split-3 scheme has to reproduce fp32-class accuracy (~1e-5 relative)."""
import ctypes

import pytest
import torch

from coslam_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _run(a, b, k, n, mode):
    c = torch.empty((128, n), device=DEV)
    _lib.check(_lib.load().cb200_selftest_mma(_lib.ptr(a), _lib.ptr(b), k, n, mode, _lib.ptr(c), _lib.stream()), 'selftest_mma')
    torch.cuda.synchronize()
    return c


@pytest.mark.parametrize('k,n', [(16, 16), (32, 16), (32, 32), (64, 32), (80, 32), (96, 32), (32, 64), (32, 80), (32, 96)])
def test_k_major_product(k, n):
    torch.manual_seed(k * 100 + n)
    a = torch.randn(128, k, device=DEV) * torch.logspace(-3, 1, k, device=DEV)
    b = torch.randn(n, k, device=DEV)
    c = _run(a, b, k, n, 0)
    ref = a.double() @ b.double().t()
    err = float((c.double() - ref).abs().max() / ref.abs().max())
    assert err < 3e-5, err


@pytest.mark.parametrize('n', [16, 32])
def test_mn_major_product(n):
    torch.manual_seed(n)
    x = torch.randn(128, 128, device=DEV)
    d = torch.randn(128, n, device=DEV) * 1e-3
    c = _run(x, d, 128, n, 1)
    ref = x.double().t() @ d.double()
    err = float((c.double() - ref).abs().max() / ref.abs().max())
    assert err < 3e-5, err
