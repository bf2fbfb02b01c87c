"""vk_logsumexp_static: the no-allocation variant of vk_logsumexp_array (csrc/k_static.cu; scratch in a static
__device__ array, no cudaMalloc on the path).  Not part of the reference.
Checked through the C ABI against the existing entry points (vk_logsumexp_array, vk_logsumexp of a
particle handle) and numpy, on calls in a row that share the scratch array -- a large input, then a small one (stale
records of the first call must not leak into the second), then the edge cases of lite/utils.py:77-91.

This file sorts last on purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-12
MAX_N = 1 << 24


def _static(x_dev, n, out_dev, torch):
    from venturecxx_b200 import _lib
    stream = torch.cuda.current_stream().cuda_stream
    _lib.call("vk_logsumexp_static", ctypes.c_void_p(x_dev.data_ptr() if n else 0), n, ctypes.c_void_p(out_dev.data_ptr()),
              ctypes.c_void_p(stream))


def _array(x):
    from venturecxx_b200 import _lib
    out = np.zeros(1)
    x = np.ascontiguousarray(x, dtype=np.float64)
    _lib.call("vk_logsumexp_array", _lib.ptr(x), x.size, 0, _lib.ptr(out))
    return float(out[0])


def _same(a, b):
    if np.isnan(b):
        return np.isnan(a)
    if np.isinf(b):
        return a == b
    return abs(a - b) <= RTOL * max(1.0, abs(b))


def test_static_logsumexp_equals_the_allocating_entry_points():
    import torch
    import venturecxx_b200 as vb
    torch.cuda.set_device(0)
    rng = np.random.default_rng(5)
    out = torch.full((1,), 123.0, dtype=torch.float64, device="cuda")
    big = rng.normal(-700.0, 30.0, size=3_000_001)          # 733 tiles, a ragged last one
    small = rng.normal(5.0, 2.0, size=777)                    # one tile: 732 stale records stay behind it
    for x in (big, small, big[:4096], small[:1]):             # calls in a row on the shared scratch
        xd = torch.from_numpy(x).cuda()
        _static(xd, x.size, out, torch)
        torch.cuda.synchronize()
        got = float(out.item())
        m = x.max()
        assert _same(got, _array(x)) and _same(got, m + np.log(np.exp(x - m).sum())), (x.size, got)
    # the weights of a particle handle: vk_logsumexp on the handle vs the static variant on the same numbers
    ts = vb.B200TraceSet(vb.NormalNormalModel(0.0, 2.0, 1.0, 4), 50_000, seed=3)
    ts.observe(1.5)
    ts.incorporate()
    lw = np.asarray(ts.log_weights)
    _static(torch.from_numpy(lw).cuda(), lw.size, out, torch)
    torch.cuda.synchronize()
    assert _same(float(out.item()), ts.logsumexp())
    ts.close()


def test_static_logsumexp_edge_cases_and_limit():
    import torch
    from venturecxx_b200 import _lib
    from venturecxx_b200.exception import VentureException
    torch.cuda.set_device(0)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    inf = np.inf
    cases = [np.full(5000, -inf), np.array([1.0, np.nan, 2.0]), np.array([inf, 1.0]), np.array([inf, -inf]),
             np.array([-inf, 3.0, -inf]), np.concatenate([np.full(4096, -inf), [2.0, 2.0]]), np.zeros(0)]
    for x in cases:
        xd = torch.from_numpy(np.ascontiguousarray(x)).cuda() if x.size else torch.zeros(1, dtype=torch.float64, device="cuda")
        _static(xd, x.size, out, torch)
        torch.cuda.synchronize()
        assert _same(float(out.item()), _array(x)), (x[:4], float(out.item()), _array(x))
    # sizes beyond the static scratch are refused with the limit in the message; nothing is launched
    xd = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(VentureException) as e:
        _static(xd, MAX_N + 1, out, torch)
    assert str(MAX_N) in str(e.value) and "exceeds" in str(e.value)
    # the largest accepted size
    full = torch.full((MAX_N,), -3.0, dtype=torch.float64, device="cuda")
    _static(full, MAX_N, out, torch)
    torch.cuda.synchronize()
    assert _same(float(out.item()), -3.0 + np.log(float(MAX_N)))
    assert _lib is not None
