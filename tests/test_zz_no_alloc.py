"""dg1d_error_norm_noalloc: the error norm for callers that are not able to allocate GPU memory.

Two kinds of user are meant: a monitoring hook that checks the error while a long run is under way and the
state buffers already occupy the card (an allocation at that point may fail and, while it runs, holds up every
stream on the device), and an application that has committed all device memory to its own pool, so that
nothing outside the pool may allocate.  For them the library keeps the partial sums of the reduction in a small
array of its own instead of asking for a scratch buffer.

Checked here, through the C ABI only: the result equals that of the ordinary entry point dg1d_error_norm bit
for bit, on two calls in a row whose sizes differ (the second must not see what the first left in the shared
array); a request larger than the built-in array is refused with a message and runs nothing.

This file sorts last on purpose: a run that stops at the first failure has then already run every other test.
It holds no state of its own, so it passes alone in a fresh process as well as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _norm_both_ways(torch, lib, check, dev, u, exact, V, w, dx):
    """Runs the ordinary and the allocation-free entry point on the same device data; returns the two results."""
    nc, nd = u.shape
    dev.pack(u.reshape(1, 1, nc, nd))
    ex = torch.from_numpy(np.ascontiguousarray(exact)).cuda()
    Vc = np.ascontiguousarray(V); wc = np.ascontiguousarray(w)
    P = C.POINTER(C.c_double)
    scratch = torch.empty(int(lib.dg1d_error_norm_scratch_doubles(dev.h)), dtype=torch.float64, device='cuda')
    out_a = torch.full((1,), -1.0, dtype=torch.float64, device='cuda')
    out_b = torch.full((1,), -2.0, dtype=torch.float64, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib.dg1d_error_norm(dev.h, dev.bufs[0].data_ptr(), ex.data_ptr(), Vc.ctypes.data_as(P), wc.ctypes.data_as(P),
                              len(wc), dx, scratch.data_ptr(), out_a.data_ptr(), stream))
    check(lib.dg1d_error_norm_noalloc(dev.h, dev.bufs[0].data_ptr(), ex.data_ptr(), Vc.ctypes.data_as(P),
                                      wc.ctypes.data_as(P), len(wc), dx, out_b.data_ptr(), stream))
    torch.cuda.synchronize()
    return float(out_a[0]), float(out_b[0])


def test_error_norm_noalloc_matches_error_norm_twice_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device; there is no CPU fallback')
    from fem_b200 import _lib
    from fem_b200.basis import Vandermonde
    from fem_b200.device import Dg1dDevice
    lib = _lib.load()
    assert lib.dg1d_error_norm_noalloc_max_doubles() == 8192
    rng = np.random.default_rng(11)
    results = []
    # first a mesh that fills many slots of the built-in array, then a small one that uses a few: stale
    # partial sums from the first call would show up in the second if they were ever read
    for nc, k in ((700_000, 2), (1_000, 3)):
        dev = Dg1dDevice('linear', 'upwind', k, nc, bc=('periodic', 'periodic'))
        xg, wg = np.polynomial.legendre.leggauss(k + 3)
        u = rng.standard_normal((nc, k + 1))
        exact = rng.standard_normal((1, nc, k + 3))
        a, b = _norm_both_ways(torch, lib, _lib.check, dev, u, exact, Vandermonde(k, xg), wg, 1.0 / nc)
        assert np.isfinite(a) and a > 0.0
        assert a == b, (nc, a, b)          # same kernels, same order of summation: not merely close, identical
        results.append(a)
    assert results[0] != results[1]


def test_error_norm_noalloc_rejects_what_does_not_fit():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device; there is no CPU fallback')
    from fem_b200 import _lib
    from fem_b200.device import Dg1dDevice
    lib = _lib.load()
    # 64 problems of 40000 cells: 157 blocks each -> 10048 partial sums > 8192
    dev = Dg1dDevice('linear', 'upwind', 1, 40_000, bc=('periodic', 'periodic'), nprob=64)
    assert lib.dg1d_error_norm_scratch_doubles(dev.h) > lib.dg1d_error_norm_noalloc_max_doubles()
    ex = torch.zeros(64 * 40_000 * 4, dtype=torch.float64, device='cuda')
    out = torch.full((64,), 7.0, dtype=torch.float64, device='cuda')
    V = np.ones((4, 2)); w = np.ones(4)
    P = C.POINTER(C.c_double)
    rc = lib.dg1d_error_norm_noalloc(dev.h, dev.bufs[0].data_ptr(), ex.data_ptr(), V.ctypes.data_as(P), w.ctypes.data_as(P),
                                     4, 1.0, out.data_ptr(), None)
    assert rc == -1
    msg = lib.dg1d_last_error().decode()
    assert '8192' in msg and 'dg1d_error_norm' in msg
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())        # nothing ran
