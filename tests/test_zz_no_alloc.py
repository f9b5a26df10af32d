"""-m gpu: the allocation-free entry point `dcvc_quad_step_enc_static` (static device scratch
inside the library) against the existing `dcvc_quad_step_enc` (caller workspace), through the C
ABI, on two calls in a row.  Collected last on purpose (file name).
Integer outputs (symbols, CDF indexes) must be bit-equal; y_hat / y_out bit-equal;
the bit estimate is the same fixed-order reduction, so it must be bit-equal too."""
import ctypes

import pytest
import torch

from dcvc_b200 import _lib
from dcvc_b200 import entropy_models as em

pytestmark = pytest.mark.gpu


def _view(t, c, coff=0):
    """NHWC fp32 tensor [h, w, cstride] -> dcvc_view of channels coff..coff+c"""
    h, w, cs = t.shape
    return _lib.View(t.data_ptr(), h, w, c, cs, coff)


def _inputs(C, h, w, seed, dev):
    g = torch.Generator().manual_seed(seed)
    y = (torch.randn(h, w, C, generator=g) * 6).to(dev)
    params = torch.cat((torch.rand(h, w, C, generator=g) * 2,
                        torch.exp(torch.rand(h, w, C, generator=g) * 9 - 5),
                        torch.randn(h, w, C, generator=g)), 2).contiguous().to(dev)
    prior = torch.cat((torch.exp(torch.rand(h, w, C, generator=g) * 9 - 5),
                       torch.randn(h, w, C, generator=g)), 2).contiguous().to(dev)
    return y, params, prior


def _run(static, y, params, prior, edges, step, dist):
    """One encoder step through the C ABI; every buffer is allocated HERE, by the caller, before
    the call -- the static entry point itself allocates nothing."""
    L = _lib.lib()
    h, w, C = y.shape
    dev = y.device
    n = C // 4 * h * w
    y_hat = torch.zeros(h, w, C, device=dev)
    y_out = torch.zeros(h, w, C, device=dev)
    sym = torch.full((n,), -7, dtype=torch.int16, device=dev)
    idx = torch.full((n,), 255, dtype=torch.uint8, device=dev)
    bits = torch.zeros(1, device=dev)
    q = _lib.Quad()
    q.y, q.params = _view(y, C), _view(params, 3 * C)
    q.prior = _view(prior, 2 * C) if step else _lib.View(None, 0, 0, 0, 0, 0)
    q.y_hat, q.y_out = _view(y_hat, C), _view(y_out, C)
    q.edges, q.step, q.distribution = edges.data_ptr(), step, dist
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        rc = L.dcvc_quad_step_enc_static(ctypes.byref(q), sym.data_ptr(), idx.data_ptr(),
                                         bits.data_ptr(), stream)
    else:
        ws = torch.empty(L.dcvc_quad_bits_ws_bytes(n) // 4 + 1, device=dev)
        rc = L.dcvc_quad_step_enc(ctypes.byref(q), sym.data_ptr(), idx.data_ptr(),
                                  bits.data_ptr(), ws.data_ptr(), stream)
    _lib.check(rc)
    torch.cuda.synchronize()
    return sym, idx, y_hat, y_out, bits


def test_static_scratch_matches_workspace_entry_twice():
    dev = torch.device("cuda:0")
    edges = em.index_edges("laplace").to(dev)
    # two calls in a row with different sizes and steps: the second (smaller) call finds the
    # first call's partials still in the shared array and must not be affected by them
    for C, h, w, step, seed in ((128, 68, 120, 1, 3), (64, 10, 14, 0, 4)):
        y, params, prior = _inputs(C, h, w, seed, dev)
        ref = _run(False, y, params, prior, edges, step, em.DIST_ID["laplace"])
        got = _run(True, y, params, prior, edges, step, em.DIST_ID["laplace"])
        for name, a, b in zip(("symbols", "indexes", "y_hat", "y_out", "bits"), ref, got):
            assert torch.equal(a, b), f"{name} differ (C={C}, {h}x{w}, step {step})"
        assert float(ref[4]) > 0


def test_static_scratch_rejects_oversize_and_does_not_allocate():
    L = _lib.lib()
    dev = torch.device("cuda:0")
    limit = L.dcvc_quad_step_enc_static_max()
    assert limit == 1 << 22
    # a step with more positions than the limit is refused before any launch; the views only
    # need plausible extents for that (no kernel runs, nothing is dereferenced)
    C, h, w = 256, 300, 300                       # C/4*h*w = 5.76 M > 4.19 M
    assert C // 4 * h * w > limit
    dummy = torch.zeros(16, device=dev)
    q = _lib.Quad()
    mk = lambda c: _lib.View(dummy.data_ptr(), h, w, c, c, 0)
    q.y, q.params, q.y_hat, q.y_out = mk(C), mk(3 * C), mk(C), mk(C)
    q.prior = _lib.View(None, 0, 0, 0, 0, 0)
    q.edges, q.step, q.distribution = dummy.data_ptr(), 0, 0
    before = L.dcvc_launch_count()
    rc = L.dcvc_quad_step_enc_static(ctypes.byref(q), None, None, dummy.data_ptr(), None)
    assert rc == -1 and L.dcvc_launch_count() == before
    assert b"static scratch limit" in L.dcvc_last_error()
    # no allocation by the call: device memory in use (driver view) is unchanged across a call
    edges = em.index_edges("laplace").to(dev)
    y, params, prior = _inputs(64, 8, 8, 0, dev)
    _run(True, y, params, prior, edges, 0, 0)      # set-up (symbol address) happens here
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    h2, w2, C2 = y.shape
    n = C2 // 4 * h2 * w2
    bufs = [torch.zeros(h2, w2, C2, device=dev) for _ in range(2)]
    sym = torch.zeros(n, dtype=torch.int16, device=dev)
    idx = torch.zeros(n, dtype=torch.uint8, device=dev)
    bits = torch.zeros(1, device=dev)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    q = _lib.Quad()
    q.y, q.params = _view(y, C2), _view(params, 3 * C2)
    q.prior = _lib.View(None, 0, 0, 0, 0, 0)
    q.y_hat, q.y_out = _view(bufs[0], C2), _view(bufs[1], C2)
    q.edges, q.step, q.distribution = edges.data_ptr(), 0, 0
    _lib.check(L.dcvc_quad_step_enc_static(ctypes.byref(q), sym.data_ptr(), idx.data_ptr(),
                                           bits.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    free2, _ = torch.cuda.mem_get_info()
    assert free2 == free1, (free0, free1, free2)
