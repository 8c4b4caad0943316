"""The allocation-free gradient entry point, exercised through the C ABI.

``dapb_lr_grad_noalloc`` is meant for hosts that cannot allocate device memory at the moment they need a
gradient (their own allocator owns the whole GPU, they sit in a loop that must not trigger cudaMalloc's
device-wide synchronisation, or they are a callback that is not allowed to allocate).  Its grid-reduction
scratch is a fixed array inside the library, shared by all calls and never cleared between them, so this
test makes two different calls in a row and checks each against the ordinary entry point
(``dapb_lr_step`` with caller-provided scratch) -- a stale ticket or stale partial sums from the first call
would corrupt the second.  The file sorts last on purpose: if it fails, ``pytest -x`` has already run
everything else.  It must pass both in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from distributed_algorithm_pyspark_b200 import _native

pytestmark = pytest.mark.gpu


def _bind(lib):
    fn = lib.dapb_lr_grad_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_uint64,
                   C.c_uint32, C.c_uint32, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dapb_lr_grad_noalloc_scratch_bytes.restype = C.c_uint64
    return fn


def _reference_entry(lib, X, y, w, D, dtype_id, seed, iteration, thr, eps, row_base):
    """The existing entry point with explicit scratch, same grid as the built-in one."""
    dev = X.device
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    grid = min(2 * sms, 296)
    partials = torch.empty(grid, D, dtype=torch.float32, device=dev)
    counts = torch.zeros(grid, dtype=torch.int32, device=dev)
    ticket = torch.zeros(1, dtype=torch.int32, device=dev)
    g = torch.zeros(D, dtype=torch.float32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    a = _native.DapbLrStepArgs()
    a.X, a.y, a.n_rows, a.row_base, a.D, a.dtype = X.data_ptr(), y.data_ptr(), X.shape[0], row_base, D, dtype_id
    a.mask, a.seed, a.iteration, a.threshold24, a.eps = 0, seed, iteration, thr, eps
    a.w_fwd, a.partials, a.counts, a.ticket, a.grid = w.data_ptr(), partials.data_ptr(), counts.data_ptr(), ticket.data_ptr(), grid
    a.fin, a.g_out, a.cnt_out = _native.FIN_GRAD_ONLY, g.data_ptr(), cnt.data_ptr()
    comm = _native.DapbComm()
    comm.world = 1
    _native.check(lib.dapb_lr_step(C.byref(a), C.byref(comm), _native.stream_ptr()), "lr_step")
    torch.cuda.synchronize()
    return g, int(cnt.item())


def test_noalloc_gradient_matches_existing_entry_point_twice_in_a_row():
    lib = _native.load()
    fn = _bind(lib)
    assert 1_000_000 < lib.dapb_lr_grad_noalloc_scratch_bytes() < 2_000_000      # the stated, modest size
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(123)
    thr = int(0.2 * (1 << 24))
    # two calls back to back with different shapes / dtypes / iterations: the second reuses the dirty scratch
    cases = [(torch.bfloat16, 1, 1024, 40000, 5, 0.0, 777), (torch.float32, 0, 200, 9001, 6, 1e-6, 0)]
    outs = []
    for dtype, dtype_id, D, n, it, eps, row_base in cases:
        X = (torch.randn(n, D, device=dev) * 0.5).to(dtype).contiguous()
        y = (torch.rand(n, device=dev) < 0.4).float()
        w = (torch.randn(D, device=dev) * 0.05).contiguous()
        g = torch.full((D,), float("nan"), dtype=torch.float32, device=dev)
        cnt = torch.full((1,), -1, dtype=torch.int32, device=dev)
        rc = fn(X.data_ptr(), y.data_ptr(), n, row_base, D, dtype_id, None, 42, it, thr, eps, w.data_ptr(),
                g.data_ptr(), cnt.data_ptr(), _native.stream_ptr())
        assert rc == 0, lib.dapb_last_error().decode()
        outs.append((X, y, w, D, dtype_id, it, eps, row_base, g, cnt))
    torch.cuda.synchronize()
    for X, y, w, D, dtype_id, it, eps, row_base, g, cnt in outs:
        g_ref, cnt_ref = _reference_entry(lib, X, y, w, D, dtype_id, 42, it, thr, eps, row_base)
        assert int(cnt.item()) == cnt_ref and cnt_ref > 0
        np.testing.assert_allclose(g.cpu().numpy(), g_ref.cpu().numpy(), rtol=2e-3,
                                   atol=2e-3 * float(g_ref.abs().max()))
        # and against plain PyTorch fp32 math of the same op
        from distributed_algorithm_pyspark_b200.utils.philox import philox_row_mask_torch
        keep = philox_row_mask_torch(X.shape[0], 0.2, 42, it, row_base, device=dev).float()
        Xf = X.float()
        r = (1.0 / (torch.exp(-(Xf @ w)) + 1.0 + eps) - y) * keep
        g_t = r @ Xf
        np.testing.assert_allclose(g.cpu().numpy(), g_t.cpu().numpy(), rtol=2e-3, atol=2e-3 * float(g_t.abs().max()))


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch():
    lib = _native.load()
    fn = _bind(lib)
    dev = torch.device("cuda", torch.cuda.current_device())
    D = 1025
    X = torch.zeros(8, D, device=dev)
    y = torch.zeros(8, device=dev)
    w = torch.zeros(D, device=dev)
    g = torch.zeros(D, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = fn(X.data_ptr(), y.data_ptr(), 8, 0, D, 0, None, 0, 0, 1 << 24, 0.0, w.data_ptr(), g.data_ptr(), cnt.data_ptr(),
            _native.stream_ptr())
    assert rc != 0
    assert "1024" in lib.dapb_last_error().decode()
