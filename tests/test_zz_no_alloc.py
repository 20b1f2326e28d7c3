"""fwi_trace_misfit_noalloc (csrc/misfit.cu): the allocation-free variant of fwi_trace_misfit, whose partials live in a
file-scope `static __device__` array instead of a caller-provided workspace.  No counterpart in the reference.
Called through the C ABI and checked against the existing entry point, two calls
in a row (the second on other data of another shape: the shared array keeps the first call's contents).

This file is collected last on purpose (a failure here hides nothing behind it)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _case(nt, S, R, seed, with_direct):
    g = torch.Generator().manual_seed(seed)
    pred = (torch.randn(nt, S, R, generator=g) * torch.rand(1, S, R, generator=g) * 3.0).to(DEV)
    direct = (0.3 * torch.randn(nt, S, R, generator=g)).to(DEV) if with_direct else None
    obs = torch.randn(nt, S, R, generator=g).clamp(-1, 1).to(DEV)
    return pred, direct, obs


def _both(L, pred, direct, obs, kind, normalize):
    """(loss of fwi_trace_misfit with a workspace of the caller's, loss of fwi_trace_misfit_noalloc)"""
    from physicsbasedfwi2_b200 import _lib
    nt, M = int(pred.shape[0]), pred[0].numel()
    d = _lib.MisfitDesc(nt, M, kind, 1 if normalize else 0, 1e-10)
    st = C.c_void_p(torch.cuda.current_stream(pred.device).cuda_stream)
    ws = torch.empty(int(L.fwi_trace_misfit_workspace_bytes(C.byref(d))), dtype=torch.uint8, device=DEV)
    ref = torch.full((1,), float("nan"), device=DEV)
    out = torch.full((1,), float("nan"), device=DEV)
    assert L.fwi_trace_misfit(C.byref(d), _p(pred), _p(direct), _p(obs), _p(ref), _p(ws), st) == 0, L.fwi_last_error()
    assert L.fwi_trace_misfit_noalloc(C.byref(d), _p(pred), _p(direct), _p(obs), _p(out), st) == 0, L.fwi_last_error()
    torch.cuda.synchronize()
    return float(ref), float(out)


def test_noalloc_misfit_matches_the_workspace_entry_point_on_two_calls_in_a_row(cuda_lib):
    L = cuda_lib
    # first call: C2-shaped traces (6 shots x 576 receivers), L1 of normalised traces with direct-wave subtraction
    a = _both(L, *_case(3000, 6, 576, 1, True), kind=0, normalize=True)
    # second call straight after it: fewer, shorter traces, MSE, no direct wave - reads the array the first call left
    b = _both(L, *_case(211, 3, 37, 2, False), kind=1, normalize=True)
    c = _both(L, *_case(500, 2, 101, 3, True), kind=0, normalize=False)
    for ref, out in (a, b, c):
        assert ref == ref and ref > 0.0                         # a loss was written
        assert abs(out - ref) <= 2e-6 * abs(ref), (out, ref)    # the tolerance of tests/test_misfit_gpu.py (same kernels:
        #                                                         the two are expected to agree to the bit)


def test_noalloc_misfit_rejects_sizes_beyond_the_static_array(cuda_lib):
    from physicsbasedfwi2_b200 import _lib
    L = cuda_lib
    d = _lib.MisfitDesc(4000, 400000, 0, 1, 1e-10)               # 1 chunk x 400000 traces x 16 B = 6.4 MB > 4 MiB
    assert L.fwi_trace_misfit_workspace_bytes(C.byref(d)) > 4 * 1024 * 1024
    # rejected on the host before any pointer is touched
    assert L.fwi_trace_misfit_noalloc(C.byref(d), C.c_void_p(256), None, C.c_void_p(256), C.c_void_p(256), None) == -1
    msg = L.fwi_last_error()
    assert b"4194304" in msg and b"fwi_trace_misfit_noalloc" in msg, msg
