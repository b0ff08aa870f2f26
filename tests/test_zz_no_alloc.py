"""qsv_inner_static_scratch (include/qsv.h; csrc/qsv_noalloc.cu): the allocation-free variant of qsv_inner
whose per-block partial sums live in a file-scope `static __device__` array.  No counterpart in the
reference.

Collected last on purpose: a failure here hides no other test.
Called through the C-ABI with raw device pointers (the operands live in torch tensors: torch is only the
owner of the device memory here) and compared with the existing entry point qsv_inner on handles holding
the same data, within the amplitude / expectation-value tolerance of the other GPU tests (1e-10), on
calls in a row that differ in operands AND in grid size, so a call that read the previous call's partial
sums or found a stale arrival ticket would fail.
"""
import ctypes as C

import numpy as np
import pytest

from qiskit_b200 import _lib

pytestmark = pytest.mark.gpu
ATOL = 1e-10


def _random_state(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.standard_normal(1 << n) + 1j * rng.standard_normal(1 << n)
    return np.ascontiguousarray(v / np.linalg.norm(v), dtype=np.complex128)


def _inner_by_handles(lib, a, b, n):
    """The existing entry point: <a|b> of two handles (statevector.py:399-413)."""
    ha, hb = C.c_void_p(), C.c_void_p()
    _lib.check(lib.qsv_create(C.byref(ha), n, 0))
    _lib.check(lib.qsv_create(C.byref(hb), n, 0))
    try:
        for h, v in ((ha, a), (hb, b)):
            _lib.check(lib.qsv_upload(h, v.ctypes.data_as(C.POINTER(C.c_double)), 0, v.size))
        out = (C.c_double * 2)()
        _lib.check(lib.qsv_inner(ha, hb, out))
        return complex(out[0], out[1])
    finally:
        lib.qsv_destroy(ha)
        lib.qsv_destroy(hb)


def test_static_scratch_inner_matches_qsv_inner_on_calls_in_a_row():
    import torch

    lib = _lib.load()
    _lib.drain_pool()
    dev = torch.device("cuda", 0)
    # (qubits, seeds): a large grid, the same operands swapped (conjugate result), then a SMALLER grid
    # (16 blocks after 1024: stale partials of the earlier calls sit right behind the live ones), then large again
    cases = [(20, 1, 2, False), (20, 1, 2, True), (12, 3, 4, False), (21, 5, 6, False), (1, 7, 8, False)]
    out = torch.full((2,), float("nan"), dtype=torch.float64, device=dev)
    for n, sa, sb, swap in cases:
        a, b = _random_state(n, sa), _random_state(n, sb)
        if swap:
            a, b = b, a
        want = _inner_by_handles(lib, a, b, n)
        assert abs(want - np.vdot(a, b)) < ATOL          # the existing entry point against NumPy
        ta = torch.from_numpy(a.view(np.float64)).to(dev)
        tb = torch.from_numpy(b.view(np.float64)).to(dev)
        torch.cuda.synchronize(dev)
        for repeat in range(2):                          # the same call twice in a row as well
            out.fill_(float("nan"))
            rc = lib.qsv_inner_static_scratch(0, ta.data_ptr(), tb.data_ptr(), a.size, out.data_ptr(), None)
            assert rc == _lib.QSV_OK, _lib.last_error()
            torch.cuda.synchronize(dev)
            got = out.cpu().numpy()
            got = complex(got[0], got[1])
            assert abs(got - want) < ATOL, (n, swap, repeat, got, want)
        # <a|a> = 1 through the same path (both operands the same pointer)
        rc = lib.qsv_inner_static_scratch(0, ta.data_ptr(), ta.data_ptr(), a.size, out.data_ptr(), None)
        assert rc == _lib.QSV_OK, _lib.last_error()
        torch.cuda.synchronize(dev)
        got = out.cpu().numpy()
        assert abs(got[0] - 1.0) < ATOL and abs(got[1]) < ATOL
        del ta, tb


def test_static_scratch_inner_on_a_side_stream_and_empty_input():
    import torch

    lib = _lib.load()
    dev = torch.device("cuda", 0)
    n = 18
    a, b = _random_state(n, 11), _random_state(n, 12)
    want = np.vdot(a, b)
    ta = torch.from_numpy(a.view(np.float64)).to(dev)
    tb = torch.from_numpy(b.view(np.float64)).to(dev)
    out = torch.zeros(2, dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    side = torch.cuda.Stream(device=dev)
    for _ in range(2):
        rc = lib.qsv_inner_static_scratch(0, ta.data_ptr(), tb.data_ptr(), a.size, out.data_ptr(),
                                          C.c_void_p(side.cuda_stream))
        assert rc == _lib.QSV_OK, _lib.last_error()
        side.synchronize()
        got = out.cpu().numpy()
        assert abs(complex(got[0], got[1]) - want) < ATOL
    # count = 0: the sum over nothing
    rc = lib.qsv_inner_static_scratch(0, ta.data_ptr(), tb.data_ptr(), 0, out.data_ptr(), None)
    assert rc == _lib.QSV_OK, _lib.last_error()
    torch.cuda.synchronize(dev)
    assert out.cpu().numpy().tolist() == [0.0, 0.0]


def test_static_scratch_inner_rejects_sizes_beyond_its_maximum():
    import torch

    lib = _lib.load()
    dev = torch.device("cuda", 0)
    t = torch.zeros(4, dtype=torch.float64, device=dev)
    out = torch.full((2,), 7.0, dtype=torch.float64, device=dev)
    rc = lib.qsv_inner_static_scratch(0, t.data_ptr(), t.data_ptr(), (1 << 30) + 1, out.data_ptr(), None)
    assert rc == _lib.ERR_INVALID
    msg = _lib.last_error()
    assert "maximum" in msg and "2^30" in msg, msg
    torch.cuda.synchronize(dev)
    assert out.cpu().numpy().tolist() == [7.0, 7.0]        # nothing was launched
    # NULL and misaligned pointers are refused before any launch as well
    assert lib.qsv_inner_static_scratch(0, None, t.data_ptr(), 2, out.data_ptr(), None) == _lib.ERR_INVALID
    assert lib.qsv_inner_static_scratch(0, t.data_ptr() + 8, t.data_ptr(), 1, out.data_ptr(), None) == _lib.ERR_INVALID
