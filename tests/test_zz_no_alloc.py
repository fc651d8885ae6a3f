"""mcb_linalg_udt_noalloc: the allocation-free variant of mcb_linalg_udt (static device scratch, device pointers in and
out; no counterpart in the reference). Through the C ABI, against the
existing entry point mcb_linalg_udt on the same inputs, on consecutive calls that share the static scratch.
Collected last on purpose (pytest -x at the end of a round must not hide other tests behind a failure here).
Device buffers come from torch, which is plumbing only: the library itself allocates nothing in this entry point."""
import ctypes as C

import numpy as np
import pytest

from montecarlo_jl_b200 import _lib

pytestmark = pytest.mark.gpu
TOL = 1e-12  # the kernels are the same ones mcb_linalg_udt launches at these sizes; the tests' UDT tolerance


def _cm(a):  # batch of numpy matrices -> [batch][col][row] buffer
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).transpose(0, 2, 1)).reshape(-1)


def _uncm(buf, b, n):
    return np.asarray(buf).reshape(b, n, n).transpose(0, 2, 1).copy()


def _existing(L, X, apply_pivot):
    b, n = X.shape[0], X.shape[1]
    U, T = np.empty(b * n * n), np.empty(b * n * n)
    D, piv = np.empty(b * n), np.empty(b * n, dtype=np.int32)
    _lib.check(L.mcb_linalg_udt(n, b, _cm(X), U, D, T, piv, apply_pivot, 0))
    return _uncm(U, b, n), D.reshape(b, n), _uncm(T, b, n), piv.reshape(b, n)


def _noalloc(L, torch, X, apply_pivot, check_free_memory=False):
    b, n = X.shape[0], X.shape[1]
    dev = torch.device("cuda", 0)
    Xd = torch.from_numpy(_cm(X)).to(dev)
    Ud = torch.full((b * n * n,), float("nan"), dtype=torch.float64, device=dev)
    Td = torch.full((b * n * n,), float("nan"), dtype=torch.float64, device=dev)
    Dd = torch.full((b * n,), float("nan"), dtype=torch.float64, device=dev)
    Pd = torch.full((b * n,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info(0)[0]   # free bytes on the device as the driver sees them
    rc = L.mcb_linalg_udt_noalloc(n, b, C.c_void_p(Xd.data_ptr()), C.c_void_p(Ud.data_ptr()),
                                  C.c_void_p(Dd.data_ptr()), C.c_void_p(Td.data_ptr()), C.c_void_p(Pd.data_ptr()),
                                  apply_pivot, 0, None)
    _lib.check(rc)
    torch.cuda.synchronize()
    if check_free_memory:   # kernels already loaded by the earlier calls: any cudaMalloc in the call would show here
        assert torch.cuda.mem_get_info(0)[0] == before
    return (_uncm(Ud.cpu().numpy(), b, n), Dd.cpu().numpy().reshape(b, n), _uncm(Td.cpu().numpy(), b, n),
            Pd.cpu().numpy().reshape(b, n))


def test_noalloc_udt_matches_the_existing_entry_point_on_consecutive_calls():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    L = _lib.lib()
    L.mcb_linalg_force_general(0)
    rng = np.random.default_rng(2024)
    # consecutive calls of different shape and content: the second and third run on scratch the earlier ones left
    cases = [(144, 3, 1), (400, 2, 0), (144, 8, 1), (144, 3, 1)]
    first = None
    for idx, (n, b, apply_pivot) in enumerate(cases):
        X = rng.standard_normal((b, n, n))
        X[-1] = X[-1] @ np.diag(np.logspace(0, -12, n))   # graded columns, like a long chain of slice matrices
        if (n, b) == (144, 3):
            if first is None:
                first = X.copy()
            else:
                X = first.copy()                           # the same input again after other calls used the scratch
        U0, D0, T0, p0 = _existing(L, X, apply_pivot)
        U1, D1, T1, p1 = _noalloc(L, torch, X, apply_pivot, check_free_memory=(idx == len(cases) - 1))
        assert np.array_equal(p0, p1), (n, b)
        assert np.abs(D1 - D0).max() <= TOL * np.abs(D0).max(), (n, b)
        assert np.abs(U1 - U0).max() <= TOL, (n, b)
        Tt0, Tt1 = (T0, T1) if apply_pivot else (np.triu(T0), np.triu(T1))
        assert np.abs(Tt1 - Tt0).max() <= TOL * max(1.0, np.abs(Tt0).max()), (n, b)
        for k in range(b):                                 # and it is a factorisation of X in its own right
            P = np.eye(n)
            if not apply_pivot:
                P = np.zeros((n, n))
                P[np.arange(n), p1[k]] = 1.0
            assert np.abs(U1[k] @ np.diag(D1[k]) @ Tt1[k] @ P - X[k]).max() < 1e-12 * np.abs(X[k]).max() * n


def test_noalloc_udt_rejects_sizes_beyond_the_static_scratch():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    L = _lib.lib()
    buf = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    p = C.c_void_p(buf.data_ptr())
    q = C.c_void_p(buf.data_ptr() + 8)
    r = C.c_void_p(buf.data_ptr() + 16)
    for n, b, what in ((144, 9, "batch"), (416, 1, "n must"), (100, 1, "n must"), (64, 1, "n must")):
        rc = L.mcb_linalg_udt_noalloc(n, b, p, q, r, r, r, 1, 0, None)
        assert rc == -1
        assert what in L.mcb_last_error().decode()
    assert L.mcb_linalg_udt_noalloc(144, 1, None, q, r, r, r, 1, 0, None) == -1
