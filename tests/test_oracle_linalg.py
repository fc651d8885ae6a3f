"""Pins the CPU oracle's linear algebra against numpy/scipy.

Restates the reference's own property tests on random inputs:
  test/slice_matrices.jl:141-235  (vmul! variants, udt_AVX_pivot!, rdivp!)
  test/flavortests_DQMC.jl:44-69  (stack Green's function vs independent computation)
  test/slice_matrices.jl:36-82    (slice-matrix multiplies)
"""
import numpy as np
import pytest
import scipy.linalg

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc

RNG = np.random.default_rng(1234)


@pytest.mark.parametrize("n", [4, 16, 33])
def test_vmul_variants(n):
    A, B = RNG.standard_normal((n, n)), RNG.standard_normal((n, n))
    d = RNG.standard_normal(n)
    assert np.allclose(orc.vmul(A, B, 0), A @ B, rtol=1e-13, atol=1e-13)
    assert np.allclose(orc.vmul(A, B, 1), A @ B.T, rtol=1e-13, atol=1e-13)
    assert np.allclose(orc.vmul(A, B, 2), A.T @ B, rtol=1e-13, atol=1e-13)
    assert np.allclose(orc.vmul(A, d, 3), A * d[None, :])
    assert np.allclose(orc.vmul(d, B, 4), d[:, None] * B)


@pytest.mark.parametrize("n", [4, 16, 64])
def test_udt_pivot_reconstruction(n):
    # test/slice_matrices.jl:202-234
    X = RNG.standard_normal((n, n))
    U, D, T, piv = orc.udt_pivot(X, apply_pivot=True)
    assert np.allclose(U @ U.T, np.eye(n), atol=1e-13)
    assert np.allclose(U @ np.diag(D) @ T, X, atol=1e-12)
    assert np.all(np.diff(D) <= 1e-12 * D[0])  # graded, rank revealing

    U2, D2, T2, piv2 = orc.udt_pivot(X, apply_pivot=False)
    P = np.zeros((n, n))
    P[np.arange(n), piv2] = 1.0
    assert np.allclose(U2 @ np.diag(D2) @ np.triu(T2) @ P, X, atol=1e-12)
    # same factorisation as LAPACK's pivoted QR up to column signs of Q
    Q, R, p = scipy.linalg.qr(X, pivoting=True)
    assert np.array_equal(p, piv2)
    assert np.allclose(np.abs(np.diag(R)), D2, rtol=1e-11)

    u = RNG.standard_normal((n, n))
    got = orc.rdivp(u, T2, piv2)
    want = u @ P.T @ np.linalg.inv(np.triu(T2))
    assert np.allclose(got, want, rtol=1e-9, atol=1e-9)


def _setup(L=4, U=1.0, beta=1.0, dtau=0.1, safe_mult=5, seed=7, mu=0.0):
    model = mcb.HubbardModelAttractive(L=L, U=U, mu=mu)
    M = int(round(beta / dtau))
    hop = model.hopping_exponentials(dtau)
    conf = model.rand_conf(M, seed, 0)
    o = orc.OracleDQMC(L * L, M, safe_mult, dtau, U, hop, conf, seed=seed)
    return model, hop, conf, o, M


def _B(model, hop, conf, l, dtau):
    lam = model.lambda_(dtau)
    return hop[2] @ np.diag(np.exp(lam * conf[:, l - 1]))


def _greens_direct(model, hop, conf, slice_, dtau):
    """G_l = [1 + B_{l-1}...B_1 B_M...B_l]^-1  (test/testfunctions.jl:79-80), plain inverse."""
    M = conf.shape[1]
    N = conf.shape[0]
    P = np.eye(N)
    for l in list(range(slice_, M + 1)) + list(range(1, slice_)):
        P = _B(model, hop, conf, l, dtau) @ P
    return np.linalg.inv(np.eye(N) + P)


def test_slice_matrix_multiplies():
    model, hop, conf, o, M = _setup()
    N = conf.shape[0]
    dtau = 0.1
    X = RNG.standard_normal((N, N))
    for l in (1, 3, M):
        B = _B(model, hop, conf, l, dtau)
        Binv = np.diag(np.exp(-model.lambda_(dtau) * conf[:, l - 1])) @ hop[3]
        assert np.allclose(B @ Binv, np.eye(N), atol=1e-12)
        assert np.allclose(o.multiply_slice(l, X, 0), B @ X, atol=1e-12)
        assert np.allclose(o.multiply_slice(l, X, 1), X @ B, atol=1e-12)
        assert np.allclose(o.multiply_slice(l, X, 2), Binv @ X, atol=1e-12)
        assert np.allclose(o.multiply_slice(l, X, 3), X @ Binv, atol=1e-12)
        assert np.allclose(o.multiply_slice(l, X, 4), B.T @ X, atol=1e-12)


def test_stack_greens_matches_direct_inverse():
    # test/flavortests_DQMC.jl:44-69: after build_stack + propagate the stack G equals an
    # independently computed G; successive wraps stay close; from-scratch G at every slice.
    model, hop, conf, o, M = _setup()
    o.build_stack()
    st = o.state()
    assert st["current_slice"] == M and st["direction"] == -1
    assert np.allclose(o.greens(), _greens_direct(model, hop, conf, M, 0.1), atol=1e-11)
    for k in range(0, M + 1):
        want = _greens_direct(model, hop, conf, k + 1 if k < M else 1, 0.1) if k >= 0 else None
        # calculate_greens(mc, k) = [1 + B_k...B_1 B_M...B_{k+1}]^-1 = G_{k+1}
        assert np.allclose(o.greens_at(k), want, atol=1e-11), k


def test_sweep_keeps_greens_consistent_with_conf():
    model, hop, conf0, o, M = _setup(beta=2.0, safe_mult=5)
    o.build_stack()
    seen = []
    for u in range(2 * M):
        o.update(1)
        st = o.state()
        seen.append(st["current_slice"])
        G = o.greens()
        want = _greens_direct(model, hop, o.conf(), st["current_slice"], 0.1)
        assert np.allclose(G, want, atol=1e-9), (u, st)
    # schedule of §3.2: M-1..1, 1..M, M
    assert seen == list(range(M - 1, 0, -1)) + list(range(1, M + 1)) + [M]
    st = o.state()
    assert st["prop_local"] == 2 * M * 16 and 0 < st["acc_local"] < st["prop_local"]
    assert not np.array_equal(o.conf(), conf0)


def test_physical_greens_transform():
    # test/measurements.jl:162-184
    model, hop, conf, o, M = _setup()
    o.build_stack()
    G = o.greens()
    assert np.allclose(o.greens(physical=True), hop[1] @ G @ hop[0], atol=1e-13)


def test_injected_uniforms_control_acceptance():
    model, hop, conf, o, M = _setup()
    o.set_uniforms(np.array([2.0]))  # never below p unless p > 1
    o.build_stack()
    o.update(3)
    a_hi = o.state()["acc_local"]
    model, hop, conf, o2, M = _setup()
    o2.set_uniforms(np.array([0.0]))  # always accepted
    o2.build_stack()
    o2.update(3)
    assert o2.state()["acc_local"] == 3 * 16 and a_hi < 3 * 16
