"""Pins the oracle's unequal-time Green's function (oracle/dqmc_oracle.cpp, ut_*) the way the reference's
own tests do (test/flavortests_DQMC.jl:73-116): equal-time limit against calculate_greens(mc, slice),
antiperiodicity G(t, 0) = -G(t, beta), plus the defining inverse evaluated with numpy where it is stable."""
import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc


def _setup(L=4, U=1.0, beta=1.0, dtau=0.1, safe_mult=5, seed=7):
    model = mcb.HubbardModelAttractive(L=L, U=U, mu=0.0)
    M = int(round(beta / dtau))
    hop = model.hopping_exponentials(dtau)
    conf = model.rand_conf(M, seed, 0)
    o = orc.OracleDQMC(L * L, M, safe_mult, dtau, U, hop, conf, seed=seed)
    return model, hop, conf, o, M


def _prod(model, hop, conf, hi, lo, dtau):
    """B_hi ... B_lo (identity when hi < lo)."""
    lam = model.lambda_(dtau)
    P = np.eye(conf.shape[0])
    for l in range(lo, hi + 1):
        P = (hop[2] @ np.diag(np.exp(lam * conf[:, l - 1]))) @ P
    return P


def direct_unequal(model, hop, conf, k, l, dtau):
    """unequal_time_stack.jl:449-450 (k >= l) and :540-541 (k <= l), plain inverses."""
    M = conf.shape[1]
    if k >= l:
        A = np.linalg.inv(_prod(model, hop, conf, k, l + 1, dtau))
        B = _prod(model, hop, conf, l, 1, dtau) @ _prod(model, hop, conf, M, k + 1, dtau)
        return np.linalg.inv(A + B)
    A = _prod(model, hop, conf, l, k + 1, dtau)
    B = np.linalg.inv(_prod(model, hop, conf, k, 1, dtau) @ _prod(model, hop, conf, M, l + 1, dtau))
    return -np.linalg.inv(A + B)


@pytest.mark.parametrize("safe_mult", [5, 2])
def test_unequal_matches_direct_inverse(safe_mult):
    model, hop, conf, o, M = _setup(beta=1.0, safe_mult=safe_mult)
    for k, l in [(0, 0), (M, M), (3, 3), (7, 2), (M, 0), (6, 5), (10, 4), (2, 7), (0, M), (4, 5), (0, 3), (5, 10)]:
        got = o.greens_unequal(k, l)
        want = direct_unequal(model, hop, conf, k, l, 0.1)
        assert np.allclose(got, want, atol=1e-11), (k, l, np.abs(got - want).max())


def test_equal_time_limit_and_antiperiodicity():
    # test/flavortests_DQMC.jl:103-115 (there: 6 sites, beta = 15, safe_mult = 5, 1e-14 / 1e-13)
    model, hop, conf, o, M = _setup(beta=15.0, safe_mult=5, seed=11)
    for s in (0, 1, 4, 5, 6, 74, 75, 149, 150):
        G1 = o.greens_at(s)
        G2 = o.greens_unequal(s, s)
        assert np.abs(G1 - G2).max() < 1e-12, s
    for s in (0, 1, 5, 37, 100, 149):
        G1 = o.greens_unequal(s, 0)
        G2 = o.greens_unequal(s, M)
        assert np.allclose(G1, -G2, atol=1e-12), s


def test_physical_transform_applies():
    model, hop, conf, o, M = _setup()
    Ge = o.greens_unequal(6, 2)
    Gp = o.greens_unequal(6, 2, physical=True)
    assert np.allclose(Gp, hop[1] @ Ge @ hop[0], atol=1e-12)


def _chain_setup(n_sites=6, beta=15.0, dtau=0.1, safe_mult=5, seed=1):
    """HubbardModelAttractive(6, 1): the 6-site periodic chain of test/flavortests_DQMC.jl:74-75."""
    import scipy.linalg as sl
    T = np.zeros((n_sites, n_sites))
    for i in range(n_sites):
        T[i, (i + 1) % n_sites] -= 1.0
        T[(i + 1) % n_sites, i] -= 1.0
    hop = [sl.expm(-0.5 * dtau * T), sl.expm(0.5 * dtau * T), sl.expm(-dtau * T), sl.expm(dtau * T)]
    M = int(round(beta / dtau))
    conf = np.random.default_rng(seed).choice([-1, 1], size=(n_sites, M)).astype(np.int8)
    return hop, conf, orc.OracleDQMC(n_sites, M, safe_mult, dtau, 1.0, hop, conf, seed=seed), M


def _stabilised_triples(o, M, N):
    want = np.empty((M, 3, N, N))
    for l in range(1, M + 1):
        want[l - 1, 0] = o.greens_unequal(0, l, physical=True)
        want[l - 1, 1] = o.greens_unequal(l, 0, physical=True)
        want[l - 1, 2] = o.greens_unequal(l, l, physical=True)
    return want


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_combined_greens_iterator_reference_case(seed):
    # test/flavortests_DQMC.jl:121-160 on its own lattice: recalculate = safe_mult ("high precision", 1e-14 there)
    # and the default 4 safe_mult ("low precision", 1e-10), each step against greens(mc, 0, l), greens(mc, l, 0),
    # greens(mc, l, l)
    hop, conf, o, M = _chain_setup(seed=seed)
    o.build_stack()
    o.propagate_to_measurement_point()
    want = _stabilised_triples(o, M, 6)
    o.build_stack()
    o.propagate_to_measurement_point()
    assert np.abs(o.combined_greens(recalculate=5) - want).max() < 1e-12
    o.build_stack()
    o.propagate_to_measurement_point()
    assert np.abs(o.combined_greens() - want).max() < 1e-10


@pytest.mark.parametrize("L,beta,safe_mult", [(2, 15.0, 5), (4, 2.0, 4)])
def test_combined_greens_iterator_square_lattice(L, beta, safe_mult):
    # the same on the square lattices of the other tests. G(l <- l) between two recalculations is a similarity
    # transform B G B^-1 that amplifies rounding (to ~1e-9 at beta = 15 after 19 slices); the reference's own 1e-10
    # holds on its 6-site chain (test above) and for G(0 <- l), G(l <- 0) here.
    model, hop, conf, o, M = _setup(L=L, beta=beta, safe_mult=safe_mult, seed=3)
    N = L * L
    o.build_stack()
    o.propagate_to_measurement_point()
    st = o.state()
    assert st["current_slice"] == 1 and st["direction"] == 1
    want = _stabilised_triples(o, M, N)
    o.build_stack()
    o.propagate_to_measurement_point()
    assert np.abs(o.combined_greens(recalculate=safe_mult) - want).max() < 1e-12
    o.build_stack()
    o.propagate_to_measurement_point()
    lo = o.combined_greens()
    assert np.abs(lo[:, :2] - want[:, :2]).max() < 1e-10
    assert np.abs(lo[:, 2] - want[:, 2]).max() < 1e-8
    eff = o.combined_greens(physical=False)
    assert np.allclose(hop[1] @ eff[5, 1] @ hop[0], lo[5, 1], atol=1e-13)


def test_combined_greens_iterator_needs_measurement_point():
    model, hop, conf, o, M = _setup()
    o.build_stack()   # leaves the handle at current_slice == M on the way down
    with pytest.raises(RuntimeError):
        o.combined_greens()
