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
