"""Pins the oracle's HubbardModelRepulsive path (oracle/dqmc_oracle.cpp: sweep_spatial_repulsive, orc_update_repulsive)
against plain numpy: the block Green's functions against the defining inverse with the spin-dependent slice
matrices (the reference's flavortests_DQMC.jl:44-69 idea), propose_local's determinant ratio against brute-force
determinants, accept_local! against a from-scratch Green's function of the flipped field, and the physics that tells
the repulsive from the attractive model (half filling, suppressed double occupancy)."""
import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc


def _setup(L=4, U=4.0, beta=1.0, dtau=0.1, safe_mult=5, seed=11):
    model = mcb.HubbardModelRepulsive(L=L, U=U)
    M = int(round(beta / dtau))
    hop = model.hopping_exponentials(dtau)
    conf = model.rand_conf(M, seed, 0)
    o = orc.OracleDQMCRepulsive(L * L, M, safe_mult, dtau, U, hop, conf, seed=seed)
    return model, hop, conf, o, M


def _direct_greens(model, hop, conf, slice_, dtau, sigma):
    """G_sigma(slice) = [1 + B_slice ... B_1 B_M ... B_{slice+1}]^-1 with
    B_l = e^{-dtau T} diag(exp(sigma lambda s_l))  (interaction_matrix_exp!, HubbardModelRepulsive.jl:112-124);
    the effective G of the stack lacks the outer e^{-+dtau T/2} like the attractive one."""
    lam = model.lambda_(dtau)
    N, M = conf.shape
    Bs = [hop[2] @ np.diag(np.exp(sigma * lam * conf[:, l])) for l in range(M)]
    P = np.eye(N)
    for l in list(range(slice_, M)) + list(range(0, slice_)):
        P = Bs[l] @ P
    return np.linalg.inv(np.eye(N) + P)


def _logdet_weight(model, hop, conf, dtau):
    lam = model.lambda_(dtau)
    N, M = conf.shape
    w = 0.0
    for sigma in (+1.0, -1.0):
        P = np.eye(N)
        for l in range(M):
            P = (hop[2] @ np.diag(np.exp(sigma * lam * conf[:, l]))) @ P
        sign, ld = np.linalg.slogdet(np.eye(N) + P)
        assert sign > 0
        w += ld
    return w


def test_block_greens_match_the_defining_inverse():
    model, hop, conf, o, M = _setup()
    for s in (0, 3, M):
        G = o.greens_at(s)
        for b, sigma in enumerate((+1.0, -1.0)):
            want = _direct_greens(model, hop, conf, s % M, 0.1, sigma)
            assert np.abs(G[b] - want).max() < 1e-9


def test_detratio_and_accept_local_against_brute_force():
    """One update() = propagate to slice M + sweep_spatial over its sites. Every accepted flip must change the
    weight prod_sigma det(1 + B_M..B_1) by the factor the acceptance used, and the Green's functions after the
    sweep must be the from-scratch ones of the new field."""
    model, hop, conf, o, M = _setup(L=4, U=4.0, beta=0.5, safe_mult=5)
    o.build_stack()
    before = o.conf()
    w0 = _logdet_weight(model, hop, before, 0.1)
    o.update(1)
    after = o.conf()
    st = o.state()
    sl = st["current_slice"]
    flipped = np.nonzero(before[:, sl - 1] != after[:, sl - 1])[0]
    assert st["acc_local"] == len(flipped) and len(flipped) > 0
    assert np.array_equal(np.delete(before, sl - 1, axis=1), np.delete(after, sl - 1, axis=1))
    # accept_local!: G after the sweep = from-scratch G of the new field. While slice sl is swept, B_sl is the
    # rightmost factor (the flip multiplies it from the right): [1 + B_{sl-1} ... B_1 B_M ... B_sl]^-1
    G = o.greens()
    for b, sigma in enumerate((+1.0, -1.0)):
        want = _direct_greens(model, hop, after, sl - 1, 0.1, sigma)
        assert np.abs(G[b] - want).max() < 1e-8
    # propose_local: replay the flips one by one; R_up R_down from the running G equals the weight ratio
    lam = model.lambda_(0.1)
    cur = before.copy()
    w = w0
    for i in flipped:
        Gs = [_direct_greens(model, hop, cur, sl - 1, 0.1, sg) for sg in (+1.0, -1.0)]
        dE = -2.0 * lam * cur[i, sl - 1]
        r = (1.0 + (np.exp(dE) - 1.0) * (1.0 - Gs[0][i, i])) * (1.0 + (np.exp(-dE) - 1.0) * (1.0 - Gs[1][i, i]))
        cur[i, sl - 1] *= -1
        w_new = _logdet_weight(model, hop, cur, 0.1)
        assert abs(np.exp(w_new - w) - r) < 1e-8 * max(1.0, r)
        w = w_new


def test_half_filling_and_suppressed_double_occupancy():
    """mu = 0 is half filling for both sectors; U > 0 pushes <n_up n_down> below the uncorrelated 1/4 (the
    attractive model pushes it above)."""
    model, hop, conf, o, M = _setup(L=4, U=4.0, beta=2.0, safe_mult=10, seed=5)
    o.build_stack()
    o.sweep(20)
    nu, nd, docc, cnt = 0.0, 0.0, 0.0, 0
    for _ in range(60):
        o.sweep(1)
        G = o.greens(physical=True)
        a, b = 1.0 - np.diag(G[0]), 1.0 - np.diag(G[1])
        nu += a.mean(); nd += b.mean(); docc += (a * b).mean(); cnt += 1
    nu, nd, docc = nu / cnt, nd / cnt, docc / cnt
    assert abs(nu + nd - 1.0) < 0.03
    assert docc < 0.2
    st = o.state()
    assert 0.05 < st["acc_local"] / st["prop_local"] < 0.95
