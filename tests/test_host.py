"""CPU-only tests: host-side logic and the C-ABI library's exported surface (no compute calls)."""
import os
import re

import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "mcb200.h")).read()
    declared = set(re.findall(r"\b(mcb_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"mcb_dqmc", "mcb_ising"}
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    _lib.build()
    L = _lib.lib()  # raises if a symbol is missing
    assert b"sm_100a" in L.mcb_version()


def test_no_cpu_fallback_without_device():
    import ctypes as C
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = _lib.lib()
    prm = _lib.DqmcParams(16, 10, 5, 1, 0.1, 1.0, 1, 0, 0, 0, 0)
    h = C.c_void_p()
    assert L.mcb_dqmc_create(C.byref(prm), C.byref(h)) == -2
    assert b"no CPU fallback" in L.mcb_last_error()
    with pytest.raises(mcb.McbError):
        mcb.MC(mcb.IsingModel(8), beta=0.3)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "montecarlo.jl_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(d, f)).read()
                assert "oracle" not in src.lower() or f == "philox.py", f
    assert "oracle" not in open(os.path.join(pkg, "philox.py")).read().replace("the oracle", "")


def test_dqmc_parameters_rules():
    # test/flavortests_DQMC.jl:2-14
    p = mcb.DQMCParameters(beta=5.0)
    assert (p.delta_tau, p.slices) == (0.1, 50)
    p = mcb.DQMCParameters(beta=5.0, delta_tau=0.05)
    assert p.slices == 100
    p = mcb.DQMCParameters(beta=50.0, slices=20)
    assert p.delta_tau == 2.5
    p = mcb.DQMCParameters(delta_tau=0.1, slices=50)
    assert p.beta == 5.0
    with pytest.raises(ValueError):
        mcb.DQMCParameters(beta=5.0, delta_tau=0.1, slices=20)
    with pytest.raises(ValueError):
        mcb.DQMCParameters(slices=20)


def test_square_lattice_tables():
    l = mcb.SquareLattice(4)
    # site 1 (1-based) has neighbours up=2, right=5, down=4, left=13 in the reference (square.jl:47-52)
    assert list(l.neighs[:, 0] + 1) == [2, 5, 4, 13]
    m = mcb.HubbardModelAttractive(L=4, U=1.0, mu=0.5, t=1.0)
    T = m.hopping_matrix()
    assert np.allclose(T, T.T) and np.allclose(np.diag(T), -0.5) and np.isclose(T.sum(), -0.5 * 16 - 64)
    # L=2: every bond counted twice, consistent with the reference's ED test
    assert mcb.HubbardModelAttractive(L=2).hopping_matrix()[0, 1] == -2.0
    eTm, eTp, eT2m, eT2p = m.hopping_exponentials(0.1)
    assert np.allclose(eTm @ eTp, np.eye(16)) and np.allclose(eT2m @ eT2p, np.eye(16))
    trg_of, dirs = l.direction_targets()
    assert trg_of.shape == (16, 16) and len(dirs) == 16
    assert all(sorted(row) == list(range(16)) for row in trg_of.tolist())   # one target per (direction, source)
    assert np.array_equal(trg_of[0], np.arange(16)) and np.allclose(dirs[0], 0)
    assert np.array_equal(trg_of[1], l.neighs[2])   # direction (1, 0) = pos[src] - pos[trg]: trg is the "down" neighbour
    nrm = np.hypot(dirs[:, 0], dirs[:, 1])
    assert np.all(np.diff(nrm) >= -1e-9)


def test_rand_conf_is_keyed_by_global_chain():
    m = mcb.HubbardModelAttractive(L=4)
    a, b = m.rand_conf(20, 123, 3), m.rand_conf(20, 123, 3)
    assert a.shape == (16, 20) and np.array_equal(a, b) and set(np.unique(a)) == {-1, 1}
    assert not np.array_equal(a, m.rand_conf(20, 123, 4))
    assert abs(m.rand_conf(500, 1, 0).mean()) < 0.05
