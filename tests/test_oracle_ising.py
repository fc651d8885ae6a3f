"""Pins the Ising oracle and the Philox stream.

Known answers from the reference's own tests (RNG-free):
  test/modeltests_IsingModel.jl:21  energy(reshape(1:27,(3,3,3))) == -18333.0
  test/modeltests_IsingModel.jl:27  energy(reshape(1:64,(8,8)))   == -164320.0
  test/modeltests_IsingModel.jl:34  propose_local(mc, m, 13, conf) == (1352.0, nothing)
Philox4x32-10 known answers: Random123 kat_vectors.
Statistical goldens: test/integration_tests.jl:9-23 (Ising L=8, beta=0.35).
"""
import ctypes as C

import numpy as np

import montecarlo_jl_b200 as mcb
from montecarlo_jl_b200 import philox
from oracle import oracle as orc


def test_reference_known_answers():
    L = orc.lib()
    assert L.ising_energy_cubic(np.arange(1, 28, dtype=np.int64), 3, 3) == -18333.0
    conf = np.arange(1, 65, dtype=np.int64)  # reshape(1:64, (8,8)) is column-major = flat order
    assert L.ising_energy_square(conf, 8) == -164320.0
    assert L.ising_energy_cubic(conf, 8, 2) == -164320.0
    assert L.ising_delta_e_square(conf, 8, 13 - 1) == 1352.0


def test_philox_known_answers():
    L = orc.lib()
    kats = [
        ([0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
        ([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
        ([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0],
         [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
    ]
    for ctr, key, want in kats:
        out = np.zeros(4, dtype=np.uint32)
        L.orc_philox(np.array(ctr, dtype=np.uint32), np.array(key, dtype=np.uint32), out)
        assert [int(x) for x in out] == want
        got = philox.philox4x32_10(np.array(ctr, dtype=np.uint32), np.array(key, dtype=np.uint32))
        assert [int(x) for x in got] == want


def test_uniform_streams_agree_between_c_and_numpy():
    L = orc.lib()
    sites = np.arange(50, dtype=np.uint32)
    u_np = philox.dqmc_uniform(123, 7, 3, 11, sites)
    u_c = np.array([L.orc_uniform(123, 7, 3, 11, int(s)) for s in sites])
    assert np.array_equal(u_np, u_c)
    assert np.all((u_c >= 0) & (u_c < 1))
    big = philox.dqmc_uniform(5, 0, 0, 0, np.arange(200000, dtype=np.uint32))
    assert abs(big.mean() - 0.5) < 5e-3 and abs(big.var() - 1 / 12) < 5e-3


def test_random_spins_c_equals_numpy():
    L = orc.lib()
    for Ls in (8, 12, 64):
        a = np.zeros(Ls * Ls, dtype=np.int8)
        L.ising_random_spins(a, Ls, 123, 5)
        b = mcb.IsingModel(Ls).rand_conf(123, 5)
        assert np.array_equal(a, b.reshape(-1, order="F"))
        assert set(np.unique(a)) == {-1, 1}


def _em(L, spins, Ls):
    E, M = C.c_int64(), C.c_int64()
    L.ising_energy_magnetization(spins, Ls, C.byref(E), C.byref(M))
    return E.value, M.value


def test_checkerboard_matches_python_statement():
    """The C oracle against a literal pure-Python statement of Appendix D (small L)."""
    L = orc.lib()
    Ls, beta, seed, rep = 6, 0.3, 99, 2
    t4, t8 = C.c_uint32(), C.c_uint32()
    L.ising_thresholds(beta, C.byref(t4), C.byref(t8))
    assert t4.value == int(np.floor(np.exp(-4 * beta) * 2.0**32))
    s = np.zeros(Ls * Ls, dtype=np.int8)
    L.ising_random_spins(s, Ls, seed, rep)
    py = s.copy().astype(int)
    L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, seed, rep, 0, 3)
    W = (Ls // 2 + 63) // 64
    for sw in range(3):
        for colour in range(2):
            for y in range(Ls):
                for x in range(Ls):
                    if (x + y) % 2 != colour:
                        continue
                    nb = (py[(x + 1) % Ls + y * Ls] + py[x + ((y + 1) % Ls) * Ls]
                          + py[(x - 1) % Ls + y * Ls] + py[x + ((y - 1) % Ls) * Ls])
                    dE = 2 * py[x + y * Ls] * nb
                    if dE <= 0:
                        acc = True
                    else:
                        xh = x // 2
                        grp = (y * W + xh // 64) * 16 + (xh % 64) // 4
                        w = philox.philox4x32_10(np.array([sw, colour, grp, 0], dtype=np.uint32),
                                                 philox.seed_key(seed, rep))
                        acc = int(w[xh % 4]) < (t4.value if dE == 4 else t8.value)
                    if acc:
                        py[x + y * Ls] *= -1
    assert np.array_equal(py.astype(np.int8), s)


def test_ising_statistics_match_reference_goldens():
    """test/integration_tests.jl:9-23: L=8, beta=0.35; both schedules, loose (statistical)."""
    L = orc.lib()
    Ls, beta, N = 8, 0.35, 64
    t4, t8 = C.c_uint32(), C.c_uint32()
    L.ising_thresholds(beta, C.byref(t4), C.byref(t8))
    Es, Ms = [], []
    for rep in range(16):
        s = np.zeros(N, dtype=np.int8)
        L.ising_random_spins(s, Ls, 123, rep)
        L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, 123, rep, 0, 100)
        for k in range(1000):
            L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, 123, rep, 100 + k, 1)
            e, m = _em(L, s, Ls)
            Es.append(e)
            Ms.append(abs(m))
    Es, Ms = np.array(Es, float), np.array(Ms, float)
    assert abs(Ms.mean() - 25.47) < 1.0
    assert abs(Es.mean() - (-59.10)) < 1.0
    assert abs(beta**2 * (np.mean(Es**2) - Es.mean() ** 2) / N - 0.585) < 0.06
    # sequential (reference-style) sweep gives the same stationary distribution
    Es2 = []
    for rep in range(8):
        s = np.zeros(N, dtype=np.int8)
        L.ising_random_spins(s, Ls, 123, rep)
        e0, _ = _em(L, s, Ls)
        en = C.c_double(float(e0))
        L.ising_sequential_sweeps(s, Ls, beta, 123, rep, 0, 100, C.byref(en))
        for k in range(1000):
            L.ising_sequential_sweeps(s, Ls, beta, 123, rep, 100 + k, 1, C.byref(en))
            Es2.append(en.value)
        assert en.value == _em(L, s, Ls)[0]  # cached energy stays exact (IsingModel.jl:83-89)
    assert abs(np.mean(Es2) - (-59.10)) < 1.2
