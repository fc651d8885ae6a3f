"""GPU parity tests of the Ising path: bit-exact spins against the CPU oracle."""
import ctypes as C

import numpy as np
import pytest

import montecarlo_jl_b200 as mcb
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def _oracle_sweeps(s0_xy, Ls, beta, seed, rep, sweep0, n):
    L = orc.lib()
    t4, t8 = C.c_uint32(), C.c_uint32()
    L.ising_thresholds(beta, C.byref(t4), C.byref(t8))
    s = np.ascontiguousarray(s0_xy.T).reshape(-1).copy()  # [y][x]
    L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, seed, rep, sweep0, n)
    E, M = C.c_int64(), C.c_int64()
    L.ising_energy_magnetization(s, Ls, C.byref(E), C.byref(M))
    return s.reshape(Ls, Ls).T, E.value, M.value


@pytest.mark.parametrize("Ls", [4, 8, 30, 128, 256, 384])
def test_spins_bit_exact(Ls):
    beta, seed, R = 0.4406867935097715, 123, 3
    sim = mcb.MC(mcb.IsingModel(Ls), beta=beta, n_replicas=R, seed=seed, replica_offset=5)
    s0 = sim.conf()
    assert np.array_equal(s0[1], mcb.IsingModel(Ls).rand_conf(seed, 6))  # pack/unpack round trip
    done = 0
    for upto in (1, 2, 10):
        sim.sweep(upto - done)
        done = upto
        got = sim.conf()
        E, M = sim.energy_magnetization()
        for r in range(R):
            want, Eo, Mo = _oracle_sweeps(s0[r], Ls, beta, seed, 5 + r, 0, upto)
            assert np.array_equal(got[r], want), (Ls, upto, r)
            assert (E[r], M[r]) == (Eo, Mo)
    sim.close()


def test_L1024_100_sweeps_bit_exact():
    Ls, beta, seed = 1024, 0.4406867935097715, 7
    sim = mcb.MC(mcb.IsingModel(Ls), beta=beta, n_replicas=2, seed=seed)
    s0 = sim.conf()
    sim.sweep(100)
    got = sim.conf()
    E, M = sim.energy_magnetization()
    for r in range(2):
        want, Eo, Mo = _oracle_sweeps(s0[r], Ls, beta, seed, r, 0, 100)
        assert np.array_equal(got[r], want)
        assert (E[r], M[r]) == (Eo, Mo)
    sim.close()


def test_replica_partition_independence():
    """Replica streams are keyed by global id: 4 replicas at once == 2 + 2 with offsets."""
    Ls, beta = 64, 0.44
    full = mcb.MC(mcb.IsingModel(Ls), beta=beta, n_replicas=4, seed=9)
    a = mcb.MC(mcb.IsingModel(Ls), beta=beta, n_replicas=2, seed=9, replica_offset=0)
    b = mcb.MC(mcb.IsingModel(Ls), beta=beta, n_replicas=2, seed=9, replica_offset=2)
    for s in (full, a, b):
        s.sweep(7)
    assert np.array_equal(full.conf()[:2], a.conf()) and np.array_equal(full.conf()[2:], b.conf())
    for s in (full, a, b):
        s.close()


def test_statistics_match_reference_goldens():
    """test/integration_tests.jl:9-23 (L=8, beta=0.35), statistical: 64 replicas x 2000 sweeps."""
    sim = mcb.MC(mcb.IsingModel(8), beta=0.35, n_replicas=64, seed=123, thermalization=200, sweeps=2000)
    r = sim.run()
    assert abs(r["M"].mean() - 25.47) < 0.5
    assert abs(r["E"].mean() - (-59.10)) < 0.5
    assert abs(r["C"].mean() - 0.585) < 0.03
    assert abs(r["chi"].mean() - 1.300) < 0.08
    sim.close()


def test_errors():
    with pytest.raises(mcb.McbError):
        mcb.MC(mcb.IsingModel(7), beta=0.4)
    with pytest.raises(mcb.McbError):
        mcb.MC(mcb.IsingModel(130), beta=0.4)


def test_full_size_L4096_bit_exact_and_conserved_counts():
    """BASELINE.json configs[4] at full lattice size: 2 of the replicas against the oracle (bit-exact),
    and the integer observables against a direct count on the downloaded spins."""
    Ls, R = 4096, 2
    sim = mcb.MC(mcb.IsingModel(Ls), beta=0.5 * np.log(1 + np.sqrt(2)), n_replicas=R, seed=123, replica_offset=5)
    s0 = sim.conf()
    sim.sweep(2)
    got = sim.conf()
    E, M = sim.energy_magnetization()
    L = orc.lib()
    import ctypes as C
    t4, t8 = C.c_uint32(), C.c_uint32()
    L.ising_thresholds(sim.beta, C.byref(t4), C.byref(t8))
    for r in range(R):
        s = np.ascontiguousarray(s0[r].T).reshape(-1).copy()
        L.ising_checkerboard_sweeps(s, Ls, t4.value, t8.value, 123, 5 + r, 0, 2)
        assert np.array_equal(s.reshape(Ls, Ls).T, got[r])
        g = got[r].astype(np.int64)
        assert M[r] == g.sum()
        assert E[r] == -(g * (np.roll(g, -1, 0) + np.roll(g, -1, 1))).sum()
    sim.close()


@pytest.mark.parametrize("L", [4, 6, 16, 128, 256])
def test_bit_packed_wire_format_round_trip(L):
    """mcb_ising_get/set_spins_bits: BitArray(conf .== 1), bit s = x + y*L, against numpy's packing of the same
    configuration; uploading the bits reproduces the spins, and a sweep from either upload gives the same state."""
    R = 3
    sim = mcb.MC(mcb.IsingModel(L), beta=0.4, n_replicas=R, seed=9)
    conf = sim.conf()                                     # (R, x, y)
    bits = sim.conf_bits()
    for r in range(R):
        lin = (conf[r].T.reshape(-1) == 1).astype(np.uint8)   # linear site index x + y*L
        pad = (-len(lin)) % 64
        want = np.packbits(np.concatenate([lin, np.zeros(pad, np.uint8)]), bitorder="little").view("<u8")
        assert np.array_equal(bits[r], want)
    sim.sweep(3)
    after = sim.conf()
    other = mcb.MC(mcb.IsingModel(L), beta=0.4, n_replicas=R, seed=9, conf=-conf)
    other.set_conf_bits(bits)
    assert np.array_equal(other.conf(), conf)
    other.sweep(3)
    assert np.array_equal(other.conf(), after)
    assert np.array_equal(other.conf_bits(), sim.conf_bits())
    sim.close(); other.close()
