"""ctypes binding of the CPU oracle (oracle/liboracle.so).

ORACLE = test infrastructure. Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module. The product package
(montecarlo.jl_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}

_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(dtype=np.int8, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(native=False, force=False):
    """Compile the oracle with the committed Makefile. Returns the .so path."""
    out = "liboracle_native.so" if native else "liboracle.so"
    path = os.path.join(_HERE, out)
    srcs = [os.path.join(_HERE, f) for f in
            ("dqmc_oracle.cpp", "ising_oracle.cpp", "measure_oracle.cpp", "linalg_ref.h", "philox_ref.h",
             "measure_ref.h")]
    stale = (not os.path.exists(path)) or any(
        os.path.getmtime(s) > os.path.getmtime(path) for s in srcs)
    if stale or force:
        cmd = ["make", "-C", _HERE, "OUT=" + out]
        if native:
            cmd.append("ARCH=native")
        subprocess.run(cmd, check=True, capture_output=True)
    return path


def lib(native=False):
    key = bool(native)
    if key in _LIBS:
        return _LIBS[key]
    L = C.CDLL(build(native))
    L.orc_create.restype = C.c_void_p
    L.orc_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _f64p, _f64p, _f64p,
                             _f64p, _i8p, C.c_uint64, C.c_uint32, C.c_int]
    L.orc_destroy.argtypes = [C.c_void_p]
    L.orc_set_conf.argtypes = [C.c_void_p, _i8p]
    L.orc_get_conf.argtypes = [C.c_void_p, _i8p]
    L.orc_set_uniforms.argtypes = [C.c_void_p, _f64p, C.c_int64]
    L.orc_build_stack.argtypes = [C.c_void_p]
    L.orc_update.argtypes = [C.c_void_p, C.c_int64]
    L.orc_sweep.argtypes = [C.c_void_p, C.c_int64]
    L.orc_propagate_only.argtypes = [C.c_void_p]
    L.orc_get_greens.argtypes = [C.c_void_p, _f64p, C.c_int]
    L.orc_get_state.argtypes = [C.c_void_p, _i64p, _f64p]
    L.orc_calculate_greens_at.argtypes = [C.c_void_p, C.c_int, _f64p]
    L.orc_greens_unequal.argtypes = [C.c_void_p, C.c_int, C.c_int, _f64p, C.c_int]
    L.orc_cgi.restype = C.c_int
    L.orc_cgi.argtypes = [C.c_void_p, C.c_int, _f64p, C.c_int]
    L.orc_vmul.argtypes = [_f64p, _f64p, _f64p, C.c_int, C.c_int]
    L.orc_udt_pivot.argtypes = [_f64p, _f64p, _f64p, _i64p, C.c_int, C.c_int]
    L.orc_rdivp.argtypes = [_f64p, _f64p, _i64p, C.c_int]
    L.orc_calculate_greens_udt.argtypes = [_f64p] * 7 + [C.c_int]
    L.orc_multiply_slice.argtypes = [C.c_void_p, C.c_int, _f64p, C.c_int]
    L.orc_uniform.restype = C.c_double
    L.orc_uniform.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    L.orc_philox.argtypes = [_u32p, _u32p, _u32p]
    L.orc_sweep_many.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int]
    L.orc_update_many.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int64, C.c_int]
    L.orc_update_repulsive.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    # measurements (measure_oracle.cpp)
    L.orc_meas_pairs.argtypes = [_f64p, C.c_int, C.c_int, C.c_int, _i32p, C.c_int64, _f64p, C.c_int]
    L.orc_meas_quads.argtypes = [_f64p, C.c_int, C.c_int, _i32p, C.c_int64, _f64p, C.c_int]
    L.orc_meas_sites.argtypes = [_f64p, C.c_int, C.c_int, C.c_int, _f64p]
    L.orc_meas_o5.argtypes = [_f64p, C.c_int, _f64p, C.c_double, _f64p]
    L.orc_run_measured.restype = C.c_int64
    L.orc_run_measured.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int64,
                                   C.c_int, _i32p, C.c_int64, C.c_int, _i32p, C.c_int64, C.c_int, _f64p,
                                   C.c_double] + [C.c_void_p] * 7
    # Ising
    L.ising_energy_cubic.restype = C.c_double
    L.ising_energy_cubic.argtypes = [_i64p, C.c_int, C.c_int]
    L.ising_energy_square.restype = C.c_double
    L.ising_energy_square.argtypes = [_i64p, C.c_int]
    L.ising_delta_e_square.restype = C.c_double
    L.ising_delta_e_square.argtypes = [_i64p, C.c_int, C.c_int64]
    L.ising_thresholds.argtypes = [C.c_double, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.ising_site_random.restype = C.c_uint32
    L.ising_site_random.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                    C.c_int]
    L.ising_checkerboard_sweeps.argtypes = [_i8p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint64,
                                            C.c_uint32, C.c_uint32, C.c_int]
    L.ising_sequential_sweeps.restype = C.c_int64
    L.ising_sequential_sweeps.argtypes = [_i8p, C.c_int, C.c_double, C.c_uint64, C.c_uint32,
                                          C.c_uint32, C.c_int, C.POINTER(C.c_double)]
    L.ising_energy_magnetization.argtypes = [_i8p, C.c_int, C.POINTER(C.c_int64),
                                             C.POINTER(C.c_int64)]
    L.ising_random_spins.argtypes = [_i8p, C.c_int, C.c_uint64, C.c_uint32]
    _LIBS[key] = L
    return L


class OracleDQMCRepulsive:
    """One Markov chain of DQMC{HubbardModelRepulsive, CheckerboardFalse} on the CPU: the two diagonal blocks of
    the reference's BlockDiagonal matrices as two sectors, `up` on the field s and `dn` on -s
    (HubbardModelRepulsive.jl:112-124), with the joint sweep_spatial of dqmc_oracle.cpp."""

    def __init__(self, N, M, safe_mult, dtau, U, hop, conf, seed=123, chain=0, check_propagation=False):
        conf = np.asarray(conf, dtype=np.int8)
        self.up = OracleDQMC(N, M, safe_mult, dtau, U, hop, conf, seed, chain, check_propagation)
        self.dn = OracleDQMC(N, M, safe_mult, dtau, U, hop, -conf, seed, chain, check_propagation)
        self.N, self.M = N, M

    def build_stack(self):
        self.up.build_stack()
        self.dn.build_stack()

    def update(self, n=1):
        self.up.L.orc_update_repulsive(self.up.h, self.dn.h, n)

    def sweep(self, n=1):
        self.update(n * 2 * self.M)

    def greens(self, physical=False):
        """(2, N, N): spin-up and spin-down block."""
        return np.stack([self.up.greens(physical), self.dn.greens(physical)])

    def greens_at(self, slice_):
        return np.stack([self.up.greens_at(slice_), self.dn.greens_at(slice_)])

    def conf(self):
        c = self.up.conf()
        assert np.array_equal(c, -self.dn.conf())
        return c

    def state(self):
        return self.up.state()


def fmat(a):
    """Julia (column-major) matrix -> flat C buffer. a[i, j] ends up at i + j*n."""
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).T).reshape(-1)


def unfmat(buf, n):
    return np.asarray(buf, dtype=np.float64).reshape(n, n).T.copy()


class OracleDQMC:
    """One Markov chain of DQMC{HubbardModelAttractive, CheckerboardFalse} on the CPU.

    conf has shape (N, M) like the reference's HubbardConf (HubbardModel.jl:4).
    """

    def __init__(self, N, M, safe_mult, dtau, U, hop, conf, seed=123, chain=0,
                 check_propagation=False, native=False):
        self.L = lib(native)
        self.N, self.M, self.dtau = N, M, dtau
        eTm, eTp, eT2m, eT2p = hop
        cbuf = np.ascontiguousarray(np.asarray(conf, dtype=np.int8).T).reshape(-1)
        self.h = self.L.orc_create(N, M, safe_mult, dtau, U, fmat(eTm), fmat(eTp), fmat(eT2m),
                                   fmat(eT2p), cbuf, seed, chain, int(check_propagation))
        if not self.h:
            raise ValueError("slices must be a multiple of safe_mult (stack.jl:115)")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def build_stack(self):
        self.L.orc_build_stack(self.h)

    def update(self, n=1):
        self.L.orc_update(self.h, n)

    def sweep(self, n=1):
        self.L.orc_sweep(self.h, n)

    def greens(self, physical=False):
        out = np.empty(self.N * self.N)
        self.L.orc_get_greens(self.h, out, int(physical))
        return unfmat(out, self.N)

    def greens_at(self, slice_):
        out = np.empty(self.N * self.N)
        self.L.orc_calculate_greens_at(self.h, slice_, out)
        return unfmat(out, self.N)

    def propagate_only(self):
        """propagate(mc) without the local updates (stack.jl:502-620)."""
        self.L.orc_propagate_only(self.h)

    def propagate_to_measurement_point(self):
        """Propagate until current_slice == 1 on the way up, where G = G(tau = 0) (the measurement point of
        DQMC.jl's run loop and the precondition of the CombinedGreensIterator)."""
        for _ in range(4 * self.M + 4):
            st = self.state()
            if st["current_slice"] == 1 and st["direction"] == 1:
                return
            self.propagate_only()
        raise RuntimeError("measurement point not reached")

    def greens_unequal(self, slice1, slice2, physical=False):
        """G(slice1 <- slice2), 0 <= slice <= M (unequal_time_stack.jl:291-305). Clobbers the work matrices."""
        out = np.empty(self.N * self.N)
        self.L.orc_greens_unequal(self.h, slice1, slice2, out, 1 if physical else 0)
        return unfmat(out, self.N)

    def combined_greens(self, recalculate=0, physical=True):
        """CombinedGreensIterator (unequal_time_stack.jl:746-883) run to the end: array (M, 3, N, N) holding
        G(0 <- l), G(l <- 0), G(l <- l) for l = 1..M. The handle has to be at current_slice == 1."""
        out = np.empty(self.M * 3 * self.N * self.N)
        if self.L.orc_cgi(self.h, recalculate, out, 1 if physical else 0) != 0:
            raise RuntimeError("CombinedGreensIterator needs current_slice == 1")
        return out.reshape(self.M, 3, self.N, self.N).transpose(0, 1, 3, 2).copy()

    def conf(self):
        out = np.empty(self.N * self.M, dtype=np.int8)
        self.L.orc_get_conf(self.h, out)
        return out.reshape(self.M, self.N).T.copy()

    def set_conf(self, conf):
        self.L.orc_set_conf(self.h, np.ascontiguousarray(np.asarray(conf, dtype=np.int8).T).reshape(-1))

    def set_uniforms(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        self.L.orc_set_uniforms(self.h, u, u.size)

    def state(self):
        st = np.zeros(5, dtype=np.int64)
        err = np.zeros(1)
        self.L.orc_get_state(self.h, st, err)
        return dict(current_slice=int(st[0]), direction=int(st[1]), prop_local=int(st[2]),
                    acc_local=int(st[3]), updates=int(st[4]), max_prop_err=float(err[0]))

    def multiply_slice(self, slice_, Mx, mode):
        buf = fmat(Mx)
        self.L.orc_multiply_slice(self.h, slice_, buf, mode)
        return unfmat(buf, self.N)


def sweep_many(chains, n_sweeps, n_threads):
    arr = (C.c_void_p * len(chains))(*[c.h for c in chains])
    chains[0].L.orc_sweep_many(arr, len(chains), n_sweeps, n_threads)


def update_many(chains, n_updates, n_threads):
    arr = (C.c_void_p * len(chains))(*[c.h for c in chains])
    chains[0].L.orc_update_many(arr, len(chains), n_updates, n_threads)


def vmul(A, B, mode):
    n = A.shape[-1]
    a = fmat(A) if A.ndim == 2 else np.ascontiguousarray(A, dtype=np.float64)
    b = fmat(B) if B.ndim == 2 else np.ascontiguousarray(B, dtype=np.float64)
    c = np.empty(n * n)
    lib().orc_vmul(c, a, b, n, mode)
    return unfmat(c, n)


def udt_pivot(X, apply_pivot=True):
    """Returns U, D, T, pivot (0-based) of udt_AVX_pivot! (UDT.jl:192-281)."""
    n = X.shape[0]
    a = fmat(X)
    u = np.empty(n * n)
    d = np.empty(n)
    p = np.empty(n, dtype=np.int64)
    lib().orc_udt_pivot(u, d, a, p, n, int(apply_pivot))
    return unfmat(u, n), d, unfmat(a, n), p


def rdivp(A, T, pivot):
    n = A.shape[0]
    a = fmat(A)
    lib().orc_rdivp(a, fmat(T), np.ascontiguousarray(pivot, dtype=np.int64), n)
    return unfmat(a, n)


def calculate_greens_udt(Ul, Dl, Tl, Ur, Dr, Tr):
    n = Ul.shape[0]
    g = np.empty(n * n)
    lib().orc_calculate_greens_udt(fmat(Ul), np.array(Dl, dtype=np.float64), fmat(Tl), fmat(Ur),
                                   np.array(Dr, dtype=np.float64), fmat(Tr), g, n)
    return unfmat(g, n)


# ---------------------------------------------------------------- measurements (measure_oracle.cpp)
class MeasurementTables:
    """The lattice iterators of a periodic L x L square lattice as the tuple lists the reference iterates
    (oracle/lattice_iterators.py): EachSitePairByDistance and EachLocalQuadByDistance{K}."""

    def __init__(self, L, K=5):
        from . import lattice_iterators as li
        self.L, self.N, self.K = L, L * L, K
        self.pairs, self.directions = li.each_site_pair_by_distance(L)
        self.n_dirs = len(self.pairs)
        self.trip = np.ascontiguousarray(li.pair_triplets(self.pairs))
        self.implied = li.each_local_quad_by_distance(self.pairs, self.N, K)
        self.quint = np.ascontiguousarray(li.quad_quintuplets(self.implied, K))
        self.n_pc = self.n_dirs * K * K
        self.synced = np.ascontiguousarray(li.synced_quad_quintuplets(self.pairs, self.N, K))


def measure(G, tables, flavors=1, T=None, U=0.0):
    """All equal-time measurements of one physical Green's function G ((fN, fN) numpy, Julia semantics G[i, j]).
    Returns a dict in the reference's output layouts: occupation (fN), cdc / sdc_x / sdc_y / sdc_z (n_dirs),
    pc (n_dirs, K, K), mx / my / mz (N), and for one flavor the O5 scalars."""
    L_ = lib()
    N, nd, K = tables.N, tables.n_dirs, tables.K
    g = fmat(G)
    out = {}
    o = np.empty(flavors * N)
    L_.orc_meas_sites(g, N, flavors, 0, o)
    out["occupation"] = o.copy()
    for k, name in ((1, "mx"), (2, "my"), (3, "mz")):
        o = np.empty(N)
        L_.orc_meas_sites(g, N, flavors, k, o)
        out[name] = o.copy()
    for k, name in enumerate(("cdc", "sdc_x", "sdc_y", "sdc_z")):
        o = np.empty(nd)
        L_.orc_meas_pairs(g, N, flavors, k, tables.trip, len(tables.trip), o, nd)
        out[name] = o
    o = np.empty(tables.n_pc)
    L_.orc_meas_quads(g, N, flavors, tables.quint, len(tables.quint), o, tables.n_pc)
    out["pc"] = o.reshape(K, K, nd).transpose(2, 1, 0).copy()  # column-major (n_dirs, K, K)
    if flavors == 1 and T is not None:
        o = np.empty(4)
        L_.orc_meas_o5(g, N, fmat(T), U, o)
        out.update(n=o[0], double_occupancy=o[1], e_kin=o[2], e_int=o[3])
    return out


def run_measured(chain, tables, thermalization, sweeps, measure_rate=10, T=None, U=0.0, first=1):
    """run!(mc) after build_stack (DQMC.jl:420-438) on an OracleDQMC or OracleDQMCRepulsive chain: sweeps
    first .. thermalization + sweeps with the reference's measurement schedule. Returns the means over
    measurements in the layouts of measure() plus G and the count."""
    rep = isinstance(chain, OracleDQMCRepulsive)
    hup, hdn = (chain.up.h, chain.dn.h) if rep else (chain.h, None)
    N, M, nd, K = chain.N, chain.M, tables.n_dirs, tables.K
    f = 2 if rep else 1
    acc = dict(G=np.zeros((f * N) ** 2), occupation=np.zeros(f * N), cdc=np.zeros(nd), sdc=np.zeros(3 * nd),
               pc=np.zeros(tables.n_pc), mz=np.zeros(N), o5=np.zeros(4))
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    Tm = fmat(T) if T is not None else np.zeros(N * N)
    cnt = lib().orc_run_measured(hup, hdn, N, M, first, thermalization + sweeps, thermalization, measure_rate,
                                 tables.trip, len(tables.trip), nd, tables.quint, len(tables.quint), tables.n_pc,
                                 Tm, U, ptr(acc["G"]), ptr(acc["occupation"]), ptr(acc["cdc"]), ptr(acc["sdc"]),
                                 ptr(acc["pc"]), ptr(acc["mz"]), ptr(acc["o5"]) if (T is not None and not rep) else None)
    k = max(cnt, 1)
    out = dict(count=cnt, G=unfmat(acc["G"], f * N) / k, occupation=acc["occupation"] / k, cdc=acc["cdc"] / k,
               sdc_x=acc["sdc"][:nd] / k, sdc_y=acc["sdc"][nd:2 * nd] / k, sdc_z=acc["sdc"][2 * nd:] / k,
               pc=acc["pc"].reshape(K, K, nd).transpose(2, 1, 0) / k, mz=acc["mz"] / k)
    if T is not None and not rep:
        out.update(n=acc["o5"][0] / k, double_occupancy=acc["o5"][1] / k, e_kin=acc["o5"][2] / k,
                   e_int=acc["o5"][3] / k)
    return out


def susceptibilities(chain, tables, recalculate=0, T=None):
    """One time-displaced measurement of a chain at its measurement point: apply!(::CombinedGreensIterator, ...)
    (src/flavors/DQMC/measurements/generic.jl:227-243) for charge_density_susceptibility, spin_density_susceptibility and
    pairing_susceptibility (measurements.jl:56-58, 144-146, 207), summed over l = 1..M, then finish!(..., delta_tau):
    times delta_tau / N (generic.jl:272-286). Kernels for packed Green's functions (G00, G0l, Gl0, Gll):
      OracleDQMC (HubbardModelAttractive.jl:225-229, 239-241, 246-248)
        cdc  4 (1 - Gll[i,i]) (1 - G00[j,j]) - 2 G0l[j,i] Gl0[i,j]      over EachSitePairByDistance
        sdc  -2 G0l[j,i] Gl0[i,j]                                        (x = y = z)
        pc   Gl0[src1,src2] Gl0[trg1,trg2]                               over EachLocalQuadByDistance{K}
      OracleDQMCRepulsive: the general kernels (measurements.jl:75-90, 157-196, 215-219) evaluated literally on the
        2N x 2N block-diagonal matrices.
    current_current_susceptibility (measurements.jl:254-262) over EachLocalQuadBySyncedDistance{K} needs the hopping matrix
    T (mc.s.hopping_matrix) and is evaluated when T is given: cc_kernel of HubbardModelAttractive.jl:249-266 / the general
    one of measurements.jl:268-300.
    Plain numpy over the iterator tuples (small lattices). Returns dict cds, sds_x, sds_y, sds_z (n_dirs; also `sds` for
    one flavor), ps (n_dirs, K, K) and, with T, ccs (n_dirs, K)."""
    N, nd, K = tables.N, tables.n_dirs, tables.K
    d_, i_, j_ = tables.trip[:, 0], tables.trip[:, 1], tables.trip[:, 2]
    lin, s1, t1, s2, t2 = (tables.quint[:, k] for k in range(5))
    cds, sx, sy, sz, ps = np.zeros(nd), np.zeros(nd), np.zeros(nd), np.zeros(nd), np.zeros(tables.n_pc)
    cc = np.zeros(nd * K)
    clin, c1, e1, c2, e2 = (tables.synced[:, k] for k in range(5))   # (lin, src1, trg1, src2, trg2)
    two = isinstance(chain, (OracleDQMCRepulsive, dict))
    if two:
        def full(u, d):
            g = np.zeros((2 * N, 2 * N))
            g[:N, :N], g[N:, N:] = u, d
            return g
        if isinstance(chain, dict):   # Green's functions from elsewhere: G00 (2, N, N), cu / cd (M, 3, N, N), dtau
            (gu, gd), cu, cd = chain["G00"], chain["cu"], chain["cd"]
        else:
            gu, gd = chain.greens(physical=True)
            cu = chain.up.combined_greens(recalculate, physical=True)
            cd = chain.dn.combined_greens(recalculate, physical=True)
        G00 = full(gu, gd)
        iN, jN = i_ + N, j_ + N
        for l in range(cu.shape[0]):
            G0l, Gl0, Gll = (full(cu[l, w], cd[l, w]) for w in range(3))
            np.add.at(cds, d_, (1 - Gll[i_, i_]) * (1 - G00[j_, j_]) - G0l[j_, i_] * Gl0[i_, j_]
                      + (1 - Gll[i_, i_]) * (1 - G00[jN, jN]) - G0l[jN, i_] * Gl0[i_, jN]
                      + (1 - Gll[iN, iN]) * (1 - G00[j_, j_]) - G0l[j_, iN] * Gl0[iN, j_]
                      + (1 - Gll[iN, iN]) * (1 - G00[jN, jN]) - G0l[jN, iN] * Gl0[iN, jN])
            np.add.at(sx, d_, Gll[iN, i_] * G00[jN, j_] - G0l[jN, i_] * Gl0[iN, j_]
                      + Gll[iN, i_] * G00[j_, jN] - G0l[j_, i_] * Gl0[iN, jN]
                      + Gll[i_, iN] * G00[jN, j_] - G0l[jN, iN] * Gl0[i_, j_]
                      + Gll[i_, iN] * G00[j_, jN] - G0l[j_, iN] * Gl0[i_, jN])
            np.add.at(sy, d_, -Gll[iN, i_] * G00[jN, j_] + G0l[jN, i_] * Gl0[iN, j_]
                      + Gll[iN, i_] * G00[j_, jN] - G0l[j_, i_] * Gl0[iN, jN]
                      + Gll[i_, iN] * G00[jN, j_] - G0l[jN, iN] * Gl0[i_, j_]
                      - Gll[i_, iN] * G00[j_, jN] + G0l[j_, iN] * Gl0[i_, jN])
            np.add.at(sz, d_, (1 - Gll[i_, i_]) * (1 - G00[j_, j_]) - G0l[j_, i_] * Gl0[i_, j_]
                      - (1 - Gll[i_, i_]) * (1 - G00[jN, jN]) + G0l[jN, i_] * Gl0[i_, jN]
                      - (1 - Gll[iN, iN]) * (1 - G00[j_, j_]) + G0l[j_, iN] * Gl0[iN, j_]
                      + (1 - Gll[iN, iN]) * (1 - G00[jN, jN]) - G0l[jN, iN] * Gl0[iN, jN])
            np.add.at(ps, lin, Gl0[s1, s2] * Gl0[t1 + N, t2 + N] - Gl0[s1, t2 + N] * Gl0[t1 + N, s2])
            if T is not None:
                Tf = full(T, T)
                for o1 in (0, N):
                    for o2 in (0, N):
                        a1, b1, a2, b2 = c1 + o1, e1 + o1, c2 + o2, e2 + o2       # s1, t1, s2, t2
                        np.add.at(cc, clin, (Tf[a1, b1] * Gll[b1, a1] - Tf[b1, a1] * Gll[a1, b1])
                                  * (Tf[a2, b2] * G00[b2, a2] - Tf[b2, a2] * G00[a2, b2])
                                  + Tf[b1, a1] * Tf[b2, a2] * (-G0l[a2, b1]) * Gl0[a1, b2]
                                  - Tf[a1, b1] * Tf[b2, a2] * (-G0l[a2, a1]) * Gl0[b1, b2]
                                  - Tf[b1, a1] * Tf[a2, b2] * (-G0l[b2, b1]) * Gl0[a1, a2]
                                  + Tf[a1, b1] * Tf[a2, b2] * (-G0l[b2, a1]) * Gl0[b1, a2])
        dtau = chain["dtau"] if isinstance(chain, dict) else chain.up.dtau
    else:
        if isinstance(chain, tuple):   # (G00, cg, dtau): Green's functions from elsewhere, e.g. stabilised evaluations
            G00, cg, dtau_given = chain
        else:
            G00 = chain.greens(physical=True)
            cg = chain.combined_greens(recalculate, physical=True)   # (M, 3, N, N): G0l, Gl0, Gll
            dtau_given = chain.dtau
        for l in range(cg.shape[0]):
            G0l, Gl0, Gll = cg[l]
            ex = -2.0 * G0l[j_, i_] * Gl0[i_, j_]
            np.add.at(sx, d_, ex)
            np.add.at(cds, d_, 4.0 * (1.0 - Gll[i_, i_]) * (1.0 - G00[j_, j_]) + ex)
            np.add.at(ps, lin, Gl0[s1, s2] * Gl0[t1, t2])
            if T is not None:
                np.add.at(cc, clin, 4.0 * (T[c1, e1] * Gll[e1, c1] - T[e1, c1] * Gll[c1, e1])
                          * (T[c2, e2] * G00[e2, c2] - T[e2, c2] * G00[c2, e2])
                          + 2.0 * T[e1, c1] * T[e2, c2] * (-G0l[c2, e1]) * Gl0[c1, e2]
                          - 2.0 * T[c1, e1] * T[e2, c2] * (-G0l[c2, c1]) * Gl0[e1, e2]
                          - 2.0 * T[e1, c1] * T[c2, e2] * (-G0l[e2, e1]) * Gl0[c1, c2]
                          + 2.0 * T[c1, e1] * T[c2, e2] * (-G0l[e2, c1]) * Gl0[e1, c2])
        sy, sz = sx, sx
        dtau = dtau_given
    f = dtau / N
    out = dict(cds=cds * f, sds_x=sx * f, sds_y=sy * f, sds_z=sz * f,
               ps=(ps * f).reshape(K, K, nd).transpose(2, 1, 0).copy())
    if not two:
        out["sds"] = out["sds_x"]
    if T is not None:
        out["ccs"] = (cc * f).reshape(K, nd).T.copy()   # column-major (n_dirs, K)
    return out
