"""DQMC flavor (host side): mirrors `DQMC(model; kwargs...)`, `run!`, `greens`, `replay!`-style
from-scratch G and the measurement accessors of src/flavors/DQMC/DQMC.jl, for a *batch* of
independent Markov chains that live on one B200. All numerics happen in libmcb200.so.
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from .models import HubbardModelAttractive, HubbardModelRepulsive


class DQMCParameters:
    """DQMCParameters (DQMC.jl:52-125): beta / delta_tau / slices consistency rules included."""

    def __init__(self, thermalization=100, sweeps=100, safe_mult=10, measure_rate=10,
                 check_propagation_error=True, check_sign_problem=True, silent=True, **kw):
        keys = set(kw)
        if keys == {"beta"}:
            kw["delta_tau"] = 0.1
            keys = set(kw)
        if keys == {"delta_tau", "beta", "slices"}:
            delta_tau, beta = kw["delta_tau"], kw["beta"]
            slices = int(round(beta / delta_tau))
            if slices != kw["slices"]:
                raise ValueError(f"Given slices ({kw['slices']}) does not match calculated slices "
                                 f"beta/delta_tau ~ {slices}")
        elif keys == {"beta", "slices"}:
            beta, slices = kw["beta"], int(kw["slices"])
            delta_tau = beta / slices
        elif keys == {"delta_tau", "slices"}:
            delta_tau, slices = kw["delta_tau"], int(kw["slices"])
            beta = delta_tau * slices
        elif keys == {"delta_tau", "beta"}:
            delta_tau, beta = kw["delta_tau"], kw["beta"]
            slices = int(round(beta / delta_tau))
        else:
            raise ValueError(f"Invalid keyword arguments to DQMCParameters {sorted(keys)}")
        self.thermalization, self.sweeps = int(thermalization), int(sweeps)
        self.safe_mult, self.measure_rate = int(safe_mult), int(measure_rate)
        self.check_propagation_error = bool(check_propagation_error)
        self.check_sign_problem, self.silent = bool(check_sign_problem), bool(silent)
        self.delta_tau, self.beta, self.slices = float(delta_tau), float(beta), slices


class DQMC:
    """`DQMC(model; beta, ...)` for a batch of `n_chains` chains (global ids chain_offset + c).

    conf has shape (n_chains, N, M): conf[c] is the reference's HubbardConf (HubbardModel.jl:4).

    With a HubbardModelRepulsive the Green's functions have shape (n_chains, 2, N, N): the spin-up and spin-down
    diagonal blocks of the reference's BlockDiagonal mc.s.greens. That model runs the sweep (build_stack, update,
    sweep, greens, greens_at, greens_unequal, conf, conf_bits, checkpoint / resume, state, stats) and the
    two-flavor measurement kernels (measurements.jl:41-207) incl. replay.
    """

    def __init__(self, model, n_chains=1, seed=123, chain_offset=0, device=0, conf=None, general_path=False,
                 **kwargs):
        if not isinstance(model, HubbardModelAttractive):
            raise TypeError("the B200 engine implements DQMC{HubbardModelAttractive, CheckerboardFalse}")
        self.model = model
        self.rep = isinstance(model, HubbardModelRepulsive)
        self._pfx = "mcb_dqmc_rep_" if self.rep else "mcb_dqmc_"
        self.p = DQMCParameters(**kwargs)
        self.n_chains, self.seed, self.chain_offset, self.device = int(n_chains), int(seed), int(chain_offset), int(device)
        self.general_path = bool(general_path)
        self.N, self.M = len(model.l), self.p.slices
        if self.M % self.p.safe_mult != 0:
            raise ValueError("slices must be a multiple of safe_mult (stack.jl:115)")
        self.L = _lib.lib()
        prm = _lib.DqmcParams(self.N, self.M, self.p.safe_mult, self.n_chains, self.p.delta_tau, model.U,
                              self.seed, self.chain_offset, self.device,
                              int(self.p.check_propagation_error), int(general_path))
        self.h = C.c_void_p()
        _lib.check(self._f("create")(C.byref(prm), C.byref(self.h)))
        self.hop = model.hopping_exponentials(self.p.delta_tau)
        _lib.check(self._f("set_hopping")(self.h, *[self._fm(a) for a in self.hop]))
        if conf is None:
            conf = np.stack([model.rand_conf(self.M, self.seed, self.chain_offset + c)
                             for c in range(self.n_chains)])
        self.set_conf(conf)
        self.last_sweep = 0
        self._tables = False

    def _f(self, name):
        """The C-ABI entry point `name` of this model's handle type."""
        return getattr(self.L, self._pfx + name)

    # -- layout helpers: numpy [i, j] (row-major view of Julia semantics) <-> column-major buffers
    @staticmethod
    def _fm(a):
        return np.ascontiguousarray(np.asarray(a, dtype=np.float64).T).reshape(-1)

    def _unbatch(self, buf):
        if buf.size == 2 * self.n_chains * self.N * self.N:  # two sectors per chain
            return buf.reshape(self.n_chains, 2, self.N, self.N).transpose(0, 1, 3, 2).copy()
        return buf.reshape(self.n_chains, self.N, self.N).transpose(0, 2, 1).copy()

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self._f("destroy")(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_conf(self, conf):
        conf = np.asarray(conf, dtype=np.int8)
        if conf.shape != (self.n_chains, self.N, self.M):
            raise ValueError(f"conf must have shape {(self.n_chains, self.N, self.M)}")
        buf = np.ascontiguousarray(conf.transpose(0, 2, 1)).reshape(-1)
        _lib.check(self._f("set_conf")(self.h, buf))

    def conf(self):
        buf = np.empty(self.n_chains * self.N * self.M, dtype=np.int8)
        _lib.check(self._f("get_conf")(self.h, buf))
        return buf.reshape(self.n_chains, self.M, self.N).transpose(0, 2, 1).copy()

    # -- recorded configurations (ConfigRecorder wire format) ------------------------------------
    def conf_bits(self):
        """compress(mc, model, conf) for every chain: (C, ceil(N*M/64)) uint64, bit b = spin i + N*l."""
        buf = np.empty(self.n_chains * self._f("conf_chunks")(self.h), dtype=np.uint64)
        _lib.check(self._f("get_conf_bits")(self.h, buf))
        return buf.reshape(self.n_chains, -1)

    def set_conf_bits(self, bits):
        bits = np.ascontiguousarray(bits, dtype=np.uint64)
        assert bits.shape == (self.n_chains, self._f("conf_chunks")(self.h))
        _lib.check(self._f("set_conf_bits")(self.h, bits.reshape(-1)))

    def replay(self, configurations):
        """The measurement loop of replay!(mc, configurations) (DQMC.jl:649-660): for every recorded configuration
        decompress, calculate_greens(mc, 0), measure. `configurations`: (n, ceil(N*M/64)) uint64 as produced by
        compress / conf_bits. One configuration per chain and batch; a remainder smaller than n_chains goes through
        a handle of that size. Returns the per-configuration-slot observables of the full batches and, if any,
        of the remainder (their moments combine like chains do, sharding.mean_and_error_from_moments)."""
        cfg = np.ascontiguousarray(configurations, dtype=np.uint64)
        n, C_ = cfg.shape[0], self.n_chains
        self.enable_measurements()
        out = []
        full = (n // C_) * C_
        if full:
            _lib.check(self._f("replay")(self.h, cfg[:full].reshape(-1), full))
            out.append(self.observables())
        if n > full:
            rest = DQMC(self.model, n_chains=n - full, seed=self.seed, beta=self.p.beta, delta_tau=self.p.delta_tau,
                        safe_mult=self.p.safe_mult, general_path=self.general_path, device=self.device)
            rest.enable_measurements(self.K)
            _lib.check(rest._f("replay")(rest.h, cfg[full:].reshape(-1), n - full))
            out.append(rest.observables())
            rest.close()
        return out

    # -- checkpoint / resume (FileIO.jl:38-156, DQMC.jl:463-477, 517-521) -----------------------
    def checkpoint(self):
        """What save(filename, mc) keeps of the Markov chain: the configuration (bit-packed), the number of complete
        sweeps, and what identifies the random streams. Call between sweeps."""
        st = self.state()
        if st["n_updates"] % (2 * self.M) != 0:
            raise ValueError("checkpoints are taken between sweeps")
        return dict(conf_bits=self.conf_bits(), last_sweep=st["n_updates"] // (2 * self.M), seed=self.seed,
                    chain_offset=self.chain_offset, n_chains=self.n_chains)

    def resume(self, ckpt):
        """resume!(mc): upload the saved configuration, rebuild the stack, continue the counters."""
        if (ckpt["seed"], ckpt["chain_offset"], ckpt["n_chains"]) != (self.seed, self.chain_offset, self.n_chains):
            raise ValueError("checkpoint belongs to different chains")
        self.set_conf_bits(ckpt["conf_bits"])
        _lib.check(self._f("resume")(self.h, int(ckpt["last_sweep"])))
        self.last_sweep = int(ckpt["last_sweep"])

    def build_stack(self):
        """init! + build_stack + propagate (DQMC.jl:412-414)."""
        _lib.check(self._f("build_stack")(self.h))

    def update(self, n=1, measure_every=0):
        if measure_every:
            self.enable_measurements()
        _lib.check(self._f("update")(self.h, n, measure_every))

    def sweep(self, n=1, measure_every=0):
        self.update(n * 2 * self.M, measure_every)

    def greens(self, physical=True):
        """greens(mc) (DQMC.jl:711-730); physical=False returns mc.s.greens. Shape (C, N, N)."""
        buf = np.empty(self.n_chains * self.model.flv * self.N * self.N)
        _lib.check(self._f("get_greens")(self.h, buf, int(physical)))
        return self._unbatch(buf)

    def greens_at(self, slice_):
        """calculate_greens(mc, slice) (stack.jl:422-480). Shape (C, N, N)."""
        buf = np.empty(self.n_chains * self.model.flv * self.N * self.N)
        _lib.check(self._f("calculate_greens_at")(self.h, slice_, buf))
        return self._unbatch(buf)

    def greens_unequal(self, slice1, slice2, physical=True):
        """greens(mc, k, l) = G(k <- l), 0 <= k, l <= M (unequal_time_stack.jl:275-305); physical=False returns the
        effective G of calculate_greens(mc, k, l). Leaves the sweep state untouched. Shape (C, N, N)."""
        buf = np.empty(self.n_chains * self.model.flv * self.N * self.N)
        _lib.check(self._f("greens_unequal")(self.h, slice1, slice2, buf, int(physical)))
        return self._unbatch(buf)

    def state(self):
        st = _lib.DqmcState()
        _lib.check(self._f("get_state")(self.h, C.byref(st)))
        return dict(current_slice=st.current_slice, direction=st.direction, n_updates=st.n_updates,
                    n_measurements=st.n_measurements)

    def stats(self):
        arr = (_lib.DqmcChainStats * self.n_chains)()
        _lib.check(self._f("get_stats")(self.h, arr))
        return dict(prop_local=np.array([a.prop_local for a in arr]),
                    acc_local=np.array([a.acc_local for a in arr]),
                    max_propagation_error=np.array([a.max_propagation_error for a in arr]),
                    n_nonfinite=np.array([a.n_nonfinite for a in arr]))

    # -- measurements ---------------------------------------------------------------------
    def enable_measurements(self, K=None):
        """Upload the lattice-iterator tables: EachSitePairByDistance and EachLocalQuadByDistance{K}
        (lattice_iterators.jl:131-355). K defaults to 1 + number of nearest-neighbour directions like
        pairing(...) does (measurements.jl:197-201), capped by the number of directions of a tiny lattice."""
        if self._tables and (K is None or K == self.K):
            return
        trg_of, dirs = self.model.l.direction_targets()
        self.directions = dirs
        self.n_dirs = len(dirs)
        if K is None:
            K = min(5, self.n_dirs)
        self.K = int(K)
        tab = np.ascontiguousarray(trg_of.astype(np.int32)).reshape(-1)
        _lib.check(self._f("set_measurement_tables")(self.h, tab, self.n_dirs, self.K,
                                                     self._fm(self.model.hopping_matrix())))
        self._tables = True

    def measure(self):
        self.enable_measurements()
        _lib.check(self._f("measure")(self.h))

    def reset_observables(self):
        _lib.check(self._f("reset_observables")(self.h))

    def observables(self):
        """Per-chain means over measurements, in the reference's output layouts (leading chain axis):
        G (C, N, N) [(C, 2, N, N) for two flavors], occupation (C, flv*N), cdc / sdc_x / sdc_y / sdc_z (C, n_dirs),
        pc (C, n_dirs, K, K), swave_pairing = pc[:, :, 0, 0], and the scalars n, double_occupancy, e_kin, e_int,
        energy; two flavors add mz (C, N)."""
        n = self._f("observable_len")(self.h)
        buf = np.empty(self.n_chains * n)
        cnt = C.c_int64()
        _lib.check(self._f("get_observables")(self.h, buf, C.byref(cnt)))
        gs = np.empty(self.n_chains * self.model.flv * self.N * self.N)
        _lib.check(self._f("get_greens_sum")(self.h, gs))
        k = max(cnt.value, 1)
        rec = buf.reshape(self.n_chains, n) / k
        nd, K, N, f = self.n_dirs, self.K, self.N, self.model.flv
        o_pc = 4 + 3 * nd
        o_occ = o_pc + nd * K * K
        pc = rec[:, o_pc:o_occ].reshape(self.n_chains, K, K, nd).transpose(0, 3, 2, 1)  # column-major (nd, K, K)
        out = dict(count=cnt.value, n=rec[:, 0], occupation_mean=rec[:, 0], double_occupancy=rec[:, 1],
                   e_kin=rec[:, 2], e_int=rec[:, 3], energy=rec[:, 2] + rec[:, 3],
                   cdc=rec[:, 4:4 + nd], sdc_x=rec[:, 4 + nd:4 + 2 * nd], sdc_y=rec[:, 4 + nd:4 + 2 * nd],
                   sdc_z=rec[:, 4 + 2 * nd:4 + 3 * nd], pc=pc, swave_pairing=pc[:, :, 0, 0],
                   occupation=rec[:, o_occ:o_occ + f * N], G=self._unbatch(gs) / k)
        if f == 2:
            out["mz"] = rec[:, o_occ + 2 * N:o_occ + 3 * N]
        return out

    def measure_susceptibilities(self, recalculate=0):
        """One time-displaced measurement (charge_density_susceptibility, spin_density_susceptibility,
        pairing_susceptibility: measurements.jl:56-58, 144-146, 207 through apply!(::CombinedGreensIterator, ...),
        generic.jl:227-243), accumulated on the device. The handle has to be at its measurement point, as for
        CombinedGreensIterator."""
        self.enable_measurements()
        _lib.check(self._f("measure_susceptibilities")(self.h, int(recalculate)))

    def schedule_susceptibilities(self, enable=True, recalculate=0):
        """Let every scheduled measurement of update / sweep / run (measure_every > 0) also take one time-displaced
        measurement, as run! does for a measurement group on the CombinedGreensIterator (and, for the one-flavor handle,
        every configuration of replay)."""
        self.enable_measurements()
        _lib.check(self._f("schedule_susceptibilities")(self.h, int(bool(enable)), int(recalculate)))

    def susceptibilities(self):
        """Per-chain means over the time-displaced measurements taken so far: cds (C, n_dirs), sds_x = sds_y and sds_z
        (C, n_dirs; all equal, also as `sds`, for a one-flavor model) — the charge and spin density susceptibilities by
        direction —, ps (C, n_dirs, K, K), the pairing susceptibility in the layout of pc, and ccs (C, n_dirs, K), the
        current-current susceptibility over EachLocalQuadBySyncedDistance{K}."""
        n = self._f("susceptibility_len")(self.h)
        if n <= 0:
            raise RuntimeError("enable_measurements first")
        buf = np.empty(self.n_chains * n)
        cnt = C.c_int64()
        _lib.check(self._f("get_susceptibilities")(self.h, buf, C.byref(cnt)))
        rec = buf.reshape(self.n_chains, n) / max(cnt.value, 1)
        nd, K = self.n_dirs, self.K
        npair = 3 if self.model.flv == 2 else 2
        ps = rec[:, npair * nd:(npair + K * K) * nd].reshape(self.n_chains, K, K, nd).transpose(0, 3, 2, 1)
        ccs = rec[:, (npair + K * K) * nd:].reshape(self.n_chains, K, nd).transpose(0, 2, 1)
        sx = rec[:, nd:2 * nd]
        sz = rec[:, 2 * nd:3 * nd] if npair == 3 else sx
        out = dict(count=cnt.value, cds=rec[:, :nd], sds_x=sx, sds_y=sx, sds_z=sz, ps=ps, ccs=ccs)
        if npair == 2:
            out["sds"] = sx
        return out

    def allreduce_moments(self, comm=None):
        """Moments of the per-chain observable means summed over the chains of all ranks of `comm` (an
        `sharding.NcclComm`; None = this handle alone): the C-ABI collective mcb_dqmc_allreduce_observables.
        Returns (record_sum, record_sumsq, G_sum, G_sumsq, n_chains) with the record in the layout of
        mcb_dqmc_get_observables and G as (N, N)."""
        if self.model.flv != 1:
            raise NotImplementedError("allreduce_moments: one-flavor handles only")
        L = self.L
        n = L.mcb_dqmc_moments_len(self.h)
        if n <= 0:
            raise RuntimeError("enable_measurements first")
        buf = np.empty(n)
        _lib.check(L.mcb_dqmc_allreduce_observables(self.h, comm.h if comm is not None else None, buf))
        R, NN = self._f("observable_len")(self.h), self.N * self.N
        g = lambda a: a.reshape(self.N, self.N).T.copy()
        return (buf[:R].copy(), buf[R:2 * R].copy(), g(buf[2 * R:2 * R + NN]), g(buf[2 * R + NN:2 * R + 2 * NN]),
                int(round(buf[-1])))

    def allreduce_susceptibility_moments(self, comm=None):
        """The same collective for the time-displaced records (mcb_dqmc_allreduce_susceptibilities): returns
        (record_sum, record_sumsq, n_chains), the record in the layout of mcb_dqmc_get_susceptibilities."""
        if self.model.flv != 1:
            raise NotImplementedError("allreduce_susceptibility_moments: one-flavor handles only")
        n = self.L.mcb_dqmc_susceptibility_len(self.h)
        if n <= 0:
            raise RuntimeError("enable_measurements first")
        buf = np.empty(2 * n + 1)
        _lib.check(self.L.mcb_dqmc_allreduce_susceptibilities(self.h, comm.h if comm is not None else None, buf))
        return buf[:n].copy(), buf[n:2 * n].copy(), int(round(buf[-1]))

    @staticmethod
    def observable_shapes(model, K=None):
        """Trailing shapes (without the chain axis) of everything observables() returns as an array; lets a rank
        without chains take part in the reduction with correctly shaped zero moments."""
        N, f = len(model.l), model.flv
        nd = N  # a periodic Bravais lattice has one direction per site
        K = min(5, nd) if K is None else K
        g = (f, N, N) if f == 2 else (N, N)
        shapes = dict(n=(), occupation_mean=(), double_occupancy=(), e_kin=(), e_int=(), energy=(), cdc=(nd,),
                      sdc_x=(nd,), sdc_y=(nd,), sdc_z=(nd,), pc=(nd, K, K), swave_pairing=(nd,),
                      occupation=(f * N,), G=g)
        if f == 2:
            shapes["mz"] = (N,)
        return shapes

    def run(self, thermalization=None, sweeps=None):
        """run!(mc) (DQMC.jl:369-515): thermalise, then sweep with measurements every measure_rate."""
        th = self.p.thermalization if thermalization is None else thermalization
        sw = self.p.sweeps if sweeps is None else sweeps
        self.enable_measurements()
        self.build_stack()
        if th:
            self.sweep(th, 0)
        if sw:
            self.sweep(sw, self.p.measure_rate)
        self.last_sweep += th + sw
        return self.observables()


def compress(conf):
    """compress(mc, model, conf) = BitArray(conf .== 1) (HubbardModel.jl:56): conf (N, M) of +-1 -> uint64 chunks in
    Julia's BitArray layout (column-major bit index, 64 bits per chunk, least significant bit first)."""
    bits = (np.asarray(conf).T.reshape(-1) == 1).astype(np.uint8)  # linear index i + N*l
    pad = (-len(bits)) % 64
    packed = np.packbits(np.concatenate([bits, np.zeros(pad, np.uint8)]), bitorder="little")
    return packed.view("<u8").astype(np.uint64)


def decompress(chunks, N, M):
    """decompress(mc, model, c) = 2c .- 1 (HubbardModel.jl:57-59)."""
    bits = np.unpackbits(np.ascontiguousarray(chunks, dtype="<u8").view(np.uint8), bitorder="little")[:N * M]
    return (2 * bits.astype(np.int8) - 1).reshape(M, N).T.copy()


class CombinedGreensIterator:
    """CombinedGreensIterator(mc, recalculate=4 safe_mult) (unequal_time_stack.jl:746-885): yields, for l = 1..M, the
    Green's functions (G(0 <- l), G(l <- 0), G(l <- l)), each of shape (C, N, N), for time-displaced measurements.
    The three UDT triples live on the device and advance one slice per step (mcb_dqmc_cgi_begin / _next); they are
    re-factorised every safe_mult slices and recomputed from the UnequalTimeStack every `recalculate` slices, as in the
    reference. The handle has to be at its measurement point (current_slice == 1 on the way up). The reference checks
    the iterator against greens(mc, l, 0) etc. to 1e-14 (recalculate = safe_mult) and 1e-10 (default),
    test/flavortests_DQMC.jl:143-160."""

    def __init__(self, mc, recalculate=0, physical=True):
        self.mc = mc
        self.recalculate = int(recalculate)
        self.physical = physical

    def __len__(self):
        return self.mc.M

    def __iter__(self):
        mc = self.mc
        _lib.check(mc._f("cgi_begin")(mc.h, self.recalculate))
        size = mc.n_chains * mc.model.flv * mc.N * mc.N
        nxt = mc._f("cgi_next")
        sl = C.c_int32(0)
        while True:
            bufs = [np.empty(size) for _ in range(3)]
            _lib.check(nxt(mc.h, int(self.physical), C.byref(sl), *[b.ctypes.data_as(C.c_void_p) for b in bufs]))
            if sl.value == 0:
                return
            yield tuple(mc._unbatch(b) for b in bufs)


def mean_and_error(per_chain):
    """Mean over chains and its standard error (chains are independent Markov chains)."""
    a = np.asarray(per_chain, dtype=np.float64)
    m = a.mean(axis=0)
    e = a.std(axis=0, ddof=1) / math.sqrt(a.shape[0]) if a.shape[0] > 1 else np.zeros_like(m)
    return m, e
