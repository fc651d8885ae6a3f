"""MC flavor (host side) for the classical Ising model: mirrors `MC(model; beta, ...)`, `run!`,
and the E/M measurements of src/flavors/MC/MC.jl and src/models/Ising/measurements.jl for a batch
of independent replicas on one B200.
"""
import ctypes as C

import numpy as np

from . import _lib
from .models import IsingModel


class MC:
    """`MC(IsingModel(L); beta, thermalization, sweeps, measure_rate)` x n_replicas.

    conf has shape (R, L, L) indexed [replica, x, y] (the reference's IsingConf is conf[x, y]).
    """

    def __init__(self, model, beta, n_replicas=1, seed=123, replica_offset=0, device=0, conf=None,
                 thermalization=0, sweeps=1000, measure_rate=1):
        if not isinstance(model, IsingModel):
            raise TypeError("the B200 engine implements MC{IsingModel{SquareLattice}}")
        self.model, self.beta = model, float(beta)
        self.R, self.Ls = int(n_replicas), model.L
        self.seed, self.replica_offset, self.device = int(seed), int(replica_offset), int(device)
        self.thermalization, self.sweeps, self.measure_rate = int(thermalization), int(sweeps), int(measure_rate)
        self.L = _lib.lib()
        self.h = C.c_void_p()
        _lib.check(self.L.mcb_ising_create(self.Ls, self.beta, self.R, self.seed, self.replica_offset,
                                           self.device, C.byref(self.h)))
        if conf is None:
            conf = np.stack([model.rand_conf(self.seed, self.replica_offset + r) for r in range(self.R)])
        self.set_conf(conf)

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.L.mcb_ising_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_conf(self, conf):
        conf = np.asarray(conf, dtype=np.int8)
        if conf.shape != (self.R, self.Ls, self.Ls):
            raise ValueError(f"conf must have shape {(self.R, self.Ls, self.Ls)}")
        # [replica][y][x]
        _lib.check(self.L.mcb_ising_set_spins(self.h, np.ascontiguousarray(conf.transpose(0, 2, 1)).reshape(-1)))

    def conf(self):
        buf = np.empty(self.R * self.Ls * self.Ls, dtype=np.int8)
        _lib.check(self.L.mcb_ising_get_spins(self.h, buf))
        return buf.reshape(self.R, self.Ls, self.Ls).transpose(0, 2, 1).copy()

    def conf_bits(self):
        """The configurations in the recorder's wire format BitArray(conf .== 1): (R, ceil(L*L/64)) uint64, bit s of a
        replica = site x + y*L (Julia's column-major linear index), least significant bit first."""
        buf = np.empty(self.R * self.L.mcb_ising_spin_chunks(self.h), dtype=np.uint64)
        _lib.check(self.L.mcb_ising_get_spins_bits(self.h, buf))
        return buf.reshape(self.R, -1)

    def set_conf_bits(self, bits):
        bits = np.ascontiguousarray(bits, dtype=np.uint64)
        if bits.shape != (self.R, self.L.mcb_ising_spin_chunks(self.h)):
            raise ValueError("bits must have shape (replicas, ceil(L*L/64))")
        _lib.check(self.L.mcb_ising_set_spins_bits(self.h, bits.reshape(-1)))

    def sweep(self, n=1):
        _lib.check(self.L.mcb_ising_sweep(self.h, n))

    def energy_magnetization(self):
        """(E, sum(conf)) per replica as exact integers (IsingModel.jl:177-185, measurements.jl:72)."""
        E = np.empty(self.R, dtype=np.int64)
        M = np.empty(self.R, dtype=np.int64)
        _lib.check(self.L.mcb_ising_measure(self.h, E, M))
        return E, M

    def run(self):
        """run!(mc) (MC.jl:190-309) + finish! of both measurements (Ising/measurements.jl:38-44,80-86)."""
        if self.thermalization:
            self.sweep(self.thermalization)
        Es, Ms = [], []
        for i in range(1, self.sweeps + 1):
            self.sweep(1)
            if i % self.measure_rate == 0:
                E, M = self.energy_magnetization()
                Es.append(E)
                Ms.append(np.abs(M))
        Es, Ms = np.array(Es, dtype=np.float64), np.array(Ms, dtype=np.float64)
        invN = 1.0 / (self.Ls * self.Ls)
        E, E2, M, M2 = Es.mean(0), (Es**2).mean(0), Ms.mean(0), (Ms**2).mean(0)
        return dict(E=E, E2=E2, e=E * invN, C=self.beta**2 * invN * (E2 - E**2),
                    M=M, M2=M2, m=M * invN, chi=self.beta * invN * (M2 - M**2))
