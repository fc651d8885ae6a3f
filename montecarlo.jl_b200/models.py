"""Model definitions (host side) mirroring the reference's model plug-ins.

HubbardModelAttractive : src/models/HubbardModel/HubbardModelAttractive.jl:27-40, :78-110
HubbardModelRepulsive  : src/models/HubbardModel/HubbardModelRepulsive.jl:27-124
IsingModel             : src/models/Ising/IsingModel.jl:1-60
Only what the engine needs as *inputs* lives here: the hopping matrix, its
exponentials, lambda, and random initial configurations.
"""
import math

import numpy as np

from . import philox
from .lattices import SquareLattice

IsingTc = 1.0 / (0.5 * math.log(1.0 + math.sqrt(2.0)))  # IsingModel.jl:7


class HubbardModelAttractive:
    def __init__(self, l=None, L=None, dims=2, U=1.0, t=1.0, mu=0.0):
        if l is None:
            if dims != 2:
                raise NotImplementedError("the B200 engine covers square lattices (SURVEY.md §2 #16)")
            l = SquareLattice(L)
        if U < 0:
            raise AssertionError("U must be positive.")  # HubbardModelAttractive.jl:31
        self.l, self.U, self.t, self.mu = l, float(U), float(t), float(mu)
        self.flv = 1

    def hopping_matrix(self):
        """T = -mu*I, T[trg, src] += -t over directed nn pairs (HubbardModelAttractive.jl:78-91)."""
        N = len(self.l)
        T = np.diag(np.full(N, -self.mu))
        for src, trg in self.l.neighbors_directed():
            T[trg, src] += -self.t
        return T

    def lambda_(self, delta_tau):
        return math.acosh(math.exp(0.5 * self.U * delta_tau))  # HubbardModelAttractive.jl:103

    def hopping_exponentials(self, delta_tau):
        """(e^{-dtau T/2}, e^{+dtau T/2}, and the products of each with itself) — stack.jl:176-179.

        The reference calls Julia's exp(::Matrix); T is real symmetric so a symmetric
        eigendecomposition gives the same matrices to rounding. They are *inputs* of the engine.
        """
        T = self.hopping_matrix()
        if not np.allclose(T, T.T):
            raise ValueError("hopping matrix must be symmetric")
        w, V = np.linalg.eigh(T)
        eTm = (V * np.exp(-0.5 * delta_tau * w)) @ V.T
        eTp = (V * np.exp(+0.5 * delta_tau * w)) @ V.T
        return eTm, eTp, eTm @ eTm, eTp @ eTp

    def rand_conf(self, nslices, seed, chain):
        """rand(DQMC, m, nslices) (HubbardModel.jl:46-48): iid +-1, shape (N, nslices)."""
        N = len(self.l)
        bits = philox.random_bits(N * nslices, seed, chain, domain=0xC0DF1E1D)
        return (2 * bits.astype(np.int8) - 1).reshape(nslices, N).T.copy()


class HubbardModelRepulsive(HubbardModelAttractive):
    """HubbardModelRepulsive (HubbardModelRepulsive.jl:27-70): two flavors. The hopping matrix of the reference is
    BlockDiagonal(T, copy(T)) (:84-100) with the T of the attractive model; the engine is handed the one block.
    The interaction exponential is exp(+-sign(power) lambda s) for spin up / down (:112-124)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.flv = 2


class IsingModel:
    def __init__(self, L, dims=2):
        if dims != 2:
            raise NotImplementedError("the B200 engine covers the square lattice (SURVEY.md §2 #14)")
        self.L, self.dims = int(L), 2
        self.l = SquareLattice(self.L)

    def rand_conf(self, seed, replica):
        """rand(MC, m) (IsingModel.jl:72): iid +-1, shape (L, L) indexed [x, y]."""
        n = self.L * self.L
        bits = philox.random_bits(n, seed, replica, domain=0xFFFFFFFF)
        return (2 * bits.astype(np.int8) - 1).reshape(self.L, self.L, order="F")
