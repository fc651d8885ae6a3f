"""Square lattice as *input definition* for the engine (host side).

Mirrors src/lattices/square.jl:4-59 of the reference: column-major site index
(0-based here: i = x + y*L), periodic, neighbour order up=(x+1,y), right=(x,y+1),
down=(x-1,y), left=(x,y-1). Index tables for distance-resolved observables follow
src/lattices/lattice_iterators.jl:131-221 (EachSitePairByDistance).
"""
import numpy as np


class SquareLattice:
    def __init__(self, L):
        self.L = int(L)
        self.sites = self.L * self.L
        lat = np.arange(self.sites).reshape(self.L, self.L, order="F")  # lat[x, y]
        self.lattice = lat
        up = np.roll(lat, -1, axis=0)
        right = np.roll(lat, -1, axis=1)
        down = np.roll(lat, 1, axis=0)
        left = np.roll(lat, 1, axis=1)
        # neighs[k, site], k = up, right, down, left       (square.jl:47-52)
        self.neighs = np.stack([a.reshape(-1, order="F") for a in (up, right, down, left)])

    def __len__(self):
        return self.sites

    def neighbors_directed(self):
        """All (src, trg) pairs, every bond in both directions (neighbors(l, Val(true)))."""
        for src in range(self.sites):
            for k in range(4):
                yield src, int(self.neighs[k, src])

    def positions(self):
        """positions(l): (x, y) per site, 1-based coordinates like the reference (square.jl:71)."""
        x = np.arange(self.sites) % self.L + 1
        y = np.arange(self.sites) // self.L + 1
        return np.stack([x, y], axis=1).astype(np.float64)

    def pair_direction_table(self, eps=1e-6):
        """dir_of[src, trg] -> direction index, ordered like EachSitePairByDistance.

        Restates lattice_iterators.jl:153-189: the displacement pos(src) - pos(trg) is
        wrapped to the shortest periodic image (ties broken by the angle-weighted
        `directed_norm`), unique displacements are numbered in order of first
        appearance and then sorted by directed_norm.
        """
        L = self.L
        pos = self.positions()
        shifts = [np.zeros(2)]
        for v in (np.array([L, 0.0]), np.array([0.0, L])):
            shifts = [e - v for e in shifts] + shifts + [e + v for e in shifts]

        def dnorm(v):
            l = float(np.hypot(v[0], v[1]))
            if l > eps:
                ang = np.arccos(v[0] / l)
                if v[1] < 0:
                    ang = 2 * np.pi - ang
                return l + eps * ang
            return l

        dirs = []
        dir_of = np.zeros((self.sites, self.sites), dtype=np.int32)
        for o in range(self.sites):
            for t in range(self.sites):
                d = pos[o] - pos[t] + shifts[0]
                for v in shifts[1:]:
                    nd = pos[o] - pos[t] + v
                    if dnorm(nd) + eps < dnorm(d):
                        d = nd
                idx = None
                for k, e in enumerate(dirs):
                    if np.allclose(e, d, atol=eps, rtol=0):
                        idx = k
                        break
                if idx is None:
                    dirs.append(d.copy())
                    idx = len(dirs) - 1
                dir_of[o, t] = idx
        order = np.argsort([dnorm(d) for d in dirs], kind="stable")
        rank = np.empty(len(dirs), dtype=np.int32)
        rank[order] = np.arange(len(dirs), dtype=np.int32)
        return rank[dir_of], np.array(dirs)[order]
