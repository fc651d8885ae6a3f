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

    def direction_targets(self, eps=1e-6):
        """trg_of[dir, src] = the target of the one pair (src, trg) that EachSitePairByDistance files under
        direction `dir` for source `src` (a periodic Bravais lattice has exactly one; -1 would mean none), and the
        displacement vectors in the iterator's order.

        Restates lattice_iterators.jl:153-189 array-wise: the displacement pos(src) - pos(trg) is moved to the
        periodic image the reference picks (scan of the 9 box shifts, replace when `directed_norm` is smaller by
        more than eps), unique displacements are numbered in order of first appearance (src-major scan) and then
        stably sorted by `directed_norm` = |d| + eps * angle(d, e_x). Inside a direction the reference lists the
        pairs by ascending src, which is the second index here; the first K rows are also the (dir, trg) lists
        EachLocalQuadByDistance{K} builds per source (lattice_iterators.jl:268-274)."""
        L, N = self.L, self.sites
        pos = self.positions()
        shifts = [np.zeros(2)]
        for v in (np.array([float(L), 0.0]), np.array([0.0, float(L)])):
            shifts = [e - v for e in shifts] + shifts + [e + v for e in shifts]

        def dnorm(d):
            l = np.hypot(d[..., 0], d[..., 1])
            safe = np.where(l > eps, l, 1.0)
            ang = np.arccos(np.clip(d[..., 0] / safe, -1.0, 1.0))
            ang = np.where(d[..., 1] < 0, 2 * np.pi - ang, ang)
            return np.where(l > eps, l + eps * ang, l)

        base = pos[:, None, :] - pos[None, :, :]          # [src, trg]
        d = base + shifts[0]
        for v in shifts[1:]:
            nd = base + v
            take = dnorm(nd) + eps < dnorm(d)
            d = np.where(take[..., None], nd, d)
        key = np.rint(d).astype(np.int64)
        code = (key[..., 0] + 4 * L) * (8 * L) + (key[..., 1] + 4 * L)
        flat = code.reshape(-1)                             # src-major scan = order of first appearance
        uniq, first = np.unique(flat, return_index=True)
        appearance = uniq[np.argsort(first, kind="stable")]
        dirs = np.stack([appearance // (8 * L) - 4 * L, appearance % (8 * L) - 4 * L], axis=1).astype(np.float64)
        order = np.argsort(dnorm(dirs), kind="stable")
        dirs = dirs[order]
        rank_of_code = {int(c): r for r, c in enumerate(appearance[order])}
        trg_of = np.full((len(dirs), N), -1, dtype=np.int32)
        for src in range(N):
            for trg in range(N):
                trg_of[rank_of_code[int(code[src, trg])], src] = trg
        return trg_of, dirs
