"""ORACLE (test infrastructure): the reference's lattice iterators as explicit tuple lists.

Restates, independently of montecarlo.jl_b200/lattices.py (the product's table builder),
  SquareLattice            src/lattices/square.jl:25-59, 71        (site numbering, neighbours, positions)
  EachSitePairByDistance   src/lattices/lattice_iterators.jl:131-189
  EachLocalQuadByDistance  src/lattices/lattice_iterators.jl:258-318
in the form the reference iterates them: `pairs[dir] = [(src, trg), ...]` and
`implied[dir12][dir1][dir2] = [(src1, trg1, src2, trg2), ...]`, everything 0-based here.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may import this module.
"""
import math

import numpy as np


def square_positions(L):
    """positions(l) for SquareLattice (square.jl:71): site n (0-based) = x + y*L sits at (x+1, y+1)."""
    return [np.array([float(n % L + 1), float(n // L + 1)]) for n in range(L * L)]


def square_lattice_vectors(L):
    """lattice_vectors(l) (square.jl:72): the periodic box."""
    return [np.array([float(L), 0.0]), np.array([0.0, float(L)])]


def generate_combinations(vs):
    """lattice_iterators.jl:138-144: all sums of {-v, 0, +v} over the box vectors, in the reference's order."""
    out = [np.zeros(len(vs[0]))]
    for v in vs:
        out = [e - v for e in out] + out + [e + v for e in out]
    return out


def directed_norm(v, eps):
    """lattice_iterators.jl:147-156: |v| + eps * angle(v, e_x), angle in [0, 2pi)."""
    l = math.sqrt(float(v[0]) ** 2 + float(v[1]) ** 2)
    if len(v) == 2 and l > eps:
        # clamp: acos of 1 + 1ulp would be NaN in Julia too, but x / hypot(x, y) never exceeds 1 in magnitude
        angle = math.acos(max(-1.0, min(1.0, float(v[0]) / l)))
        if v[1] < 0:
            angle = 2 * math.pi - angle
        return l + eps * angle
    return l


def each_site_pair_by_distance(L, eps=1e-6):
    """EachSitePairByDistance(lattice) (lattice_iterators.jl:158-189).

    Returns (pairs, directions): pairs[dir] is the list of (src, trg) in the order the reference pushes them
    (origin-major), directions[dir] the displacement pos[src] - pos[trg] wrapped to its shortest image."""
    pos = square_positions(L)
    wrap = generate_combinations(square_lattice_vectors(L))
    directions = []
    bonds = []
    for origin in range(L * L):
        for trg, p in enumerate(pos):
            d = pos[origin] - p + wrap[0]
            for v in wrap[1:]:
                new_d = pos[origin] - p + v
                if directed_norm(new_d, eps) + eps < directed_norm(d, eps):
                    d = new_d.copy()
            idx = None
            for k, e in enumerate(directions):
                if np.all(np.abs(e - d) <= eps):   # isapprox(dir, d, atol = eps) on vectors uses the 2-norm;
                    idx = k                        # displacements are integer vectors, so both agree
                    break
            if idx is None:
                directions.append(d.copy())
                bonds.append([])
                idx = len(directions) - 1
            bonds[idx].append((origin, trg))
    # sortperm(directions, by = directed_norm): stable
    order = sorted(range(len(directions)), key=lambda k: directed_norm(directions[k], eps))
    return [bonds[k] for k in order], [directions[k] for k in order]


def pair_triplets(pairs):
    """The iteration of EachSitePairByDistance (lattice_iterators.jl:195-218): (dir, src, trg), dir-major."""
    return np.array([(d, s, t) for d, lst in enumerate(pairs) for (s, t) in lst], dtype=np.int32).reshape(-1, 3)


def each_local_quad_by_distance(pairs, n_sites, K):
    """EachLocalQuadByDistance{K}(lattice) (lattice_iterators.jl:258-318): implied[dir12][dir1][dir2] =
    list of (src1, trg1, src2, trg2)."""
    trg_from_src = [[] for _ in range(n_sites)]
    for d in range(K):
        for (src, trg) in pairs[d]:
            trg_from_src[src].append((d, trg))
    nd = len(pairs)
    implied = [[[[] for _ in range(K)] for _ in range(K)] for _ in range(nd)]
    for dir12 in range(nd):
        for (src1, src2) in pairs[dir12]:
            for (dir1, trg1) in trg_from_src[src1]:
                for (dir2, trg2) in trg_from_src[src2]:
                    implied[dir12][dir1][dir2].append((src1, trg1, src2, trg2))
    return implied


def quad_quintuplets(implied, K):
    """The iteration of EachLocalQuadByDistance (lattice_iterators.jl:328-351): (lin, src1, trg1, src2, trg2)
    with lin the column-major linear index of (dir12, dir1, dir2) in an array of size (n_dirs, K, K)."""
    nd = len(implied)
    rows = []
    for dir2 in range(K):
        for dir1 in range(K):
            for dir12 in range(nd):
                lin = dir12 + nd * (dir1 + K * dir2)
                for q in implied[dir12][dir1][dir2]:
                    rows.append((lin,) + q)
    return np.array(rows, dtype=np.int32).reshape(-1, 5)


def synced_quad_quintuplets(pairs, n_sites, K):
    """EachLocalQuadBySyncedDistance{K}(lattice) and its iteration (lattice_iterators.jl:390-464): the quadruples
    (src1, trg1, src2, trg2) with src2 - src1 in direction dir12 and BOTH targets in the same direction dir1 < K of their
    source, as rows (lin, src1, trg1, src2, trg2), lin = the column-major linear index of (dir12, dir1) in an array of
    size (n_dirs, K). Order: dir1-major, then dir12, then the reference's push order (pair, trg1, trg2)."""
    trg_from_src = [[] for _ in range(n_sites)]
    for d in range(K):
        for (src, trg) in pairs[d]:
            trg_from_src[src].append((d, trg))
    nd = len(pairs)
    implied = [[[] for _ in range(K)] for _ in range(nd)]
    for dir12 in range(nd):
        for (src1, src2) in pairs[dir12]:
            for (dir1, trg1) in trg_from_src[src1]:
                for (dir2, trg2) in trg_from_src[src2]:
                    if dir1 == dir2:
                        implied[dir12][dir1].append((src1, trg1, src2, trg2))
    rows = []
    for dir1 in range(K):
        for dir12 in range(nd):
            for q in implied[dir12][dir1]:
                rows.append((dir12 + nd * dir1,) + q)
    return np.array(rows, dtype=np.int32).reshape(-1, 5)
