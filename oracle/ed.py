"""ORACLE (test infrastructure): exact diagonalisation of small Hubbard clusters.

Restates the reference's ED test helper test/ED/ED.jl in matrix form (numpy):
  HamiltonMatrix                 test/ED/ED.jl:68-121   H = -t sum_{directed nn} c+_trg c_src
                                                           + U sum_i (n_up - 1/2)(n_dn - 1/2) - mu sum_i n_i
                                                        with U -> -|U| for HubbardModelAttractive, mu = 0 for
                                                        HubbardModelRepulsive
  expectation_value              test/ED/ED.jl:441-475  Tr(e^{-beta H} O) / Z
  Greens / calculate_Greens_matrix  :146-155, 479-500   G[i, j] = <c_j c+_i>  (note the index order)
  charge_density_correlation, m_z, spin_density_correlation_{x,z}, pairing_correlation  :180-400
  expectation_value_integrated   test/ED/ED.jl:625-675  step * sum_{tau = step .. beta} <A(tau) B(0)>
Fermion operators are Jordan-Wigner matrices over the mode order lin = 2*site + substate (ED.jl:44-46).
Only tests/ may import this module.
"""
import numpy as np

UP, DOWN = 0, 1


class HubbardED:
    def __init__(self, neighs, U_signed, mu, t=1.0, beta=1.0):
        """neighs: (4, n_sites) neighbour table, column = source (square.jl:47-52)."""
        self.ns = neighs.shape[1]
        self.nm = 2 * self.ns
        self.dim = 1 << self.nm
        self.beta = beta
        # annihilators: c_m |s> = (-1)^{popcount(s & (bit_m - 1))} |s ^ bit_m>  if bit m set   (ED.jl:52-59)
        states = np.arange(self.dim)
        self.c = []
        for m in range(self.nm):
            bit = 1 << m
            occ = (states & bit) != 0
            par = np.array([bin(int(s) & (bit - 1)).count("1") & 1 for s in states])
            A = np.zeros((self.dim, self.dim))
            src = states[occ]
            A[src ^ bit, src] = 1.0 - 2.0 * par[occ]
            self.c.append(A)
        self.cd = [A.T.copy() for A in self.c]
        H = np.zeros((self.dim, self.dim))
        for sub in (UP, DOWN):
            for source in range(self.ns):
                for target in neighs[:, source]:
                    H -= t * self.cd[self.mode(int(target), sub)] @ self.c[self.mode(source, sub)]
        for p in range(self.ns):
            nu, nd = self.n(p, UP), self.n(p, DOWN)
            one = np.eye(self.dim)
            H += U_signed * (nu - 0.5 * one) @ (nd - 0.5 * one)
            H -= mu * (nu + nd)
        assert np.allclose(H, H.T)
        w, V = np.linalg.eigh(H)
        wts = np.exp(-beta * (w - w.min()))
        self.rho = (V * (wts / wts.sum())) @ V.T
        self.w, self.V = w - w.min(), V

    def mode(self, site, sub):
        return 2 * site + sub

    def n(self, site, sub):
        m = self.mode(site, sub)
        return self.cd[m] @ self.c[m]

    def ev(self, O):
        return float(np.trace(self.rho @ O))

    # ---- observables in the reference's conventions
    def greens_matrix(self):
        """calculate_Greens_matrix (ED.jl:479-500): G[s1*ns + i, s2*ns + j] = <c_{j,s2} c+_{i,s1}>."""
        G = np.zeros((self.nm, self.nm))
        for s1 in (UP, DOWN):
            for s2 in (UP, DOWN):
                for i in range(self.ns):
                    for j in range(self.ns):
                        G[s1 * self.ns + i, s2 * self.ns + j] = self.ev(self.c[self.mode(j, s2)] @ self.cd[self.mode(i, s1)])
        return G

    def charge_density_correlation(self, site1, site2):
        """ED.jl:180-205: n_site1 n_site2."""
        n1 = self.n(site1, UP) + self.n(site1, DOWN)
        n2 = self.n(site2, UP) + self.n(site2, DOWN)
        return self.ev(n1 @ n2)

    def m_z(self, site):
        return self.ev(self.n(site, UP) - self.n(site, DOWN))

    def spin_density_correlation_z(self, site1, site2):
        m1 = self.n(site1, UP) - self.n(site1, DOWN)
        m2 = self.n(site2, UP) - self.n(site2, DOWN)
        return self.ev(m1 @ m2)

    def spin_density_correlation_x(self, site1, site2):
        """ED.jl:287-316: (c+_up c_dn + c+_dn c_up)_site1 (same)_site2."""
        def sx(s):
            return self.cd[self.mode(s, UP)] @ self.c[self.mode(s, DOWN)] + self.cd[self.mode(s, DOWN)] @ self.c[self.mode(s, UP)]
        return self.ev(sx(site1) @ sx(site2))

    def pairing_correlation(self, i, j, k, l):
        """ED.jl:389-398: c_{i,up} c_{j,dn} c+_{l,dn} c+_{k,up}."""
        return self.ev(self.c[self.mode(i, UP)] @ self.c[self.mode(j, DOWN)] @ self.cd[self.mode(l, DOWN)] @ self.cd[self.mode(k, UP)])

    def double_occupancy(self, site):
        return self.ev(self.n(site, UP) @ self.n(site, DOWN))

    def occupation(self, site):
        return self.ev(self.n(site, UP) + self.n(site, DOWN))

    def integrated(self, A, B, step):
        """expectation_value_integrated (ED.jl:625-675): step * sum over tau = beta, beta - step, .., step of
        Tr(e^{-(beta - tau) H} A e^{-tau H} B) / Z = <A(tau) B(0)>, A(tau) = e^{tau H} A e^{-tau H}."""
        Ae, Be = self.V.T @ A @ self.V, self.V.T @ B @ self.V
        Z = np.exp(-self.beta * self.w).sum()
        total = 0.0
        for l in range(1, int(round(self.beta / step)) + 1):
            tau = l * step
            total += step * float(np.sum(np.exp(-(self.beta - tau) * self.w)[:, None] * Ae
                                         * np.exp(-tau * self.w)[None, :] * Be.T)) / Z
        return total

    def number(self, site):
        return self.n(site, UP) + self.n(site, DOWN)

    def mz_operator(self, site):
        return self.n(site, UP) - self.n(site, DOWN)
