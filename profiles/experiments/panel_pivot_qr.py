"""Numerical experiment (CPU, numpy): does a *panel-pivoted* blocked Householder QR (the nb largest remaining columns
form the next panel, full column pivoting inside the panel, compact-WY trailing update) give the same Green's functions
as the reference's step-by-step column pivoting (oracle = udt_AVX_pivot!)? G is unique, so the question is only one of
rounding: max|G_variant - G_oracle| / max|G_oracle| against the 1e-10 parity gate.

    python profiles/experiments/panel_pivot_qr.py L U beta dtau nb
"""
import sys
import time

import numpy as np
import scipy.linalg

sys.path.insert(0, ".")
import montecarlo_jl_b200 as mcb  # noqa: E402  (host-side model definitions only; no GPU needed)
from oracle import oracle as orc  # noqa: E402


def qr_panel_pivot(X, nb):
    """X[:, perm] = Q R with panel-wise pivoting. Returns Q, R (upper triangular), perm."""
    A = X.copy()
    n = A.shape[0]
    perm = np.arange(n)
    Q = np.eye(n)
    j = 0
    while j < n:
        b = min(nb, n - j)
        # norms of the trailing columns below row j, choose the b largest
        nrm = np.sqrt((A[j:, j:] ** 2).sum(0))
        order = np.argsort(-nrm, kind="stable")
        sel = j + order
        A[:, j:] = A[:, sel]
        perm[j:] = perm[sel]
        # panel factorisation with pivoting inside the panel (LAPACK on the panel)
        Qp, Rp, pp = scipy.linalg.qr(A[j:, j:j + b], pivoting=True, mode="full")
        A[:, j:j + b] = A[:, j + pp]
        perm[j:j + b] = perm[j + pp]
        A[j:, j:] = Qp.T @ A[j:, j:]
        Q[:, j:] = Q[:, j:] @ Qp
        A[j + b:, j:j + b] = 0.0
        j += b
    return Q, np.triu(A), perm


def udt(X, mode, nb):
    n = X.shape[0]
    if mode == "qrcp":
        Q, R, p = scipy.linalg.qr(X, pivoting=True)
    elif mode == "prepivot":
        p = np.argsort(-np.sqrt((X ** 2).sum(0)), kind="stable")
        Q, R = np.linalg.qr(X[:, p])
    else:
        Q, R, p = qr_panel_pivot(X, nb)
    D = np.abs(np.diag(R))
    T = np.empty_like(R)
    T[:, p] = R / D[:, None]
    return Q, D, T


def greens_from_stack(model, conf, dtau, safe_mult, mode, nb):
    """build_stack (stack.jl:242-255) + the first calculate_greens (G at slice M going down): returns G_eff."""
    N, M = conf.shape
    lam = model.lambda_(dtau)
    hop = model.hopping_exponentials(dtau)
    eT2 = hop[2]
    U, D, T = np.eye(N), np.ones(N), np.eye(N)
    # G_M = [1 + B_{M-1} ... B_1 B_M]^-1 is what the stack holds after build_stack + propagate (current slice M, down)
    order = [M - 1] + list(range(0, M - 1))
    for lo in range(0, M, safe_mult):
        for l in order[lo:lo + safe_mult]:
            U = eT2 @ (np.exp(lam * conf[:, l])[:, None] * U)
        U, D, T1 = udt(U * D[None, :], mode, nb)
        T = T1 @ T
    # G = [1 + U D T]^-1 = T^-1 [U' T^-1 + D]^-1 U'   (calculate_greens with identity on the right)
    Ti = np.linalg.inv(T)
    Q2, D2, T2 = udt(U.T @ Ti + np.diag(D), mode, nb)
    return Ti @ np.linalg.inv(T2) @ np.diag(1.0 / D2) @ Q2.T @ U.T, D


def main():
    L, U_, beta, dtau, nb = int(sys.argv[1]), float(sys.argv[2]), float(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5])
    model = mcb.HubbardModelAttractive(L=L, U=U_)
    M = int(round(beta / dtau))
    conf = model.rand_conf(M, 123, 0)
    t = time.time()
    o = orc.OracleDQMC(L * L, M, 10, dtau, U_, model.hopping_exponentials(dtau), conf, seed=123, chain=0)
    o.build_stack()
    Go = o.greens()
    print(f"oracle build_stack {time.time() - t:.1f} s; max|G| = {np.abs(Go).max():.3f}")
    for mode in ("qrcp", "panel", "prepivot"):
        t = time.time()
        G, D = greens_from_stack(model, conf, dtau, 10, mode, nb)
        print(f"{mode:9s} nb={nb:3d}: rel max-norm diff vs oracle {np.abs(G - Go).max() / np.abs(Go).max():.3e}   "
              f"(log10 D range {np.log10(D.max() / D.min()):.1f}, {time.time() - t:.1f} s)")


if __name__ == "__main__":
    main()
