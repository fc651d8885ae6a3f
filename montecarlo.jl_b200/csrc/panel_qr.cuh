// Chain-resident UDT building blocks on the register layout of panel.cuh (one CTA per chain):
//   gemm_regB   C = opA(A) * X with X read out of the registers of the owning warp by shuffles
//   qr_pivot    column-pivoted Householder QR   (udt_AVX_pivot!, src/linalg/UDT.jl:192-249)
//   form_q      explicit Q by backward accumulation (UDT.jl:250-266)
//   make_t      D = |diag R|, T = D^-1 R P'      (UDT.jl:271-297)
//   rsolve_t    X' = T^-T B'  i.e.  X = B T^-1   (rdivp!, src/linalg/general.jl:138-166)
//
// Columns are never moved: the pivoting is a map position <-> physical column kept in shared
// memory, so T = D^-1 R P' is simply the factored matrix with everything below each column's
// position zeroed, and rdivp!(A, T_dirty, pivot) == A * T^-1 with that same T.
#pragma once
#include "panel.cuh"

namespace mcb {
namespace panel {

constexpr unsigned FULL = 0xffffffffu;

template <int N>
struct QrSmem {
    static constexpr int NW = Geo<N>::NW;
    double cand[2][NW][N]; // per warp: its best remaining column (double buffered over steps)
    double cval[2][NW];    // ... the squared norm of that column's trailing part
    int ckey[2][NW];       // ... (current position << 16) | physical column
    double tau[N];
    double D[N];
    double Dinv[N];
    int piv[N]; // physical column at position j
    int pos[N]; // position of physical column k
};

// out = opA(A) * X. A in shared memory (TA: stored transposed), X = this thread block's register
// matrix; a warp's B fragments are its own column panel, fetched with two shuffles per k-step.
template <int N, bool TA>
__device__ __forceinline__ void gemm_regB(double (&out)[Geo<N>::MT][2], const double* As,
                                          const double (&X)[Geo<N>::MT][2], const Lane& t)
{
    constexpr int LD = Geo<N>::LD, MT = Geo<N>::MT;
#pragma unroll
    for (int m = 0; m < MT; ++m) out[m][0] = out[m][1] = 0.0;
    const double* ap = TA ? As + t.q + t.r * LD : As + t.r + t.q * LD;
    const bool odd = t.r & 1;
#pragma unroll
    for (int k0 = 0; k0 < N; k0 += 4) {
        // B[k0 + q][8w + r] lives in lane ((k0 % 8 + q), r >> 1), register X[k0 / 8][r & 1]
        const int src = (((k0 & 7) + t.q) << 2) | (t.r >> 1);
        const double v0 = __shfl_sync(FULL, X[k0 / 8][0], src);
        const double v1 = __shfl_sync(FULL, X[k0 / 8][1], src);
        const double b = odd ? v1 : v0;
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            const double a = TA ? ap[k0 + 8 * m * LD] : ap[8 * m + k0 * LD];
            dmma884(out[m][0], out[m][1], a, b);
        }
    }
}

// sum over the 8 lanes that share a column (same q), result in all of them
__device__ __forceinline__ double col_sum(double v)
{
    v += __shfl_xor_sync(FULL, v, 4);
    v += __shfl_xor_sync(FULL, v, 8);
    v += __shfl_xor_sync(FULL, v, 16);
    return v;
}

// Column-pivoted Householder QR of the register matrix A (in place). On return the physical column
// k = qs.piv[j] holds R[0..j, j] in rows <= j and the Householder vector of step j in rows > j
// (implicit leading 1), qs.tau[j] its factor. Pivot rule of indmaxcolumn (UDT.jl:151-168): first
// strict maximum of the freshly recomputed trailing column norms, in the column order the
// reference's swaps would have produced (tracked as `curpos`). One __syncthreads per step.
template <int N>
__device__ __forceinline__ void qr_pivot(double (&A)[Geo<N>::MT][2], QrSmem<N>& qs, const Lane& t)
{
    constexpr int MT = Geo<N>::MT, NW = Geo<N>::NW;
    const int col0 = 8 * t.w + 2 * t.q;
    int curpos[2] = {col0, col0 + 1};
    bool done[2] = {false, false};
    double nrm[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        double s = 0.0;
#pragma unroll
        for (int m = 0; m < MT; ++m) s += A[m][c] * A[m][c];
        nrm[c] = col_sum(s);
    }
#pragma unroll
    for (int mj = 0; mj < MT; ++mj) { // row block of the step: rows below 8*mj are finished (compile time)
#pragma unroll 1
        for (int rj = 0; rj < 8; ++rj) {
            const int j = 8 * mj + rj;
            const int buf = j & 1;
            // ---- best remaining column of this warp
            double bv = done[0] ? -1.0 : nrm[0];
            int bk = (curpos[0] << 16) | col0;
            {
                const double v1 = done[1] ? -1.0 : nrm[1];
                const int k1 = (curpos[1] << 16) | (col0 + 1);
                if (v1 > bv || (v1 == bv && k1 < bk)) { bv = v1; bk = k1; }
            }
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                const double ov = __shfl_xor_sync(FULL, bv, o);
                const int ok = __shfl_xor_sync(FULL, bk, o);
                if (ov > bv || (ov == bv && ok < bk)) { bv = ov; bk = ok; }
            }
            {
                const int bc = bk & 0xffff;
                double* cd = qs.cand[buf][t.w] + t.r;
                if (bc == col0) {
#pragma unroll
                    for (int m = mj; m < MT; ++m) cd[8 * m] = A[m][0];
                } else if (bc == col0 + 1) {
#pragma unroll
                    for (int m = mj; m < MT; ++m) cd[8 * m] = A[m][1];
                }
                if (t.l == 0) { qs.cval[buf][t.w] = bv; qs.ckey[buf][t.w] = bk; }
            }
            __syncthreads();
            // ---- global winner (every thread evaluates the same NW candidates)
            double maxval = qs.cval[buf][0];
            int key = qs.ckey[buf][0], wb = 0;
#pragma unroll
            for (int w = 1; w < NW; ++w) {
                const double v = qs.cval[buf][w];
                const int k = qs.ckey[buf][w];
                if (v > maxval || (v == maxval && k < key)) { maxval = v; key = k; wb = w; }
            }
            const int pj = key & 0xffff, jm = key >> 16;
            // ---- reflector! (UDT.jl:133-148)
            const double* x = qs.cand[buf][wb];
            const double xj = x[j];
            double nu = 0.0, tj = 0.0, inv = 0.0;
            const bool live = maxval > 0.0;
            if (live) {
                // nu = copysign(||x||, x_j), tau = (x_j + nu) / nu, inv = 1 / (x_j + nu), from one
                // reciprocal square root and one reciprocal (Newton-refined hardware seeds)
                const double rs = fast_rsqrt(maxval);
                nu = copysign(maxval * rs, xj);
                tj = (xj + nu) * copysign(rs, xj);
                inv = fast_rcp(xj + nu);
            }
            // v = inv * xt with xt = (0 .. 0, x_j + nu, x[j+1:]); xt is read from the candidate buffer
            // where it is needed (no register copy), inv is folded into the scalars. Only the row
            // blocks m >= mj are touched.
            const double xjn = xj + nu;
            const double xd = t.r < rj ? 0.0 : (t.r == rj ? (live ? xjn : 0.0) : x[8 * mj + t.r]);
            // ---- reflectorApply! on the remaining columns (UDT.jl:32-50) + fresh trailing norms
            double dot0 = xd * A[mj][0], dot1 = xd * A[mj][1];
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double xm = x[8 * m + t.r];
                dot0 += xm * A[m][0];
                dot1 += xm * A[m][1];
            }
            dot0 = col_sum(dot0);
            dot1 = col_sum(dot1);
            const bool piv0 = (col0 == pj), piv1 = (col0 + 1 == pj);
            const double tinv2 = tj * inv * inv; // A -= xt * (tau inv^2 (xt . A))
            const double s0 = (done[0] || piv0) ? 0.0 : tinv2 * dot0;
            const double s1 = (done[1] || piv1) ? 0.0 : tinv2 * dot1;
            A[mj][0] -= s0 * xd;
            A[mj][1] -= s1 * xd;
            double nn0 = t.r > rj ? A[mj][0] * A[mj][0] : 0.0;
            double nn1 = t.r > rj ? A[mj][1] * A[mj][1] : 0.0;
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double xm = x[8 * m + t.r];
                A[m][0] -= s0 * xm;
                A[m][1] -= s1 * xm;
                nn0 += A[m][0] * A[m][0];
                nn1 += A[m][1] * A[m][1];
            }
            nrm[0] = col_sum(nn0);
            nrm[1] = col_sum(nn1);
            if (piv0) done[0] = true;
            else if (!done[0] && curpos[0] == j) curpos[0] = jm; // the swap j <-> jm (UDT.jl:219-231)
            if (piv1) done[1] = true;
            else if (!done[1] && curpos[1] == j) curpos[1] = jm;
            // the pivot column keeps R above the diagonal, -nu on it and the Householder vector below
            if (live && piv0) {
                if (t.r >= rj) A[mj][0] = t.r == rj ? -nu : xd * inv;
#pragma unroll
                for (int m = mj + 1; m < MT; ++m) A[m][0] = x[8 * m + t.r] * inv;
            } else if (live && piv1) {
                if (t.r >= rj) A[mj][1] = t.r == rj ? -nu : xd * inv;
#pragma unroll
                for (int m = mj + 1; m < MT; ++m) A[m][1] = x[8 * m + t.r] * inv;
            }
            if (t.tid == 0) { qs.tau[j] = tj; qs.piv[j] = pj; qs.pos[pj] = j; }
        }
    }
    __syncthreads();
}

// Q = H_0 H_1 ... H_{N-1} (explicit). F = the factored matrix in shared memory (ld = LD), written
// there by the caller from qr_pivot's result; every warp builds its own 8 columns independently.
template <int N>
__device__ __forceinline__ void form_q(double (&Q)[Geo<N>::MT][2], const double* F, const QrSmem<N>& qs,
                                       const Lane& t)
{
    constexpr int MT = Geo<N>::MT, LD = Geo<N>::LD;
    const int col0 = 8 * t.w + 2 * t.q;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        Q[m][0] = (8 * m + t.r == col0) ? 1.0 : 0.0;
        Q[m][1] = (8 * m + t.r == col0 + 1) ? 1.0 : 0.0;
    }
#pragma unroll
    for (int mk = MT - 1; mk >= 0; --mk) {
#pragma unroll 1
        for (int rk = 7; rk >= 0; --rk) {
            const int k = 8 * mk + rk;
            if (8 * t.w + 7 < k) continue; // columns left of k are still unit vectors: H_k leaves them alone
            const double* x = F + (size_t)qs.piv[k] * LD;
            const double tk = qs.tau[k];
            const double vd = t.r < rk ? 0.0 : (t.r == rk ? 1.0 : x[8 * mk + t.r]);
            double dot0 = vd * Q[mk][0], dot1 = vd * Q[mk][1];
#pragma unroll
            for (int m = mk + 1; m < MT; ++m) {
                const double vm = x[8 * m + t.r];
                dot0 += vm * Q[m][0];
                dot1 += vm * Q[m][1];
            }
            const double s0 = tk * col_sum(dot0), s1 = tk * col_sum(dot1);
            Q[mk][0] -= s0 * vd;
            Q[mk][1] -= s1 * vd;
#pragma unroll
            for (int m = mk + 1; m < MT; ++m) {
                const double vm = x[8 * m + t.r];
                Q[m][0] -= s0 * vm;
                Q[m][1] -= s1 * vm;
            }
        }
    }
}

// D_i = |R_ii| and 1/D_i into qs (from the shared-memory copy F of the factored matrix)
template <int N>
__device__ __forceinline__ void make_d(const double* F, QrSmem<N>& qs, const Lane& t)
{
    constexpr int LD = Geo<N>::LD;
    for (int i = t.tid; i < N; i += Geo<N>::THREADS) {
        const double d = fabs(F[i + (size_t)qs.piv[i] * LD]);
        qs.D[i] = d;
        qs.Dinv[i] = 1.0 / d;
    }
}

// A (factored, registers) -> T = D^-1 R P' in place (UDT.jl:283-297). Needs qs.Dinv, qs.pos.
template <int N>
__device__ __forceinline__ void make_t(double (&A)[Geo<N>::MT][2], const QrSmem<N>& qs, const Lane& t)
{
    constexpr int MT = Geo<N>::MT;
    const int col0 = 8 * t.w + 2 * t.q;
    const int p0 = qs.pos[col0], p1 = qs.pos[col0 + 1];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const int i = 8 * m + t.r;
        const double di = qs.Dinv[i];
        A[m][0] = i <= p0 ? A[m][0] * di : 0.0;
        A[m][1] = i <= p1 ? A[m][1] * di : 0.0;
    }
}

// Y = T^-T B (both register matrices hold TRANSPOSES: B = b', Y = x' with x = b T^-1).
// Tt = T stored transposed in shared memory: T[j][k] at Tt[k + j*LD]. Forward substitution over
// positions j; each warp solves its own 8 columns, no block-level synchronisation.
template <int N>
__device__ __forceinline__ void rsolve_t(double (&B)[Geo<N>::MT][2], double (&Y)[Geo<N>::MT][2],
                                         const double* Tt, const QrSmem<N>& qs, const Lane& t)
{
    constexpr int MT = Geo<N>::MT, LD = Geo<N>::LD;
#pragma unroll
    for (int mj = 0; mj < MT; ++mj) {
#pragma unroll
        for (int rj = 0; rj < 8; ++rj) {
            const int j = 8 * mj + rj;
            const int p = qs.piv[j];
            const int pm = p >> 3;
            double s0 = B[0][0], s1 = B[0][1];
#pragma unroll
            for (int m = 1; m < MT; ++m)
                if (pm == m) { s0 = B[m][0]; s1 = B[m][1]; }
            const int src = ((p & 7) << 2) | t.q;
            const double* trow = Tt + (size_t)j * LD;
            const double rinv = fast_rcp(trow[p]);
            const double y0 = __shfl_sync(FULL, s0, src) * rinv;
            const double y1 = __shfl_sync(FULL, s1, src) * rinv;
            if (t.r == rj) { Y[mj][0] = y0; Y[mj][1] = y1; }
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const double tv = trow[8 * m + t.r];
                B[m][0] -= tv * y0;
                B[m][1] -= tv * y1;
            }
        }
    }
}

// global (column-major, ld = N) -> registers holding the TRANSPOSE: acc(i, j) = g[j + i*N]
template <int N>
__device__ __forceinline__ void g2r_t(double (&acc)[Geo<N>::MT][2], const double* __restrict__ g, const Lane& t)
{
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m) {
        const double2 v = *reinterpret_cast<const double2*>(g + (8 * t.w + 2 * t.q) + (size_t)(8 * m + t.r) * N);
        acc[m][0] = v.x; acc[m][1] = v.y;
    }
}

// shared (ld = LD) -> registers holding the TRANSPOSE: acc(i, j) = S[j + i*LD]
template <int N>
__device__ __forceinline__ void s2r_t(double (&acc)[Geo<N>::MT][2], const double* S, const Lane& t)
{
    constexpr int LD = Geo<N>::LD;
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m) {
        const double2 v = *reinterpret_cast<const double2*>(S + (8 * t.w + 2 * t.q) + (8 * m + t.r) * LD);
        acc[m][0] = v.x; acc[m][1] = v.y;
    }
}

} // namespace panel
} // namespace mcb
