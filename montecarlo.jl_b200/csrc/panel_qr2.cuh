// Column-pivoted Householder QR (udt_AVX_pivot!, src/linalg/UDT.jl:192-266) of a register-resident matrix:
// a CTA's N / (8 CW) warps hold CW 8-column blocks each,
//     lane l (r = l>>2, q = l&3): A[m][cw][c] = X[8m + r][8 CW w + 8 cw + 2q + c].
// The per-step work that does not shrink with the warp count (pivot search, reflector scalars, barrier)
// is replicated in every warp, the per-column work grows with CW: CW = 2 is the measured optimum at
// N = 64 (CW = 1: 8.8 ms, CW = 2: 6.05 ms, CW = 4: 7.6 ms of factorisations per dqmc_A512 sweep).
// The warps synchronise on named barrier 1; pivoting state (tau, piv, pos) lives in QrSmem.
#pragma once
#include "panel_qr.cuh"

namespace mcb {
namespace panel {

template <int N, int CW_ = 2>
struct Geo2 {
    static_assert(N % (8 * CW_) == 0, "needs N to be a multiple of the warp's column count");
    static constexpr int CW = CW_;         // 8-column blocks per warp
    static constexpr int KC = 2 * CW_;     // columns per lane
    static constexpr int NW = N / (8 * CW_); // participating warps
    static constexpr int NT = NW * 32;     // participating threads
    static constexpr int MT = N / 8;
    static constexpr int LD = Geo<N>::LD;
};

__device__ __forceinline__ void bar2(int nthreads)
{
    asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
}

struct Lane2 {
    int w, l, r, q, col0;
    __device__ __forceinline__ Lane2(int cw = 2)
    {
        w = threadIdx.x >> 5; l = threadIdx.x & 31; r = l >> 2; q = l & 3; col0 = 8 * cw * w + 2 * q;
    }
};

template <int N, int CW = 2>
__device__ __forceinline__ void r2s2(const double (&A)[Geo2<N>::MT][CW][2], double* S, const Lane2& t)
{
    constexpr int LD = Geo2<N>::LD;
#pragma unroll
    for (int m = 0; m < Geo2<N>::MT; ++m)
#pragma unroll
        for (int cw = 0; cw < CW; ++cw)
#pragma unroll
            for (int c = 0; c < 2; ++c) S[(8 * m + t.r) + (t.col0 + 8 * cw + c) * LD] = A[m][cw][c];
}

// Column-pivoted Householder QR, in place. On return the physical column k = qs.piv[j] holds R[0..j, j] in
// rows <= j and the Householder vector of step j in rows > j (implicit leading 1), qs.tau[j] its factor.
// Pivot rule of indmaxcolumn (UDT.jl:151-168): first strict maximum of the freshly recomputed trailing
// column norms, in the column order the reference's swaps would have produced (tracked as `curpos`).
// A Householder step is one long dependency chain (pivot search -> barrier -> reflector scalars -> dot
// products -> update -> fresh norms), so the code is arranged for latency rather than instruction count:
//  * squared norms are non-negative doubles, so their bit patterns order like integers: the pivot
//    search compares (hi, lo, key) words on the integer pipe and reduces over the warp with REDUX;
//  * the pivot column stays untouched in registers; -nu and 1/(x_j + nu) are kept per step
//    (qs.D / qs.Dinv, overwritten by make_d later) and applied once after the last step.
// EARLY (large N, F != nullptr): a pivot column is scaled the moment it is eliminated and every block of 8 rows goes
// to the shared-memory matrix F as soon as its 8 steps are over, so that its registers are free for the rest of
// the factorisation (at N = 144 the 144 matrix registers per thread do not fit next to the working set).
template <int N, int CW = 2, bool EARLY = false>
__device__ __forceinline__ void qr_pivot2(double (&A)[Geo2<N>::MT][CW][2], QrSmem<N>& qs, const Lane2& t,
                                          double* F = nullptr)
{
    constexpr int MT = Geo2<N, CW>::MT, NW = Geo2<N, CW>::NW, NT = Geo2<N, CW>::NT, KC = 2 * CW;
    constexpr int LDF = Geo2<N, CW>::LD;
    int curpos[KC];
    bool done[KC];
    double nrm[KC];
#pragma unroll
    for (int k = 0; k < KC; ++k) {
        curpos[k] = t.col0 + 8 * (k >> 1) + (k & 1);
        done[k] = false;
        double s = 0.0;
#pragma unroll
        for (int m = 0; m < MT; ++m) s += A[m][k >> 1][k & 1] * A[m][k >> 1][k & 1];
        nrm[k] = col_sum(s);
    }
#pragma unroll
    for (int mj = 0; mj < MT; ++mj) {
#pragma unroll 1
        for (int rj = 0; rj < 8; ++rj) {
            const int j = 8 * mj + rj;
            const int buf = j & 1;
            // ---- best remaining column of this warp: max norm, ties to the lowest position
            int hi[KC], key[KC];
            unsigned lo[KC];
#pragma unroll
            for (int k = 0; k < KC; ++k) {
                hi[k] = done[k] ? -1 : __double2hiint(nrm[k]);
                lo[k] = (unsigned)__double2loint(nrm[k]);
                key[k] = (curpos[k] << 16) | (t.col0 + 8 * (k >> 1) + (k & 1));
            }
            int mh = hi[0];
#pragma unroll
            for (int k = 1; k < KC; ++k) mh = max(mh, hi[k]);
            mh = __reduce_max_sync(FULL, mh);
            unsigned ml = 0u;
#pragma unroll
            for (int k = 0; k < KC; ++k) ml = max(ml, hi[k] == mh ? lo[k] : 0u);
            ml = __reduce_max_sync(FULL, ml);
            int bk = 0x7fffffff;
#pragma unroll
            for (int k = 0; k < KC; ++k) bk = min(bk, (hi[k] == mh && lo[k] == ml) ? key[k] : 0x7fffffff);
            bk = __reduce_min_sync(FULL, bk);
            {
                // the warp's candidate column (bk is uniform over the warp)
                const int bc = bk & 0xffff;
                double* cd = qs.cand[buf][t.w] + t.r;
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    const bool mine = bc == t.col0 + 8 * (k >> 1) + (k & 1);
#pragma unroll
                    for (int m = mj; m < MT; ++m)
                        if (mine) cd[8 * m] = A[m][k >> 1][k & 1];
                }
                if (t.l == 0) { qs.cval[buf][t.w] = __hiloint2double(mh, (int)ml); qs.ckey[buf][t.w] = bk; }
            }
            bar2(NT);
            // ---- global winner
            int wh[NW], wk[NW];
            unsigned wl[NW];
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                const double v = qs.cval[buf][w];
                wh[w] = __double2hiint(v); wl[w] = (unsigned)__double2loint(v); wk[w] = qs.ckey[buf][w];
            }
            int gh = wh[0];
#pragma unroll
            for (int w = 1; w < NW; ++w) gh = max(gh, wh[w]);
            unsigned gl = 0u;
#pragma unroll
            for (int w = 0; w < NW; ++w) gl = max(gl, wh[w] == gh ? wl[w] : 0u);
            int gkey = 0x7fffffff;
#pragma unroll
            for (int w = 0; w < NW; ++w) gkey = min(gkey, (wh[w] == gh && wl[w] == gl) ? wk[w] : 0x7fffffff);
            const double maxval = __hiloint2double(gh, (int)gl);
            const int pj = gkey & 0xffff, jm = gkey >> 16;
            const double* x = qs.cand[buf][pj / (8 * CW)];
            const double xlow = t.r > rj ? x[8 * mj + t.r] : 0.0;
            // (large N: the vector is read again in the update loop instead of being held in MT registers)
            constexpr bool RELOAD = N > 64;
            double xm[RELOAD ? 1 : MT];
            if (!RELOAD) {
#pragma unroll
                for (int m = mj + 1; m < MT; ++m) xm[m] = x[8 * m + t.r];
            }
            const double xj = x[j];
            const bool live = maxval > 0.0;
            const double rs = live ? fast_rsqrt(maxval) : 0.0;
            const double nu = copysign(maxval * rs, xj);
            const double xjn = xj + nu;
            const double tj = xjn * copysign(rs, xj);
            const double inv = live ? fast_rcp(xjn) : 0.0;
            const double tinv2 = tj * inv * inv;
            const double xd = t.r == rj ? xjn : xlow;
            double dot[KC];
#pragma unroll
            for (int k = 0; k < KC; ++k) dot[k] = xd * A[mj][k >> 1][k & 1];
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double xv = RELOAD ? x[8 * m + t.r] : xm[RELOAD ? 0 : m];
#pragma unroll
                for (int k = 0; k < KC; ++k) dot[k] += xv * A[m][k >> 1][k & 1];
            }
#pragma unroll
            for (int k = 0; k < KC; ++k) dot[k] = col_sum(dot[k]);
            // ---- reflectorApply! (UDT.jl:32-50) + fresh trailing norms
            if (RELOAD) asm volatile("" ::: "memory");
            double s[KC], nn[KC];
#pragma unroll
            for (int k = 0; k < KC; ++k) {
                const bool is_piv = (t.col0 + 8 * (k >> 1) + (k & 1) == pj);
                s[k] = (done[k] || is_piv) ? 0.0 : tinv2 * dot[k];
                A[mj][k >> 1][k & 1] -= s[k] * xd;
                nn[k] = t.r > rj ? A[mj][k >> 1][k & 1] * A[mj][k >> 1][k & 1] : 0.0;
            }
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double xv = RELOAD ? x[8 * m + t.r] : xm[RELOAD ? 0 : m];
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    A[m][k >> 1][k & 1] -= s[k] * xv;
                    nn[k] += A[m][k >> 1][k & 1] * A[m][k >> 1][k & 1];
                }
            }
#pragma unroll
            for (int k = 0; k < KC; ++k) nrm[k] = col_sum(nn[k]);
#pragma unroll
            for (int k = 0; k < KC; ++k) {
                const bool is_piv = (t.col0 + 8 * (k >> 1) + (k & 1) == pj);
                // the swap j <-> jm (UDT.jl:219-231); a pivot column remembers its final position
                curpos[k] = is_piv ? j : ((!done[k] && curpos[k] == j) ? jm : curpos[k]);
                done[k] = done[k] || is_piv;
            }
            if (threadIdx.x == 0) {
                qs.tau[j] = tj; qs.piv[j] = pj; qs.pos[pj] = j;
                qs.D[j] = -nu; qs.Dinv[j] = inv;
            }
            if (EARLY && pj / (8 * CW) == t.w) {   // my warp owns the pivot column: -nu on row j, vector below
#pragma unroll
                for (int k = 0; k < KC; ++k) {
                    const bool is_piv = (t.col0 + 8 * (k >> 1) + (k & 1) == pj);
                    const double a = A[mj][k >> 1][k & 1];
                    A[mj][k >> 1][k & 1] = !is_piv ? a : (t.r < rj ? a : (t.r == rj ? -nu : a * inv));
#pragma unroll
                    for (int m = mj + 1; m < MT; ++m) A[m][k >> 1][k & 1] *= is_piv ? inv : 1.0;
                }
            }
        }
        if (EARLY) {   // rows 8 mj .. 8 mj + 7 are final
#pragma unroll
            for (int k = 0; k < KC; ++k)
                F[(8 * mj + t.r) + (size_t)(t.col0 + 8 * (k >> 1) + (k & 1)) * LDF] = A[mj][k >> 1][k & 1];
        }
    }
    bar2(NT);
    if (EARLY) return;
    // every column keeps R above its position, -nu on it and the Householder vector below
#pragma unroll
    for (int k = 0; k < KC; ++k) {
        const int pos = curpos[k];
        const double dg = qs.D[pos], iv = qs.Dinv[pos];
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            const int i = 8 * m + t.r;
            const double a = A[m][k >> 1][k & 1];
            A[m][k >> 1][k & 1] = i < pos ? a : (i == pos ? dg : a * iv);
        }
    }
    bar2(NT); // qs.D / qs.Dinv are free again
}

// Q = H_0 H_1 ... H_{N-1} (explicit, UDT.jl:250-266). F = the factored matrix in shared memory (ld = LD);
// every warp builds its own columns independently.
template <int N, int CW = 2>
__device__ __forceinline__ void form_q2(double (&Q)[Geo2<N>::MT][CW][2], const double* F, const QrSmem<N>& qs,
                                        const Lane2& t)
{
    constexpr int MT = Geo2<N>::MT, LD = Geo2<N>::LD, KC = 2 * CW;
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int k = 0; k < KC; ++k) Q[m][k >> 1][k & 1] = (8 * m + t.r == t.col0 + 8 * (k >> 1) + (k & 1)) ? 1.0 : 0.0;
#pragma unroll
    for (int mk = MT - 1; mk >= 0; --mk) {
#pragma unroll 1
        for (int rk = 7; rk >= 0; --rk) {
            const int k = 8 * mk + rk;
            if (8 * CW * t.w + 8 * CW - 1 < k) continue; // columns left of k are still unit vectors
            const double* x = F + (size_t)qs.piv[k] * LD;
            const double tk = qs.tau[k];
            const double vd = t.r < rk ? 0.0 : (t.r == rk ? 1.0 : x[8 * mk + t.r]);
            double dot[KC];
#pragma unroll
            for (int c = 0; c < KC; ++c) dot[c] = vd * Q[mk][c >> 1][c & 1];
#pragma unroll
            for (int m = mk + 1; m < MT; ++m) {
                const double vm = x[8 * m + t.r];
#pragma unroll
                for (int c = 0; c < KC; ++c) dot[c] += vm * Q[m][c >> 1][c & 1];
            }
            double s[KC];
#pragma unroll
            for (int c = 0; c < KC; ++c) {
                s[c] = tk * col_sum(dot[c]);
                Q[mk][c >> 1][c & 1] -= s[c] * vd;
            }
#pragma unroll
            for (int m = mk + 1; m < MT; ++m) {
                const double vm = x[8 * m + t.r];
#pragma unroll
                for (int c = 0; c < KC; ++c) Q[m][c >> 1][c & 1] -= s[c] * vm;
            }
        }
    }
}

} // namespace panel
} // namespace mcb
