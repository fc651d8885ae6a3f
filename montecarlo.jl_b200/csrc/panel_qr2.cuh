// The Householder steps of panel_qr.cuh on HALF the warps: the first N/16 warps of the CTA hold the
// matrix with 16 columns per warp,
//     lane l (r = l>>2, q = l&3): A[m][cw][c] = X[8m + r][16 w + 8 cw + 2q + c],
// the other warps wait. The per-step work that does not shrink with the warp count (pivot search,
// reflector scalars, barrier) is replicated in every participating warp, so fewer, fatter warps
// execute about half the instructions per Householder step. The participating warps synchronise on
// named barrier 1; everything else (pivot rule, reflector, storage of R / v / tau / piv / pos in
// QrSmem) is identical to qr_pivot / form_q.
#pragma once
#include "panel_qr.cuh"

namespace mcb {
namespace panel {

template <int N>
struct Geo2 {
    static_assert(N % 16 == 0, "needs N to be a multiple of 16");
    static constexpr int CW = 2;
    static constexpr int NW = N / 16;      // participating warps
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
    __device__ __forceinline__ Lane2()
    {
        w = threadIdx.x >> 5; l = threadIdx.x & 31; r = l >> 2; q = l & 3; col0 = 16 * w + 2 * q;
    }
};

template <int N>
__device__ __forceinline__ void s2r2(double (&A)[Geo2<N>::MT][2][2], const double* S, const Lane2& t)
{
    constexpr int LD = Geo2<N>::LD;
#pragma unroll
    for (int m = 0; m < Geo2<N>::MT; ++m)
#pragma unroll
        for (int cw = 0; cw < 2; ++cw)
#pragma unroll
            for (int c = 0; c < 2; ++c) A[m][cw][c] = S[(8 * m + t.r) + (t.col0 + 8 * cw + c) * LD];
}

template <int N>
__device__ __forceinline__ void r2s2(const double (&A)[Geo2<N>::MT][2][2], double* S, const Lane2& t)
{
    constexpr int LD = Geo2<N>::LD;
#pragma unroll
    for (int m = 0; m < Geo2<N>::MT; ++m)
#pragma unroll
        for (int cw = 0; cw < 2; ++cw)
#pragma unroll
            for (int c = 0; c < 2; ++c) S[(8 * m + t.r) + (t.col0 + 8 * cw + c) * LD] = A[m][cw][c];
}

// Column-pivoted Householder QR (see qr_pivot). Called by the first Geo2<N>::NW warps only.
template <int N>
__device__ __forceinline__ void qr_pivot2(double (&A)[Geo2<N>::MT][2][2], QrSmem<N>& qs, const Lane2& t)
{
    constexpr int MT = Geo2<N>::MT, NW = Geo2<N>::NW, NT = Geo2<N>::NT;
    int curpos[4];
    bool done[4];
    double nrm[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
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
            // ---- best remaining column of this warp
            double bv = -2.0;
            int bk = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double v = done[k] ? -1.0 : nrm[k];
                const int key = (curpos[k] << 16) | (t.col0 + 8 * (k >> 1) + (k & 1));
                if (v > bv || (v == bv && key < bk)) { bv = v; bk = key; }
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
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (bc == t.col0 + 8 * (k >> 1) + (k & 1)) {
#pragma unroll
                        for (int m = mj; m < MT; ++m) cd[8 * m] = A[m][k >> 1][k & 1];
                    }
                if (t.l == 0) { qs.cval[buf][t.w] = bv; qs.ckey[buf][t.w] = bk; }
            }
            bar2(NT);
            // ---- global winner
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
                const double rs = fast_rsqrt(maxval);
                nu = copysign(maxval * rs, xj);
                tj = (xj + nu) * copysign(rs, xj);
                inv = fast_rcp(xj + nu);
            }
            const double xjn = xj + nu;
            const double xd = t.r < rj ? 0.0 : (t.r == rj ? (live ? xjn : 0.0) : x[8 * mj + t.r]);
            // ---- reflectorApply! (UDT.jl:32-50) + fresh trailing norms
            double dot[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) dot[k] = xd * A[mj][k >> 1][k & 1];
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double xm = x[8 * m + t.r];
#pragma unroll
                for (int k = 0; k < 4; ++k) dot[k] += xm * A[m][k >> 1][k & 1];
            }
            const double tinv2 = tj * inv * inv;
            double s[4], nn[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const bool is_piv = (t.col0 + 8 * (k >> 1) + (k & 1) == pj);
                const double d = col_sum(dot[k]);
                s[k] = (done[k] || is_piv) ? 0.0 : tinv2 * d;
                A[mj][k >> 1][k & 1] -= s[k] * xd;
                nn[k] = t.r > rj ? A[mj][k >> 1][k & 1] * A[mj][k >> 1][k & 1] : 0.0;
            }
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double xm = x[8 * m + t.r];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    A[m][k >> 1][k & 1] -= s[k] * xm;
                    nn[k] += A[m][k >> 1][k & 1] * A[m][k >> 1][k & 1];
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const bool is_piv = (t.col0 + 8 * (k >> 1) + (k & 1) == pj);
                nrm[k] = col_sum(nn[k]);
                if (is_piv) done[k] = true;
                else if (!done[k] && curpos[k] == j) curpos[k] = jm; // the swap j <-> jm (UDT.jl:219-231)
                // the pivot column keeps R above the diagonal, -nu on it and the Householder vector below
                if (live && is_piv) {
                    if (t.r >= rj) A[mj][k >> 1][k & 1] = t.r == rj ? -nu : xd * inv;
#pragma unroll
                    for (int m = mj + 1; m < MT; ++m) A[m][k >> 1][k & 1] = x[8 * m + t.r] * inv;
                }
            }
            if (threadIdx.x == 0) { qs.tau[j] = tj; qs.piv[j] = pj; qs.pos[pj] = j; }
        }
    }
    bar2(NT);
}

// Explicit Q (see form_q); every participating warp builds its own 16 columns.
template <int N>
__device__ __forceinline__ void form_q2(double (&Q)[Geo2<N>::MT][2][2], const double* F, const QrSmem<N>& qs,
                                        const Lane2& t)
{
    constexpr int MT = Geo2<N>::MT, LD = Geo2<N>::LD;
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int k = 0; k < 4; ++k) Q[m][k >> 1][k & 1] = (8 * m + t.r == t.col0 + 8 * (k >> 1) + (k & 1)) ? 1.0 : 0.0;
#pragma unroll
    for (int mk = MT - 1; mk >= 0; --mk) {
#pragma unroll 1
        for (int rk = 7; rk >= 0; --rk) {
            const int k = 8 * mk + rk;
            if (16 * t.w + 15 < k) continue; // columns left of k are still unit vectors
            const double* x = F + (size_t)qs.piv[k] * LD;
            const double tk = qs.tau[k];
            const double vd = t.r < rk ? 0.0 : (t.r == rk ? 1.0 : x[8 * mk + t.r]);
            double dot[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) dot[c] = vd * Q[mk][c >> 1][c & 1];
#pragma unroll
            for (int m = mk + 1; m < MT; ++m) {
                const double vm = x[8 * m + t.r];
#pragma unroll
                for (int c = 0; c < 4; ++c) dot[c] += vm * Q[m][c >> 1][c & 1];
            }
            double s[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                s[c] = tk * col_sum(dot[c]);
                Q[mk][c >> 1][c & 1] -= s[c] * vd;
            }
#pragma unroll
            for (int m = mk + 1; m < MT; ++m) {
                const double vm = x[8 * m + t.r];
#pragma unroll
                for (int c = 0; c < 4; ++c) Q[m][c >> 1][c & 1] -= s[c] * vm;
            }
        }
    }
}

} // namespace panel
} // namespace mcb
