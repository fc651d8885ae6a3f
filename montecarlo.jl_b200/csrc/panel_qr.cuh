// Chain-resident UDT building blocks on the register layout of panel.cuh (one CTA per chain):
//   gemm_regB   C = opA(A) * X with X read out of the registers of the owning warp by shuffles
//   QrSmem      pivoting state of a factorisation (candidate columns, tau, pivot map)
//   col_sum     reduction over the 8 lanes that hold one column
// The Householder steps themselves are in panel_qr2.cuh (qr_pivot2 / form_q2).
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
} // namespace panel
} // namespace mcb
