// Chain-resident building blocks: one CTA per Markov chain, one N x N matrix held in the REGISTERS
// of the CTA in the accumulator layout of the FP64 tensor-core instruction (DMMA m8n8k4).
//
//   warps  NW = N/8 ; warp w owns the column panel [8w, 8w+8) of the matrix, all N rows
//   lane l: r = l>>2, q = l&3 ; element acc[m][c] = X[8m + r][8w + 2q + c],  m < N/8, c < 2
//
// A GEMM leaves its result in exactly this layout, the Householder QR wants whole columns inside one
// warp (column dots are 3 shuffles), and the rank-1 Green's-function update touches every element
// once with 8 + 2 operand loads per 16 FMAs -- so wrap, local updates, UDT and the pivoted solve all
// run without the matrix ever leaving the SM. Operands of a GEMM are staged in shared memory,
// column-major with leading dimension LD = N + 4 (LD % 16 == 4 makes every DMMA fragment load
// conflict-free: a half-warp reads rows r = 0..3 of columns q = 0..3 -> banks r + 4q).
#pragma once
#include "common.cuh"

namespace mcb {
namespace panel {

template <int N>
struct Geo {
    static_assert(N % 16 == 0, "panel kernels need N to be a multiple of 16");
    static constexpr int NW = N / 8;
    static constexpr int MT = N / 8;
    static constexpr int THREADS = NW * 32;
    static constexpr int LD = N + 4;
    static constexpr int MAT = N * LD; // doubles per shared-memory matrix
};

struct Lane {
    int tid, w, l, r, q;
    __device__ __forceinline__ Lane()
    {
        tid = threadIdx.x; w = tid >> 5; l = tid & 31; r = l >> 2; q = l & 3;
    }
};

// global (column-major, ld = N) -> shared (ld = LD), optionally transposed
template <int N, bool TRANS>
__device__ __forceinline__ void g2s(double* __restrict__ S, const double* __restrict__ g, int tid)
{
    constexpr int LD = Geo<N>::LD, TH = Geo<N>::THREADS;
#pragma unroll 4
    for (int e = tid; e < N * N; e += TH) {
        const int i = e % N, j = e / N;
        if (TRANS) S[j + i * LD] = g[e]; else S[i + j * LD] = g[e];
    }
}

template <int N>
__device__ __forceinline__ void g2r(double (&acc)[Geo<N>::MT][2], const double* __restrict__ g, const Lane& t)
{
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m)
#pragma unroll
        for (int c = 0; c < 2; ++c) acc[m][c] = g[(8 * m + t.r) + (size_t)(8 * t.w + 2 * t.q + c) * N];
}

template <int N>
__device__ __forceinline__ void r2g(const double (&acc)[Geo<N>::MT][2], double* __restrict__ g, const Lane& t)
{
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m)
#pragma unroll
        for (int c = 0; c < 2; ++c) g[(8 * m + t.r) + (size_t)(8 * t.w + 2 * t.q + c) * N] = acc[m][c];
}

// registers -> shared; element (i, j) lands at S[i + j*LD], or at S[j + i*LD] when TRANS
template <int N, bool TRANS>
__device__ __forceinline__ void r2s(const double (&acc)[Geo<N>::MT][2], double* S, const Lane& t)
{
    constexpr int LD = Geo<N>::LD;
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int i = 8 * m + t.r, j = 8 * t.w + 2 * t.q + c;
            if (TRANS) S[j + i * LD] = acc[m][c]; else S[i + j * LD] = acc[m][c];
        }
}

template <int N>
__device__ __forceinline__ void s2r(double (&acc)[Geo<N>::MT][2], const double* S, const Lane& t)
{
    constexpr int LD = Geo<N>::LD;
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m)
#pragma unroll
        for (int c = 0; c < 2; ++c) acc[m][c] = S[(8 * m + t.r) + (8 * t.w + 2 * t.q + c) * LD];
}

// acc[m][c] *= rowv[8m + r]
template <int N>
__device__ __forceinline__ void scale_rows(double (&acc)[Geo<N>::MT][2], const double* rowv, const Lane& t)
{
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m) {
        const double s = rowv[8 * m + t.r];
        acc[m][0] *= s; acc[m][1] *= s;
    }
}
template <int N>
__device__ __forceinline__ void scale_cols(double (&acc)[Geo<N>::MT][2], const double* colv, const Lane& t)
{
    const double s0 = colv[8 * t.w + 2 * t.q], s1 = colv[8 * t.w + 2 * t.q + 1];
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m) { acc[m][0] *= s0; acc[m][1] *= s1; }
}

// acc = opA(A) * opB(B), both operands in shared memory (ld = LD).
//   TA: A operand is stored transposed (element A[i][k] read from As[k + i*LD])
//   TB: B operand is stored transposed (element B[k][j] read from Bs[j + k*LD])
template <int N, bool TA, bool TB>
__device__ __forceinline__ void gemm(double (&acc)[Geo<N>::MT][2], const double* As, const double* Bs,
                                     const Lane& t)
{
    constexpr int LD = Geo<N>::LD, MT = Geo<N>::MT;
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[m][0] = acc[m][1] = 0.0;
    const double* bp = TB ? Bs + (8 * t.w + t.r) + t.q * LD : Bs + t.q + (8 * t.w + t.r) * LD;
    const double* ap = TA ? As + t.q + t.r * LD : As + t.r + t.q * LD;
#pragma unroll 4
    for (int k0 = 0; k0 < N; k0 += 4) {
        const double b = TB ? bp[k0 * LD] : bp[k0];
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            const double a = TA ? ap[k0 + 8 * m * LD] : ap[8 * m + k0 * LD];
            dmma884(acc[m][0], acc[m][1], a, b);
        }
    }
}

} // namespace panel
} // namespace mcb
