// Chain-resident stabilisation kernels: see stab.cuh. One CTA per chain; the N x N matrices of a step live
// in registers (accumulator layout of DMMA m8n8k4) or in shared-memory operand buffers, global memory is
// touched for the inputs and the results of a kernel. Products (FP64-pipe bound) and factorisations
// (latency bound) are separate kernels on purpose: a dependent DFMA that queues behind another warp's
// DMMAs takes several times its latency (profiles/microbench), so the two kinds must not share an SM.
#include "panel_qr2.cuh"
#include "stab.cuh"

namespace mcb {

namespace {

using namespace panel;

template <int N>
__device__ __forceinline__ void load_vec(double* dst, const double* __restrict__ src, int tid)
{
    for (int i = tid; i < N; i += Geo<N>::THREADS) dst[i] = src[i];
}

// exp(+lambda s_i) of one slice for the 8 rows this lane owns
template <int N>
__device__ __forceinline__ void load_ev(double (&ev)[Geo<N>::MT], const int8_t* __restrict__ row, double ep,
                                        double em, const Lane& t)
{
#pragma unroll
    for (int m = 0; m < Geo<N>::MT; ++m) ev[m] = __ldg(row + 8 * m + t.r) > 0 ? ep : em;
}

// ------------------------------------------------------------------ add_slice_sequence_*
// The slice-matrix products of add_slice_sequence_* alone: Xout = B_hi ... B_lo Usrc diag(Dsrc) (left) or
// the daggered descending sequence (right). Used when the factorisation runs as a kernel of its own
// (udt2_kernel, more chains per SM than the fused advance_kernel allows).
template <int N>
struct MultSmem {
    double S0[N * Geo<N>::LD];
    double d0[N];
};

template <int N>
__global__ void __launch_bounds__(Geo<N>::THREADS) slice_products_kernel(AdvanceArgs a, double* Xout)
{
    constexpr int MT = Geo<N>::MT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    MultSmem<N>& sm = *reinterpret_cast<MultSmem<N>*>(smem_raw);
    const Lane t;
    const int chain = blockIdx.x;
    const size_t nn = (size_t)N * N;
    const int8_t* cg = a.conf + (size_t)chain * a.sConf;
    double X[MT][2], Y[MT][2];
    g2r<N>(X, a.Usrc + chain * nn, t);
    g2s<N, false>(sm.S0, a.eT2, t.tid);
    load_vec<N>(sm.d0, a.Dsrc + (size_t)chain * N, t.tid);
    const int cnt = a.hi - a.lo + 1;
    double ev[MT];
    load_ev<N>(ev, cg + (size_t)((a.left ? a.lo : a.hi) - 1) * N, a.ep, a.em, t);
    __syncthreads();
    for (int s = 0; s < cnt; ++s) {
        double evn[MT];
        if (s + 1 < cnt) {
            const int nxt = a.left ? a.lo + s + 1 : a.hi - s - 1;
            load_ev<N>(evn, cg + (size_t)(nxt - 1) * N, a.ep, a.em, t);
        }
        if (a.left) {
#pragma unroll
            for (int m = 0; m < MT; ++m) { X[m][0] *= ev[m]; X[m][1] *= ev[m]; }
            gemm_regB<N, false>(Y, sm.S0, X, t);
#pragma unroll
            for (int m = 0; m < MT; ++m) { X[m][0] = Y[m][0]; X[m][1] = Y[m][1]; }
        } else {
            gemm_regB<N, true>(Y, sm.S0, X, t);
#pragma unroll
            for (int m = 0; m < MT; ++m) { X[m][0] = Y[m][0] * ev[m]; X[m][1] = Y[m][1] * ev[m]; }
        }
        if (s + 1 < cnt) {
#pragma unroll
            for (int m = 0; m < MT; ++m) ev[m] = evn[m];
        }
    }
    scale_cols<N>(X, sm.d0, t);
    r2g<N>(X, Xout + chain * nn, t);
}

// ------------------------------------------------------------------ udt_AVX_pivot!, calculate_greens_AVX!
// udt_AVX_pivot! as a kernel of its own: N/16 warps with 16 columns each (panel_qr2.cuh), ONE
// shared-memory matrix next to the pivoting buffers, so that several chains share an SM (N = 64:
// 45 KB and 128 threads per chain) and N = 144 fits at all (9 warps, 144 registers of matrix per thread).
// udt2_kernel arguments. Input: X, or (X == nullptr) the scaled product diag(rs) PA PB' diag(cs).
// Output: U, D, pivot and T, or (Tsrc != nullptr, apply_pivot form) T * Tsrc in T's place.
struct UdtFused {
    const double* X = nullptr;
    const double *PA = nullptr, *PB = nullptr, *rs = nullptr, *cs = nullptr;
    double *U = nullptr, *D = nullptr, *T = nullptr;
    int* piv = nullptr;
    int apply_pivot = 1;
    const double* Tsrc = nullptr;
};

template <int N>
struct UdtSmem {
    double F[N * Geo<N>::LD];
    QrSmem<N> qs;
};

template <int N, int CW>
__global__ void __launch_bounds__(Geo2<N, CW>::NT, (N <= 64 ? 4 : 1))
udt2_kernel(UdtFused a)
{
    constexpr int MT = Geo2<N, CW>::MT, LD = Geo2<N, CW>::LD, NT = Geo2<N, CW>::NT, KC = 2 * CW;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    UdtSmem<N>& sm = *reinterpret_cast<UdtSmem<N>*>(smem_raw);
    const Lane2 t(CW);
    const int tid = threadIdx.x;
    const int chain = blockIdx.x;
    const size_t nn = (size_t)N * N;
    {
        double A[MT][CW][2];
        if (a.X) {
            const double* X = a.X + chain * nn;
#pragma unroll
            for (int m = 0; m < MT; ++m)
#pragma unroll
                for (int k = 0; k < KC; ++k)
                    A[m][k >> 1][k & 1] = X[(8 * m + t.r) + (size_t)(t.col0 + 8 * (k >> 1) + (k & 1)) * N];
        } else {
            // X = diag(rs) (PA PB') diag(cs) computed here (calculate_greens, stack.jl:346-348): PB staged in
            // F, PA streamed from global memory as DMMA fragments
            const double* PB = a.PB + chain * nn;
            for (int e = tid; e < N * N; e += NT) sm.F[(e % N) + (e / N) * LD] = PB[e];
            __syncthreads();
            const double* PA = a.PA + chain * nn + t.r + (size_t)t.q * N;
            const double* bp = sm.F + (8 * CW * t.w + t.r) + t.q * LD;
#pragma unroll
            for (int m = 0; m < MT; ++m)
#pragma unroll
                for (int cw = 0; cw < CW; ++cw) A[m][cw][0] = A[m][cw][1] = 0.0;
#pragma unroll 2
            for (int k0 = 0; k0 < N; k0 += 4) {
                double av[MT], b[CW];
#pragma unroll
                for (int m = 0; m < MT; ++m) av[m] = __ldg(PA + 8 * m + (size_t)k0 * N);
#pragma unroll
                for (int cw = 0; cw < CW; ++cw) b[cw] = bp[8 * cw + k0 * LD];
#pragma unroll
                for (int m = 0; m < MT; ++m)
#pragma unroll
                    for (int cw = 0; cw < CW; ++cw) dmma884(A[m][cw][0], A[m][cw][1], av[m], b[cw]);
            }
            const double* rs = a.rs + (size_t)chain * N;
            const double* cs = a.cs + (size_t)chain * N;
#pragma unroll
            for (int k = 0; k < KC; ++k) {
                const double c = __ldg(cs + t.col0 + 8 * (k >> 1) + (k & 1));
#pragma unroll
                for (int m = 0; m < MT; ++m) A[m][k >> 1][k & 1] *= c;
            }
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const double r = __ldg(rs + 8 * m + t.r);
#pragma unroll
                for (int k = 0; k < KC; ++k) A[m][k >> 1][k & 1] *= r;
            }
        }
        if (N > 64) {
            // (the product prologue staged PB in F: everybody is done with it before the first rows arrive)
            if (!a.X) __syncthreads();
            qr_pivot2<N, CW, true>(A, sm.qs, t, sm.F);
        } else {
            qr_pivot2<N, CW>(A, sm.qs, t);
            r2s2<N, CW>(A, sm.F, t);
        }
    }
    __syncthreads();
    for (int i = tid; i < N; i += NT) {
        const double d = fabs(sm.F[i + (size_t)sm.qs.piv[i] * LD]);
        sm.qs.D[i] = d;
        sm.qs.Dinv[i] = 1.0 / d;
        a.D[(size_t)chain * N + i] = d;
        a.piv[(size_t)chain * N + i] = sm.qs.piv[i];
    }
    {
        double Q[MT][CW][2];
        form_q2<N, CW>(Q, sm.F, sm.qs, t);
        double* U = a.U + chain * nn;
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int k = 0; k < KC; ++k)
                U[(8 * m + t.r) + (size_t)(t.col0 + 8 * (k >> 1) + (k & 1)) * N] = Q[m][k >> 1][k & 1];
    }
    __syncthreads();
    if (a.Tsrc) {
        // T = D^-1 R P' materialised in F, then T_out = T * Tsrc (t_stack update, stack.jl:285, 309)
        for (int e = tid; e < N * N; e += NT) {
            const int i = e % N, j = e / N;
            const double v = sm.F[i + j * LD];
            sm.F[i + j * LD] = i <= sm.qs.pos[j] ? v * sm.qs.Dinv[i] : 0.0;
        }
        __syncthreads();
        double C[MT][CW][2];
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int cw = 0; cw < CW; ++cw) C[m][cw][0] = C[m][cw][1] = 0.0;
        const double* ap = sm.F + t.r + t.q * LD;
        const double* Bg = a.Tsrc + chain * nn + t.q + (size_t)(8 * CW * t.w + t.r) * N;
#pragma unroll 2
        for (int k0 = 0; k0 < N; k0 += 4) {
            double b[CW];
#pragma unroll
            for (int cw = 0; cw < CW; ++cw) b[cw] = __ldg(Bg + k0 + (size_t)(8 * cw) * N);
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const double av = ap[8 * m + k0 * LD];
#pragma unroll
                for (int cw = 0; cw < CW; ++cw) dmma884(C[m][cw][0], C[m][cw][1], av, b[cw]);
            }
        }
        double* To = a.T + chain * nn;
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int k = 0; k < KC; ++k)
                To[(8 * m + t.r) + (size_t)(t.col0 + 8 * (k >> 1) + (k & 1)) * N] = C[m][k >> 1][k & 1];
        return;
    }
    // T from the shared-memory copy of the factored matrix (UDT.jl:283-306)
    for (int e = tid; e < N * N; e += NT) {
        const int i = e % N, j = e / N;
        if (a.apply_pivot) { // T[:, p] = D^-1 R[:, pos(p)] masked below the position
            a.T[chain * nn + e] = i <= sm.qs.pos[j] ? sm.F[i + j * LD] * sm.qs.Dinv[i] : 0.0;
        } else {             // columns in position order
            a.T[chain * nn + e] = i <= j ? sm.F[i + sm.qs.piv[j] * LD] * sm.qs.Dinv[i] : 0.0;
        }
    }
}

// warp width of the stand-alone factorisation per matrix size: fewer, fatter warps execute fewer
// instructions per Householder step (the per-step scalar work is replicated in every warp)
template <int N> struct UdtCw { static constexpr int value = 2; };
template <> struct UdtCw<64> { static constexpr int value = 2; };

// rdivp!(A, T, O, pivot) (src/linalg/general.jl:138-166) for the same large size: out = A[:, p] triu(T)^-1.
// Solved in the transposed domain: each warp owns 8 rows of A (= 8 columns of A'), loaded already
// permuted, and runs the forward substitution over the positions j without block-level barriers;
// triu(T) is staged transposed in shared memory so that row j is contiguous; x_j goes straight to
// global memory.
template <int N>
__global__ void __launch_bounds__(Geo<N>::THREADS, 1)
rdivp_big_kernel(const double* Ag, double* Og, const double* Tg, const int* pivg)
{
    constexpr int MT = Geo<N>::MT, LD = Geo<N>::LD;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    double* Tt = reinterpret_cast<double*>(smem_raw);       // T[j][k] at Tt[k + j*LD]
    int* piv = reinterpret_cast<int*>(Tt + N * LD);
    const Lane t;
    const int chain = blockIdx.x;
    const size_t nn = (size_t)N * N;
    const double* T = Tg + chain * nn;
    for (int e = t.tid; e < N * N; e += Geo<N>::THREADS) {
        const int j = e % N, k = e / N; // T[j, k], coalesced over j
        Tt[k + j * LD] = T[e];
    }
    for (int i = t.tid; i < N; i += Geo<N>::THREADS) piv[i] = pivg[(size_t)chain * N + i];
    __syncthreads();
    // B(i, c) = A[c, p[i]] for the 2 columns c of the transposed problem this lane holds
    double B[MT][2];
    const double* A = Ag + chain * nn;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const double2 v = *reinterpret_cast<const double2*>(A + (8 * t.w + 2 * t.q) + (size_t)piv[8 * m + t.r] * N);
        B[m][0] = v.x; B[m][1] = v.y;
    }
    double* O = Og + chain * nn + (8 * t.w + 2 * t.q);
#pragma unroll
    for (int mj = 0; mj < MT; ++mj) {
#pragma unroll
        for (int rj = 0; rj < 8; ++rj) {
            const int j = 8 * mj + rj;
            const double* trow = Tt + j * LD;
            const double rinv = fast_rcp(trow[j]);
            const double y0 = __shfl_sync(FULL, B[mj][0], rj * 4 + t.q) * rinv;
            const double y1 = __shfl_sync(FULL, B[mj][1], rj * 4 + t.q) * rinv;
            if (t.r == rj) *reinterpret_cast<double2*>(O + (size_t)j * N) = make_double2(y0, y1);
            // positions j' > j: b[j'] -= T[j, j'] x_j  (the diagonal block masks r <= rj)
            {
                const double tv = t.r > rj ? trow[8 * mj + t.r] : 0.0;
                B[mj][0] -= tv * y0;
                B[mj][1] -= tv * y1;
            }
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double tv = trow[8 * m + t.r];
                B[m][0] -= tv * y0;
                B[m][1] -= tv * y1;
            }
        }
    }
}

// The two halves of calculate_greens_AVX! that sit between / after the factorisations when those run
// as kernels of their own (udt2_kernel): products and triangular solves of one chain in one CTA, the
// intermediates in registers / shared memory. T comes "dirty" (position order, apply_pivot = 0).
template <int N>
struct GlueSmem {
    static constexpr int LD = Geo<N>::LD;
    double S0[N * LD];
    double S1[N * LD];
    double Tt[N * LD]; // T[j][k] at Tt[k + j*LD]
    double d[N];
    int piv[N];
};

// x = b[:, p] triu(T)^-1 in the transposed domain (see rdivp_big_kernel); column j of the result, times
// colscale[j] if given, goes to S[.. + j*LD] and, if Og != nullptr, unscaled to global memory.
template <int N, bool SCALE>
__device__ __forceinline__ void solve_perm(const double* __restrict__ A, const GlueSmem<N>& sm, double* S,
                                           double* Og, const Lane& t)
{
    constexpr int MT = Geo<N>::MT, LD = Geo<N>::LD;
    double B[MT][2];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const double2 v = *reinterpret_cast<const double2*>(A + (8 * t.w + 2 * t.q) + (size_t)sm.piv[8 * m + t.r] * N);
        B[m][0] = v.x; B[m][1] = v.y;
    }
    const int c0 = 8 * t.w + 2 * t.q;
#pragma unroll
    for (int mj = 0; mj < MT; ++mj) {
#pragma unroll
        for (int rj = 0; rj < 8; ++rj) {
            const int j = 8 * mj + rj;
            const double* trow = sm.Tt + j * LD;
            const double rinv = fast_rcp(trow[j]);
            const double y0 = __shfl_sync(FULL, B[mj][0], rj * 4 + t.q) * rinv;
            const double y1 = __shfl_sync(FULL, B[mj][1], rj * 4 + t.q) * rinv;
            if (t.r == rj) {
                if (Og) *reinterpret_cast<double2*>(Og + c0 + (size_t)j * N) = make_double2(y0, y1);
                const double sc = SCALE ? sm.d[j] : 1.0;
                *reinterpret_cast<double2*>(S + c0 + j * LD) = make_double2(y0 * sc, y1 * sc);
            }
            {
                const double tv = t.r > rj ? trow[8 * mj + t.r] : 0.0;
                B[mj][0] -= tv * y0;
                B[mj][1] -= tv * y1;
            }
#pragma unroll
            for (int m = mj + 1; m < MT; ++m) {
                const double tv = trow[8 * m + t.r];
                B[m][0] -= tv * y0;
                B[m][1] -= tv * y1;
            }
        }
    }
}

template <int N>
__device__ __forceinline__ void load_tt(GlueSmem<N>& sm, const double* __restrict__ T, const int* __restrict__ piv,
                                        const Lane& t)
{
    constexpr int LD = Geo<N>::LD;
    for (int e = t.tid; e < N * N; e += Geo<N>::THREADS) {
        const int j = e % N, k = e / N;
        sm.Tt[k + j * LD] = T[e];
    }
    for (int i = t.tid; i < N; i += Geo<N>::THREADS) sm.piv[i] = piv[i];
}

// W3 = Ul Q1 ; W4 = Ur[:, p1] triu(T1)^-1 ; A2 = W3' W4 + diag(D1)            (stack.jl:360-368)
template <int N>
__global__ void __launch_bounds__(Geo<N>::THREADS, (N <= 64 ? 2 : 1))
greens_mid_kernel(const double* Ulg, const double* Q1g, const double* T1g, const int* piv1g, const double* D1g,
                  const double* Urg, double* W3g, double* W4g, double* A2g)
{
    constexpr int MT = Geo<N>::MT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    GlueSmem<N>& sm = *reinterpret_cast<GlueSmem<N>*>(smem_raw);
    const Lane t;
    const int chain = blockIdx.x;
    const size_t nn = (size_t)N * N;
    double Q[MT][2], W[MT][2];
    g2s<N, false>(sm.S0, Ulg + chain * nn, t.tid);
    load_tt<N>(sm, T1g + chain * nn, piv1g + (size_t)chain * N, t);
    load_vec<N>(sm.d, D1g + (size_t)chain * N, t.tid);
    g2r<N>(Q, Q1g + chain * nn, t);
    __syncthreads();
    gemm_regB<N, false>(W, sm.S0, Q, t);
    r2g<N>(W, W3g + chain * nn, t);
    solve_perm<N, false>(Urg + chain * nn, sm, sm.S1, W4g + chain * nn, t);
    __syncthreads();
    r2s<N, false>(W, sm.S0, t);
    __syncthreads();
    gemm<N, true, false>(Q, sm.S0, sm.S1, t);
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int c = 0; c < 2; ++c)
            if (8 * m + t.r == 8 * t.w + 2 * t.q + c) Q[m][c] += sm.d[8 * m + t.r];
    r2g<N>(Q, A2g + chain * nn, t);
}

// W0 = W3 Q2 ; W5 = W4[:, p2] triu(T2)^-1 ; G = (W5 diag(1/D2)) W0'            (stack.jl:377-391)
template <int N>
__global__ void __launch_bounds__(Geo<N>::THREADS, (N <= 64 ? 2 : 1))
greens_end_kernel(const double* W3g, const double* Q2g, const double* T2g, const int* piv2g, const double* D2g,
                  const double* W4g, double* Gg)
{
    constexpr int MT = Geo<N>::MT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    GlueSmem<N>& sm = *reinterpret_cast<GlueSmem<N>*>(smem_raw);
    const Lane t;
    const int chain = blockIdx.x;
    const size_t nn = (size_t)N * N;
    double Q[MT][2], W[MT][2];
    g2s<N, false>(sm.S0, W3g + chain * nn, t.tid);
    load_tt<N>(sm, T2g + chain * nn, piv2g + (size_t)chain * N, t);
    for (int i = t.tid; i < N; i += Geo<N>::THREADS) sm.d[i] = 1.0 / D2g[(size_t)chain * N + i];
    g2r<N>(Q, Q2g + chain * nn, t);
    __syncthreads();
    gemm_regB<N, false>(W, sm.S0, Q, t);
    solve_perm<N, true>(W4g + chain * nn, sm, sm.S1, nullptr, t);
    __syncthreads();
    r2s<N, false>(W, sm.S0, t);
    __syncthreads();
    gemm<N, false, true>(Q, sm.S1, sm.S0, t);
    r2g<N>(Q, Gg + chain * nn, t);
}

// rdivp! for matrices too large for one CTA: the rows of A are independent, so a CTA takes 32 of them
// (4 warps x 8 rows, held in registers in the transposed layout) and streams triu(T) through shared
// memory 8 rows at a time (double buffered). grid = (ceil(N / 32), chains).
template <int N>
struct RowsGeo {
    static constexpr int MT = N / 8, LDT = N + 4, WARPS = 4, THREADS = 128;
    static constexpr size_t SMEM = 2 * 8 * LDT * sizeof(double) + N * sizeof(int);
};

// positions 8 MJ .. 8 MJ + 7 (compile-time recursion keeps every register index static)
template <int N, int MJ>
__device__ __forceinline__ void rows_block(double (&B)[N / 8][2], double (*Tt)[8][N + 4], const double* T,
                                           double* O, int tid, int r, int q, bool live)
{
    constexpr int MT = N / 8;
    if constexpr (MJ + 1 < MT) {
        // prefetch rows 8 (MJ+1) .. of triu(T): 8 consecutive doubles per column k >= 8 (MJ+1)
        for (int e = tid; e < 8 * (N - 8 * (MJ + 1)); e += 128) {
            const int rr = e & 7, k = 8 * (MJ + 1) + (e >> 3);
            Tt[(MJ + 1) & 1][rr][k] = T[(8 * (MJ + 1) + rr) + (size_t)k * N];
        }
    }
#pragma unroll
    for (int rj = 0; rj < 8; ++rj) {
        const int j = 8 * MJ + rj;
        const double* trow = Tt[MJ & 1][rj];
        const double rinv = fast_rcp(trow[j]);
        const double y0 = __shfl_sync(FULL, B[MJ][0], rj * 4 + q) * rinv;
        const double y1 = __shfl_sync(FULL, B[MJ][1], rj * 4 + q) * rinv;
        if (r == rj && live) *reinterpret_cast<double2*>(O + (size_t)j * N) = make_double2(y0, y1);
        {
            const double tv = r > rj ? trow[8 * MJ + r] : 0.0;
            B[MJ][0] -= tv * y0;
            B[MJ][1] -= tv * y1;
        }
#pragma unroll
        for (int m = MJ + 1; m < MT; ++m) {
            const double tv = trow[8 * m + r];
            B[m][0] -= tv * y0;
            B[m][1] -= tv * y1;
        }
    }
    __syncthreads();
    if constexpr (MJ + 1 < MT) rows_block<N, MJ + 1>(B, Tt, T, O, tid, r, q, live);
}

template <int N>
__global__ void __launch_bounds__(128, 1)
rdivp_rows_kernel(const double* Ag, double* Og, const double* Tg, const int* pivg)
{
    constexpr int MT = N / 8, LDT = N + 4, WARPS = 4;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    double (*Tt)[8][LDT] = reinterpret_cast<double (*)[8][LDT]>(smem_raw); // Tt[buffer][row in block][k]
    int* piv = reinterpret_cast<int*>(smem_raw + 2 * 8 * LDT * sizeof(double));
    const int tid = threadIdx.x, wl = tid >> 5, l = tid & 31, r = l >> 2, q = l & 3;
    const int chain = blockIdx.y;
    const int row0 = (blockIdx.x * WARPS + wl) * 8 + 2 * q; // first of this lane's two rows of A
    const bool live = row0 < N;                               // (N % 8 == 0: both rows in range)
    const size_t nn = (size_t)N * N;
    const double* T = Tg + chain * nn;
    for (int i = tid; i < N; i += 128) piv[i] = pivg[(size_t)chain * N + i];
    for (int e = tid; e < 8 * N; e += 128) Tt[0][e & 7][e >> 3] = T[(e & 7) + (size_t)(e >> 3) * N];
    __syncthreads();
    double B[MT][2];
    const double* A = Ag + chain * nn;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        double2 v = make_double2(0.0, 0.0);
        if (live) v = *reinterpret_cast<const double2*>(A + row0 + (size_t)piv[8 * m + r] * N);
        B[m][0] = v.x; B[m][1] = v.y;
    }
    rows_block<N, 0>(B, Tt, T, Og + chain * nn + row0, tid, r, q, live);
}

template <int N>
int set_attrs()
{
    MCB_CUDA(cudaFuncSetAttribute(slice_products_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MultSmem<N>)));
    MCB_CUDA(cudaFuncSetAttribute(rdivp_big_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(N * Geo<N>::LD * sizeof(double) + N * sizeof(int))));
    MCB_CUDA(cudaFuncSetAttribute(udt2_kernel<N, UdtCw<N>::value>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UdtSmem<N>)));
    MCB_CUDA(cudaFuncSetAttribute(greens_mid_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GlueSmem<N>)));
    MCB_CUDA(cudaFuncSetAttribute(greens_end_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GlueSmem<N>)));
    return 0;
}

} // namespace

bool stab_supported(int n) { return n == 16 || n == 32 || n == 64; }
bool udt_panel_supported(int n) { return stab_supported(n) || n == 144; }
bool rdivp_panel_supported(int n) { return stab_supported(n) || n == 144 || n == 400; }

int init_stab_kernels()
{
    if (set_attrs<16>() || set_attrs<32>() || set_attrs<64>()) return -2;
    MCB_CUDA(cudaFuncSetAttribute(udt2_kernel<144, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(UdtSmem<144>)));
    MCB_CUDA(cudaFuncSetAttribute(rdivp_rows_kernel<400>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * 8 * 404 * sizeof(double) + 400 * sizeof(int))));
    MCB_CUDA(cudaFuncSetAttribute(rdivp_big_kernel<144>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(144 * Geo<144>::LD * sizeof(double) + 144 * sizeof(int))));
    return 0;
}

#define STAB_DISPATCH(n, CALL)                                                         \
    switch (n) {                                                                       \
    case 16: { constexpr int N = 16; CALL; break; }                                    \
    case 32: { constexpr int N = 32; CALL; break; }                                    \
    case 64: { constexpr int N = 64; CALL; break; }                                    \
    default: set_error("chain-resident kernels: unsupported matrix size"); return -1;  \
    }

int launch_slice_products(const AdvanceArgs& a, double* Xout, cudaStream_t st)
{
    if (a.batch <= 0) return 0;
    const double n3 = (double)a.n * a.n * a.n;
    LaunchScope scope(LC_ADVANCE, st, 2.0 * (a.hi - a.lo + 1) * n3 * a.batch);
    STAB_DISPATCH(a.n, MCB_CUDA(launch_pdl(slice_products_kernel<N>, dim3(a.batch), dim3(Geo<N>::THREADS), sizeof(MultSmem<N>), st, a, Xout)));
    MCB_CUDA(cudaGetLastError());
    return 0;
}

int launch_greens_mid(int n, int batch, const double* Ul, const double* Q1, const double* T1, const int* piv1,
                      const double* D1, const double* Ur, double* W3, double* W4, double* A2, cudaStream_t st)
{
    if (batch <= 0) return 0;
    LaunchScope scope(LC_GREENS, st, 5.0 * n * (double)n * n * batch);
    STAB_DISPATCH(n, MCB_CUDA(launch_pdl(greens_mid_kernel<N>, dim3(batch), dim3(Geo<N>::THREADS), sizeof(GlueSmem<N>), st, Ul, Q1, T1, piv1, D1, Ur, W3, W4, A2)));
    MCB_CUDA(cudaGetLastError());
    return 0;
}

int launch_greens_end(int n, int batch, const double* W3, const double* Q2, const double* T2, const int* piv2,
                      const double* D2, const double* W4, double* G, cudaStream_t st)
{
    if (batch <= 0) return 0;
    LaunchScope scope(LC_GREENS, st, 5.0 * n * (double)n * n * batch);
    STAB_DISPATCH(n, MCB_CUDA(launch_pdl(greens_end_kernel<N>, dim3(batch), dim3(Geo<N>::THREADS), sizeof(GlueSmem<N>), st, W3, Q2, T2, piv2, D2, W4, G)));
    MCB_CUDA(cudaGetLastError());
    return 0;
}

static int launch_udt2(int n, int batch, const UdtFused& a, double flops, cudaStream_t st)
{
    if (batch <= 0) return 0;
    LaunchScope scope(LC_UDT, st, flops * batch);
    if (n == 144) {
        MCB_CUDA(launch_pdl(udt2_kernel<144, 2>, dim3(batch), dim3(Geo2<144>::NT), sizeof(UdtSmem<144>), st, a));
        return 0;
    }
    STAB_DISPATCH(n, MCB_CUDA(launch_pdl(udt2_kernel<N, UdtCw<N>::value>, dim3(batch), dim3(Geo2<N, UdtCw<N>::value>::NT), sizeof(UdtSmem<N>), st, a)));
    MCB_CUDA(cudaGetLastError());
    return 0;
}

int launch_udt_panel(int n, int batch, const double* X, double* U, double* D, double* T, int* pivot,
                     int apply_pivot, cudaStream_t st)
{
    UdtFused a;
    a.X = X; a.U = U; a.D = D; a.T = T; a.piv = pivot; a.apply_pivot = apply_pivot;
    return launch_udt2(n, batch, a, 8.0 / 3.0 * n * (double)n * n, st);
}

int launch_udt_of_product(int n, int batch, const double* PA, const double* PB, const double* rs, const double* cs,
                          double* U, double* D, double* T, int* pivot, cudaStream_t st)
{
    UdtFused a;
    a.PA = PA; a.PB = PB; a.rs = rs; a.cs = cs; a.U = U; a.D = D; a.T = T; a.piv = pivot; a.apply_pivot = 0;
    return launch_udt2(n, batch, a, (2.0 + 8.0 / 3.0) * n * (double)n * n, st);
}

int launch_udt_times(int n, int batch, const double* X, double* U, double* D, const double* Tsrc, double* Tout,
                     int* pivot, cudaStream_t st)
{
    UdtFused a;
    a.X = X; a.U = U; a.D = D; a.T = Tout; a.piv = pivot; a.apply_pivot = 1; a.Tsrc = Tsrc;
    return launch_udt2(n, batch, a, (2.0 + 8.0 / 3.0) * n * (double)n * n, st);
}

// out = in[:, pivot] * triu(T)^-1 for the sizes with a register-resident solver (udt_panel_supported && n > 64)
int launch_rdivp_panel(int n, int batch, const double* in, double* out, const double* T, const int* pivot,
                       cudaStream_t st)
{
    if (batch <= 0) return 0;
    LaunchScope scope(LC_RDIVP, st, 1.0 * n * (double)n * n * batch);
    if (n == 400) {
        rdivp_rows_kernel<400><<<dim3((400 + 31) / 32, batch), 128, 2 * 8 * 404 * sizeof(double) + 400 * sizeof(int), st>>>(in, out, T, pivot);
        MCB_CUDA(cudaGetLastError());
        return 0;
    }
#define RDIVP_CASE(NN)                                                                                  \
    case NN:                                                                                            \
        MCB_CUDA(launch_pdl(rdivp_big_kernel<NN>, dim3(batch), dim3(Geo<NN>::THREADS),                  \
                            NN * Geo<NN>::LD * sizeof(double) + NN * sizeof(int), st, in, out, T, pivot));   \
        break;
    switch (n) {
        RDIVP_CASE(16) RDIVP_CASE(32) RDIVP_CASE(64) RDIVP_CASE(144)
    default: set_error("rdivp panel kernel: unsupported matrix size"); return -1;
    }
#undef RDIVP_CASE
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mcb
