// ORACLE (test infrastructure only — never linked into the product library).
//
// CPU restatement of MonteCarlo.jl's hand-written real-Float64 linear algebra.
// All matrices are square n×n, column-major (Julia layout): A[i,j] = A[i + j*n].
// Each routine cites the reference function it restates (paths relative to
// /root/reference). The reference vectorises with LoopVectorization's @avx,
// which re-associates sums, so bit-equality with Julia is not defined; parity
// is 1e-10 relative max-norm.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#define IDX(i, j, n) ((size_t)(i) + (size_t)(j) * (size_t)(n))

// C = A * B                                   src/linalg/general.jl:7-15
static inline void ref_vmul_nn(double* C, const double* A, const double* B, int n)
{
    for (int j = 0; j < n; ++j) {
        double* c = C + IDX(0, j, n);
        for (int i = 0; i < n; ++i) c[i] = 0.0;
        for (int k = 0; k < n; ++k) {
            const double b = B[IDX(k, j, n)];
            const double* a = A + IDX(0, k, n);
#pragma omp simd
            for (int i = 0; i < n; ++i) c[i] += a[i] * b;
        }
    }
}
// C = A * diag(d)                             src/linalg/general.jl:16-20
static inline void ref_vmul_nd(double* C, const double* A, const double* d, int n)
{
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i) C[IDX(i, j, n)] = A[IDX(i, j, n)] * d[j];
}
// C = diag(d) * B                             src/linalg/general.jl:21-25
static inline void ref_vmul_dn(double* C, const double* d, const double* B, int n)
{
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i) C[IDX(i, j, n)] = d[i] * B[IDX(i, j, n)];
}
// C = A * B'                                  src/linalg/general.jl:26-35
static inline void ref_vmul_nt(double* C, const double* A, const double* B, int n)
{
    for (int j = 0; j < n; ++j) {
        double* c = C + IDX(0, j, n);
        for (int i = 0; i < n; ++i) c[i] = 0.0;
        for (int k = 0; k < n; ++k) {
            const double b = B[IDX(j, k, n)];
            const double* a = A + IDX(0, k, n);
#pragma omp simd
            for (int i = 0; i < n; ++i) c[i] += a[i] * b;
        }
    }
}
// C = A' * B                                  src/linalg/general.jl:36-45
static inline void ref_vmul_tn(double* C, const double* A, const double* B, int n)
{
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i) {
            const double* a = A + IDX(0, i, n);
            const double* b = B + IDX(0, j, n);
            double acc = 0.0;
#pragma omp simd reduction(+ : acc)
            for (int k = 0; k < n; ++k) acc += a[k] * b[k];
            C[IDX(i, j, n)] = acc;
        }
}
// A[i,i] += d[i]                              src/linalg/general.jl:67-71
static inline void ref_rvadd_diag(double* A, const double* d, int n)
{
    for (int i = 0; i < n; ++i) A[IDX(i, i, n)] += d[i];
}
static inline void ref_eye(double* A, int n)
{
    memset(A, 0, sizeof(double) * (size_t)n * n);
    for (int i = 0; i < n; ++i) A[IDX(i, i, n)] = 1.0;
}
static inline void ref_copy(double* dst, const double* src, int n)
{
    memcpy(dst, src, sizeof(double) * (size_t)n * n);
}

// A <- A[:,pivot] * triu(T)^-1                src/linalg/general.jl:138-166
// O is n×n scratch; T's strict lower triangle is ignored ("dirty" T).
static inline void ref_rdivp(double* A, const double* T, double* O, const int64_t* pivot, int n)
{
    for (int j = 0; j < n; ++j) {
        const double* src = A + IDX(0, pivot[j], n);
        double* dst = O + IDX(0, j, n);
        for (int i = 0; i < n; ++i) dst[i] = src[i];
    }
    for (int j = 0; j < n; ++j) {
        double* aj = A + IDX(0, j, n);
        const double* oj = O + IDX(0, j, n);
        for (int i = 0; i < n; ++i) aj[i] = oj[i];
        for (int k = 0; k < j; ++k) {
            const double t = T[IDX(k, j, n)];
            const double* ak = A + IDX(0, k, n);
#pragma omp simd
            for (int i = 0; i < n; ++i) aj[i] -= ak[i] * t;
        }
        const double tjj = T[IDX(j, j, n)];
        for (int i = 0; i < n; ++i) aj[i] /= tjj;
    }
}

// First strict maximum of trailing column norms, recomputed from scratch.
//                                             src/linalg/UDT.jl:151-168
static inline int ref_indmaxcolumn(const double* A, int j, int n, double* maxval)
{
    double mx = 0.0;
    for (int k = j; k < n; ++k) mx += A[IDX(k, j, n)] * A[IDX(k, j, n)];
    int ii = j;
    for (int i = j + 1; i < n; ++i) {
        double mi = 0.0;
        const double* a = A + IDX(0, i, n);
#pragma omp simd reduction(+ : mi)
        for (int k = j; k < n; ++k) mi += a[k] * a[k];
        if (fabs(mi) > mx) { mx = mi; ii = i; }
    }
    *maxval = mx;
    return ii;
}

// Householder reflector for column j using the squared norm `normu`.
//                                             src/linalg/UDT.jl:133-148
static inline double ref_reflector(double* A, double normu, int j, int n)
{
    double xi1 = A[IDX(j, j, n)];
    if (normu == 0.0) return 0.0;
    normu = sqrt(normu);
    const double nu = copysign(normu, xi1);
    xi1 += nu;
    A[IDX(j, j, n)] = -nu;
    for (int i = j + 1; i < n; ++i) A[IDX(i, j, n)] /= xi1;
    return xi1 / nu;
}

// Column-pivoted Householder QR -> U (explicit), D = |diag R|, T = D^-1 R P'
// (apply_pivot != 0) or "dirty" D^-1 R left in the upper triangle with the
// reflector vectors below (apply_pivot == 0).
//                                             src/linalg/UDT.jl:192-306, :32-50
static inline void ref_udt_pivot(double* U, double* D, double* A, int64_t* pivot, double* temp,
                                 int n, int apply_pivot)
{
    for (int i = 0; i < n; ++i) pivot[i] = i;
    for (int j = 0; j < n; ++j) {
        double maxval;
        const int jm = ref_indmaxcolumn(A, j, n, &maxval);
        if (jm != j) {
            int64_t tp = pivot[jm]; pivot[jm] = pivot[j]; pivot[j] = tp;
            for (int i = 0; i < n; ++i) {
                double t = A[IDX(i, jm, n)];
                A[IDX(i, jm, n)] = A[IDX(i, j, n)];
                A[IDX(i, j, n)] = t;
            }
        }
        const double tau = ref_reflector(A, maxval, j, n);
        temp[j] = tau;
        // reflectorApply! on the trailing block (rows j.., cols j+1..)
        const double* x = A + IDX(0, j, n);
        for (int c = j + 1; c < n; ++c) {
            double* a = A + IDX(0, c, n);
            double v = a[j];
#pragma omp simd reduction(+ : v)
            for (int i = j + 1; i < n; ++i) v += x[i] * a[i];
            v = tau * v;
            a[j] -= v;
#pragma omp simd
            for (int i = j + 1; i < n; ++i) a[i] -= x[i] * v;
        }
    }
    // explicit Q by backward accumulation        UDT.jl:250-266
    ref_eye(U, n);
    U[IDX(n - 1, n - 1, n)] -= temp[n - 1];
    for (int k = n - 2; k >= 0; --k) {
        const double* x = A + IDX(0, k, n);
        for (int j = k; j < n; ++j) {
            double* u = U + IDX(0, j, n);
            double v = u[k];
#pragma omp simd reduction(+ : v)
            for (int i = k + 1; i < n; ++i) v += x[i] * u[i];
            v = temp[k] * v;
            u[k] -= v;
#pragma omp simd
            for (int i = k + 1; i < n; ++i) u[i] -= x[i] * v;
        }
    }
    for (int i = 0; i < n; ++i) D[i] = fabs(A[IDX(i, i, n)]);
    if (apply_pivot) {                          // UDT.jl:283-297
        for (int i = 0; i < n; ++i) {
            const double d = 1.0 / D[i];
            for (int j = 0; j < i; ++j) temp[pivot[j]] = 0.0;
            for (int j = i; j < n; ++j) temp[pivot[j]] = d * A[IDX(i, j, n)];
            for (int j = 0; j < n; ++j) A[IDX(i, j, n)] = temp[j];
        }
    } else {                                    // UDT.jl:298-306
        for (int i = 0; i < n; ++i) {
            const double d = 1.0 / D[i];
            for (int j = i; j < n; ++j) A[IDX(i, j, n)] *= d;
        }
    }
}

// G = [1 + Ul Dl Tl Tr' Dr Ur']^-1 ; all inputs clobbered.
//                                             src/flavors/DQMC/stack.jl:337-393
static inline void ref_calculate_greens(double* Ul, double* Dl, double* Tl, double* Ur, double* Dr,
                                        double* Tr, double* G, int64_t* pivot, double* temp, int n)
{
    ref_vmul_nt(G, Tl, Tr, n);
    ref_vmul_nd(Tr, G, Dr, n);
    ref_vmul_dn(G, Dl, Tr, n);
    ref_udt_pivot(Tr, Dr, G, pivot, temp, n, 0);

    ref_vmul_nn(Tl, Ul, Tr, n);
    ref_rdivp(Ur, G, Ul, pivot, n);
    ref_vmul_tn(Tr, Tl, Ur, n);

    ref_rvadd_diag(Tr, Dr, n);

    ref_udt_pivot(Ul, Dr, Tr, pivot, temp, n, 0);
    ref_rdivp(Ur, Tr, G, pivot, n);
    ref_vmul_nn(Tr, Tl, Ul, n);

    for (int i = 0; i < n; ++i) Dl[i] = 1.0 / Dr[i];

    ref_vmul_nd(Ul, Ur, Dl, n);
    ref_vmul_nt(G, Ul, Tr, n);
}
