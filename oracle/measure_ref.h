// ORACLE (test infrastructure, never linked into libmcb200.so): the reference's DQMC measurement kernels
// and their lattice-iterator drivers, restated loop for loop. G is column-major, G[i + j*ld] = G[i, j],
// all indices 0-based here (the reference's are 1-based).
//
//   greens_kernel, occupation_kernel, cdc_kernel, m{x,y,z}_kernel, sdc_{x,y,z}_kernel, pc_kernel (general,
//   two flavors in one 2N x 2N matrix)                 src/flavors/DQMC/measurements/measurements.jl:31-207
//   their HubbardModelAttractive overloads (one flavor) src/models/HubbardModel/HubbardModelAttractive.jl:221-245
//   apply!(EachSiteAndFlavor / EachSitePairByDistance / EachLocalQuadByDistance), prepare!, finish! (the 1/N)
//                                                       src/flavors/DQMC/measurements/generic.jl:279-347
#pragma once
#include <cstddef>
#include <cstdint>

namespace measure_ref {

inline double I(int j, int i) { return j == i ? 1.0 : 0.0; } // LinearAlgebra.I[j, i]

// ---- one-flavor kernels (HubbardModelAttractive.jl:221-245); G is N x N
inline double cdc_attractive(const double* G, int N, int i, int j)
{
    // 4 * (1 - G[i, i]) * (1 - G[j, j]) + 2 * (I[j, i] - G[j, i]) * G[i, j]
    return 4 * (1 - G[i + (size_t)i * N]) * (1 - G[j + (size_t)j * N]) +
           2 * (I(j, i) - G[j + (size_t)i * N]) * G[i + (size_t)j * N];
}
inline double sdc_attractive(const double* G, int N, int i, int j)
{
    // 2(I[j,i] - G[j,i]) * G[i,j]   (x, y and z are the same expression)
    return 2 * (I(j, i) - G[j + (size_t)i * N]) * G[i + (size_t)j * N];
}
inline double pc_attractive(const double* G, int N, int src1, int trg1, int src2, int trg2)
{
    // G[src1, src2] * G[trg1, trg2]
    return G[src1 + (size_t)src2 * N] * G[trg1 + (size_t)trg2 * N];
}

// ---- two-flavor kernels (measurements.jl:58-207); G is 2N x 2N, flavor-major (up block first)
struct G2 {
    const double* g;
    int N;
    double operator()(int i, int j) const { return g[i + (size_t)j * 2 * N]; }
};
inline double cdc_general(G2 G, int i, int j)
{
    const int N = G.N;
    return (1 - G(i, i)) * (1 - G(j, j)) + (I(j, i) - G(j, i)) * G(i, j) +          // <n_up n_up>
           (1 - G(i, i)) * (1 - G(j + N, j + N)) - G(j + N, i) * G(i, j + N) +        // <n_up n_dn>
           (1 - G(i + N, i + N)) * (1 - G(j, j)) - G(j, i + N) * G(i + N, j) +        // <n_dn n_up>
           (1 - G(i + N, i + N)) * (1 - G(j + N, j + N)) + (I(j, i) - G(j + N, i + N)) * G(i + N, j + N);
}
inline double mx_general(G2 G, int i) { return -G(i + G.N, i) - G(i, i + G.N); }
inline double my_general(G2 G, int i) { return G(i + G.N, i) - G(i, i + G.N); }
inline double mz_general(G2 G, int i) { return G(i + G.N, i + G.N) - G(i, i); }
inline double sdc_x_general(G2 G, int i, int j)
{
    const int N = G.N;
    return G(i + N, i) * G(j + N, j) - G(j + N, i) * G(i + N, j) +
           G(i + N, i) * G(j, j + N) + (I(j, i) - G(j, i)) * G(i + N, j + N) +
           G(i, i + N) * G(j + N, j) + (I(j, i) - G(j + N, i + N)) * G(i, j) +
           G(i, i + N) * G(j, j + N) - G(j, i + N) * G(i, j + N);
}
inline double sdc_y_general(G2 G, int i, int j)
{
    const int N = G.N;
    return -G(i + N, i) * G(j + N, j) + G(j + N, i) * G(i + N, j) +
           G(i + N, i) * G(j, j + N) + (I(j, i) - G(j, i)) * G(i + N, j + N) +
           G(i, i + N) * G(j + N, j) + (I(j, i) - G(j + N, i + N)) * G(i, j) -
           G(i, i + N) * G(j, j + N) + G(j, i + N) * G(i, j + N);
}
inline double sdc_z_general(G2 G, int i, int j)
{
    const int N = G.N;
    return (1 - G(i, i)) * (1 - G(j, j)) + (I(j, i) - G(j, i)) * G(i, j) -
           (1 - G(i, i)) * (1 - G(j + N, j + N)) + G(j + N, i) * G(i, j + N) -
           (1 - G(i + N, i + N)) * (1 - G(j, j)) + G(j, i + N) * G(i + N, j) +
           (1 - G(i + N, i + N)) * (1 - G(j + N, j + N)) + (I(j, i) - G(j + N, i + N)) * G(i + N, j + N);
}
inline double pc_general(G2 G, int src1, int trg1, int src2, int trg2)
{
    const int N = G.N;
    return G(src1, src2) * G(trg1 + N, trg2 + N) - G(src1, trg2 + N) * G(trg1 + N, src2);
}

} // namespace measure_ref
