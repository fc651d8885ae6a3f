// Chain-resident stabilisation kernels (one CTA per Markov chain, matrices in registers / shared
// memory, see panel.cuh and panel_qr2.cuh). They replace, for the matrix sizes the register layout
// supports, the general-path sequences of dqmc_engine.cu:
//   slice products + udt(_times)        = add_slice_sequence_left / _right   src/flavors/DQMC/stack.jl:272-311
//   udt_of_product, greens_mid, udt, greens_end = calculate_greens_AVX!      src/flavors/DQMC/stack.jl:337-393
//   udt                                 = udt_AVX_pivot!                      src/linalg/UDT.jl:192-306
//   rdivp                               = rdivp!                              src/linalg/general.jl:138-166
#pragma once
#include "common.cuh"

namespace mcb {

struct AdvanceArgs {
    int n = 0, batch = 0;
    const double* Usrc = nullptr; const double* Dsrc = nullptr; const double* Tsrc = nullptr;
    double* Udst = nullptr; double* Ddst = nullptr; double* Tdst = nullptr; // strides n*n / n per chain
    const double* eT2 = nullptr;       // e^{-dtau T}, shared by all chains
    const int8_t* conf = nullptr; long long sConf = 0; // [chain][slice][site]
    int lo = 0, hi = 0;                // 1-based slice range of the sequence
    int left = 1;                      // 1: B_hi ... B_lo (ascending) ; 0: B'_lo ... B'_hi (descending, daggered)
    double ep = 1.0, em = 1.0;         // exp(+lambda s) for s = +1 / -1
};

bool stab_supported(int n);
bool udt_panel_supported(int n);   // launch_udt_panel exists for this size
bool rdivp_panel_supported(int n); // launch_rdivp_panel exists for this size
// only the products of add_slice_sequence_*: Xout = (B chain) * Usrc * diag(Dsrc)
int launch_slice_products(const AdvanceArgs& a, double* Xout, cudaStream_t st);
// X -> U, D, T (= D^-1 R P' when apply_pivot, else the reference's "dirty" upper triangle in
// position order) and the 0-based pivot vector; same contract as launch_udt (udt.cuh)
int launch_udt_panel(int n, int batch, const double* X, double* U, double* D, double* T, int* pivot,
                     int apply_pivot, cudaStream_t st);
// the same factorisation with the product that feeds it / consumes it done by the same CTA (stab_supported sizes):
//   _of_product: X = diag(rs) (PA PB') diag(cs), T "dirty"            calculate_greens, stack.jl:346-349
//   _times:      T_out = (D^-1 R P') * Tsrc instead of T              add_slice_sequence_*, stack.jl:285, 309
int launch_udt_of_product(int n, int batch, const double* PA, const double* PB, const double* rs, const double* cs,
                          double* U, double* D, double* T, int* pivot, cudaStream_t st);
int launch_udt_times(int n, int batch, const double* X, double* U, double* D, const double* Tsrc, double* Tout,
                     int* pivot, cudaStream_t st);
int launch_rdivp_panel(int n, int batch, const double* in, double* out, const double* T, const int* pivot,
                       cudaStream_t st);
// calculate_greens between / after its two factorisations (T1, T2 "dirty", i.e. apply_pivot = 0):
//   mid: W3 = Ul Q1, W4 = Ur[:, p1] triu(T1)^-1, A2 = W3' W4 + diag(D1)     stack.jl:360-368
//   end: G = (W4[:, p2] triu(T2)^-1 diag(1/D2)) (W3 Q2)'                     stack.jl:377-391
int launch_greens_mid(int n, int batch, const double* Ul, const double* Q1, const double* T1, const int* piv1,
                      const double* D1, const double* Ur, double* W3, double* W4, double* A2, cudaStream_t st);
int launch_greens_end(int n, int batch, const double* W3, const double* Q2, const double* T2, const int* piv2,
                      const double* D2, const double* W4, double* G, cudaStream_t st);
int init_stab_kernels();

} // namespace mcb
