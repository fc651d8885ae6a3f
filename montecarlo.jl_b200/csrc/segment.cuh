// Chain-resident slice-visit kernel: for a run of consecutive time slices, wrap_greens! followed by
// sweep_spatial, with the Green's function living in registers for the whole run (panel.cuh).
#pragma once
#include "common.cuh"

namespace mcb {

struct SegmentArgs {
    int n = 0, batch = 0;
    double* G = nullptr; long long sG = 0;
    int8_t* conf = nullptr; long long sConf = 0; // [chain][slice][site], per chain stride
    const double* eT2 = nullptr;    // e^{-dtau T}
    const double* eTinv2 = nullptr; // e^{+dtau T}
    const double* eT2T = nullptr;   // their transposes (segment_resident.cu streams rows of the exponentials as columns)
    const double* eTinv2T = nullptr;
    int n_slices = 0;
    int start_slice = 0; // 1-based slice of the first sweep_spatial
    int direction = 1;
    int n_visits = 0;
    int first_wrap = 0;  // wrap onto start_slice before its sweep_spatial
    double eVp[2] = {1, 1}, eVm[2] = {1, 1}; // exp(+lambda s), exp(-lambda s) for s = +1, -1
    double gamma_p = 0, gamma_m = 0, efac_p = 0, efac_m = 0;
    uint32_t key0 = 0, chain_offset = 0;
    long long n_updates0 = 0; // update() counter of the first visit (Philox counter words 0, 1)
    long long* acc = nullptr;
    long long* prop = nullptr;
    long long* nonfinite = nullptr;
    int phase_sync = 0; // CTA-wide barriers around the wrap (all chains of the CTA present: batch % 4 == 0)
};

// true if a chain-resident instantiation exists for this matrix size
bool segment_supported(int n);
int launch_segment(const SegmentArgs& a, cudaStream_t st);
int init_segment_kernels();
// the products of add_slice_sequence_left / _right with the running product resident in shared memory (segment_resident.cu)
struct SliceChainArgs {
    int n = 0, batch = 0;
    const double* Usrc = nullptr;   // [batch][n][n]
    const double* Dsrc = nullptr;   // [batch][n] column scaling applied at the end (null: none)
    double* dst = nullptr;          // [batch][n][n]
    const int8_t* conf = nullptr; long long sConf = 0;
    const double* eT2 = nullptr;    // e^{-dtau T}
    const double* eT2T = nullptr;   // its transpose
    int lo = 0, hi = 0, left = 1;   // slices lo .. hi (1-based); left: B_hi ... B_lo, else B_lo' ... B_hi'
    double ep = 1, em = 1;          // exp(+lambda s) for s = +1 / -1
};
int launch_slice_chain_resident(const SliceChainArgs& a, cudaStream_t st);
// the same run of slice visits with G resident in one CTA's shared memory (segment_resident.cu): N = 144
bool segment_resident_supported(int n);
int launch_segment_resident(const SegmentArgs& a, cudaStream_t st);
int init_segment_resident_kernels();

} // namespace mcb
