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
// the same run of slice visits with G resident in one CTA's shared memory (segment_resident.cu): N = 144
bool segment_resident_supported(int n);
int launch_segment_resident(const SegmentArgs& a, cudaStream_t st);
int init_segment_resident_kernels();

} // namespace mcb
