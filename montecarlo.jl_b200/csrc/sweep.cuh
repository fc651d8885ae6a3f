// Local-update kernels: sweep_spatial of one time slice for a batch of chains.
#pragma once
#include "common.cuh"

namespace mcb {

struct SweepArgs {
    int n = 0, batch = 0;
    double* G = nullptr; long long sG = 0;
    int8_t* conf = nullptr; long long sConf = 0; // row of the current slice, per chain stride
    // constants of propose_local (HubbardModelAttractive.jl:113-127), computed on the host:
    // for s = +1 / -1 : gamma = exp(dE) - 1, efac = exp(-dE), dE = -2 lambda s
    double gamma_p = 0, gamma_m = 0, efac_p = 0, efac_m = 0;
    uint32_t key0 = 0, chain_offset = 0, sweep = 0, visit = 0;
    long long* acc = nullptr;   // [batch]
    long long* prop = nullptr;  // [batch]
    long long* nonfinite = nullptr; // [batch]
    // HubbardModelRepulsive (HubbardModelRepulsive.jl:126-232): two spin sectors share the field with opposite
    // sign. The second sector is kept as an attractive-like chain on the negated field (G2, conf2 = -conf);
    // the two sectors only meet in the acceptance probability p = R_up * R_down (no bosonic factor).
    double* G2 = nullptr;
    int8_t* conf2 = nullptr;
};

int launch_sweep_spatial(const SweepArgs& a, cudaStream_t st);
int init_sweep_kernels();
// same contract, delayed rank-k updates flushed on the FP64 tensor pipe (sweep_delayed.cu)
int launch_sweep_delayed(const SweepArgs& a, cudaStream_t st);
int init_sweep_delayed_kernels();

} // namespace mcb
