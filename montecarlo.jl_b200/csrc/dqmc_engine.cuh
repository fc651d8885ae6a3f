// Engine state of one DQMC handle: a batch of independent Markov chains on one device.
// Host code replays the reference's (chain-independent) slice schedule and launches batched
// kernels; see dqmc_engine.cu for the mapping to src/flavors/DQMC/stack.jl.
#pragma once
#include "../../include/mcb200.h"
#include "common.cuh"
#include "gemm.cuh"
#include "sweep.cuh"
#include "segment.cuh"
#include "stab.cuh"
#include "udt.cuh"
#include <vector>

struct mcb_dqmc {
    mcb_dqmc_params p{};
    int N = 0, M = 0, s = 0, C = 0, ne = 0;
    double lambda = 0;
    double eVp[2] = {1, 1}; // exp(+lambda * (+1)), exp(+lambda * (-1))   power > 0
    double eVm[2] = {1, 1}; // exp(-lambda * (+1)), exp(-lambda * (-1))   power < 0
    cudaStream_t st = nullptr;
    size_t nn = 0;      // N*N
    size_t mat = 0;     // C*N*N   (one batched matrix)
    size_t vec = 0;     // C*N
    // device memory
    double* hop[4] = {nullptr, nullptr, nullptr, nullptr}; // eT, eTinv, eT2, eTinv2
    int8_t* conf = nullptr;                                 // [C][M][N]
    double* G = nullptr;
    double* Gtmp = nullptr;
    double* W[8] = {};
    double* Dw[2] = {};
    double* Dx = nullptr;
    int* pivot = nullptr;
    std::vector<double*> bufU, bufT, bufD; // ne + 1 physical slot buffers
    std::vector<int> slot;                 // logical slot -> physical buffer
    int spare = 0;
    long long *acc = nullptr, *prop = nullptr, *nonfinite = nullptr;
    double* prop_err = nullptr; // [C]
    // measurement
    int32_t* dir_of_pair = nullptr;
    double* hopT = nullptr;
    int n_dirs = 0;
    double* obs = nullptr;  // [C][obs_len]
    double* Gsum = nullptr; // [C][N*N]
    int64_t n_meas = 0;
    // schedule state
    int current_slice = 0, direction = 1;
    int64_t n_updates = 0;
    mcb::Profiler prof;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    bool use_panel = false; // chain-resident kernels (panel.cuh) exist for this N
    bool hopping_set = false, conf_set = false, stack_built = false;

    int obs_len() const { return 4 + 2 * n_dirs; }
};

namespace mcb {
int engine_alloc(mcb_dqmc* h);
void engine_free(mcb_dqmc* h);
int engine_build_stack(mcb_dqmc* h);
int engine_propagate(mcb_dqmc* h, bool defer_wrap, bool* wrap_pending);
int engine_update(mcb_dqmc* h);
int engine_update_batch(mcb_dqmc* h, int64_t max_updates, bool stop_at_measure_point, int64_t* done);
int engine_calculate_greens_at(mcb_dqmc* h, int slice);
int engine_physical_greens(mcb_dqmc* h, double* out_dev);
int engine_measure(mcb_dqmc* h);
// batched calculate_greens on raw device buffers (shared with the linalg entry point)
int calc_greens_raw(int n, int batch, const double* Ul, const double* Dl, const double* Tl,
                    const double* Ur, const double* Dr, const double* Tr, double* G, double* const W[6],
                    double* const Dw[2], int* pivot, cudaStream_t st);
} // namespace mcb
