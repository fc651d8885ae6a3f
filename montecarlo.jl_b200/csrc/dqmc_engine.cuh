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
#include "measure.cuh"
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
    uint64_t* conf_bits = nullptr;                          // staging for the bit-packed wire format
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
    // measurement tables and accumulators (measure.cu)
    mcb::MeasureState ms;
    // schedule state
    int current_slice = 0, direction = 1;
    int64_t n_updates = 0;
    mcb::Profiler prof;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    bool use_panel = false; // chain-resident kernels (panel.cuh) exist for this N
    bool td_scheduled = false;         // time-displaced measurements ride along with the scheduled equal-time ones
    int td_recalculate = 0;
    bool use_resident_segment = false; // wrap + sweep with G in shared memory (segment_resident.cu), general stabilisation
    double* hopT[2] = {nullptr, nullptr}; // transposes of hop[2], hop[3] for that kernel
    bool hopping_set = false, conf_set = false, stack_built = false;
    struct UtStack* ut = nullptr; // unequal-time stacks and work matrices, allocated on first use (unequal.cu)
    int64_t conf_version = 0;     // bumped whenever the configuration may have changed
};

namespace mcb {
// exp(+-lambda s) of slice l (1-based) as a Scale; power > 0 -> e^{+lambda s}
inline Scale conf_scale(const mcb_dqmc* h, int slice, bool positive)
{
    Scale sc;
    sc.s = h->conf + (size_t)(slice - 1) * h->N;
    sc.stride_s = (long long)h->M * h->N;
    sc.ep = positive ? h->eVp[0] : h->eVm[0];
    sc.em = positive ? h->eVp[1] : h->eVm[1];
    return sc;
}
inline Scale vec_scale(const double* d, long long stride, bool inv = false)
{
    Scale sc;
    sc.d = d; sc.stride_d = stride; sc.inv = inv ? 1 : 0;
    return sc;
}

inline GemmArgs base_gemm(const mcb_dqmc* h, const double* A, bool sharedA, const double* B, bool sharedB,
                   double* C)
{
    GemmArgs g;
    g.n = h->N; g.batch = h->C;
    g.A = A; g.strideA = sharedA ? 0 : (long long)h->nn;
    g.B = B; g.strideB = sharedB ? 0 : (long long)h->nn;
    g.C = C; g.strideC = (long long)h->nn;
    return g;
}

int engine_alloc(mcb_dqmc* h);
void engine_free(mcb_dqmc* h);
int engine_build_stack(mcb_dqmc* h);
int engine_propagate(mcb_dqmc* h, bool defer_wrap, bool* wrap_pending);
int engine_update(mcb_dqmc* h);
int engine_update_batch(mcb_dqmc* h, int64_t max_updates, bool stop_at_measure_point, int64_t* done);
int engine_calculate_greens_at(mcb_dqmc* h, int slice);
int engine_physical_greens(mcb_dqmc* h, double* out_dev);
int engine_measure(mcb_dqmc* h);
// unequal-time Green's function G(k <- l) of the current configuration into h->G (unequal.cu)
int engine_greens_unequal(mcb_dqmc* h, int slice1, int slice2);
void unequal_free(mcb_dqmc* h);
double* unequal_result(mcb_dqmc* h);
// CombinedGreensIterator on the device (unequal.cu)
int engine_measure_td(mcb_dqmc* h, int recalculate, bool from_g0 = false);
int engine_cgi_begin_from_g0(mcb_dqmc* h, int recalculate);
int engine_cgi_begin(mcb_dqmc* h, int recalculate);
int engine_cgi_next(mcb_dqmc* h, int physical, int* l_out, const double* results[3]);
int engine_cgi_last(mcb_dqmc* h, const double* results[3]);
// add_slice_sequence_left/right on raw (U, D, T) triples
int advance_raw(mcb_dqmc* h, bool left, const double* Usrc, const double* Dsrc, const double* Tsrc, double* Udst,
                double* Ddst, double* Tdst, int lo, int hi);
// batched calculate_greens on raw device buffers (shared with the linalg entry point); resident: use the
// chain-resident kernels of stab.cuh where they exist for n (the handle's use_panel / the linalg force flag)
int calc_greens_raw(int n, int batch, const double* Ul, const double* Dl, const double* Tl,
                    const double* Ur, const double* Dr, const double* Tr, double* G, double* const W[6],
                    double* const Dw[2], int* pivot, cudaStream_t st, bool resident);
} // namespace mcb
