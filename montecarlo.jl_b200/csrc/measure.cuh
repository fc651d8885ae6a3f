// Equal-time observables of a batch of chains (measure.cu): the reference's measurement kernels driven by its
// lattice iterators, one deterministic pass per measurement.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mcb {

// Tables + accumulators of one handle. flavors = 1: HubbardModelAttractive (one N x N block, both spins);
// flavors = 2: HubbardModelRepulsive (spin-up and spin-down blocks of the BlockDiagonal G).
struct MeasureState {
    int N = 0, C = 0, flavors = 1;
    int n_dirs = 0, K = 0;
    double U_signed = 0.0;       // coefficient of (n_up - 1/2)(n_dn - 1/2): -|U| attractive, +|U| repulsive
    int32_t* trg_of = nullptr;   // [n_dirs][N]: target of the pair (src, trg) in direction dir, -1 = none
    double* hopT = nullptr;      // hopping matrix T, N x N column-major
    double* obs = nullptr;       // [C][obs_len()] sums over measurements
    double* Gsum = nullptr;      // [C][flavors][N*N] sums of the physical G
    int64_t n_meas = 0;
    // time-displaced measurements, per chain sums over measurements:
    //   one flavor   [cds[n_dirs] | sds[n_dirs] (x = y = z) | ps[n_dirs*K*K] | ccs[n_dirs*K]]
    //   two flavors  [cds[n_dirs] | sds_x[n_dirs] (= y) | sds_z[n_dirs] | ps[n_dirs*K*K] | ccs[n_dirs*K]]
    //   (ps column-major (dir12, dir1, dir2), ccs column-major (dir12, dir1))
    double* td = nullptr;        // [C][td_len()]
    double* G00 = nullptr;       // [flavors][C][N*N] the equal-time physical G of the measurement in progress
    int64_t n_td = 0;
    int td_pairs() const { return flavors == 2 ? 3 : 2; }
    int td_len() const { return n_dirs * (td_pairs() + K * K + K); }

    // record of one chain: [n, double occupancy, E_kin/N, E_int/N | cdc[n_dirs] | sdc_x[n_dirs] | sdc_z[n_dirs] |
    //                       pc[n_dirs*K*K] (column-major (dir12, dir1, dir2)) | occupation[flavors*N] | mz[N] (2 flavors)]
    int off_cdc() const { return 4; }
    int off_sdcx() const { return 4 + n_dirs; }
    int off_sdcz() const { return 4 + 2 * n_dirs; }
    int off_pc() const { return 4 + 3 * n_dirs; }
    int off_occ() const { return off_pc() + n_dirs * K * K; }
    int off_mz() const { return off_occ() + flavors * N; }
    int obs_len() const { return off_mz() + (flavors == 2 ? N : 0); }
    bool ready() const { return obs != nullptr; }
};

int measure_setup(MeasureState& m, int N, int C, int flavors, const int32_t* trg_of_host, int n_dirs, int K,
                  const double* hop_host, double U_signed, cudaStream_t st);
void measure_free(MeasureState& m);
int measure_reset(MeasureState& m, cudaStream_t st);
// accumulate the observables of the physical Green's functions Gu (and Gd; pass Gu for one flavor), [C][N*N]
int measure_accumulate(MeasureState& m, const double* Gu, const double* Gd, cudaStream_t st);
// one slice l of a time-displaced measurement: td += factor * kernel(G00, G0l, Gl0, Gll) per output bin (measure.cu)
// up[4] / dn[4] = {G00, G0l, Gl0, Gll} of the spin-up / spin-down block, [C][N*N] each (dn = up for one flavor)
int measure_td_accumulate(MeasureState& m, const double* const up[4], const double* const dn[4], double factor, cudaStream_t st);
// the buffer(s) for G00 of the measurement in progress, allocated on first use: g00[f] for f < flavors
int measure_td_g00(MeasureState& m, double* g00[2]);

} // namespace mcb
