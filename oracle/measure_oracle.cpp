// ORACLE (test infrastructure, never linked into libmcb200.so).
// Drivers for the measurement kernels of measure_ref.h: the reference's prepare! / apply! / finish! sequence per
// lattice iterator (src/flavors/DQMC/measurements/generic.jl:279-347) and the measuring part of run!
// (src/flavors/DQMC/DQMC.jl:420-438). The iterator tuple lists come from oracle/lattice_iterators.py.
#include "measure_ref.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

using namespace measure_ref;

extern "C" {
// from dqmc_oracle.cpp
void orc_update(void* h, int64_t n);
void orc_update_repulsive(void* hup, void* hdn, int64_t n);
void orc_get_greens(void* h, double* out, int physical);
void orc_get_state(void* h, int64_t* out, double* max_prop_err);

// apply!(::EachSitePairByDistance, ...) (generic.jl:331-336) between prepare! (:283) and finish! (:292-295).
// trip: n x (dir, src, trg). kind: 0 cdc, 1 sdc_x, 2 sdc_y, 3 sdc_z. flavors 1: G is N x N and the
// HubbardModelAttractive overloads apply; flavors 2: G is 2N x 2N.
void orc_meas_pairs(const double* G, int N, int flavors, int kind, const int32_t* trip, int64_t n, double* out,
                    int n_out)
{
    for (int d = 0; d < n_out; ++d) out[d] = 0.0;
    for (int64_t e = 0; e < n; ++e) {
        const int dir = trip[3 * e], src = trip[3 * e + 1], trg = trip[3 * e + 2];
        double v;
        if (flavors == 1) v = kind == 0 ? cdc_attractive(G, N, src, trg) : sdc_attractive(G, N, src, trg);
        else {
            const G2 g{G, N};
            v = kind == 0 ? cdc_general(g, src, trg) : kind == 1 ? sdc_x_general(g, src, trg)
              : kind == 2 ? sdc_y_general(g, src, trg) : sdc_z_general(g, src, trg);
        }
        out[dir] += v;
    }
    for (int d = 0; d < n_out; ++d) out[d] *= 1.0 / N;
}

// apply!(::EachLocalQuadByDistance, ...) (generic.jl:341-347) with pc_kernel; quint: n x (lin, src1, trg1, src2, trg2)
void orc_meas_quads(const double* G, int N, int flavors, const int32_t* quint, int64_t n, double* out, int n_out)
{
    for (int d = 0; d < n_out; ++d) out[d] = 0.0;
    for (int64_t e = 0; e < n; ++e) {
        const int32_t* q = quint + 5 * e;
        out[q[0]] += flavors == 1 ? pc_attractive(G, N, q[1], q[2], q[3], q[4])
                                  : pc_general(G2{G, N}, q[1], q[2], q[3], q[4]);
    }
    for (int d = 0; d < n_out; ++d) out[d] *= 1.0 / N;
}

// apply!(::EachSiteAndFlavor / ::EachSite, ...) (generic.jl:299-311). kind 0: occupation_kernel over all
// flavors*N diagonal entries (measurements.jl:41-44); 1, 2, 3: mx, my, mz over the N sites (0 for one flavor,
// HubbardModelAttractive.jl:231-233)
void orc_meas_sites(const double* G, int N, int flavors, int kind, double* out)
{
    if (kind == 0) {
        const int n = flavors * N;
        for (int i = 0; i < n; ++i) out[i] = 1 - G[i + (size_t)i * n];
        return;
    }
    for (int i = 0; i < N; ++i) {
        if (flavors == 1) { out[i] = 0.0; continue; }
        const G2 g{G, N};
        out[i] = kind == 1 ? mx_general(g, i) : kind == 2 ? my_general(g, i) : mz_general(g, i);
    }
}

// SURVEY.md 8a O5 (no reference kernel; definitions from the Hamiltonian docs/src/models/hubbardattractive.md:6-8,
// H = -t sum c+c - U sum (n_up - 1/2)(n_dn - 1/2) - mu sum n, spin sectors share G given the field):
//   out[0] = <n>      = (1/N) sum_i 2 (1 - G_ii)
//   out[1] = <n_up n_dn> = (1/N) sum_i (1 - G_ii)^2
//   out[2] = E_kin/N  = (2/N) sum_ij T_ij (delta_ij - G_ji)        (T carries -mu on its diagonal)
//   out[3] = E_int/N  = -(U/N) sum_i (D_i - n_i/2 + 1/4)
void orc_meas_o5(const double* G, int N, const double* T, double U, double* out)
{
    double occ = 0, docc = 0, ekin = 0;
    for (int i = 0; i < N; ++i) {
        const double ni = 1 - G[i + (size_t)i * N];
        occ += 2 * ni;
        docc += ni * ni;
    }
    for (int j = 0; j < N; ++j)
        for (int i = 0; i < N; ++i) ekin += T[i + (size_t)j * N] * (I(i, j) - G[j + (size_t)i * N]);
    out[0] = occ / N;
    out[1] = docc / N;
    out[2] = 2 * ekin / N;
    out[3] = -U * (docc - 0.5 * occ + 0.25 * N) / N;
}

// The sweep loop of run! (DQMC.jl:420-438) for sweeps first..last (1-based, inclusive): 2M updates per sweep,
// after each one `current_slice == 1 && direction == 1 && i > thermalization && i % measure_rate == 0` triggers
// the measurements (greens!(mc) = the physical G, generic.jl:207-215). Sums over measurements go to the non-null
// accumulators; returns the number of measurements taken. hdn != null: HubbardModelRepulsive (two sectors, G
// assembled as the BlockDiagonal 2N x 2N matrix).
int64_t orc_run_measured(void* hup, void* hdn, int N, int M, int64_t first, int64_t last, int64_t thermalization,
                         int measure_rate, const int32_t* trip, int64_t n_trip, int n_dirs, const int32_t* quint,
                         int64_t n_quint, int n_pc, const double* T, double U, double* accG, double* acc_occ,
                         double* acc_cdc, double* acc_sdc /*[3][n_dirs]*/, double* acc_pc, double* acc_mz,
                         double* acc_o5)
{
    const int fl = hdn ? 2 : 1, n = fl * N;
    std::vector<double> G((size_t)n * n, 0.0), blk((size_t)N * N), tmp(std::max<size_t>({(size_t)n, (size_t)n_dirs, (size_t)n_pc, 4}));
    int64_t count = 0;
    for (int64_t i = first; i <= last; ++i) {
        for (int u = 0; u < 2 * M; ++u) {
            if (hdn) orc_update_repulsive(hup, hdn, 1); else orc_update(hup, 1);
            int64_t st[5]; double pe;
            orc_get_state(hup, st, &pe);
            if (!(st[0] == 1 && st[1] == 1 && i > thermalization && i % measure_rate == 0)) continue;
            if (!hdn) orc_get_greens(hup, G.data(), 1);
            else {
                for (int s = 0; s < 2; ++s) {
                    orc_get_greens(s ? hdn : hup, blk.data(), 1);
                    for (int c = 0; c < N; ++c)
                        memcpy(&G[(size_t)(s * N) + (size_t)(s * N + c) * n], &blk[(size_t)c * N], N * sizeof(double));
                }
            }
            ++count;
            if (accG) for (size_t e = 0; e < G.size(); ++e) accG[e] += G[e];
            if (acc_occ) { orc_meas_sites(G.data(), N, fl, 0, tmp.data()); for (int k = 0; k < n; ++k) acc_occ[k] += tmp[k]; }
            if (acc_cdc) { orc_meas_pairs(G.data(), N, fl, 0, trip, n_trip, tmp.data(), n_dirs); for (int k = 0; k < n_dirs; ++k) acc_cdc[k] += tmp[k]; }
            if (acc_sdc) for (int x = 0; x < 3; ++x) {
                orc_meas_pairs(G.data(), N, fl, 1 + x, trip, n_trip, tmp.data(), n_dirs);
                for (int k = 0; k < n_dirs; ++k) acc_sdc[(size_t)x * n_dirs + k] += tmp[k];
            }
            if (acc_pc) { orc_meas_quads(G.data(), N, fl, quint, n_quint, tmp.data(), n_pc); for (int k = 0; k < n_pc; ++k) acc_pc[k] += tmp[k]; }
            if (acc_mz) { orc_meas_sites(G.data(), N, fl, 3, tmp.data()); for (int k = 0; k < N; ++k) acc_mz[k] += tmp[k]; }
            if (acc_o5 && !hdn) { orc_meas_o5(G.data(), N, T, U, tmp.data()); for (int k = 0; k < 4; ++k) acc_o5[k] += tmp[k]; }
        }
    }
    return count;
}

} // extern "C"
