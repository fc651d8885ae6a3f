// ORACLE (test infrastructure only — never linked into the product library).
//
// CPU restatement of MonteCarlo.jl's classical Ising Metropolis path
// (paths relative to /root/reference):
//   local rule   src/models/Ising/IsingModel.jl:92-101 (propose_local, square lattice)
//                src/flavors/MC/MC.jl:316-332        (sweep: accept iff dE<=0 || rand()<exp(-beta dE))
//                src/models/Ising/IsingModel.jl:83-89 (accept_local!: E += dE; s = -s)
//   energy       src/models/Ising/IsingModel.jl:149-185
//   neighbours   src/lattices/square.jl:47-52 : up=(x+1,y) right=(x,y+1) down=(x-1,y) left=(x,y-1),
//                site index x + y*L (0-based here), periodic.
//
// The reference sweeps sites sequentially with Julia's MersenneTwister. The
// checkerboard schedule + counter RNG below are this project's specification
// (SURVEY.md Appendix D); this file is its plain per-site statement, written
// without any bit tricks so it is independent of the CUDA kernel's packing.
// Pinned by the reference's RNG-free known answers
// (test/modeltests_IsingModel.jl:21,27,34) in tests/test_oracle_ising.py.
#include "philox_ref.h"
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

extern "C" {

// Generic D-dimensional periodic cubic-lattice energy over "upright" bonds
// (IsingModel.jl:149-170 with neighbors(l, Val(false)); lattices/cubic.jl:44-57).
// conf holds arbitrary integers (the reference KAT uses 1..L^D).
double ising_energy_cubic(const int64_t* conf, int L, int D)
{
    int64_t sites = 1;
    for (int d = 0; d < D; ++d) sites *= L;
    double E = 0.0;
    for (int64_t src = 0; src < sites; ++src) {
        int64_t stride = 1;
        for (int d = 0; d < D; ++d) {
            const int64_t coord = (src / stride) % L;
            const int64_t trg = src + (((coord + 1) % L) - coord) * stride;
            E -= (double)(conf[src] * conf[trg]);
            stride *= L;
        }
    }
    return E;
}

// Square-lattice specialisation (IsingModel.jl:177-185): E = -sum_i s_i (s_up + s_right)
double ising_energy_square(const int64_t* conf, int L)
{
    double E = 0.0;
    for (int y = 0; y < L; ++y)
        for (int x = 0; x < L; ++x) {
            const int64_t s = conf[x + (int64_t)y * L];
            const int64_t up = conf[(x + 1) % L + (int64_t)y * L];
            const int64_t right = conf[x + (int64_t)((y + 1) % L) * L];
            E -= (double)(s * up + s * right);
        }
    return E;
}

// propose_local for the square lattice (IsingModel.jl:92-101); `site` 0-based.
double ising_delta_e_square(const int64_t* conf, int L, int64_t site)
{
    const int x = (int)(site % L), y = (int)(site / L);
    const int64_t up = conf[(x + 1) % L + (int64_t)y * L];
    const int64_t right = conf[x + (int64_t)((y + 1) % L) * L];
    const int64_t down = conf[(x + L - 1) % L + (int64_t)y * L];
    const int64_t left = conf[x + (int64_t)((y + L - 1) % L) * L];
    return 2.0 * (double)conf[site] * (double)(up + right + down + left);
}

// Integer acceptance thresholds, computed once on the host in double precision.
void ising_thresholds(double beta, uint32_t* thr4, uint32_t* thr8)
{
    // saturated so that beta = 0 (product 2^32) stays representable; beta < 0 is rejected by the callers
    *thr4 = (uint32_t)fmin(floor(exp(-4.0 * beta) * 4294967296.0), 4294967295.0);
    *thr8 = (uint32_t)fmin(floor(exp(-8.0 * beta) * 4294967296.0), 4294967295.0);
}

static inline int words_per_row(int L) { return (L / 2 + 63) / 64; }

// 32-bit acceptance word of site (x,y) in (sweep, colour) for replica `replica`.
// Mapping: xh = x/2 is the site's position inside its colour plane row; 64 plane
// positions form one word, 4 consecutive positions share one Philox call.
uint32_t ising_site_random(uint64_t seed, uint32_t replica, uint32_t sweep, int colour, int L, int x,
                           int y)
{
    const int W = words_per_row(L);
    const int xh = x >> 1;
    const uint32_t group = (uint32_t)(((uint64_t)y * W + (xh >> 6)) * 16 + ((xh & 63) >> 2));
    uint32_t ctr[4] = {sweep, (uint32_t)colour, group, 0u};
    uint32_t key[2] = {(uint32_t)seed ^ (uint32_t)(seed >> 32), replica};
    uint32_t o[4];
    philox4x32_10_ref(ctr, key, o);
    return o[xh & 3];
}

// n checkerboard sweeps on int8 spins (+1/-1), spins[x + y*L]. sweep0 = index of the first sweep.
void ising_checkerboard_sweeps(int8_t* spins, int L, uint32_t thr4, uint32_t thr8, uint64_t seed,
                               uint32_t replica, uint32_t sweep0, int n_sweeps)
{
    for (int sw = 0; sw < n_sweeps; ++sw)
        for (int colour = 0; colour < 2; ++colour)
            for (int y = 0; y < L; ++y)
                for (int x = (y + colour) & 1; x < L; x += 2) {
                    const int64_t i = x + (int64_t)y * L;
                    const int nb = spins[(x + 1) % L + (int64_t)y * L] + spins[x + (int64_t)((y + 1) % L) * L] +
                                   spins[(x + L - 1) % L + (int64_t)y * L] +
                                   spins[x + (int64_t)((y + L - 1) % L) * L];
                    const int dE = 2 * spins[i] * nb;
                    bool acc = dE <= 0;
                    if (!acc) {
                        const uint32_t r = ising_site_random(seed, replica, sweep0 + sw, colour, L, x, y);
                        acc = r < (dE == 4 ? thr4 : thr8);
                    }
                    if (acc) spins[i] = (int8_t)-spins[i];
                }
}

// Reference-style sequential sweep (MC.jl:316-332) with the double-precision rule
// rand() < exp(-beta dE). The uniform of proposal (sweep, site) comes from the
// DQMC-style 53-bit Philox stream; drawn only when dE > 0 like the reference
// (irrelevant for an indexed stream). Returns accepted count; *energy is updated.
int64_t ising_sequential_sweeps(int8_t* spins, int L, double beta, uint64_t seed, uint32_t replica,
                                uint32_t sweep0, int n_sweeps, double* energy)
{
    int64_t acc = 0;
    const double w4 = exp(-beta * 4.0), w8 = exp(-beta * 8.0);
    for (int sw = 0; sw < n_sweeps; ++sw)
        for (int y = 0; y < L; ++y)
            for (int x = 0; x < L; ++x) {
                const int64_t i = x + (int64_t)y * L;
                const int nb = spins[(x + 1) % L + (int64_t)y * L] + spins[x + (int64_t)((y + 1) % L) * L] +
                               spins[(x + L - 1) % L + (int64_t)y * L] +
                               spins[x + (int64_t)((y + L - 1) % L) * L];
                const int dE = 2 * spins[i] * nb;
                bool a = dE <= 0;
                if (!a) {
                    const double u = dqmc_uniform_ref(seed, replica, sweep0 + sw, (uint32_t)y, (uint32_t)x);
                    a = u < (dE == 4 ? w4 : w8);
                }
                if (a) { spins[i] = (int8_t)-spins[i]; *energy += dE; ++acc; }
            }
    return acc;
}

// E = -sum_i s_i (s_up + s_right), M = sum_i s_i  on int8 spins (exact integers).
void ising_energy_magnetization(const int8_t* spins, int L, int64_t* E, int64_t* Msum)
{
    int64_t e = 0, m = 0;
    for (int y = 0; y < L; ++y)
        for (int x = 0; x < L; ++x) {
            const int s = spins[x + (int64_t)y * L];
            e -= s * (spins[(x + 1) % L + (int64_t)y * L] + spins[x + (int64_t)((y + 1) % L) * L]);
            m += s;
        }
    *E = e; *Msum = m;
}

// Hot start: spin(x,y) of replica from Philox, counter = (0xFFFFFFFF, 0, site/4 low, site/4 high)
void ising_random_spins(int8_t* spins, int L, uint64_t seed, uint32_t replica)
{
    const int64_t n = (int64_t)L * L;
    uint32_t key[2] = {(uint32_t)seed ^ (uint32_t)(seed >> 32), replica};
    for (int64_t g = 0; g * 128 < n; ++g) {
        uint32_t ctr[4] = {0xFFFFFFFFu, 0u, (uint32_t)g, (uint32_t)(g >> 32)};
        uint32_t o[4];
        philox4x32_10_ref(ctr, key, o);
        for (int b = 0; b < 128 && g * 128 + b < n; ++b)
            spins[g * 128 + b] = ((o[b >> 5] >> (b & 31)) & 1u) ? 1 : -1;
    }
}

} // extern "C"
