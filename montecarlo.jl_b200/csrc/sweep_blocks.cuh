// Block-wise local updates shared by the delayed sweep kernels (sweep_delayed.cu: G in L2; segment_resident.cu: G in
// shared memory): the Metropolis decisions of a block of K sites on its K x K diagonal block, held in the registers of
// one warp, and the two triangular recurrences that turn the decisions into the update vectors of the block.
// Reference: sweep_spatial / propose_local / accept_local! (src/flavors/DQMC/DQMC.jl:546-582,
// src/models/HubbardModel/HubbardModelAttractive.jl:113-155) with the rank-1 updates re-associated, not changed.
#pragma once
#include "sweep.cuh"

namespace mcb {
namespace swb {

// ---- the sites of a block without a block-wide barrier per site (one spin sector) -------------------
// The decisions of a block only need its K x K diagonal block D = G[i0.., i0..]. Four or eight warps work on a copy DD in
// shared memory (thread = row a, every fourth / eighth column) and walk through the sites - G_ii, Metropolis decision,
// rank-1 update of the trailing part of DD, one named barrier of these warps - like a right-looking elimination:
// after site j neither row j nor column j of DD is touched again, so at the end DD holds, for every site, the row
// of D at the moment the site was visited (upper triangle) and its column (lower triangle, then scaled by x_j),
// next to x_j itself (0 if rejected). The full vectors follow from two triangular recurrences that need no
// decisions and no barriers, one element of a column / of a row per thread:
//   u_j   = x_j (e_i - [Cc_j - sum_{j'<j} u_j' DD[j'][j]])                     (column role, u_j replaces Cc_j)
//   row_j = Rc_j + sum_{j'<j} (x_j' DD[j][j']) row_j'                            (row role, frozen rows in Rc)
// which is what the site-by-site updates of the columns j' > j and rows k > j add up to.
constexpr int DLD = 33;        // row stride of DD (conflict-free column access)

// One warp, the diagonal block in registers: lane a holds row a (KMAX entries), the pivot row travels by shuffles,
// no shared-memory round trip and no barrier per site (the 8-warp shared-memory version took 670 cycles per site,
// most of it the named barrier and the LDS -> FMA -> STS chain). Same operations in the same order, hence the same bits.
#ifndef SWB_KMAX
#define SWB_KMAX 24
#endif
constexpr int KMAX = SWB_KMAX;   // block size cap (delayed_geom): entries of a row in registers (24; 32 in experiments)

struct DecideCtx {
    double gamma_p, gamma_m, efac_p, efac_m;
    int s_mine;        // lane j carries the field and the uniform of site j
    double u_mine;
    int kb, i0, lane, nacc;
    double* xs;
    int8_t* s_out;
};

// site J of the block (compile-time recursion keeps every register index static). The phase is bound by the
// shuffle pipe (2 SHFL per entry of the pivot row: 12.3 k cycles per block of 24 at N = 144); starting the next
// site's reciprocal chain under the rest of this site's update was measured and is no faster (14.4 k).
template <int KM, int J>
__device__ __forceinline__ void decide_site(double (&reg)[KM], DecideCtx& c)
{
    if constexpr (J < KM) {
        if (J < c.kb) {   // warp-uniform
            const double gii = __shfl_sync(0xffffffffu, reg[J], J);
            const int s = __shfl_sync(0xffffffffu, c.s_mine, J);
            const double u = __shfl_sync(0xffffffffu, c.u_mine, J);
            const double gamma = s > 0 ? c.gamma_p : c.gamma_m;
            const double efac = s > 0 ? c.efac_p : c.efac_m;
            const double rt = 1.0 + gamma * (1.0 - gii);
            const double x = gamma * fast_rcp(rt);          // speculative: only used if accepted
            const double p = efac * (rt * rt);
            const bool accept = (p > 1.0 || u < p);
            // -u of row `lane` (u = x (e_i - G[:, i])), zero for the rows that are final and for a rejected flip: the update
            // below is unconditional (reg + 0 * row = reg), so the shuffles of row J do not wait for the decision
            const double mu = (accept && c.lane > J) ? x * reg[J] : 0.0;
#pragma unroll
            for (int b = J + 1; b < KM; ++b) {
                const double rowjb = __shfl_sync(0xffffffffu, reg[b], J);
                reg[b] = fma(mu, rowjb, reg[b]);
            }
            c.nacc += accept ? 1 : 0;
            // column J is final: keep it scaled by x_J (0 if rejected) for the row recurrence
            if (c.lane > J) reg[J] = mu;
            if (c.lane == 0) { c.xs[J] = accept ? x : 0.0; c.s_out[c.i0 + J] = (int8_t)(accept ? -s : s); }
            decide_site<KM, J + 1>(reg, c);
        }
    }
}

// KM = entries of a row held in registers (the block size cap of the caller: KMAX for the L2-resident sweep, 16 for the
// shared-memory-resident one); gm = {gamma_p, gamma_m, efac_p, efac_m} of propose_local
struct DecideConst { double gamma_p, gamma_m, efac_p, efac_m; };

template <int KM>
__device__ __forceinline__ int decide_rows(const DecideConst& a, double (&reg)[KM], int i0, int kb, const int8_t* s_in,
                                           int8_t* s_out, const double* unif, double* DD, double* xs, int lane)
{
    DecideCtx c{a.gamma_p, a.gamma_m, a.efac_p, a.efac_m, lane < kb ? (int)s_in[i0 + lane] : 1,
                lane < kb ? unif[i0 + lane] : 2.0, kb, i0, lane, 0, xs, s_out};
    decide_site<KM, 0>(reg, c);
    {   // all 32 rows and columns of the table: the recurrences read (and discard) the padded ones
        double* myrow = DD + lane * DLD;
#pragma unroll
        for (int b = 0; b < KM; ++b) myrow[b] = reg[b];
#pragma unroll
        for (int b = KM; b < 32; ++b) myrow[b] = 0.0;
    }
    return c.nacc;
}

// Both recurrences run in sub-blocks of 8 sites: the contributions of all earlier sub-blocks are plain sums with
// eight independent accumulators (one load of the earlier vector element per eight FMAs), only the 8 x 8 triangle
// inside the sub-block is a dependent chain.
__device__ __forceinline__ void column_recurrence(double* Cc, int ld, int i0, int kb, const double* DD, const double* xs, int e)
{
    for (int j0 = 0; j0 < kb; j0 += 8) {
        double c[8], u[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) c[q] = (j0 + q < kb) ? Cc[(j0 + q) * ld + e] : 0.0;
        for (int jp = 0; jp < j0; ++jp) {
            const double up = Cc[jp * ld + e];          // u_jp, final
            const double* d = DD + jp * DLD + j0;        // row jp of the block at the moment of site jp
#pragma unroll
            for (int q = 0; q < 8; ++q) c[q] = fma(-up, d[q], c[q]);
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            double cc = c[q];
#pragma unroll
            for (int p = 0; p < q; ++p) cc = fma(-u[p], DD[(j0 + p) * DLD + j0 + q], cc);
            u[q] = (j0 + q < kb) ? xs[j0 + q] * ((e == i0 + j0 + q ? 1.0 : 0.0) - cc) : 0.0;
            if (j0 + q < kb) Cc[(j0 + q) * ld + e] = u[q];
        }
    }
}

__device__ __forceinline__ void row_recurrence(double* Rc, int ld, int kb, const double* DD, int e)
{
    for (int j0 = 0; j0 < kb; j0 += 8) {
        double r[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) r[q] = (j0 + q < kb) ? Rc[(j0 + q) * ld + e] : 0.0;
        for (int jp = 0; jp < j0; ++jp) {
            const double rp = Rc[jp * ld + e];           // row_jp, final
            const double* d = DD + j0 * DLD + jp;         // x_jp times column jp of the block at the moment of site jp
#pragma unroll
            for (int q = 0; q < 8; ++q) r[q] = fma(d[q * DLD], rp, r[q]);
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
#pragma unroll
            for (int p = 0; p < q; ++p) r[q] = fma(DD[(j0 + q) * DLD + j0 + p], r[p], r[q]);
            if (j0 + q < kb) Rc[(j0 + q) * ld + e] = r[q];
        }
    }
}

} // namespace swb
} // namespace mcb
