// sweep_spatial (src/flavors/DQMC/DQMC.jl:546-582) with propose_local / accept_local!
// (src/models/HubbardModel/HubbardModelAttractive.jl:113-155) for any N, with DELAYED rank-k
// Green's-function updates: one CTA per chain, G stays in global memory (L2).
//
// The sites of a slice are processed in blocks of K. For a block starting at site i0 the CTA holds
// in shared memory the K current columns Cc = G[:, i0:i0+K) and the K current rows Rc = G[i0:i0+K, :].
// An accepted flip at site i = i0 + j is the rank-1 update G -= u v' with u = x (e_i - G[:, i]),
// v = G[i, :], x = gamma / (1 + gamma (1 - G_ii))  (accept_local!, :142-153). It is applied at once
// only to the columns j' > j of Cc and the rows k > j of Rc (2 K N flop instead of 2 N^2); u is
// recorded in place of column j of Cc (no later site reads it; the store is deferred to the next barrier
// because the current site still does), v stays frozen in Rc[j, :]. After the block the whole matrix is brought up to
// date with ONE product G -= U * Rc (N x K x N) on the FP64 tensor pipe. The Metropolis decisions see
// exactly the Green's function the immediate updates would give (the arithmetic is re-associated only).
// One barrier per site: the next site's G_ii is published by the thread that owns its row. (For N > 64 the sites
// of a block are decided on its diagonal block first and the vectors follow from recurrences, see decide_rows.)
//
// Small batches of large matrices (config D: 32 chains, N = 400) leave most SMs idle with one CTA per chain,
// and the flush is 2/3 of the kernel. There a thread-block CLUSTER of `cls` CTAs serves one chain: every CTA
// runs the (cheap, sequential) site loop redundantly on its own copy of the block's columns and rows, so
// no data is exchanged, and the flush tiles are divided among the CTAs; one cluster barrier per block
// orders the G tiles written by the peers before the next block's loads.
//
// HubbardModelRepulsive: a cluster of two CTAs per chain, one per spin sector (a.G / a.G2 on the fields
// a.conf / a.conf2 = -a.conf). Each CTA publishes the next site's G_ii to its partner through distributed
// shared memory, the per-site barrier becomes a cluster barrier, and both take the same decision from
// p = R_own * R_partner (propose_local, HubbardModelRepulsive.jl:126-154; G is block diagonal, so the 2x2
// determinant is the product of the diagonal entries).
#include "sweep.cuh"
#include "sweep_blocks.cuh"
#include "tma.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>

namespace cg = cooperative_groups;

namespace mcb {

namespace {

constexpr int SMEM_LIMIT = 227 * 1024;
constexpr size_t TABLE_BYTES = (32 * 33 + 32 + SWB_KMAX * SWB_KMAX) * sizeof(double) + 16;   // DD, xs of the block-wise decisions, the next diagonal block, tile counter

#ifdef SWD_TIMING
__device__ long long swd_clk[32];
#define SWD_T(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) swd_clk[i] = clock64(); } while (0)
#else
#define SWD_T(i) do { } while (0)
#endif

struct DelayedGeom {
    int n, np, ld, K, threads; // np = n rounded up to 8, ld = np + 4 (ld % 16 == 4 keeps DMMA fragment loads conflict-free)
    size_t smem;
    int cls;   // CTAs per chain (thread-block cluster), 1 = plain launch
    int two;   // 1: the two CTAs of a cluster are the two spin sectors of the repulsive model (cls == 2)
    int exp;   // timing experiments (MCB_SWEEP_EXP): 1 = no flush, 2 = no site loop; results invalid
};

__host__ __device__ inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

DelayedGeom delayed_geom(int n)
{
    DelayedGeom g;
    g.n = n;
    g.np = round_up(n, 8);
    g.ld = g.np + ((20 - g.np % 16) % 16 == 0 ? 16 : (20 - g.np % 16) % 16); // ld % 16 == 4, ld > np
    const size_t hdr = (size_t)(2 * n + 16) * sizeof(double) + 2 * (size_t)round_up(n, 8);
    int K = (int)((SMEM_LIMIT - 2048 - hdr - TABLE_BYTES) / (2 * (size_t)g.ld * sizeof(double)));
    K = K / 4 * 4;
    // block size: 24 measured best at N = 144 (196 against 204 us per slice with 32) and at N = 400 (1092 / 1124)
    static const int kcap = getenv("MCB_SWEEP_K") ? atoi(getenv("MCB_SWEEP_K")) : swb::KMAX;
    if (K > kcap) K = kcap;
    if (K > swb::KMAX) K = swb::KMAX;   // the deciding warp keeps a row of the diagonal block in KMAX registers
    if (K > round_up(n, 4)) K = round_up(n, 4);
    g.K = K;
    // 512 threads (128 registers each) for every N > 64: no spills in the deciding warp and the old tile values
    // prefetched in the flush outweigh the sequential column / row recurrences at N = 400 (1.35 M against 1.49 M
    // cycles per slice); MCB_SWEEP_THREADS=1024 restores the 64-register variant for comparisons
    static const int thr = getenv("MCB_SWEEP_THREADS") ? atoi(getenv("MCB_SWEEP_THREADS")) : 0;
    g.threads = n <= 32 ? 128 : (n <= 64 ? 256 : 512);
    if (thr == 1024 && n > 256) g.threads = 1024;
    g.smem = hdr + 2 * (size_t)K * g.ld * sizeof(double) + TABLE_BYTES;
    static const int exp_flags = experiment_flags("MCB_SWEEP_EXP");
    g.exp = exp_flags;
    return g;
}

using namespace swb;   // decide_rows, column_recurrence, row_recurrence, DLD, KMAX (sweep_blocks.cuh)

// BW: block-wise decisions (one spin sector only); the site-by-site loop otherwise
template <int MAXT, bool BW>
__global__ void __launch_bounds__(MAXT) sweep_delayed_kernel(SweepArgs a, DelayedGeom g)
{
    extern __shared__ __align__(16) double sm[];
    const int n = g.n, ld = g.ld, K = g.K;
    const int cls = g.cls;
    const int chain = blockIdx.x / cls, crank = blockIdx.x % cls, tid = threadIdx.x, T = blockDim.x;
    cg::cluster_group cluster = cg::this_cluster();
    double* unif = sm;                  // [n]
    double* gnext = unif + n;           // [2] G_ii of the next site (double buffered)
    double* Cc = gnext + 16;            // [K][ld]  column j of the block at Cc + j*ld
    double* Rc = Cc + (size_t)K * ld;   // [K][ld]  row k of the block at Rc + k*ld
    double* U = Cc;                     // recorded u of site j replaces column j (zero if rejected)
    int8_t* s_in = (int8_t*)(Rc + (size_t)K * ld);
    int8_t* s_out = s_in + round_up(n, 8);
    double* DD = reinterpret_cast<double*>(s_out + round_up(n, 8));   // [32][DLD] diagonal block of the decisions
    double* xs = DD + 32 * DLD;                                        // [32] x_j, 0 for rejected sites
    double* Dn = xs + 32;                                              // [KMAX][KMAX] diagonal block of the NEXT block before the flush
    int* tile_ctr = reinterpret_cast<int*>(Dn + KMAX * KMAX);          // flush tiles are handed out dynamically
    volatile int* decide_done = tile_ctr + 1;                          // set by the deciding warp
    const bool two = g.two != 0;
    const int fcls = two ? 1 : cls, frank = two ? 0 : crank;   // division of the flush / checks among the CTAs
    double* Gg = ((two && crank) ? a.G2 : a.G) + (size_t)chain * a.sG;
    int8_t* cg = ((two && crank) ? a.conf2 : a.conf) + (size_t)chain * a.sConf;
    double* gpeer = gnext + 2;          // [2] the partner sector's G_ii (written by the partner)
    double* gpeer_remote = two ? cluster.map_shared_rank(gpeer, crank ^ 1) : gpeer;

    for (int i = tid; i < n; i += T) {
        s_in[i] = cg[i];
        unif[i] = dqmc_uniform(a.key0, a.chain_offset + (uint32_t)chain, a.sweep, a.visit, (uint32_t)i);
    }
    int nacc = 0;
    bool bad = false;   // a non-finite entry of G, seen in the last block's flush (which rewrites the whole matrix)
    SWD_T(0);
    // (the host picks BW, see launch_sweep_delayed)
    constexpr bool blockwise = BW;
    const int nv = round_up(n, 32);                 // threads per group (whole warps)
    // groups of nv threads that share the per-site work; more groups shorten the loops but every
    // extra warp repeats the fixed per-site instructions, so two groups at most
    const int ngrp = T / nv >= 2 ? 2 : 1;
    const int warp = tid >> 5, lane = tid & 31, r = lane >> 2, q = lane & 3, nwarps = T >> 5;

    // thread = (element e of a length-n vector, group grp): the groups stride over the remaining
    // columns / rows of the block so that all threads have work
    const int e = tid % nv, grp = tid / nv;
    const bool worker = grp < ngrp && e < n;
    const int gstep = ngrp * ld;
    const DecideConst dconst{a.gamma_p, a.gamma_m, a.efac_p, a.efac_m};
    // Cluster of CTAs on one chain (one spin sector): the decisions stay redundant, but the block's columns and rows are
    // divided - every CTA loads elements [e0, e1) of the K columns and of the K rows, runs the two recurrences there and
    // sends its part of u_j and of the frozen rows to the peers' shared memory as bulk copies (cp.async.bulk
    // shared::cta -> shared::cluster, one per vector, peer and array), each CTA waiting on an mbarrier that counts the
    // incoming bytes. Loads and recurrences shrink by the cluster size; the flush tiles stay interleaved as before.
    // The exchange itself costs ~8 k cycles per block of 24 sites at N = 400 (115 KB out and in per CTA) whichever way it is
    // done - per-thread st.shared::cluster, these bulk copies, or a round trip through an L2 scratch with multicast TMA
    // loads were all measured within 10 % of each other - so the net gain is 8 % of the kernel, not the 25 % the divided
    // work alone would give.
    const bool split = blockwise && cls > 1 && !two;
    const int echunk = (g.np + fcls - 1) / fcls;
    const int e0 = split ? frank * echunk : 0, e1 = split ? min(g.np, e0 + echunk) : g.np, ne = e1 - e0;
    uint64_t* xbar = reinterpret_cast<uint64_t*>(tile_ctr + 2);   // counts the bytes the peers send per block
    uint32_t xphase = 0;
    if (split && tid == 0) mbar_init(xbar, 1);   // (visible to the peers after the first cluster barrier)

    if constexpr (blockwise) {
        // The decisions of block b + 1 run on one warp UNDER the flush of block b (see below); those of the first block
        // have nothing to hide under and read the diagonal block straight from global memory.
        __syncthreads();   // s_in, unif
        if (warp == 0 && !(g.exp & 2)) {
            const int kb0 = min(K, n);
            double reg[KMAX];
#pragma unroll
            for (int b = 0; b < KMAX; ++b) reg[b] = (lane < kb0 && b < kb0) ? Gg[lane + (size_t)b * n] : 0.0;
            nacc += decide_rows<KMAX>(dconst, reg, 0, kb0, s_in, s_out, unif, DD, xs, lane);
        }
    }

    for (int i0 = 0; i0 < n; i0 += K) {
        const int kb = min(K, n - i0);
        const int kb4 = round_up(kb, 4);
        // previous flush done (G up to date in every CTA of the chain, buffers free)
        // (two sectors: also before the first block, the partner must be running before its shared memory is written)
        if (cls > 1 && (i0 > 0 || two || split)) cluster.sync(); else __syncthreads();
        if (split && tid == 0)   // the peers' shares of this block's vectors: everything but elements e0 .. e1, both arrays
            mbar_expect_tx(xbar, (uint32_t)((size_t)round_up(min(K, n - i0), 4) * (g.np - ne) * 2 * sizeof(double)));
        if (i0 == K) SWD_T(1);
        // ---- load the block's columns and rows; zero the padding the flush will read. Flattened over the elements with
        // eight independent loads in flight per thread (a loop over columns per warp serialises the L2 latencies).
        if (split) {
            // this CTA's share of the columns (rows e0 .. e1) and of the rows (columns e0 .. e1) of the block
            const int ncol = kb4 * ne;
            for (int base = 0; base < ncol; base += 8 * T) {
                double v[8];
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int j = idx / ne, row = e0 + idx - j * ne;
                    v[z] = (idx < ncol && j < kb && row < n) ? Gg[(size_t)(i0 + j) * n + row] : 0.0;
                }
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int j = idx / ne, row = e0 + idx - j * ne;
                    if (idx < ncol) Cc[j * ld + row] = v[z];
                }
            }
            for (int base = 0; base < ncol; base += 8 * T) {
                double v[8];
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int l = e0 + idx / kb4, k = idx % kb4;
                    v[z] = (idx < ncol && k < kb && l < n) ? Gg[(size_t)l * n + i0 + k] : 0.0;
                }
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int l = e0 + idx / kb4, k = idx % kb4;
                    if (idx < ncol) Rc[k * ld + l] = v[z];
                }
            }
        } else {
            const int ncol = kb4 * ld;   // Cc: element (row, j) at Cc[j * ld + row], consecutive threads = consecutive rows
            for (int base = 0; base < ncol; base += 8 * T) {
                double v[8];
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int j = idx / ld, row = idx - j * ld;
                    v[z] = (idx < ncol && j < kb && row < n) ? Gg[(size_t)(i0 + j) * n + row] : 0.0;
                }
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    if (idx < ncol) Cc[idx] = v[z];
                }
            }
            // Rc: element (k, l) at Rc[k * ld + l], consecutive threads = consecutive k of one column l (contiguous in G)
            const int nrow = kb4 * ld;
            for (int base = 0; base < nrow; base += 8 * T) {
                double v[8];
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int l = idx / kb4, k = idx - l * kb4;
                    v[z] = (idx < nrow && k < kb && l < n) ? Gg[(size_t)l * n + i0 + k] : 0.0;
                }
#pragma unroll
                for (int z = 0; z < 8; ++z) {
                    const int idx = base + z * T + tid;
                    const int l = idx / kb4, k = idx - l * kb4;
                    if (idx < nrow) Rc[k * ld + l] = v[z];
                }
            }
        }
        const int i1 = i0 + K, kb1 = max(0, min(K, n - i1));   // the next block (kb1 == 0: none)
        if (tid == 0) { *tile_ctr = 0; *decide_done = 0; }
        if constexpr (blockwise) {
            // the next block's diagonal block as it is BEFORE this block's flush (G is quiescent here)
            for (int idx = tid; idx < kb1 * kb1; idx += T) {
                const int ra = idx % kb1, cb = idx / kb1;
                Dn[cb * KMAX + ra] = Gg[(i1 + ra) + (size_t)(i1 + cb) * n];
            }
            // ---- this block's decisions were taken under the previous flush (DD, xs): the vector recurrences
            __syncthreads();
            if (i0 == K) SWD_T(2);
            if (!(g.exp & 2)) {
                if (g.exp & 4) {
                } else if (split) {
                    for (int it = tid; it < 2 * ne; it += T) {
                        if (it < ne) { if (e0 + it < n) column_recurrence(Cc, ld, i0, kb, DD, xs, e0 + it); }
                        else if (e0 + it - ne < n) row_recurrence(Rc, ld, kb, DD, e0 + it - ne);
                    }
                } else if (2 * nv <= T) {
                    if (tid < nv) { if (tid < n) column_recurrence(Cc, ld, i0, kb, DD, xs, tid); }
                    else if (tid < 2 * nv) { if (tid - nv < n) row_recurrence(Rc, ld, kb, DD, tid - nv); }
                } else {
                    for (int ee = tid; ee < n; ee += T) {
                        column_recurrence(Cc, ld, i0, kb, DD, xs, ee);
                        row_recurrence(Rc, ld, kb, DD, ee);
                    }
                }
            }
            if (split) {
                // this CTA's elements of the update vectors (and the zero padding the flush reads) into the peers' copies
                fence_async_smem();
                __syncthreads();
                if (i0 == K) SWD_T(8);
                for (int w = tid; w < (fcls - 1) * 2 * kb4; w += T) {
                    const int p = w / (2 * kb4), rem = w - p * 2 * kb4, arr = rem / kb4, j = rem - arr * kb4;
                    const uint32_t peer = (uint32_t)((frank + 1 + p) % fcls);   // rank in the cluster
                    const double* src = (arr ? Rc : Cc) + j * ld + e0;
                    bulk_s2peer(mapa_u32(smem_u32(src), peer), src, (uint32_t)(ne * sizeof(double)), mapa_u32(smem_u32(xbar), peer));
                }
                if (i0 == K) SWD_T(9);
                mbar_wait(xbar, xphase);
                xphase ^= 1;
            } else {
                __syncthreads();
            }
            if (i0 == K) SWD_T(3);
            // ---- one warp: the next block's diagonal block brought up to date, D' = D - U[blk', :] Rc[:, blk'] (K^3 FMAs
            // instead of waiting for the N x K x N flush), then its decisions. The other warps start the flush at once.
            if (warp == 0 && kb1 > 0 && !(g.exp & 2)) {
                {   // KT x KT DMMA tiles of the K x K x K product, subtracted from Dn in place
                    constexpr int KT = KMAX / 8;
                    const double* ua[KT];
                    const double* rb[KT];
#pragma unroll
                    for (int x = 0; x < KT; ++x) {
                        ua[x] = U + q * ld + min(i1 + 8 * x + r, g.np - 1);
                        rb[x] = Rc + q * ld + min(i1 + 8 * x + r, g.np - 1);
                    }
                    double c[KT][KT][2];
#pragma unroll
                    for (int x = 0; x < KT; ++x)
#pragma unroll
                        for (int y = 0; y < KT; ++y) c[x][y][0] = c[x][y][1] = 0.0;
                    for (int k0 = 0; k0 < kb4; k0 += 4) {
                        double av[KT], bv[KT];
#pragma unroll
                        for (int x = 0; x < KT; ++x) { av[x] = ua[x][k0 * ld]; bv[x] = rb[x][k0 * ld]; }
#pragma unroll
                        for (int x = 0; x < KT; ++x)
#pragma unroll
                            for (int y = 0; y < KT; ++y) dmma884(c[x][y][0], c[x][y][1], av[x], bv[y]);
                    }
#pragma unroll
                    for (int x = 0; x < KT; ++x)
#pragma unroll
                        for (int y = 0; y < KT; ++y)
#pragma unroll
                            for (int cc = 0; cc < 2; ++cc) {
                                const int row = 8 * x + r, col = 8 * y + 2 * q + cc;
                                if (row < kb1 && col < kb1) Dn[col * KMAX + row] -= c[x][y][cc];
                            }
                    __syncwarp();
                }
                double reg[KMAX];
#pragma unroll
                for (int b = 0; b < KMAX; ++b) reg[b] = (lane < kb1 && b < kb1) ? Dn[b * KMAX + lane] : 0.0;
                nacc += decide_rows<KMAX>(dconst, reg, i1, kb1, s_in, s_out, unif, DD, xs, lane);
                if (i0 == K) SWD_T(4);
                __syncwarp();
                if (g.exp & 128) { if (lane == 0) { __threadfence_block(); *decide_done = 1; } }
                else if (!(g.exp & 64)) asm volatile("bar.arrive 1, %0;" ::"r"(T / 4) : "memory");
            } else if (kb1 > 0 && !(g.exp & 2) && (g.exp & 128)) {
                while (!*decide_done) __nanosleep(100);   // experiment: every warp waits for the decisions
            } else if (kb1 > 0 && !(g.exp & 2) && !(g.exp & 64) && (warp & 3) == 0) {
                // The warps that share the deciding warp's scheduler stay out of the flush until it is done: their DMMAs
                // would queue in front of its dependent FP64 chain (decisions 32 k cycles with them, 24 k without, at
                // N = 144). They park on a named barrier (a polling loop takes issue slots from the decisions).
                // Experiment bit 64 lets them in.
                asm volatile("bar.sync 1, %0;" ::"r"(T / 4) : "memory");
            }
        } else {
            // ---- site by site (two spin sectors: the partner's G_ii arrives through distributed shared memory)
            if (tid == 0) {
                const double g0 = Gg[i0 + (size_t)i0 * n];
                gnext[0] = g0;
                if (two) gpeer_remote[0] = g0;
            }
            double u_prev = 0.0;   // u of the previous site, stored once nobody reads its column any more
            for (int j = 0; j < ((g.exp & 2) ? 0 : kb); ++j) {
                if (two) cluster.sync(); else __syncthreads();
                if (tid >= ngrp * nv) continue; // warps without per-site work only keep the barrier count
                if (j > 0 && worker && grp == 0) U[(j - 1) * ld + e] = u_prev;
                u_prev = 0.0;
                const int i = i0 + j;
                const double gii = gnext[j & 1];
                const int s = s_in[i];
                const double gamma = s > 0 ? a.gamma_p : a.gamma_m;
                const double efac = s > 0 ? a.efac_p : a.efac_m;
                const double rt = 1.0 + gamma * (1.0 - gii);
                double p = efac * (rt * rt);
                if (two) { // the partner sits on the negated field
                    const double rp = 1.0 + (s > 0 ? a.gamma_m : a.gamma_p) * (1.0 - gpeer[j & 1]);
                    p = rt * rp;
                }
                const bool accept = (p > 1.0 || unif[i] < p); // uniform over the CTA
                const double* cj = Cc + j * ld;
                const double* vj = Rc + j * ld;
                if (accept) {
                    const double x = gamma * fast_rcp(rt);
                    if (worker) {
                        // columns j' > j of the block: element (row e, column j')
                        const double u = x * ((e == i ? 1.0 : 0.0) - cj[e]);
                        u_prev = u;
                        const int first = j + 1 + grp;
                        double* cp = Cc + first * ld + e;
                        const double* vp = vj + i0 + first;
                        if (grp == 0 && e == i + 1 && first < kb) {
                            const double cnew = *cp - u * *vp;
                            gnext[(j + 1) & 1] = cnew;
                            if (two) gpeer_remote[(j + 1) & 1] = cnew;
                        }
                        // (operands are fetched four iterations at a time: the compiler cannot prove that
                        // the stores do not alias the next loads)
                        const double xv = x * vj[e];
                        double* rp = Rc + first * ld + e;
                        const double* up = cj + i0 + first;
                        for (int jp = first; jp < kb; jp += 4 * ngrp) {
                            double c[4], v[4], rr[4], uu[4];
    #pragma unroll
                            for (int z = 0; z < 4; ++z) {
                                const bool ok = jp + z * ngrp < kb;
                                c[z] = ok ? cp[z * gstep] : 0.0;
                                v[z] = ok ? vp[z * ngrp] : 0.0;
                                rr[z] = ok ? rp[z * gstep] : 0.0;
                                uu[z] = ok ? up[z * ngrp] : 0.0;
                            }
    #pragma unroll
                            for (int z = 0; z < 4; ++z) {
                                if (jp + z * ngrp < kb) {
                                    cp[z * gstep] = c[z] - u * v[z];    // columns j' > j: element (row e, column j')
                                    rp[z * gstep] = rr[z] + uu[z] * xv; // rows k > j: element (row k, column e), u_k = -x G[i0+k, i]
                                }
                            }
                            cp += 4 * gstep; vp += 4 * ngrp; rp += 4 * gstep; up += 4 * ngrp;
                        }
                    }
                    ++nacc;
                } else if (j + 1 < kb) {
                    if (tid == 0) {
                        const double gn = Cc[(j + 1) * ld + i + 1];
                        gnext[(j + 1) & 1] = gn;
                        if (two) gpeer_remote[(j + 1) & 1] = gn;
                    }
                }
                if (tid == 0) s_out[i] = (int8_t)(accept ? -s : s);
            }
            __syncthreads();
            if (!(g.exp & 2) && worker && grp == 0) U[(kb - 1) * ld + e] = u_prev;
            __syncthreads();
        }
        // ---- flush: G -= U * Rc on the FP64 tensor pipe. Warp tile 32 x 16 (4 x 2 DMMA tiles), handed out by a counter:
        // the deciding warp joins when it is done.
        const int tm = (g.np + 31) / 32, tn = (g.np + 15) / 16;
        const int ntile = (g.exp & 1) ? 0 : tm * tn;
        const bool last_block = i0 + K >= n;
        for (;;) {
            int tnext = 0;
            if (lane == 0) tnext = atomicAdd(tile_ctr, 1);
            tnext = __shfl_sync(0xffffffffu, tnext, 0);
            const int tile = tnext * fcls + frank;
            if (tile >= ntile) break;
            const int row0 = (tile % tm) * 32, col0 = (tile / tm) * 16;
            // rows / columns past the padded size are clamped: their products are never stored
            const double* ua[4];
            const double* rb[2];
#pragma unroll
            for (int x = 0; x < 4; ++x) ua[x] = U + q * ld + min(row0 + 8 * x + r, g.np - 1);
#pragma unroll
            for (int y = 0; y < 2; ++y) rb[y] = Rc + q * ld + min(col0 + 8 * y + r, g.np - 1);
            double c[4][2][2];
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 2; ++y) c[x][y][0] = c[x][y][1] = 0.0;
            // read-modify-write of the tile. Where the register budget allows (512 threads: 128 registers each) the old
            // values are requested BEFORE the product, so that their L2 latency hides under the DMMAs; the 1024-thread
            // variant (64 registers) loads them afterwards.
            constexpr bool PREFETCH = MAXT <= 512;
            double gold[4][2][2];
            if ((n & 1) == 0) {
                // Even n: the product is formed TRANSPOSED (the fragments of Rc as the A operand, those of U as B), so that
                // the two accumulator elements of a thread are G(row, col), G(row + 1, col): adjacent in the column-major G,
                // one 16-byte load and store instead of two of 8 bytes. c[x][y][cc] <-> row0 + 8x + 2q + cc, col0 + 8y + r.
                auto load_old2 = [&]() {
#pragma unroll
                    for (int x = 0; x < 4; ++x)
#pragma unroll
                        for (int y = 0; y < 2; ++y) {
                            const int row = row0 + 8 * x + 2 * q, col = col0 + 8 * y + r;
                            double2 v = make_double2(0.0, 0.0);
                            if (row < n && col < n) v = __ldcg(reinterpret_cast<const double2*>(Gg + row + (size_t)col * n));
                            gold[x][y][0] = v.x; gold[x][y][1] = v.y;
                        }
                };
                if (PREFETCH) load_old2();
                for (int k0 = 0; k0 < kb4; k0 += 4) {
                    double av[4], bv[2];
#pragma unroll
                    for (int x = 0; x < 4; ++x) av[x] = ua[x][k0 * ld];
#pragma unroll
                    for (int y = 0; y < 2; ++y) bv[y] = rb[y][k0 * ld];
#pragma unroll
                    for (int x = 0; x < 4; ++x)
#pragma unroll
                        for (int y = 0; y < 2; ++y) dmma884(c[x][y][0], c[x][y][1], bv[y], av[x]);
                }
                if (!PREFETCH) load_old2();
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y) {
                        const int row = row0 + 8 * x + 2 * q, col = col0 + 8 * y + r;
                        const double2 gnew = make_double2(gold[x][y][0] - c[x][y][0], gold[x][y][1] - c[x][y][1]);
                        if (row < n && col < n) {
                            *reinterpret_cast<double2*>(Gg + row + (size_t)col * n) = gnew;
                            // exponent field all ones = Inf / NaN (integer test: the FP64 pipe is the busy one here)
                            if (last_block)
                                bad |= ((__double2hiint(gnew.x) & 0x7ff00000) == 0x7ff00000) |
                                       ((__double2hiint(gnew.y) & 0x7ff00000) == 0x7ff00000);
                        }
                    }
                continue;
            }
            auto load_old = [&]() {
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y)
#pragma unroll
                        for (int cc = 0; cc < 2; ++cc) {
                            const int row = row0 + 8 * x + r, col = col0 + 8 * y + 2 * q + cc;
                            gold[x][y][cc] = (row < n && col < n) ? __ldcg(Gg + row + (size_t)col * n) : 0.0;
                        }
            };
            if (PREFETCH) load_old();
            for (int k0 = 0; k0 < kb4; k0 += 4) {
                double av[4], bv[2];
#pragma unroll
                for (int x = 0; x < 4; ++x) av[x] = ua[x][k0 * ld];
#pragma unroll
                for (int y = 0; y < 2; ++y) bv[y] = rb[y][k0 * ld];
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y) dmma884(c[x][y][0], c[x][y][1], av[x], bv[y]);
            }
            if (!PREFETCH) load_old();
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 2; ++y)
#pragma unroll
                    for (int cc = 0; cc < 2; ++cc) {
                        const int row = row0 + 8 * x + r, col = col0 + 8 * y + 2 * q + cc;
                        const double gnew = gold[x][y][cc] - c[x][y][cc];
                        if (row < n && col < n) {
                            Gg[row + (size_t)col * n] = gnew;
                            // exponent field all ones = Inf / NaN (integer test: the FP64 pipe is the busy one here)
                            if (last_block) bad |= (__double2hiint(gnew) & 0x7ff00000) == 0x7ff00000;
                        }
                    }
        }
        if (i0 == K) { __syncthreads(); SWD_T(5); }
    }
    (void)nwarps;
    if (cls > 1) cluster.sync(); else __syncthreads();
    SWD_T(6);
    // every CTA of the chain has checked its share of G while storing it (NaN / Inf survive the later blocks' updates, so
    // the last flush sees them all; the timing experiments that skip the flush skip the check); rank 0 owns the outputs
    if (crank == 0 || two)
        for (int i = tid; i < n; i += T) cg[i] = s_out[i];
    if (__syncthreads_or(bad) && tid == 0) atomicAdd((unsigned long long*)&a.nonfinite[chain], 1ull);
    if (crank == 0 && tid == 0) { a.acc[chain] += nacc; a.prop[chain] += n; }
    SWD_T(7);
}

} // namespace

static int g_sms = 0;

int init_sweep_delayed_kernels()
{
    int dev = 0;
    MCB_CUDA(cudaGetDevice(&dev));
    MCB_CUDA(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
    MCB_CUDA(cudaFuncSetAttribute(sweep_delayed_kernel<512, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    MCB_CUDA(cudaFuncSetAttribute(sweep_delayed_kernel<1024, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    MCB_CUDA(cudaFuncSetAttribute(sweep_delayed_kernel<512, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    MCB_CUDA(cudaFuncSetAttribute(sweep_delayed_kernel<1024, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    return 0;
}

int launch_sweep_delayed(const SweepArgs& a, cudaStream_t st)
{
    DelayedGeom g = delayed_geom(a.n);
    if (g.K < 4) { set_error("delayed sweep: matrix too large for one CTA's shared memory"); return -1; }
    // CTAs per chain: as many as keep the whole batch in one wave of one CTA per SM (clusters of 4 leave a few
    // SMs of each GPC unused, hence the margin); only where the flush is large enough to be worth dividing
    static const int force_cls = getenv("MCB_SWEEP_CLS") ? atoi(getenv("MCB_SWEEP_CLS")) : 0;
    int cls = 1;
    if (a.n >= 128) {
        if (a.batch * 4 <= g_sms - 16) cls = 4;
        else if (a.batch * 2 <= g_sms - 4) cls = 2;
    }
    if (force_cls == 1 || force_cls == 2 || force_cls == 4) cls = force_cls;
    g.two = a.G2 != nullptr ? 1 : 0;
    if (g.two) {
        if (!a.conf2) { set_error("delayed sweep: second sector without its field"); return -1; }
        cls = 2;
    }
    g.cls = cls;
    LaunchScope scope(LC_SWEEP, st, 0.0);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.batch * cls, 1, 1);
    cfg.blockDim = dim3(g.threads, 1, 1);
    cfg.dynamicSmemBytes = g.smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cls; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    // block-wise decisions for one spin sector and N > 64 (measured per slice: N = 144 182 against 203 us, N = 400 794
    // against 1092 us); MCB_SWEEP_EXP bit 16 forces the site-by-site loop, bit 32 the block-wise scheme for every size
    const bool bw = !g.two && !(g.exp & 16) && (g.threads >= 512 || (g.exp & 32));
    if (g.threads <= 512) {
        if (bw) MCB_CUDA(cudaLaunchKernelEx(&cfg, sweep_delayed_kernel<512, true>, a, g));
        else MCB_CUDA(cudaLaunchKernelEx(&cfg, sweep_delayed_kernel<512, false>, a, g));
    } else {
        if (bw) MCB_CUDA(cudaLaunchKernelEx(&cfg, sweep_delayed_kernel<1024, true>, a, g));
        else MCB_CUDA(cudaLaunchKernelEx(&cfg, sweep_delayed_kernel<1024, false>, a, g));
    }
#ifdef SWD_TIMING
    {
        static int shown = 0;
        if (shown++ < 3) {
            long long h[32];
            cudaStreamSynchronize(st);
            cudaMemcpyFromSymbol(h, swd_clk, sizeof(h));
            fprintf(stderr, "sweep_delayed n=%d K=%d cls=%d cycles: whole loop %lld | tail %lld || block 1: load %lld recurrences %lld (of which push %lld, cluster barrier %lld) next diagonal + decide %lld (under the flush) flush %lld\n",
                    a.n, g.K, g.cls, h[6] - h[0], h[7] - h[6], h[2] - h[1], h[3] - h[2], h[9] - h[8], h[3] - h[9], h[4] - h[3], h[5] - h[3]);
        }
    }
#endif
    return 0;
}

} // namespace mcb
