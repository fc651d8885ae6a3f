// Chain-resident wrap_greens! + sweep_spatial for N = 144 (12 x 12 lattices, BASELINE config C): one CTA of 12 warps per
// chain keeps the whole Green's function (144 x 144 doubles, padded to 148 rows: 170 KB) in shared memory for a run of
// up to safe_mult consecutive time slices, as segment_kernel (segment.cu) does in registers for N <= 64.
//   wrap_greens!                 src/flavors/DQMC/stack.jl:491-500, slice_matrices.jl:23-76
//   sweep_spatial / propose_local / accept_local!   DQMC.jl:546-582, HubbardModelAttractive.jl:113-155
//
// Wrap: the two products of a wrap run on 12 warps (6 x 2 warp tiles of 24 x 72, the whole 144 x 144 result in DMMA
// accumulators), one operand read as fragments straight out of the resident G, the other - the hopping exponential, the
// same matrix for every chain and therefore L2-resident - streamed in panels of 8 columns by TMA bulk copies into a ring
// of 4 stages guarded by full (byte-counting) and empty (one arrival per warp) mbarriers; no CTA barrier inside a product. The
// exp(+-lambda s) diagonals are applied to the A fragments / at the write-back. The second product needs rows of the
// exponential; the host uploads its transpose so that these are contiguous columns too.
//
// Sweep: delayed rank-16 updates as in sweep_delayed.cu (decisions of block b + 1 by one warp under the flush of block
// b, triangular recurrences for the update vectors, flush G -= U Rc on the FP64 tensor pipe), except that the flush is a
// read-modify-write of shared memory: the delayed-update kernel with G in L2 moves 2 x 166 KB per block and chain through
// L2 (12 x per slice, 255 MB per slice for 128 chains - it is L2-bandwidth-bound there), this kernel moves G once per run.
#include "segment.cuh"
#include "sweep_blocks.cuh"
#include "tma.cuh"
#include "prof.cuh"
#include <stdlib.h>

namespace mcb {

namespace {

using namespace swb;

#ifdef SWD_TIMING
__device__ long long sgr_clk[32];
#define SGR_T(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) sgr_clk[i] = clock64(); } while (0)
#else
#define SGR_T(i) do { } while (0)
#endif

template <int N>
struct RGeo {
    static constexpr int LD = N + 4;          // column stride of G, of the block's columns / rows and of the ring (% 16 == 4)
    static constexpr int K = 16;              // sites per block
    static constexpr int THREADS = 384;       // 12 warps, three per scheduler: 168 registers each (a wrap product holds 108 in accumulators)
    static constexpr int BK = 8, STAGES = 4;  // ring of hopping-exponential panels; lives where Cc / Rc live during the sweep
    static constexpr int WM = N / 6, WN = N / 2, IT = WM / 8, JT = WN / 8;
    static_assert(LD % 16 == 4 && N % 48 == 0 && N % BK == 0, "unsupported size");
    static_assert(2 * K * LD == STAGES * BK * LD, "the ring is the Cc / Rc area");
    // shared memory carve-up (doubles)
    static constexpr int OFF_G = 0;
    static constexpr int OFF_CC = OFF_G + N * LD;
    static constexpr int OFF_RC = OFF_CC + K * LD;
    static constexpr int OFF_DD = OFF_RC + K * LD;
    static constexpr int OFF_XS = OFF_DD + 32 * DLD;
    static constexpr int OFF_DN = OFF_XS + 32;
    static constexpr int OFF_UNIF = OFF_DN + K * K;
    static constexpr int OFF_SC = OFF_UNIF + N;          // mid1, row1, mid2, col2
    static constexpr int OFF_BAR = OFF_SC + 4 * N;       // full[STAGES], empty[STAGES], gbar
    static constexpr int OFF_MISC = OFF_BAR + 2 * STAGES + 2;   // tile counter, decide flag
    static constexpr int OFF_S = OFF_MISC + 2;           // s_in, s_out (int8)
    static constexpr size_t SMEM = (size_t)OFF_S * sizeof(double) + 2 * ((N + 15) / 16 * 16);
    static_assert(SMEM <= 227 * 1024, "does not fit one SM");
};

// column_recurrence / row_recurrence (sweep_blocks.cuh) with the block's original columns / rows read from G itself
// (column j of the block at Gcol + j * ldg, row k at Grow + k with column stride ldg) instead of from a copy
__device__ __forceinline__ void column_recurrence_g(const double* Gcol, int ldg, double* U, int ld, int i0, int kb,
                                                    const double* DD, const double* xs, int e)
{
    for (int j0 = 0; j0 < kb; j0 += 8) {
        double c[8], u[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) c[q] = (j0 + q < kb) ? Gcol[(size_t)(j0 + q) * ldg + e] : 0.0;
        for (int jp = 0; jp < j0; ++jp) {
            const double up = U[jp * ld + e];
            const double* d = DD + jp * DLD + j0;
#pragma unroll
            for (int q = 0; q < 8; ++q) c[q] = fma(-up, d[q], c[q]);
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            double cc = c[q];
#pragma unroll
            for (int p = 0; p < q; ++p) cc = fma(-u[p], DD[(j0 + p) * DLD + j0 + q], cc);
            u[q] = (j0 + q < kb) ? xs[j0 + q] * ((e == i0 + j0 + q ? 1.0 : 0.0) - cc) : 0.0;
            if (j0 + q < kb) U[(j0 + q) * ld + e] = u[q];
        }
    }
}

__device__ __forceinline__ void row_recurrence_g(const double* Grow, int ldg, double* Rc, int ld, int kb, const double* DD, int e)
{
    for (int j0 = 0; j0 < kb; j0 += 8) {
        double r[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) r[q] = (j0 + q < kb) ? Grow[(size_t)e * ldg + j0 + q] : 0.0;
        for (int jp = 0; jp < j0; ++jp) {
            const double rp = Rc[jp * ld + e];
            const double* d = DD + j0 * DLD + jp;
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

// One product of a wrap. WHICH == 1: G <- diag(row) H (diag(mid) G)   (A = columns of H from the ring, B = G)
//                        WHICH == 2: G <- ((G diag(mid)) H) diag(col)   (A = G, B = rows of H = columns of H' from the ring)
// Hcols points at the matrix whose COLUMNS are streamed (H for 1, H' for 2). cnt counts the ring stages used so far (the
// same sequence in every thread): stage = cnt % STAGES, the phase of its barrier follows from cnt / STAGES.
template <int N, int WHICH>
__device__ __forceinline__ void wrap_product(double* __restrict__ Gs, double* __restrict__ ring, const double* __restrict__ Hcols,
                                             const double* __restrict__ mid, const double* __restrict__ outsc, uint64_t* full,
                                             uint64_t* empty, uint32_t& cnt, int tid)
{
    using R = RGeo<N>;
    constexpr int LD = R::LD, BK = R::BK, STAGES = R::STAGES, IT = R::IT, JT = R::JT, nk = N / BK;
    const int warp = tid >> 5, lane = tid & 31, r = lane >> 2, q = lane & 3;
    double acc[IT][JT][2];
    const int wm = (warp % 6) * R::WM, wn = (warp / 6) * R::WN;
    // one bulk copy per column of a panel. The first panels go out from warp 0; stage s is refilled for panel kt + STAGES by
    // warp kt % 12 once all twelve warps have released it (arrivals on empty[s]): no CTA-wide barrier inside the product.
    auto issue = [&](int kt) {
        const uint32_t s = (cnt + kt) % STAGES;
        if (lane == 0) mbar_expect_tx(full + s, BK * N * sizeof(double));
        __syncwarp();
        if (lane < BK)
            bulk_g2s(ring + (size_t)(s * BK + lane) * LD, Hcols + (size_t)(kt * BK + lane) * N, N * sizeof(double), full + s);
    };
    if (warp == 0)
        for (int kt = 0; kt < STAGES; ++kt) issue(kt);
#pragma unroll
    for (int i = 0; i < IT; ++i)
#pragma unroll
        for (int j = 0; j < JT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    for (int kt = 0; kt < nk; ++kt) {
        const uint32_t c = cnt + kt, s = c % STAGES;
        mbar_wait(full + s, (c / STAGES) & 1);
        const double* hs = ring + (size_t)s * BK * LD;
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            const int k = kt * BK + kk + q;
            const double mv = mid[k];
            double av[IT], bv[JT];
#pragma unroll
            for (int i = 0; i < IT; ++i)
                av[i] = (WHICH == 1 ? hs[(kk + q) * LD + wm + 8 * i + r] : Gs[(size_t)k * LD + wm + 8 * i + r]) * mv;
#pragma unroll
            for (int j = 0; j < JT; ++j)
                bv[j] = WHICH == 1 ? Gs[(size_t)(wn + 8 * j + r) * LD + k] : hs[(kk + q) * LD + wn + 8 * j + r];
#pragma unroll
            for (int i = 0; i < IT; ++i)
#pragma unroll
                for (int j = 0; j < JT; ++j) dmma884(acc[i][j][0], acc[i][j][1], av[i], bv[j]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + s);
        if (kt + STAGES < nk && warp == kt % 12) {
            mbar_wait(empty + s, (c / STAGES) & 1);
            issue(kt + STAGES);
        }
    }
    cnt += nk;
    __syncthreads();   // every fragment of the old G has been read
    {
#pragma unroll
        for (int i = 0; i < IT; ++i) {
            const int gm = wm + 8 * i + r;
            const double rs = WHICH == 1 ? outsc[gm] : 1.0;
#pragma unroll
            for (int j = 0; j < JT; ++j)
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int gn = wn + 8 * j + 2 * q + cc;
                    Gs[(size_t)gn * LD + gm] = acc[i][j][cc] * (WHICH == 1 ? rs : outsc[gn]);
                }
        }
    }
    __syncthreads();
}

template <int N>
__global__ void __launch_bounds__(RGeo<N>::THREADS, 1) segment_resident_kernel(SegmentArgs a)
{
    using R = RGeo<N>;
    constexpr int LD = R::LD, K = R::K, T = R::THREADS;
    extern __shared__ __align__(128) double sm[];
    double* Gs = sm + R::OFF_G;
    double* Cc = sm + R::OFF_CC;       // [K][LD] columns of the block, then the recorded u_j
    double* Rc = sm + R::OFF_RC;       // [K][LD] rows of the block, then the frozen rows
    double* ring = Cc;                 // [STAGES][BK][LD] during a wrap
    double* DD = sm + R::OFF_DD;
    double* xs = sm + R::OFF_XS;
    double* Dn = sm + R::OFF_DN;       // [K][K] diagonal block of the next block of sites
    double* unif = sm + R::OFF_UNIF;
    double* sc_mid1 = sm + R::OFF_SC;
    double* sc_row1 = sc_mid1 + N;
    double* sc_mid2 = sc_row1 + N;
    double* sc_col2 = sc_mid2 + N;
    uint64_t* full = reinterpret_cast<uint64_t*>(sm + R::OFF_BAR);
    uint64_t* empty = full + R::STAGES;
    uint64_t* gbar = empty + R::STAGES;
    int* tile_ctr = reinterpret_cast<int*>(sm + R::OFF_MISC);
    int8_t* s_in = reinterpret_cast<int8_t*>(sm + R::OFF_S);
    int8_t* s_out = s_in + (N + 15) / 16 * 16;

    const int chain = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, r = lane >> 2, q = lane & 3;
    double* Gg = a.G + (size_t)chain * a.sG;
    int8_t* cg = a.conf + (size_t)chain * a.sConf;
    const uint32_t k1 = a.chain_offset + (uint32_t)chain;
    const DecideConst dconst{a.gamma_p, a.gamma_m, a.efac_p, a.efac_m};

    if (tid == 0) {
        for (int s = 0; s < R::STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 12); }
        mbar_init(gbar, 1);
    }
    __syncthreads();
    // ---- G -> shared memory, one bulk copy per column
    if (tid == 0) mbar_expect_tx(gbar, (uint32_t)(N * N * sizeof(double)));
    if (tid < N) bulk_g2s(Gs + (size_t)tid * LD, Gg + (size_t)tid * N, N * sizeof(double), gbar);
    mbar_wait(gbar, 0);
    SGR_T(0);

    int nacc = 0;
    bool bad = false;
    uint32_t cnt = 0;   // ring stages used so far
    for (int v = 0; v < a.n_visits; ++v) {
        const int slice = a.start_slice + a.direction * v;
        const bool wrapping = v > 0 || a.first_wrap;
        {
            const long long upd = a.n_updates0 + v;
            const uint32_t sweep = (uint32_t)(upd / (2LL * a.n_slices));
            const uint32_t visit = (uint32_t)(upd % (2LL * a.n_slices));
            for (int i = tid; i < N; i += T) {
                s_in[i] = cg[(size_t)(slice - 1) * N + i];
                unif[i] = dqmc_uniform(a.key0, k1, sweep, visit, (uint32_t)i);
            }
        }
        if (wrapping) {
            // ---- wrap_greens!: up   G <- e^{-dtau T} diag(e^{+lambda s}) G diag(e^{-lambda s}) e^{+dtau T}   (slice matrix B_{slice-1})
            //                    down G <- diag(e^{-lambda s}) e^{+dtau T} G e^{-dtau T} diag(e^{+lambda s})   (slice matrix B_slice)
            const int l = a.direction > 0 ? slice - 1 : slice;
            for (int i = tid; i < N; i += T) {
                const int s = cg[(size_t)(l - 1) * N + i];
                const double ep = s > 0 ? a.eVp[0] : a.eVp[1], em = s > 0 ? a.eVm[0] : a.eVm[1];
                sc_mid1[i] = a.direction > 0 ? ep : 1.0;
                sc_row1[i] = a.direction > 0 ? 1.0 : em;
                sc_mid2[i] = a.direction > 0 ? em : 1.0;
                sc_col2[i] = a.direction > 0 ? 1.0 : ep;
            }
            fence_async_smem();   // the ring area was written through the generic proxy by the previous sweep
            __syncthreads();
            if (v == 1) SGR_T(1);
#ifdef SWD_TIMING
            if (v == 2) SGR_T(25);
            if (blockIdx.x == 0 && tid == 32 && v == 2) sgr_clk[26] = clock64();
#endif
            wrap_product<N, 1>(Gs, ring, a.direction > 0 ? a.eT2 : a.eTinv2, sc_mid1, sc_row1, full, empty, cnt, tid);
            if (v == 1) SGR_T(2);
            wrap_product<N, 2>(Gs, ring, a.direction > 0 ? a.eTinv2T : a.eT2T, sc_mid2, sc_col2, full, empty, cnt, tid);
            if (v == 1) SGR_T(3);
        } else {
            __syncthreads();   // s_in, unif
        }

        // ---- sweep_spatial of `slice`. The decisions of the first block have nothing to hide under.
        if (warp == 0) {
            double reg[K];
#pragma unroll
            for (int b = 0; b < K; ++b) reg[b] = lane < K ? Gs[(size_t)b * LD + lane] : 0.0;
            nacc += decide_rows<K>(dconst, reg, 0, K, s_in, s_out, unif, DD, xs, lane);
            if (v == 1) SGR_T(4);
        }
        const bool last_visit = v == a.n_visits - 1;
        for (int i0 = 0; i0 < N; i0 += K) {
            const int i1 = i0 + K;
            const bool has_next = i1 < N;
            __syncthreads();   // the previous flush is complete, the decisions of this block are in DD / xs
            if (v == 1 && i0 == K) SGR_T(5);
            if (v == 1) SGR_T(12 + i0 / K);
            if (v == 1 && i0 == 2 * K) SGR_T(9);
            // ---- the update vectors of the block: two triangular recurrences, one element per thread, reading the block's
            // columns and rows straight out of G (no copies) and writing u_j / the frozen rows into Cc / Rc. The threads
            // left over fetch the next block's diagonal block as it is BEFORE this flush.
            {
                constexpr int NV = (N + 31) / 32 * 32;
                if (tid < NV) { if (tid < N) column_recurrence_g(Gs + (size_t)i0 * LD, LD, Cc, LD, i0, K, DD, xs, tid); }
                else if (tid < 2 * NV) { if (tid - NV < N) row_recurrence_g(Gs + i0, LD, Rc, LD, K, DD, tid - NV); }
                else {
                    if (tid == 2 * NV) *tile_ctr = 0;
                    if (has_next)
                    for (int idx = tid - 2 * NV; idx < K * K; idx += T - 2 * NV) {
                        const int ra = idx % K, cb = idx / K;
                        Dn[cb * K + ra] = Gs[(size_t)(i1 + cb) * LD + i1 + ra];
                    }
                }
            }
            __syncthreads();
            if (v == 1 && i0 == K) SGR_T(7);
            if (v == 1 && i0 == 8 * K) SGR_T(21);
            if (v == 1 && i0 == 7 * K) SGR_T(22);
            // ---- one warp: next diagonal block D' = D - U[blk', :] Rc[:, blk'] (2 x 2 DMMA tiles), then its decisions
            if (warp == 0 && has_next) {
                double c[2][2][2];
#pragma unroll
                for (int x = 0; x < 2; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y) c[x][y][0] = c[x][y][1] = 0.0;
#pragma unroll
                for (int k0 = 0; k0 < K; k0 += 4) {
                    double av[2], bv[2];
#pragma unroll
                    for (int x = 0; x < 2; ++x) {
                        av[x] = Cc[(k0 + q) * LD + i1 + 8 * x + r];
                        bv[x] = Rc[(k0 + q) * LD + i1 + 8 * x + r];
                    }
#pragma unroll
                    for (int x = 0; x < 2; ++x)
#pragma unroll
                        for (int y = 0; y < 2; ++y) dmma884(c[x][y][0], c[x][y][1], av[x], bv[y]);
                }
#pragma unroll
                for (int x = 0; x < 2; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y)
#pragma unroll
                        for (int cc = 0; cc < 2; ++cc) Dn[(8 * y + 2 * q + cc) * K + 8 * x + r] -= c[x][y][cc];
                __syncwarp();
                double reg[K];
#pragma unroll
                for (int b = 0; b < K; ++b) reg[b] = lane < K ? Dn[b * K + lane] : 0.0;
                nacc += decide_rows<K>(dconst, reg, i1, K, s_in, s_out, unif, DD, xs, lane);
                if (v == 1 && i0 == K) SGR_T(8);
                // the two warps on this scheduler were parked on a named barrier (no polling: a spinning warp takes issue
                // slots from the decisions); let them into the flush
                asm volatile("bar.arrive 1, 96;" ::: "memory");
            }
            // ---- flush G -= U Rc in shared memory: 18 work items of 16 rows x 72 columns (2 x 9 DMMA tiles whose accumulators
            // start as the G tile; the A operand is -U), handed out by a counter. The warps that share the deciding warp's
            // scheduler join when it is done (their DMMAs would queue in front of its dependent FP64 chain, sweep_delayed.cu).
            if (has_next && (warp & 3) == 0 && warp != 0) asm volatile("bar.sync 1, 96;" ::: "memory");
            {
                const bool check = last_visit && !has_next;
                constexpr int FR = 1;                  // row tiles (of 8) per work item
                constexpr int FI = 2 * N / (8 * FR);   // work items
                for (;;) {
                    int item = 0;
                    if (lane == 0) item = atomicAdd(tile_ctr, 1);
                    item = __shfl_sync(0xffffffffu, item, 0);
                    if (item >= FI) break;
                    const int row0 = (item >> 1) * (8 * FR), col0 = (item & 1) * (N / 2);
                    // The product is formed transposed (fragments of Rc as the A operand, of -U as B): a thread's two
                    // accumulator elements are then G(row, col), G(row + 1, col), adjacent in the column-major resident G:
                    // 16-byte shared-memory loads and stores. c[x][y][cc] <-> row0 + 8x + 2q + cc, col0 + 8y + r.
                    double c[FR][9][2];
#pragma unroll
                    for (int x = 0; x < FR; ++x)
#pragma unroll
                        for (int y = 0; y < 9; ++y) {
                            const double2 v = *reinterpret_cast<const double2*>(Gs + (size_t)(col0 + 8 * y + r) * LD + row0 + 8 * x + 2 * q);
                            c[x][y][0] = v.x; c[x][y][1] = v.y;
                        }
#pragma unroll
                    for (int k0 = 0; k0 < K; k0 += 4) {
                        double av[FR], bv[9];
#pragma unroll
                        for (int x = 0; x < FR; ++x) av[x] = -Cc[(k0 + q) * LD + row0 + 8 * x + r];
#pragma unroll
                        for (int y = 0; y < 9; ++y) bv[y] = Rc[(k0 + q) * LD + col0 + 8 * y + r];
#pragma unroll
                        for (int x = 0; x < FR; ++x)
#pragma unroll
                            for (int y = 0; y < 9; ++y) dmma884(c[x][y][0], c[x][y][1], bv[y], av[x]);
                    }
#pragma unroll
                    for (int x = 0; x < FR; ++x)
#pragma unroll
                        for (int y = 0; y < 9; ++y) {
                            *reinterpret_cast<double2*>(Gs + (size_t)(col0 + 8 * y + r) * LD + row0 + 8 * x + 2 * q) =
                                make_double2(c[x][y][0], c[x][y][1]);
                            if (check)
                                bad |= ((__double2hiint(c[x][y][0]) & 0x7ff00000) == 0x7ff00000) |
                                       ((__double2hiint(c[x][y][1]) & 0x7ff00000) == 0x7ff00000);
                        }
                }
#ifdef SWD_TIMING
                if (blockIdx.x == 0 && tid == 32 && v == 1 && i0 == 7 * K) sgr_clk[23] = clock64();
                if (blockIdx.x == 0 && tid == 32 && v == 1 && i0 == 8 * K) sgr_clk[24] = clock64();
#endif
            }
        }
        __syncthreads();   // last flush done; s_out complete
        if (v == 1) SGR_T(10);
#ifdef SWD_TIMING
        if (blockIdx.x == 0 && tid == 32 && v == 1) sgr_clk[27] = clock64();
#endif
        for (int i = tid; i < N; i += T) cg[(size_t)(slice - 1) * N + i] = s_out[i];
    }
    // ---- G back to global memory
    fence_async_smem();
    const int anybad = __syncthreads_or(bad);
    if (tid < N) {
        bulk_s2g(Gg + (size_t)tid * N, Gs + (size_t)tid * LD, N * sizeof(double));
        bulk_commit();
        bulk_wait_all();
    }
    SGR_T(11);
    if (tid == 0) {
        a.acc[chain] += nacc;
        a.prop[chain] += (long long)N * a.n_visits;
        if (anybad) a.nonfinite[chain] += 1;
    }
}

// The slice-matrix chains of add_slice_sequence_left / _right (src/flavors/DQMC/stack.jl:242-311) with the running
// product resident in shared memory: dst = (B_hi ... B_lo Usrc) diag(Dsrc) (left) or (B_lo' ... B_hi' Usrc) diag(Dsrc)
// (right), B_l = e^{-dtau T} diag(e^{+lambda s_l}) (slice_matrices.jl:42-48, 70-76). One launch replaces hi - lo + 1
// batched GEMM launches and their round trips of the product through L2.
template <int N>
__global__ void __launch_bounds__(RGeo<N>::THREADS, 1) slice_chain_resident_kernel(SliceChainArgs a)
{
    using R = RGeo<N>;
    constexpr int LD = R::LD, T = R::THREADS;
    extern __shared__ __align__(128) double sm[];
    double* Gs = sm + R::OFF_G;
    double* ring = sm + R::OFF_CC;
    double* sc_mid = sm + R::OFF_SC;
    double* sc_row = sc_mid + N;
    double* dcol = sc_row + N;
    uint64_t* full = reinterpret_cast<uint64_t*>(sm + R::OFF_BAR);
    uint64_t* empty = full + R::STAGES;
    uint64_t* gbar = empty + R::STAGES;
    const int chain = blockIdx.x, tid = threadIdx.x;
    const double* src = a.Usrc + (size_t)chain * N * N;
    double* dst = a.dst + (size_t)chain * N * N;
    const int8_t* cg = a.conf + (size_t)chain * a.sConf;

    if (tid == 0) {
        for (int s = 0; s < R::STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 12); }
        mbar_init(gbar, 1);
    }
    __syncthreads();
    if (tid == 0) mbar_expect_tx(gbar, (uint32_t)(N * N * sizeof(double)));
    if (tid < N) bulk_g2s(Gs + (size_t)tid * LD, src + (size_t)tid * N, N * sizeof(double), gbar);
    for (int i = tid; i < N; i += T) dcol[i] = a.Dsrc ? a.Dsrc[(size_t)chain * N + i] : 1.0;
    mbar_wait(gbar, 0);

    uint32_t cnt = 0;
    const int count = a.hi - a.lo + 1;
    for (int t = 0; t < count; ++t) {
        const int slice = a.left ? a.lo + t : a.hi - t;
        for (int i = tid; i < N; i += T) {
            const double ev = cg[(size_t)(slice - 1) * N + i] > 0 ? a.ep : a.em;
            sc_mid[i] = a.left ? ev : 1.0;
            sc_row[i] = a.left ? 1.0 : ev;
        }
        __syncthreads();
        wrap_product<N, 1>(Gs, ring, a.left ? a.eT2 : a.eT2T, sc_mid, sc_row, full, empty, cnt, tid);
    }
    // column scaling by Dsrc, then out (one bulk copy per column)
    for (int idx = tid; idx < N * N; idx += T) {
        const int col = idx / N, row = idx - col * N;
        Gs[(size_t)col * LD + row] *= dcol[col];
    }
    fence_async_smem();
    __syncthreads();
    if (tid < N) {
        bulk_s2g(dst + (size_t)tid * N, Gs + (size_t)tid * LD, N * sizeof(double));
        bulk_commit();
        bulk_wait_all();
    }
}

} // namespace

bool segment_resident_supported(int n) { return n == 144; }

int init_segment_resident_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(segment_resident_kernel<144>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RGeo<144>::SMEM));
    MCB_CUDA(cudaFuncSetAttribute(slice_chain_resident_kernel<144>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RGeo<144>::SMEM));
    return 0;
}

int launch_slice_chain_resident(const SliceChainArgs& a, cudaStream_t st)
{
    if (a.n != 144) { set_error("slice_chain_resident: unsupported size"); return -1; }
    if (a.hi < a.lo || a.batch <= 0) { set_error("slice_chain_resident: empty chain"); return -1; }
    LaunchScope scope(LC_GEMM, st, 2.0 * a.n * (double)a.n * a.n * (a.hi - a.lo + 1) * a.batch);
    slice_chain_resident_kernel<144><<<a.batch, RGeo<144>::THREADS, RGeo<144>::SMEM, st>>>(a);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

int launch_segment_resident(const SegmentArgs& a, cudaStream_t st)
{
    if (a.n != 144) { set_error("segment_resident: unsupported size"); return -1; }
    if (!a.eT2T || !a.eTinv2T) { set_error("segment_resident: transposed hopping exponentials missing"); return -1; }
    if (a.n_visits <= 0 || a.batch <= 0) return 0;
    const int wraps = a.n_visits - 1 + (a.first_wrap ? 1 : 0);
    LaunchScope scope(LC_FUSED, st, 4.0 * a.n * (double)a.n * a.n * wraps * a.batch);
    segment_resident_kernel<144><<<a.batch, RGeo<144>::THREADS, RGeo<144>::SMEM, st>>>(a);
    MCB_CUDA(cudaGetLastError());
#ifdef SWD_TIMING
    {
        static int shown = 0;
        if (a.n_visits >= 2 && shown++ < 3) {
            long long h[32];
            cudaStreamSynchronize(st);
            cudaMemcpyFromSymbol(h, sgr_clk, sizeof(h));
            fprintf(stderr, "segment_resident visits=%d cycles: kernel (after the G load) %lld || visit 1: product 1 %lld product 2 %lld first decisions %lld "
                    "| block 1: recurrences %lld next decisions %lld (under the flush) block total %lld | sweep of the visit %lld\n",
                    a.n_visits, h[11] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3], h[7] - h[5], h[8] - h[7], h[9] - h[5],
                    h[10] - h[3]);
            fprintf(stderr, "  block starts relative to the end of the wrap:");
            for (int b = 0; b < 9; ++b) fprintf(stderr, " %lld", h[12 + b] - h[3]);
            fprintf(stderr, " end %lld | block 7 recurrences done %lld, block 8 recurrences done %lld\n", h[10] - h[3], h[22] - h[3], h[21] - h[3]);
            fprintf(stderr, "  warp 1 leaves the flush of block 7 at %lld, of block 8 at %lld; passes the end-of-visit barrier at %lld; next visit's pre-wrap barrier: thread 0 %lld, thread 32 %lld\n", h[23] - h[3], h[24] - h[3], h[27] - h[3], h[25] - h[3], h[26] - h[3]);
        }
    }
#endif
    return 0;
}

} // namespace mcb
