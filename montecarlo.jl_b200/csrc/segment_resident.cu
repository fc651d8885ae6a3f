// Chain-resident wrap_greens! + sweep_spatial for N = 144 (12 x 12 lattices, BASELINE config C): one CTA of 12 warps per
// chain keeps the whole Green's function (144 x 144 doubles, padded to 148 rows: 170 KB) in shared memory for a run of
// up to safe_mult consecutive time slices, as segment_kernel (segment.cu) does in registers for N <= 64.
//   wrap_greens!                 src/flavors/DQMC/stack.jl:491-500, slice_matrices.jl:23-76
//   sweep_spatial / propose_local / accept_local!   DQMC.jl:546-582, HubbardModelAttractive.jl:113-155
//
// Wrap: the two products of a wrap run on 12 warps (6 x 2 warp tiles of 24 x 72, the whole 144 x 144 result in DMMA
// accumulators), one operand read as fragments straight out of the resident G, the other - the hopping exponential, the
// same matrix for every chain and therefore L2-resident - streamed in panels of 8 columns by TMA bulk copies into a ring
// of 4 stages, each guarded by an mbarrier that counts the bytes (refilled after a CTA barrier, as in gemm_tma.cu). The
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

// One product of a wrap. WHICH == 1: G <- diag(row) H (diag(mid) G)   (A = columns of H from the ring, B = G)
//                        WHICH == 2: G <- ((G diag(mid)) H) diag(col)   (A = G, B = rows of H = columns of H' from the ring)
// Hcols points at the matrix whose COLUMNS are streamed (H for 1, H' for 2). cnt counts the ring stages used so far (the
// same sequence in every thread): stage = cnt % STAGES, the phase of its barrier follows from cnt / STAGES.
template <int N, int WHICH>
__device__ __forceinline__ void wrap_product(double* __restrict__ Gs, double* __restrict__ ring, const double* __restrict__ Hcols,
                                             const double* __restrict__ mid, const double* __restrict__ outsc, uint64_t* full,
                                             uint32_t& cnt, int tid)
{
    using R = RGeo<N>;
    constexpr int LD = R::LD, BK = R::BK, STAGES = R::STAGES, IT = R::IT, JT = R::JT, nk = N / BK;
    const int warp = tid >> 5, lane = tid & 31, r = lane >> 2, q = lane & 3;
    double acc[IT][JT][2];
    const int wm = (warp % 6) * R::WM, wn = (warp / 6) * R::WN;
    // one bulk copy per column of a panel, issued by the first warp
    auto issue = [&](int kt) {
        if (warp == 0) {
            const uint32_t s = (cnt + kt) % STAGES;
            if (lane == 0) mbar_expect_tx(full + s, BK * N * sizeof(double));
            __syncwarp();
            if (lane < BK)
                bulk_g2s(ring + (size_t)(s * BK + lane) * LD, Hcols + (size_t)(kt * BK + lane) * N, N * sizeof(double), full + s);
        }
    };
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
        if (kt + STAGES < nk) {
            __syncthreads();   // every warp is done with stage s: refill it
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
    uint64_t* gbar = full + R::STAGES;
    int* tile_ctr = reinterpret_cast<int*>(sm + R::OFF_MISC);
    volatile int* decide_done = tile_ctr + 1;
    int8_t* s_in = reinterpret_cast<int8_t*>(sm + R::OFF_S);
    int8_t* s_out = s_in + (N + 15) / 16 * 16;

    const int chain = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, r = lane >> 2, q = lane & 3;
    double* Gg = a.G + (size_t)chain * a.sG;
    int8_t* cg = a.conf + (size_t)chain * a.sConf;
    const uint32_t k1 = a.chain_offset + (uint32_t)chain;
    const DecideConst dconst{a.gamma_p, a.gamma_m, a.efac_p, a.efac_m};

    if (tid == 0) {
        for (int s = 0; s < R::STAGES; ++s) mbar_init(full + s, 1);
        mbar_init(gbar, 1);
    }
    __syncthreads();
    // ---- G -> shared memory, one bulk copy per column
    if (tid == 0) mbar_expect_tx(gbar, (uint32_t)(N * N * sizeof(double)));
    if (tid < N) bulk_g2s(Gs + (size_t)tid * LD, Gg + (size_t)tid * N, N * sizeof(double), gbar);
    mbar_wait(gbar, 0);

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
            wrap_product<N, 1>(Gs, ring, a.direction > 0 ? a.eT2 : a.eTinv2, sc_mid1, sc_row1, full, cnt, tid);
            wrap_product<N, 2>(Gs, ring, a.direction > 0 ? a.eTinv2T : a.eT2T, sc_mid2, sc_col2, full, cnt, tid);
        } else {
            __syncthreads();   // s_in, unif
        }

        // ---- sweep_spatial of `slice`. The decisions of the first block have nothing to hide under.
        if (warp == 0) {
            double reg[K];
#pragma unroll
            for (int b = 0; b < K; ++b) reg[b] = lane < K ? Gs[(size_t)b * LD + lane] : 0.0;
            nacc += decide_rows<K>(dconst, reg, 0, K, s_in, s_out, unif, DD, xs, lane);
        }
        const bool last_visit = v == a.n_visits - 1;
        for (int i0 = 0; i0 < N; i0 += K) {
            const int i1 = i0 + K;
            const bool has_next = i1 < N;
            __syncthreads();   // the previous flush is complete, the decisions of this block are in DD / xs
            // ---- copies of the block's columns and rows, the next block's diagonal block (all as they are BEFORE this flush)
            for (int idx = tid; idx < K * N; idx += T) {
                const int j = idx / N, row = idx - j * N;
                Cc[j * LD + row] = Gs[(size_t)(i0 + j) * LD + row];
            }
            for (int idx = tid; idx < K * N; idx += T) {
                const int l = idx / K, k = idx - l * K;
                Rc[k * LD + l] = Gs[(size_t)l * LD + i0 + k];
            }
            if (has_next && tid < K * K) {
                const int ra = tid % K, cb = tid / K;
                Dn[cb * K + ra] = Gs[(size_t)(i1 + cb) * LD + i1 + ra];
            }
            if (tid == 0) { *tile_ctr = 0; *decide_done = 0; }
            __syncthreads();
            // ---- the update vectors of the block (two triangular recurrences, one element per thread)
            {
                constexpr int NV = (N + 31) / 32 * 32;
                if (tid < NV) { if (tid < N) column_recurrence(Cc, LD, i0, K, DD, xs, tid); }
                else if (tid < 2 * NV) { if (tid - NV < N) row_recurrence(Rc, LD, K, DD, tid - NV); }
            }
            __syncthreads();
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
                __syncwarp();
                if (lane == 0) { __threadfence_block(); *decide_done = 1; }
            } else if (has_next && (warp & 3) == 0) {
                // the warps on the deciding warp's scheduler stay out of the flush until it is done (sweep_delayed.cu)
                while (!*decide_done) __nanosleep(100);
            }
            // ---- flush G -= U Rc in shared memory; warp tiles of 32 x 16 handed out by a counter
            constexpr int TM = (N + 31) / 32, TN = N / 16;
            const bool check = last_visit && !has_next;
            for (;;) {
                int tnext = 0;
                if (lane == 0) tnext = atomicAdd(tile_ctr, 1);
                tnext = __shfl_sync(0xffffffffu, tnext, 0);
                if (tnext >= TM * TN) break;
                const int row0 = (tnext % TM) * 32, col0 = (tnext / TM) * 16;
                double c[4][2][2];
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y) c[x][y][0] = c[x][y][1] = 0.0;
                const double* ua[4];
                const double* rb[2];
#pragma unroll
                for (int x = 0; x < 4; ++x) ua[x] = Cc + q * LD + min(row0 + 8 * x + r, N - 1);
#pragma unroll
                for (int y = 0; y < 2; ++y) rb[y] = Rc + q * LD + col0 + 8 * y + r;
#pragma unroll
                for (int k0 = 0; k0 < K; k0 += 4) {
                    double av[4], bv[2];
#pragma unroll
                    for (int x = 0; x < 4; ++x) av[x] = ua[x][k0 * LD];
#pragma unroll
                    for (int y = 0; y < 2; ++y) bv[y] = rb[y][k0 * LD];
#pragma unroll
                    for (int x = 0; x < 4; ++x)
#pragma unroll
                        for (int y = 0; y < 2; ++y) dmma884(c[x][y][0], c[x][y][1], av[x], bv[y]);
                }
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 2; ++y)
#pragma unroll
                        for (int cc = 0; cc < 2; ++cc) {
                            const int row = row0 + 8 * x + r, col = col0 + 8 * y + 2 * q + cc;
                            if (row < N) {
                                double* gp = Gs + (size_t)col * LD + row;
                                const double gnew = *gp - c[x][y][cc];
                                *gp = gnew;
                                if (check) bad |= (__double2hiint(gnew) & 0x7ff00000) == 0x7ff00000;
                            }
                        }
            }
        }
        __syncthreads();   // last flush done; s_out complete
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
    if (tid == 0) {
        a.acc[chain] += nacc;
        a.prop[chain] += (long long)N * a.n_visits;
        if (anybad) a.nonfinite[chain] += 1;
    }
}

} // namespace

bool segment_resident_supported(int n) { return n == 144; }

int init_segment_resident_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(segment_resident_kernel<144>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RGeo<144>::SMEM));
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
    return 0;
}

} // namespace mcb
