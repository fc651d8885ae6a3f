// wrap_greens! (src/flavors/DQMC/stack.jl:491-500) + sweep_spatial (src/flavors/DQMC/DQMC.jl:546-582)
// with propose_local / accept_local! (src/models/HubbardModel/HubbardModelAttractive.jl:113-155),
// fused over a run of consecutive time slices.
//
// Geometry: a CTA works on CH = 4 Markov chains at once; each chain belongs to NWC = N/16 warps and its
// Green's function stays in their REGISTERS for the whole run of slices, in the accumulator layout of
// the FP64 tensor-core instruction (DMMA m8n8k4):
//     warp wl of the chain owns the 16 columns [16 wl, 16 wl + 16), all N rows;
//     lane l (r = l>>2, q = l&3): acc[m][cw][c] = G[8m + r][16 wl + 8 cw + 2q + c].
// The hopping exponentials e^{-+dtau T} are staged ONCE per CTA in shared memory and shared by its four
// chains; each chain has a private staging matrix X for the operand that has to cross warps. Chains
// synchronise on their own named barrier, never CTA-wide, so the four latency chains of the strictly
// sequential site loop interleave on one SM.
//
// Wrap: two N^3 products on the FP64 tensor pipe; the slice matrices B_l = e^{-dtau T} diag(e^{lambda s})
// are never materialised (slice_matrices.jl:23-39): the diagonal factor is applied in registers.
// Local updates: DELAYED rank-8 updates, two barriers per block of 8 sites. All warps publish the 8
// rows of the block; the ONE warp that owns its 8 columns then runs the 8 strictly sequential
// Metropolis decisions on private copies of those rows and columns (warp-local: shuffles and
// __syncwarp only), recording  -x (e_i - G[:,i])  and  G[i,:]  of every accepted flip
// (accept_local!, x = gamma / (1 + gamma (1 - G_ii))); after the second barrier every warp applies
// G += NUX * V (N x 8 x N) to its registers on the FP64 tensor pipe.
#include "segment.cuh"

namespace mcb {

namespace {

constexpr int CH = 4; // chains per CTA
constexpr int CW = 2; // 8-column blocks per warp

template <int N>
struct SGeo {
    static_assert(N % 16 == 0, "chain-resident kernels need N to be a multiple of 16");
    static constexpr int MT = N / 8;         // 8-row blocks
    static constexpr int NWC = N / (8 * CW); // warps per chain
    static constexpr int TPC = NWC * 32;     // threads per chain
    static constexpr int THREADS = CH * TPC;
    static constexpr int LD = N + 4;         // LD % 16 == 4: conflict-free DMMA fragment loads
};

template <int N>
struct alignas(16) ChainSmem {
    static constexpr int LD = SGeo<N>::LD;
    static constexpr int XSIZE = N * LD > 48 * LD + 160 ? N * LD : 48 * LD + 160;
    double X[XSIZE];   // wrap: staging matrix; site loop: 2 x (Rb | NUX | V) (8 x LD each, by block parity) | block scalars
    double unif[N];
    double sc_row[N];
    double sc_col[N];
    int flag;
    int nacc;
    int8_t s_in[N];
    int8_t s_out[N];
    __device__ double* Rb(int par) { return X + par * 24 * LD; }            // rows of the block, Rb[k*LD + col]
    __device__ double* NUX(int par) { return X + par * 24 * LD + 8 * LD; }  // -x (e_i - G[:, i]) of site j at NUX[j*LD + row]
    __device__ double* V(int par) { return X + par * 24 * LD + 16 * LD; }   // G[i, :] of site j at V[j*LD + col]
    // results of the sequential 8 x 8 decision chain of the block
    __device__ double* nuS() { return X + 48 * LD; }        // [j][k]: NUX_j at block row k (0 if rejected)
    __device__ double* rowS() { return X + 48 * LD + 64; }  // [j][k]: V_j at block column k
    __device__ double* xS() { return X + 48 * LD + 128; }   // [j]: x_j (0 if rejected)
};

template <int N>
struct SegSmem {
    double eT2[N * SGeo<N>::LD];
    double eTinv2[N * SGeo<N>::LD];
    ChainSmem<N> ch[CH];
};

__device__ __forceinline__ void chain_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

struct CLane {
    int ctid, wl, l, r, q, col0; // thread in chain, warp in chain, lane, row-in-block, quad, first column
    // rot rotates the warp roles, so that the same role of different chains sits on different SMSPs
    __device__ __forceinline__ CLane(int tid_in_chain, int rot, int nw)
    {
        ctid = tid_in_chain; wl = ((ctid >> 5) + rot) % nw; l = ctid & 31; r = l >> 2; q = l & 3;
        col0 = 8 * CW * wl + 2 * q;
    }
};

template <int N>
__device__ __forceinline__ void r2s(const double (&acc)[SGeo<N>::MT][CW][2], double* S, const CLane& t)
{
    constexpr int LD = SGeo<N>::LD;
#pragma unroll
    for (int m = 0; m < SGeo<N>::MT; ++m)
#pragma unroll
        for (int cw = 0; cw < CW; ++cw)
#pragma unroll
            for (int c = 0; c < 2; ++c) S[(8 * m + t.r) + (t.col0 + 8 * cw + c) * LD] = acc[m][cw][c];
}

// acc = A * B, both in shared memory (ld = LD), column-major
template <int N>
__device__ __forceinline__ void gemm(double (&acc)[SGeo<N>::MT][CW][2], const double* As, const double* Bs,
                                     const CLane& t)
{
    constexpr int LD = SGeo<N>::LD, MT = SGeo<N>::MT;
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int cw = 0; cw < CW; ++cw) acc[m][cw][0] = acc[m][cw][1] = 0.0;
    const double* ap = As + t.r + t.q * LD;
    const double* bp = Bs + t.q + (8 * CW * t.wl + t.r) * LD;
#pragma unroll 4
    for (int k0 = 0; k0 < N; k0 += 4) {
        double b[CW];
#pragma unroll
        for (int cw = 0; cw < CW; ++cw) b[cw] = bp[k0 + 8 * cw * LD];
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            const double a = ap[8 * m + k0 * LD];
#pragma unroll
            for (int cw = 0; cw < CW; ++cw) dmma884(acc[m][cw][0], acc[m][cw][1], a, b[cw]);
        }
    }
}

template <int N>
__device__ __forceinline__ void scale_rows(double (&acc)[SGeo<N>::MT][CW][2], const double* v, const CLane& t)
{
#pragma unroll
    for (int m = 0; m < SGeo<N>::MT; ++m) {
        const double s = v[8 * m + t.r];
#pragma unroll
        for (int cw = 0; cw < CW; ++cw) { acc[m][cw][0] *= s; acc[m][cw][1] *= s; }
    }
}
template <int N>
__device__ __forceinline__ void scale_cols(double (&acc)[SGeo<N>::MT][CW][2], const double* v, const CLane& t)
{
#pragma unroll
    for (int cw = 0; cw < CW; ++cw) {
        const double s0 = v[t.col0 + 8 * cw], s1 = v[t.col0 + 8 * cw + 1];
#pragma unroll
        for (int m = 0; m < SGeo<N>::MT; ++m) { acc[m][cw][0] *= s0; acc[m][cw][1] *= s1; }
    }
}

// sites 8 IT .. 8 IT + 7 of sweep_spatial.
// The only strictly sequential part of the 8 decisions involves the 8 x 8 diagonal block
// D = G[blk, blk]: flip j needs D_jj and updates D by a rank-1 term. Phase 2a runs that chain in the
// warp that owns the block's columns (lane r holds row r of D) and records, per accepted flip j,
// x_j, nu_j[k] = NUX_j at block row k and row_j[k] = V_j at block column k. With these scalars the
// full vectors follow from two triangular recurrences without any sequential decision (phase 2b,
// one row or one column per thread):
//     NUX_k = x_k (G0[:, k] - e_k + sum_{m<k} NUX_m row_m[k]),   V_k = G0[k, :] + sum_{m<k} nu_m[k] V_m
template <int N, int IT>
__device__ __forceinline__ void site_block(double (&acc)[SGeo<N>::MT][CW][2], ChainSmem<N>& cs,
                                           const SegmentArgs& a, const CLane& t, int bar_id, int& nacc)
{
    constexpr int MT = SGeo<N>::MT, LD = SGeo<N>::LD, TPC = SGeo<N>::TPC;
    constexpr int WI = IT / CW, CWI = IT % CW;  // owner of the block's 8 columns
    // (block parity: the next block's phase 1 may overlap this block's phase 3 in another warp)
    double* Rb = cs.Rb(IT & 1);
    double* NUX = cs.NUX(IT & 1);
    double* V = cs.V(IT & 1);
    double* nuS = cs.nuS();
    double* rowS = cs.rowS();
    double* xS = cs.xS();
    // ---- phase 1: publish the rows 8 IT .. 8 IT + 7 (everybody) and the columns (their owner; into NUX)
#pragma unroll
    for (int cw = 0; cw < CW; ++cw)
        *reinterpret_cast<double2*>(Rb + t.r * LD + t.col0 + 8 * cw) = make_double2(acc[IT][cw][0], acc[IT][cw][1]);
    if (t.wl == WI) {
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            NUX[(2 * t.q) * LD + 8 * m + t.r] = acc[m][CWI][0];
            NUX[(2 * t.q + 1) * LD + 8 * m + t.r] = acc[m][CWI][1];
        }
    }
    chain_sync(bar_id, TPC);
    // ---- phase 2a: the sequential decisions on the 8 x 8 diagonal block
    if (t.wl == WI) {
        const int r8 = t.l & 7; // lanes 8..31 mirror lanes 0..7
        double Dr[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) Dr[c] = Rb[r8 * LD + 8 * IT + c];
        // everything that does not depend on the previous decision is read up front; the chain itself is
        // branch-free (the update is computed speculatively and masked), because every FP64 instruction
        // on it queues behind the DMMAs of the other chains of the SM
        int sgn = 0; // bit j: s_j > 0
#pragma unroll
        for (int j = 0; j < 8; ++j) sgn |= (cs.s_in[8 * IT + j] > 0 ? 1 : 0) << j;
        int flips = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const double gii = __shfl_sync(0xffffffffu, Dr[j], j);
            double rowj[8];
#pragma unroll
            for (int c = j + 1; c < 8; ++c) rowj[c] = __shfl_sync(0xffffffffu, Dr[c], j);
            const bool up = (sgn >> j) & 1;
            const double gam = up ? a.gamma_p : a.gamma_m;
            const double efc = up ? a.efac_p : a.efac_m;
            const double dm = Dr[j] - (r8 == j ? 1.0 : 0.0);
            const double rt = 1.0 + gam * (1.0 - gii);          // propose_local: detratio = rt^2
            const double p = efc * (rt * rt);                   // DQMC.jl:570
            // DQMC.jl:573: p > 1 || rand() < p; the uniform is < 1, so the second test alone decides
            const bool accept = cs.unif[8 * IT + j] < p;
            const double xs = gam * fast_rcp(rt);               // accept_local!: gamma / (1 + gamma u_i)
            const double x = accept ? xs : 0.0;
            const double nu = accept ? xs * dm : 0.0;
#pragma unroll
            for (int c = j + 1; c < 8; ++c) Dr[c] += nu * rowj[c];
            flips |= (accept ? 1 : 0) << j;
            if (t.l < 8) nuS[j * 8 + t.l] = nu;
            if (t.l == 0) {
                xS[j] = x;
#pragma unroll
                for (int c = j + 1; c < 8; ++c) rowS[j * 8 + c] = rowj[c];
            }
        }
        nacc += __popc(flips);
        if (t.l < 8) {
            const int8_t s = cs.s_in[8 * IT + t.l];
            cs.s_out[8 * IT + t.l] = ((flips >> t.l) & 1) ? (int8_t)-s : s;
        }
    }
    chain_sync(bar_id, TPC);
    // ---- phase 2b: the vectors of the accepted flips, one row (NUX) or one column (V) per thread
    // (right-looking: a finished vector entry is added to all later sums at once, so that the
    // dependent chain is one FMA and one product per step; the order of the additions is unchanged)
    if (t.ctid < N) {
        const int rho = t.ctid;
        double v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = NUX[k * LD + rho] - (rho == 8 * IT + k ? 1.0 : 0.0);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            v[k] *= xS[k];
#pragma unroll
            for (int c = k + 1; c < 8; ++c) v[c] += v[k] * rowS[k * 8 + c];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) NUX[k * LD + rho] = v[k];
    } else if (t.ctid < 2 * N) {
        const int kap = t.ctid - N;
        double v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = Rb[k * LD + kap];
#pragma unroll
        for (int k = 0; k < 8; ++k)
#pragma unroll
            for (int c = k + 1; c < 8; ++c) v[c] += nuS[k * 8 + c] * v[k];
#pragma unroll
        for (int k = 0; k < 8; ++k) V[k * LD + kap] = v[k];
    }
    chain_sync(bar_id, TPC);
    // ---- phase 3: G += NUX * V on the tensor pipe (k = 8)
    {
        const double* ap = NUX + t.q * LD + t.r;
        const double* bp = V + t.q * LD + 8 * CW * t.wl + t.r;
#pragma unroll
        for (int k0 = 0; k0 < 8; k0 += 4) {
            double b[CW];
#pragma unroll
            for (int cw = 0; cw < CW; ++cw) b[cw] = bp[k0 * LD + 8 * cw];
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const double av = ap[k0 * LD + 8 * m];
#pragma unroll
                for (int cw = 0; cw < CW; ++cw) dmma884(acc[m][cw][0], acc[m][cw][1], av, b[cw]);
            }
        }
    }
    if constexpr (IT + 1 < MT) site_block<N, IT + 1>(acc, cs, a, t, bar_id, nacc);
}

template <int N>
__global__ void __launch_bounds__(SGeo<N>::THREADS, 1) segment_kernel(SegmentArgs a)
{
    griddep_wait(); griddep_launch();   // programmatic dependent launch (common.cuh)
    constexpr int MT = SGeo<N>::MT, TPC = SGeo<N>::TPC, LD = SGeo<N>::LD;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SegSmem<N>& sm = *reinterpret_cast<SegSmem<N>*>(smem_raw);
    // the hopping exponentials, once per CTA
    for (int e = threadIdx.x; e < N * N; e += SGeo<N>::THREADS) {
        const int i = e % N, j = e / N;
        sm.eT2[i + j * LD] = a.eT2[e];
        sm.eTinv2[i + j * LD] = a.eTinv2[e];
    }
    __syncthreads();
    const int cl = threadIdx.x / TPC; // chain within the CTA
    const int chain = blockIdx.x * CH + cl;
    if (chain >= a.batch) return; // whole chains leave together; they only use their own barrier below
    const CLane t(threadIdx.x % TPC, a.phase_sync ? cl : 0, TPC / 32);
    const int bar_id = 1 + cl;
    ChainSmem<N>& cs = sm.ch[cl];
    double* Gg = a.G + (size_t)chain * a.sG;
    int8_t* cg = a.conf + (size_t)chain * a.sConf;
    const uint32_t k1 = a.chain_offset + (uint32_t)chain;

    double acc[MT][CW][2];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int cw = 0; cw < CW; ++cw)
#pragma unroll
            for (int c = 0; c < 2; ++c) acc[m][cw][c] = Gg[(8 * m + t.r) + (size_t)(t.col0 + 8 * cw + c) * N];

    int nacc = 0; // accepted flips decided by this warp
    if (t.ctid == 0) { cs.flag = 0; cs.nacc = 0; }
    for (int v = 0; v < a.n_visits; ++v) {
        const int slice = a.start_slice + a.direction * v;
        // the fields and uniforms of this visit's sweep_spatial (read together with the wrap's fields,
        // one global-memory round trip per visit)
        const bool wrapping = v > 0 || a.first_wrap;
        {
            const long long upd = a.n_updates0 + v;
            const uint32_t sweep = (uint32_t)(upd / (2LL * a.n_slices));
            const uint32_t visit = (uint32_t)(upd % (2LL * a.n_slices));
            for (int i = t.ctid; i < N; i += TPC) {
                cs.s_in[i] = cg[(size_t)(slice - 1) * N + i];
                cs.unif[i] = dqmc_uniform(a.key0, k1, sweep, visit, (uint32_t)i);
            }
        }
        if (wrapping) {
            // ---- wrap_greens!
            // The products saturate the FP64 pipe, and a dependent DFMA of another chain's site loop
            // that queues behind them takes several times its latency: all chains of the CTA do their
            // products in the same window and their site loops in the other.
            if (a.phase_sync) __syncthreads();
            const int l = a.direction > 0 ? slice - 1 : slice; // slice matrix B_l
            for (int i = t.ctid; i < N; i += TPC) {
                const int s = cg[(size_t)(l - 1) * N + i];
                const double ep = s > 0 ? a.eVp[0] : a.eVp[1], em = s > 0 ? a.eVm[0] : a.eVm[1];
                cs.sc_row[i] = a.direction > 0 ? ep : em;
                cs.sc_col[i] = a.direction > 0 ? em : ep;
            }
            chain_sync(bar_id, TPC);
            if (a.direction > 0) {
                // G <- e^{-dtau T} diag(e^{+lambda s}) G diag(e^{-lambda s}) e^{+dtau T}
                scale_rows<N>(acc, cs.sc_row, t);
                r2s<N>(acc, cs.X, t);
                chain_sync(bar_id, TPC);
                gemm<N>(acc, sm.eT2, cs.X, t);
                chain_sync(bar_id, TPC);
                scale_cols<N>(acc, cs.sc_col, t);
                r2s<N>(acc, cs.X, t);
                chain_sync(bar_id, TPC);
                gemm<N>(acc, cs.X, sm.eTinv2, t);
            } else {
                // G <- diag(e^{-lambda s}) e^{+dtau T} G e^{-dtau T} diag(e^{+lambda s})
                r2s<N>(acc, cs.X, t);
                chain_sync(bar_id, TPC);
                gemm<N>(acc, sm.eTinv2, cs.X, t);
                chain_sync(bar_id, TPC);
                scale_rows<N>(acc, cs.sc_row, t);
                r2s<N>(acc, cs.X, t);
                chain_sync(bar_id, TPC);
                gemm<N>(acc, cs.X, sm.eT2, t);
                scale_cols<N>(acc, cs.sc_col, t);
            }
            if (a.phase_sync) __syncthreads();
        }
        // ---- sweep_spatial of `slice`
        // the wrap's last product reads all of cs.X as its A operand and site_block<0> overwrites columns 0..15 of
        // it: without the CTA barrier above the chain needs its own (also orders s_in / unif when nothing wrapped)
        if (!wrapping || !a.phase_sync) chain_sync(bar_id, TPC);
        site_block<N, 0>(acc, cs, a, t, bar_id, nacc);
        chain_sync(bar_id, TPC);
        for (int i = t.ctid; i < N; i += TPC) cg[(size_t)(slice - 1) * N + i] = cs.s_out[i];
    }
    bool bad = false;
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int cw = 0; cw < CW; ++cw)
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const double g = acc[m][cw][c];
                bad |= !isfinite(g);
                Gg[(8 * m + t.r) + (size_t)(t.col0 + 8 * cw + c) * N] = g;
            }
    if (bad) cs.flag = -1; // (benign race: any writer flags the chain)
    if (t.l == 0 && nacc) atomicAdd(&cs.nacc, nacc);
    chain_sync(bar_id, TPC);
    if (t.ctid == 0) {
        a.acc[chain] += cs.nacc;
        a.prop[chain] += (long long)N * a.n_visits;
        if (cs.flag == -1) a.nonfinite[chain] += 1;
    }
}

template <int N>
int launch_t(const SegmentArgs& a, cudaStream_t st)
{
    // algorithmic flop: 4 N^3 per wrap; the rank-1 updates are added by the caller from the
    // acceptance counters
    const int wraps = a.n_visits - 1 + (a.first_wrap ? 1 : 0);
    LaunchScope scope(LC_FUSED, st, 4.0 * N * (double)N * N * wraps * a.batch);
    MCB_CUDA(launch_pdl(segment_kernel<N>, dim3((a.batch + CH - 1) / CH), dim3(SGeo<N>::THREADS), sizeof(SegSmem<N>), st, a));
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace

bool segment_supported(int n) { return n == 16 || n == 32 || n == 64; }

int init_segment_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(segment_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem<16>)));
    MCB_CUDA(cudaFuncSetAttribute(segment_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem<32>)));
    MCB_CUDA(cudaFuncSetAttribute(segment_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem<64>)));
    return 0;
}

int launch_segment(const SegmentArgs& a, cudaStream_t st)
{
    if (a.n_visits <= 0 || a.batch <= 0) return 0;
    switch (a.n) {
    case 16: return launch_t<16>(a, st);
    case 32: return launch_t<32>(a, st);
    case 64: return launch_t<64>(a, st);
    default: set_error("segment kernel: unsupported matrix size"); return -1;
    }
}

} // namespace mcb
