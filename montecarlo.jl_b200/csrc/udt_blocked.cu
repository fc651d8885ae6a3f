// Blocked, panel-pivoted Householder QR with compact-WY block reflectors on the FP64 tensor pipe (DMMA):
// udt_AVX_pivot! (src/linalg/UDT.jl:192-306) for N % 16 == 0, 128 <= N <= 400 (configs C / D: N = 144, 400).
//
// The reference pivots column by column: every step applies one reflector to the whole trailing matrix (BLAS-2) and
// re-measures all trailing norms; on a GPU that is a chain of N dependent, barrier-separated steps. Here the matrix is
// factored in panels of NB = 16 columns:
//   1. qrb_panel_kernel   (one CTA per chain) the exact squared norms of the trailing columns are known from the
//      previous update; the NB largest form the next panel, sorted by decreasing norm, and are moved to positions
//      j .. j+NB-1 (the displaced columns fill the holes; all column traffic is TMA bulk copies). The panel is factored
//      by unpivoted Householder steps with the reference's reflector! convention, all warps of the CTA on registers,
//      ONE block barrier per step (qr_blocked.cuh). Output: V (unit lower trapezoid) and R in place, tau, D = |R_kk|,
//      and the NB x NB triangular factor T of  H_1 .. H_NB = I - V T V'  (Gram matrix V'V on the tensor pipe).
//   2. qrb_apply_kernel   (grid = column blocks x chains) every warp keeps an 8-column strip of the trailing matrix in
//      registers and applies  A <- A - V T' (V' A)  as three DMMA products; the strip itself serves as the tensor
//      operand of V' A (the k index of a DMMA product may be permuted freely as long as both operands agree, so the
//      accumulator layout of one product is the A-operand layout of the next). The squared norms of the updated
//      columns below the panel come out of the same registers, in a fixed order.
//   3. the same apply kernel, reflectors in reverse order, forms the explicit Q from the identity (UDT.jl:250-266).
//   4. qrb_emit_kernel    T = D^-1 R P' or the "dirty" D^-1 R + pivot (UDT.jl:283-306).
// Pivot order differs from the reference's (panel-wise instead of step-wise selection); U D T = X, the grading of D
// and the accuracy of the Green's functions built from it do not (profiles/experiments/panel_pivot_qr.py: 4e-12 /
// 8e-13 relative to the step-wise CPU restatement at configs C / D, the same distance LAPACK's dgeqp3 has).
#include "udt.cuh"
#include <map>
#include <mutex>
#include <stdlib.h>
#include <utility>
#include "prof.cuh"
#include "qr_blocked.cuh"
#include "tma.cuh"

namespace mcb {

namespace {

using qrb::NB;
using qrb::TLD;

struct QrbArgs {
    int n = 0, j = 0;
    double* A = nullptr;     // [batch][n*n] the matrix being factored (V below, R above the diagonal afterwards)
    double* D = nullptr; long long sD = 0;
    double* norms = nullptr; // [batch][n] squared norms of rows >= j of the columns at positions >= j
    int* perm = nullptr;     // [batch][n] original column index of the column at each position
    double* Tfac = nullptr;  // [batch][n/NB][NB*NB] block-reflector factors, column-major
};

// ---------------------------------------------------------------------------------------------- panel
// PW = warps that run the Householder steps, GW = warps of the Gram product
template <int PW, int GW>
size_t panel_smem(int n)
{
    const size_t ldp = n + 8; // ldp % 16 == 8 for n % 16 == 0
    return (2 * NB * ldp + NB * NB /*Ptop*/ + n /*nrm*/ + 2 * PW * NB + 2 * NB + NB * NB /*Ssm*/ +
            (size_t)GW * NB * NB /*Sred*/ + NB /*taus*/ + 2 /*mbarrier*/) * sizeof(double) + (4 * NB + 4) * sizeof(int);
}

// NT threads; the Householder steps run on warps 0 .. PW-1 (more warps only add redundant per-step work: 1425
// cycles per step with 8 warps of a 256-thread CTA, 910 with 4), everything else on all of them
#ifdef SWD_TIMING
__device__ long long qrb_clk[16];
#define QRB_T(i) do { if (blockIdx.x == 0 && threadIdx.x == 0 && a.j == 5 * NB) qrb_clk[i] = clock64(); } while (0)
#else
#define QRB_T(i) do { } while (0)
#endif

template <int NT, int PW, int TPW, int GW>
__global__ void __launch_bounds__(NT) qrb_panel_kernel(QrbArgs a)
{
    extern __shared__ __align__(16) double sm[];
    const int n = a.n, j = a.j, chain = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int T = NT, NW = NT / 32;
    const int ldp = n + 8;
    double* P = sm;                              // [NB][ldp] the panel, all n rows
    double* Dsp = P + (size_t)NB * ldp;          // [NB][ldp] staging of the displaced columns
    double* Ptop = Dsp + (size_t)NB * ldp;       // [NB][NB] raw top block of the factored panel
    double* nrm = Ptop + NB * NB;                // [n]
    double* part = nrm + n;                      // [2][PW][NB]
    double* rowk = part + 2 * PW * NB;           // [2][NB]
    double* Ssm = rowk + 2 * NB;                 // [NB][NB]
    double* Sred = Ssm + NB * NB;                // [GW][NB*NB]; also the selection's rank table (n ints)
    double* taus = Sred + (size_t)GW * NB * NB;  // [NB]
    uint64_t* mbar = reinterpret_cast<uint64_t*>(taus + NB);
    int* sel = reinterpret_cast<int*>(mbar + 2); // [NB] selected positions, [NB] displaced, [NB] holes, [NB] perm staging
    int* disp = sel + NB;
    int* hole = disp + NB;
    int* pstage = hole + NB;
    int* ndisp = pstage + NB;
    double* A = a.A + (size_t)chain * n * n;
    const double* gn = a.norms + (size_t)chain * n;
    int* perm = a.perm + (size_t)chain * n;
    const int m = n - j, mt = m / 8;

    QRB_T(0);
    if (tid == 0) mbar_init(mbar, 1);
    griddep_wait();   // the trailing update of the previous panel
    // ---- squared norms of the trailing columns (rows >= j): exact, from the previous update; computed here at j = 0
    if (j == 0) {
        for (int c = warp; c < n; c += NW) {
            const double* col = A + (size_t)c * n;
            double s = 0.0;
#pragma unroll 8
            for (int i = lane; i < n; i += 32) { const double v = __ldcg(col + i); s = fma(v, v, s); }
            s = warp_sum(s);
            if (lane == 0) nrm[c] = (s == s) ? s : INFINITY; // NaN columns sort first: the ranking needs a total order
        }
        for (int c = tid; c < n; c += T) perm[c] = c;
    } else {
        for (int c = j + tid; c < n; c += T) { const double s = __ldcg(gn + c); nrm[c] = (s == s) ? s : INFINITY; }
    }
    __syncthreads();
    QRB_T(1);
    // ---- the NB largest, by decreasing norm (ties: lower position first)
    qrb::select_by_rank<NT>(nrm, j, n, tid, Sred, sel);
    QRB_T(2);
    // ---- gather the panel (all rows: rows < j carry R entries of earlier panels) and the displaced columns
    if (tid < NB) pstage[tid] = perm[sel[tid]];
    if (warp == 0) {
        // the lists by ballots (lane = slot s / panel position), the copies issued by the lanes in parallel
        const int sv = lane < NB ? sel[lane] : 0x7fffffff;
        const bool inside = lane < NB && sv < j + NB;                       // selected and already inside the panel
        const unsigned chosen = __reduce_or_sync(0xffffffffu, inside ? 1u << (sv - j) : 0u);
        const unsigned outside = __ballot_sync(0xffffffffu, lane < NB && !inside);
        const unsigned freed = ~chosen & 0xffffu;                           // panel positions whose column is displaced
        const bool displaced = lane < NB && ((freed >> lane) & 1u);
        const int dslot = __popc(freed & ((1u << lane) - 1u));              // displaced: by increasing position
        if (displaced) disp[dslot] = j + lane;
        if (lane < NB && !inside) hole[__popc(outside & ((1u << lane) - 1u))] = sv;   // holes: by increasing slot
        const uint32_t colbytes = (uint32_t)n * sizeof(double);
        if (lane == 0) { *ndisp = __popc(freed); mbar_expect_tx(mbar, (uint32_t)(NB + __popc(freed)) * colbytes); }
        __syncwarp();
        if (lane < NB) bulk_g2s(P + (size_t)lane * ldp, A + (size_t)sv * n, colbytes, mbar);
        if (displaced) bulk_g2s(Dsp + (size_t)dslot * ldp, A + (size_t)(j + lane) * n, colbytes, mbar);
    }
    __syncthreads();
    const int nd = *ndisp;
    int moved = -1;
    if (tid < nd) moved = perm[disp[tid]];
    mbar_wait(mbar, 0);
    __syncthreads();   // every thread has read its perm entries; the staged columns are complete
    QRB_T(3);
    if (tid < nd) perm[hole[tid]] = moved;
    if (tid < NB) perm[j + tid] = pstage[tid];
    if (tid < nd) {   // one copy per thread; every issuing thread waits for its own groups before the CTA ends (below)
        bulk_s2g(A + (size_t)hole[tid] * n, Dsp + (size_t)tid * ldp, (uint32_t)n * sizeof(double));
        bulk_commit();
    }
    // ---- panel into the registers of PW warps, NB Householder steps (named barrier of these warps per step)
    const int q = lane & 3, rl = lane >> 2;
    if (warp < PW) {
        double x[TPW][4];
#pragma unroll
        for (int u = 0; u < TPW; ++u) {
            const int t = warp + PW * u;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) x[u][cc] = (t < mt) ? P[(size_t)qrb::panel_col(q, cc) * ldp + j + 8 * t + rl] : 0.0;
        }
        qrb::panel_steps<PW, TPW, 1>(x, mt, warp, lane, part, rowk, taus, a.D + (size_t)chain * a.sD + j);
        // ---- V back to P; the raw top block is kept aside while P carries the unit lower trapezoid for V'V
#pragma unroll
        for (int u = 0; u < TPW; ++u) {
            const int t = warp + PW * u;
            if (t < mt) {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) P[(size_t)qrb::panel_col(q, cc) * ldp + j + 8 * t + rl] = x[u][cc];
            }
        }
    }
    __syncthreads();
    QRB_T(4);
    griddep_launch();   // what is left takes about as long as the next kernel's launch
    for (int e = tid; e < NB * NB; e += T) {
        const int c = e / NB, r = e % NB;
        const double v = P[(size_t)c * ldp + j + r];
        Ptop[e] = v;
        P[(size_t)c * ldp + j + r] = r > c ? v : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    qrb::panel_gram<GW>(P, ldp, j, mt, warp, lane, T, tid, Sred, Ssm);
    if (warp == 0) qrb::panel_tfactor(Ssm, taus, lane, a.Tfac + ((size_t)chain * (n / NB) + j / NB) * NB * NB);
    for (int e = tid; e < NB * NB; e += T) P[(size_t)(e / NB) * ldp + j + e % NB] = Ptop[e];
    fence_async_smem();
    __syncthreads();
    QRB_T(5);
    if (tid < NB) {   // threads 0 .. nd - 1 also issued a displaced column above
        bulk_s2g(A + (size_t)(j + tid) * n, P + (size_t)tid * ldp, (uint32_t)n * sizeof(double));
        bulk_commit();
        bulk_wait_all();
    }
    QRB_T(6);
}

// ---------------------------------------------------------------------------------------------- apply
// C <- (I - V op(T) V') C for the columns [c_lo, n) of C, rows >= j; V = unit lower trapezoid stored in A[j:, j:j+NB].
//   mode 0 (factorisation): op(T) = T', C = A itself (trailing columns), epilogue = squared norms of rows >= j + NB
//   mode 1 (form Q):        op(T) = T,  C = Q
// One warp per 8-column strip (qrb::apply_strip); V arrives by TMA bulk copies while the strip is loaded.
struct ApplyArgs {
    int n = 0, j = 0, c_lo = 0, mode = 0;
    const double* A = nullptr; // reflectors
    const double* Tfac = nullptr;
    double* Cm = nullptr;      // [batch][n*n]
    double* norms = nullptr;
};

template <int RT, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32) qrb_apply_kernel(ApplyArgs a)
{
    extern __shared__ __align__(16) double sm[];
    const int n = a.n, j = a.j, chain = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = n - j, mt = m / 8;
    const int ldv = m + ((8 - (m & 15)) & 15);       // ldv % 16 == 8
    double* Vs = sm;                                  // [NB][ldv]   V[r][k] at k * ldv + r
    double* Vt = Vs + (size_t)NB * ldv;               // [m][TLD]    V[r][k] at r * TLD + k
    double* Ts = Vt + (size_t)m * TLD;                // [NB][TLD]   op(T)[k][k2] at k2 * TLD + k
    uint64_t* mbar = reinterpret_cast<uint64_t*>(Ts + (size_t)NB * TLD);
    const double* A = a.A + (size_t)chain * n * n;
    griddep_wait();
    if (tid == 0) {
        mbar_init(mbar, 1);
        const uint32_t colbytes = (uint32_t)m * sizeof(double);
        mbar_expect_tx(mbar, NB * colbytes);
        for (int k = 0; k < NB; ++k) bulk_g2s(Vs + (size_t)k * ldv, A + (size_t)(j + k) * n + j, colbytes, mbar);
    }
    {
        const double* Tg = a.Tfac + ((size_t)chain * (n / NB) + j / NB) * NB * NB;
        for (int e = tid; e < NB * NB; e += NWARPS * 32) {
            const int k = e % NB, k2 = e / NB; // Tg[k + k2 * NB] = T[k][k2]
            const double v = __ldcg(Tg + e);
            if (a.mode == 0) Ts[k2 * TLD + k] = v;   // Y' = W' T
            else Ts[k * TLD + k2] = v;               // Y' = W' T'
        }
    }
    const int c0 = a.c_lo + 8 * (blockIdx.x * NWARPS + warp);
    const bool active = c0 < n;
    const int q = lane & 3, rl = lane >> 2;
    double* Ccol = a.Cm + (size_t)chain * n * n + (size_t)(active ? c0 + rl : 0) * n + j + 2 * q;
    double c[RT][2];
#pragma unroll
    for (int t = 0; t < RT; ++t) {
        if (active && t < mt) {
            const double2 v = __ldcg(reinterpret_cast<const double2*>(Ccol + 8 * t));
            c[t][0] = v.x; c[t][1] = v.y;
        } else { c[t][0] = 0.0; c[t][1] = 0.0; }
    }
    __syncthreads();          // the barrier is initialised for everybody
    mbar_wait(mbar, 0);
    // unit lower trapezoid: the top NB x NB block still carries R on and above its diagonal
    for (int e = tid; e < NB * NB; e += NWARPS * 32) {
        const int k = e / NB, r = e % NB;
        if (r <= k) Vs[(size_t)k * ldv + r] = (r == k) ? 1.0 : 0.0;
    }
    __syncthreads();
    for (int e = tid; e < (NB / 2) * m; e += NWARPS * 32) {
        const int kp = e / m, r = e - kp * m;   // two k at a time: 16-byte stores
        *reinterpret_cast<double2*>(Vt + (size_t)r * TLD + 2 * kp) =
            make_double2(Vs[(size_t)(2 * kp) * ldv + r], Vs[(size_t)(2 * kp + 1) * ldv + r]);
    }
    __syncthreads();
    if (!active) return;
    double nrm = qrb::apply_strip<RT>(c, mt, NB / 8, Vs, ldv, Vt, Ts, lane);
    griddep_launch();
#pragma unroll
    for (int t = 0; t < RT; ++t)
        if (t < mt) *reinterpret_cast<double2*>(Ccol + 8 * t) = make_double2(c[t][0], c[t][1]);
    if (a.mode == 0) {
        nrm += __shfl_xor_sync(0xffffffffu, nrm, 1);
        nrm += __shfl_xor_sync(0xffffffffu, nrm, 2);
        if (q == 0) a.norms[(size_t)chain * n + c0 + rl] = nrm;
    }
}

// ---------------------------------------------------------------------------------------------- form Q in one launch
// Q = H_0 H_1 ... H_{npan-1} I. The columns of Q are independent, so a warp keeps its 8-column strip in registers for
// ALL block reflectors (25 at N = 400) instead of one launch of the apply kernel per reflector: no launch gaps, the
// strip never travels, and the next reflector block arrives by TMA while the current one is applied. The strip's tiles
// are counted from the last row upwards (apply_strip<RT, true>): the block of panel p acts on the rows >= 16 p, a prefix
// of the tiles that grows as p falls. Panels p > (c0 + 7) / 16 leave the strip at column c0 alone (it is still e_c there).
template <int RT, int NWARPS>
__global__ void __launch_bounds__(NWARPS * 32) qrb_formq_kernel(int n, const double* __restrict__ Afac,
                                                                const double* __restrict__ Tfac, double* __restrict__ U)
{
    extern __shared__ __align__(16) double sm[];
    const int chain = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = lane & 3, rl = lane >> 2;
    const int ldv = n + 8, npan = n / NB;            // n % 16 == 0: ldv % 16 == 8
    double* Vs = sm;                                  // [2][NB][ldv]  V[r][k] at k * ldv + r, r relative to the panel's first row
    double* Vt = Vs + (size_t)2 * NB * ldv;           // [n][TLD]      V[r][k] at r * TLD + k
    double* Ts = Vt + (size_t)n * TLD;                // [NB][TLD]     T[k][k2] at k * TLD + k2
    uint64_t* mbar = reinterpret_cast<uint64_t*>(Ts + (size_t)NB * TLD);   // [2]
    const double* A = Afac + (size_t)chain * n * n;
    const int cta_c0 = blockIdx.x * NWARPS * 8, c0 = cta_c0 + 8 * warp;
    const bool active = c0 < n;
    const int p_hi = (min(n, cta_c0 + NWARPS * 8) - 1) / NB;
    const int p_w = active ? (c0 + 7) / NB : -1;
    if (tid == 0) { mbar_init(mbar, 1); mbar_init(mbar + 1, 1); }
    griddep_wait();
    double c[RT][2];
#pragma unroll
    for (int t = 0; t < RT; ++t) {
        const int row = n - 8 * (t + 1) + 2 * q;
        c[t][0] = (active && row == c0 + rl) ? 1.0 : 0.0;
        c[t][1] = (active && row + 1 == c0 + rl) ? 1.0 : 0.0;
    }
    __syncthreads();
    auto issue = [&](int p, int buf) {
        const int j = p * NB;
        const uint32_t colbytes = (uint32_t)(n - j) * sizeof(double);
        mbar_expect_tx(mbar + buf, NB * colbytes);
        for (int k = 0; k < NB; ++k) bulk_g2s(Vs + (size_t)(buf * NB + k) * ldv, A + (size_t)(j + k) * n + j, colbytes, mbar + buf);
    };
    if (tid == 0) issue(p_hi, 0);
    int it = 0;
    for (int p = p_hi; p >= 0; --p, ++it) {
        const int buf = it & 1, m = n - p * NB, mt = m / 8;
        double* V = Vs + (size_t)buf * NB * ldv;
        {
            const double* Tg = Tfac + ((size_t)chain * npan + p) * NB * NB;   // Tg[k + k2 * NB] = T[k][k2]
            for (int e = tid; e < NB * NB; e += NWARPS * 32) Ts[(e % NB) * TLD + e / NB] = __ldcg(Tg + e);
        }
        mbar_wait(mbar + buf, (it >> 1) & 1);
        // unit lower trapezoid: the top NB x NB block still carries R on and above its diagonal
        for (int e = tid; e < NB * NB; e += NWARPS * 32) {
            const int k = e / NB, r = e % NB;
            if (r <= k) V[(size_t)k * ldv + r] = (r == k) ? 1.0 : 0.0;
        }
        __syncthreads();
        for (int e = tid; e < (NB / 2) * m; e += NWARPS * 32) {
            const int kp = e / m, r = e - kp * m;
            *reinterpret_cast<double2*>(Vt + (size_t)r * TLD + 2 * kp) =
                make_double2(V[(size_t)(2 * kp) * ldv + r], V[(size_t)(2 * kp + 1) * ldv + r]);
        }
        __syncthreads();
        // the other buffer was last read by the strips of the previous panel: free since the barriers above
        if (tid == 0 && p > 0) issue(p - 1, buf ^ 1);
        if (p <= p_w) qrb::apply_strip<RT, true>(c, mt, 1 << 30, V, ldv, Vt, Ts, lane, m);
        __syncthreads();   // Vt and Ts are free
    }
    griddep_launch();
    if (active) {
        double* Ucol = U + (size_t)chain * n * n + (size_t)(c0 + rl) * n + 2 * q;
#pragma unroll
        for (int t = 0; t < RT; ++t)
            if (8 * (t + 1) <= n) *reinterpret_cast<double2*>(Ucol + n - 8 * (t + 1)) = make_double2(c[t][0], c[t][1]);
    }
}

size_t formq_smem(int n)
{
    return ((size_t)2 * NB * (n + 8) + (size_t)n * TLD + (size_t)NB * TLD + 2) * sizeof(double);
}

__global__ void qrb_identity_kernel(double* U, int n)
{
    double* u = U + (size_t)blockIdx.y * n * n;
    for (size_t e = blockIdx.x * blockDim.x + threadIdx.x; e < (size_t)n * n; e += (size_t)gridDim.x * blockDim.x)
        u[e] = (e % n == e / n) ? 1.0 : 0.0;
}

// T = D^-1 R P' (apply_pivot) or D^-1 R in position order with a zero lower triangle; pivot = perm
__global__ void qrb_emit_kernel(const double* A, const double* D, long long sD, const int* perm, double* T, int* pivot,
                                long long sP, int n, int apply_pivot)
{
    const int chain = blockIdx.y;
    griddep_wait();
    const double* a = A + (size_t)chain * n * n;
    const double* d = D + (size_t)chain * sD;
    const int* pm = perm + (size_t)chain * n;
    double* t = T + (size_t)chain * n * n;
    for (int pos = blockIdx.x; pos < n; pos += gridDim.x) {
        const int dst = apply_pivot ? pm[pos] : pos;
        for (int i = threadIdx.x; i < n; i += blockDim.x) t[i + (size_t)dst * n] = (i <= pos) ? a[i + (size_t)pos * n] / d[i] : 0.0;
        if (threadIdx.x == 0) pivot[(size_t)chain * sP + pos] = pm[pos];
    }
}

size_t apply_smem(int n, int j)
{
    const int m = n - j;
    const int ldv = m + ((8 - (m & 15)) & 15);
    return ((size_t)NB * ldv + (size_t)m * TLD + (size_t)NB * TLD + 2) * sizeof(double);
}

template <int RT, int NWARPS>
int launch_apply(const ApplyArgs& a, int batch, cudaStream_t st)
{
    const int strips = (a.n - a.c_lo) / 8;
    if (strips <= 0) return 0;
    dim3 grid((strips + NWARPS - 1) / NWARPS, batch);
    MCB_CUDA(launch_pdl(qrb_apply_kernel<RT, NWARPS>, grid, dim3(NWARPS * 32), apply_smem(a.n, a.j), st, a));
    return 0;
}

constexpr int AW = 8; // warps (strips) per CTA of the apply kernel

// The whole factorisation on caller-provided scratch: norms [batch][n] doubles, perm [batch][n] ints,
// Tfac [batch][n/NB][NB*NB] doubles. Launches only; no allocation.
int udt_blocked_on_scratch(int n, int batch, double* X, double* U, double* D, long long sD, double* T, int* pivot,
                           long long sP, int apply_pivot, double* norms, int* perm, double* Tfac, cudaStream_t st)
{
    const int npan = n / NB;
    const bool small = n <= 192;
    QrbArgs a;
    a.n = n; a.A = X; a.D = D; a.sD = sD; a.norms = norms; a.perm = perm; a.Tfac = Tfac;
    ApplyArgs ap;
    ap.n = n; ap.A = X; ap.Tfac = Tfac; ap.norms = norms;
    for (int p = 0; p < npan; ++p) {
        a.j = p * NB;
        if (small) MCB_CUDA(launch_pdl(qrb_panel_kernel<128, 4, 6, 4>, dim3(batch), dim3(128), panel_smem<4, 4>(n), st, a));
        else MCB_CUDA(launch_pdl(qrb_panel_kernel<512, 8, 7, 8>, dim3(batch), dim3(512), panel_smem<8, 8>(n), st, a));
        ap.j = a.j; ap.c_lo = a.j + NB; ap.mode = 0; ap.Cm = X;
        if (small) { if (launch_apply<24, AW>(ap, batch, st)) return -2; }
        else { if (launch_apply<50, AW>(ap, batch, st)) return -2; }
    }
#ifdef SWD_TIMING
    {
        static int shown = 0;
        if (!small && shown++ < 2) {
            long long h[16];
            cudaStreamSynchronize(st);
            cudaMemcpyFromSymbol(h, qrb_clk, sizeof(h));
            fprintf(stderr, "qrb_panel n=%d panel 5 cycles: norms %lld select %lld lists+gather %lld steps %lld top+gram+T %lld store %lld | total %lld\n",
                    n, h[1] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3], h[5] - h[4], h[6] - h[5], h[6] - h[0]);
        }
    }
#endif
    // explicit Q: the block reflectors in reverse order on the identity; panel p leaves columns < p*NB alone
    // MCB_FORMQ_FUSED=0: one launch of the apply kernel per reflector block (comparisons)
    static const bool fused_formq = !(getenv("MCB_FORMQ_FUSED") && atoi(getenv("MCB_FORMQ_FUSED")) == 0);
    if (!small && fused_formq) {
        MCB_CUDA(launch_pdl(qrb_formq_kernel<50, AW>, dim3((n + AW * 8 - 1) / (AW * 8), batch), dim3(AW * 32), formq_smem(n), st, n,
                            (const double*)X, (const double*)Tfac, U));
    } else {
        qrb_identity_kernel<<<dim3(8, batch), 256, 0, st>>>(U, n);
        MCB_CUDA(cudaGetLastError());
        ap.mode = 1; ap.Cm = U;
        for (int p = npan - 1; p >= 0; --p) {
            ap.j = p * NB; ap.c_lo = p * NB;
            if (small) { if (launch_apply<24, AW>(ap, batch, st)) return -2; }
            else { if (launch_apply<50, AW>(ap, batch, st)) return -2; }
        }
    }
    MCB_CUDA(launch_pdl(qrb_emit_kernel, dim3(n < 64 ? n : 64, batch), dim3(128), 0, st, (const double*)X, (const double*)D, sD,
                        (const int*)perm, T, pivot, sP, n, apply_pivot));
    return 0;
}

} // namespace

bool udt_blocked_supported(int n) { return n % NB == 0 && n >= 128 && n <= 400; }

int init_udt_blocked_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(qrb_panel_kernel<128, 4, 6, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)panel_smem<4, 4>(192)));
    MCB_CUDA(cudaFuncSetAttribute(qrb_panel_kernel<512, 8, 7, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)panel_smem<8, 8>(400)));
    MCB_CUDA(cudaFuncSetAttribute(qrb_apply_kernel<24, AW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)apply_smem(192, 0)));
    MCB_CUDA(cudaFuncSetAttribute(qrb_apply_kernel<50, AW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)apply_smem(400, 0)));
    MCB_CUDA(cudaFuncSetAttribute(qrb_formq_kernel<50, AW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)formq_smem(400)));
    return 0;
}

// Auxiliary streams for launch_udt_blocked: the panel kernel is a latency chain on one CTA per matrix (42 us per panel at
// N = 400 whatever the batch), the apply kernel is throughput-bound. With the batch cut into groups that run on their own
// streams, the panel kernels of one group overlap the apply kernels of the others.
// One set of auxiliary streams and events per (device, caller's stream): two handles driven from two host threads must
// not share the fork / join events (a wait would pick up the other thread's record and let a group start before its own
// stream's earlier work is done), and sharing the streams would serialise them. A set lives until the process exits; a
// stream handle value that the runtime hands out again after a destroy simply finds its set.
namespace {
constexpr int MAX_GROUPS = 4;
struct GroupStreams {
    cudaStream_t s[MAX_GROUPS - 1] = {};
    cudaEvent_t fork = nullptr, join[MAX_GROUPS - 1] = {};
};
std::map<std::pair<int, cudaStream_t>, GroupStreams> g_groups;
std::mutex g_groups_mutex;

int group_streams(cudaStream_t st, GroupStreams** out)
{
    int dev = 0;
    MCB_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(g_groups_mutex);
    auto it = g_groups.find({dev, st});
    if (it == g_groups.end()) {
        GroupStreams g;
        for (int i = 0; i < MAX_GROUPS - 1; ++i) {
            MCB_CUDA(cudaStreamCreateWithFlags(&g.s[i], cudaStreamNonBlocking));
            MCB_CUDA(cudaEventCreateWithFlags(&g.join[i], cudaEventDisableTiming));
        }
        MCB_CUDA(cudaEventCreateWithFlags(&g.fork, cudaEventDisableTiming));
        it = g_groups.emplace(std::make_pair(dev, st), g).first;
    }
    *out = &it->second;   // std::map nodes do not move
    return 0;
}
} // namespace

int scratch_alloc(void** p, size_t bytes, cudaStream_t st)
{
    static const bool own = !(getenv("MCB_SCRATCH_POOL") && atoi(getenv("MCB_SCRATCH_POOL")) == 0);
    if (!own) { MCB_CUDA(cudaMallocAsync(p, bytes, st)); return 0; }
    static cudaMemPool_t pools[64] = {};
    static std::mutex pool_mutex;
    int dev = 0;
    MCB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("scratch pool: device index out of range"); return -1; }
    cudaMemPool_t pool = nullptr;
    {
        std::lock_guard<std::mutex> lock(pool_mutex);
        if (!pools[dev]) {
            cudaMemPoolProps props = {};
            props.allocType = cudaMemAllocationTypePinned;
            props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = dev;
            MCB_CUDA(cudaMemPoolCreate(&pools[dev], &props));
            unsigned long long keep = ~0ull;
            MCB_CUDA(cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep));
        }
        pool = pools[dev];
    }
    MCB_CUDA(cudaMallocFromPoolAsync(p, bytes, pool, st));
    return 0;
}

// same contract as launch_udt (udt.cuh); X is the workspace the factorisation happens in
int launch_udt_blocked(int n, int batch, double* X, double* U, double* D, long long sD, double* T, int* pivot,
                       long long sP, int apply_pivot, cudaStream_t st)
{
    const int npan = n / NB;
    double* norms = nullptr; int* perm = nullptr; double* Tfac = nullptr;
    if (scratch_alloc((void**)&norms, (size_t)batch * n * sizeof(double), st)) return -2;
    if (scratch_alloc((void**)&perm, (size_t)batch * n * sizeof(int), st)) return -2;
    if (scratch_alloc((void**)&Tfac, (size_t)batch * npan * NB * NB * sizeof(double), st)) return -2;
    // MCB_UDT_GROUPS=k: k groups (1 .. 4) instead of the default (comparisons)
    static const int env_groups = getenv("MCB_UDT_GROUPS") ? atoi(getenv("MCB_UDT_GROUPS")) : 0;
    int groups = 1;
    // two groups: 1558 us per batched 400 x 400 factorisation of 32 against 1750 on one stream. Three or four groups give
    // the same time as two: the factorisation is the latency chain of one group (25 x (panel + apply) + form Q) as soon as
    // the apply kernels of a group fit the SMs in one wave
    if (n > 192 && batch >= 8) groups = env_groups >= 1 && env_groups <= MAX_GROUPS ? env_groups : 2;
    int rc = 0;
    {
        LaunchScope scope(LC_UDT, st, 8.0 / 3.0 * n * (double)n * n * batch);
        GroupStreams* gs = nullptr;
        if (groups > 1 && group_streams(st, &gs)) return -2;
        if (groups > 1) MCB_CUDA(cudaEventRecord(gs->fork, st));
        const size_t nn = (size_t)n * n;
        for (int g = 0; g < groups && rc == 0; ++g) {
            const int b0 = (int)((long long)batch * g / groups), b1 = (int)((long long)batch * (g + 1) / groups);
            cudaStream_t s = g == 0 ? st : gs->s[g - 1];
            if (g > 0) MCB_CUDA(cudaStreamWaitEvent(s, gs->fork, 0));
            rc = udt_blocked_on_scratch(n, b1 - b0, X + b0 * nn, U + b0 * nn, D + (size_t)b0 * sD, sD, T + b0 * nn,
                                        pivot + (size_t)b0 * sP, sP, apply_pivot, norms + (size_t)b0 * n, perm + (size_t)b0 * n,
                                        Tfac + (size_t)b0 * npan * NB * NB, s);
            if (g > 0) {
                MCB_CUDA(cudaEventRecord(gs->join[g - 1], s));
                MCB_CUDA(cudaStreamWaitEvent(st, gs->join[g - 1], 0));
            }
        }
    }
    MCB_CUDA(cudaFreeAsync(norms, st));
    MCB_CUDA(cudaFreeAsync(perm, st));
    MCB_CUDA(cudaFreeAsync(Tfac, st));
    return rc;
}

} // namespace mcb

// ---------------------------------------------------------------------------------------------- allocation-free entry
// mcb_linalg_udt_noalloc (include/mcb200.h): the blocked factorisation above for callers that may not call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it). No counterpart in the reference.
// The scratch the factorisation needs (trailing column norms, the column permutation
// and the NB x NB block-reflector factors) lives in a file-scope device array of fixed size: n <= 400, batch <= 8.
// Every GPU that loads the library holds these 448 000 bytes until the process exits. All calls share the array, so
// two calls that overlap on the GPU (different streams, or different host threads) race; calls on one stream are fine.
// Nothing in it has to start at zero: every element that is read was written earlier in the same call.
#include "../../include/mcb200.h"
#include <mutex>

namespace mcb {
namespace {

constexpr int NOALLOC_MAX_N = 400, NOALLOC_MAX_BATCH = 8;
constexpr size_t NA_NORMS_BYTES = (size_t)NOALLOC_MAX_BATCH * NOALLOC_MAX_N * sizeof(double);                     //  25 600
constexpr size_t NA_TFAC_BYTES = (size_t)NOALLOC_MAX_BATCH * (NOALLOC_MAX_N / NB) * NB * NB * sizeof(double);     // 409 600
constexpr size_t NA_PERM_BYTES = (size_t)NOALLOC_MAX_BATCH * NOALLOC_MAX_N * sizeof(int);                         //  12 800
static __device__ __align__(256) unsigned char qrb_static_scratch[NA_NORMS_BYTES + NA_TFAC_BYTES + NA_PERM_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
unsigned char* g_na_scratch[NA_MAX_DEVICES] = {}; // device address of qrb_static_scratch per device (it is a
                                                  // host-side stand-in in host code, not an address)

// once per device, with that device current: kernel attributes and the device address of the scratch array
int noalloc_setup(int dev, unsigned char** out)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        set_error("no CUDA device: libmcb200 has no CPU fallback");
        return MCB_ERR_CUDA;
    }
    if (dev < 0 || dev >= count || dev >= NA_MAX_DEVICES) { set_error("bad device id"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(dev));
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_scratch[dev]) {
        if (init_udt_blocked_kernels() != 0) return MCB_ERR_CUDA;
        unsigned char* p = nullptr;
        MCB_CUDA(cudaGetSymbolAddress((void**)&p, qrb_static_scratch));
        if (!p) { set_error("cudaGetSymbolAddress returned a null address"); return MCB_ERR_CUDA; }
        g_na_scratch[dev] = p;
    }
    *out = g_na_scratch[dev];
    return 0;
}

} // namespace
} // namespace mcb

extern "C" int mcb_linalg_udt_noalloc(int32_t n, int32_t batch, double* X_dev, double* U_dev, double* D_dev,
                                      double* T_dev, int32_t* pivot_dev, int32_t apply_pivot, int32_t device_id,
                                      void* stream)
{
    using namespace mcb;
    if (!X_dev || !U_dev || !D_dev || !T_dev || !pivot_dev || n <= 0 || batch <= 0) { set_error("bad argument"); return MCB_ERR_ARG; }
    if (!udt_blocked_supported(n) || n > NOALLOC_MAX_N) {
        set_error("mcb_linalg_udt_noalloc: n must be a multiple of 16 with 128 <= n <= 400");
        return MCB_ERR_ARG;
    }
    if (batch > NOALLOC_MAX_BATCH) {
        set_error("mcb_linalg_udt_noalloc: batch exceeds the static scratch (at most 8 matrices per call)");
        return MCB_ERR_ARG;
    }
    if (X_dev == U_dev || X_dev == T_dev || U_dev == T_dev) { set_error("udt: buffers alias"); return MCB_ERR_ARG; }
    unsigned char* base = nullptr;
    const int rc = noalloc_setup(device_id, &base);
    if (rc != 0) return rc;
    double* norms = reinterpret_cast<double*>(base);
    double* Tfac = reinterpret_cast<double*>(base + NA_NORMS_BYTES);
    int* perm = reinterpret_cast<int*>(base + NA_NORMS_BYTES + NA_TFAC_BYTES);
    // launches only: no cudaMalloc / cudaFree / event creation, no synchronisation (the caller owns the stream)
    return udt_blocked_on_scratch(n, batch, X_dev, U_dev, D_dev, n, T_dev, pivot_dev, n, apply_pivot, norms, perm, Tfac,
                                  (cudaStream_t)stream);
}
