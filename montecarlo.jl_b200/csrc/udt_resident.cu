// udt_AVX_pivot! (src/linalg/UDT.jl:192-306) for N = 128 / 144 (config C) as ONE kernel, one CTA per chain, the whole
// matrix resident in shared memory (144 x 152 doubles = 175 KB of the 227 KB an SM offers): the blocked panel-pivoted
// factorisation of udt_blocked.cu without its 29 launches per call and without a global-memory round trip per panel.
//   load     X -> shared memory by TMA bulk copies (one per column), squared column norms
//   panel p  the NB = 16 largest remaining columns are gathered into a reflector buffer (swizzled, qr_blocked.cuh),
//            displaced columns fill the holes; NB Householder steps on registers, one block barrier each;
//            V'V and the block-reflector factor T on the tensor pipe; every warp then applies I - V T' V' to its
//            8-column strips of the trailing matrix with three DMMA products per strip, operands from shared memory,
//            and leaves the updated squared norms for the next selection
//   emit     T = D^-1 R P' (or the "dirty" D^-1 R + pivot), D
//   form Q   in place, panels in reverse order (UDT.jl:250-266 blocked): the panel's columns become unit vectors, the
//            block reflector is applied to the columns from the panel on; U leaves by TMA bulk copies
// Same pivot order and arithmetic as udt_blocked.cu (same device functions), hence the same results.
#include "udt.cuh"
#include "prof.cuh"
#include "qr_blocked.cuh"
#include "tma.cuh"

namespace mcb {

namespace {

using qrb::NB;
using qrb::TLD;
using qrb::vs_index;

#ifdef UDTR_TIMING
__device__ long long udtr_clk[64];
#define UDTR_T(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) udtr_clk[i] = clock64(); } while (0)
#else
#define UDTR_T(i) do { } while (0)
#endif

constexpr int RW = 16;              // warps per CTA
constexpr int RTH = RW * 32;
constexpr int RMAXN = 144;
constexpr int RRT = RMAXN / 8;      // row tiles of a strip
constexpr int PW = 4;               // warps that run the Householder steps of a panel (the others wait: more warps
                                    // only add redundant per-step work, 1900 against 900 cycles per step)
constexpr int RTPW = (RRT + PW - 1) / PW;
constexpr int RGW = 4;              // warps of the Gram product

struct ResArgs {
    int n = 0;
    const double* X = nullptr;
    double* U = nullptr;
    double* D = nullptr; long long sD = 0;
    double* T = nullptr;
    int* pivot = nullptr; long long sP = 0;
    int apply_pivot = 0;
    double* Tfac = nullptr;  // [batch][n/NB][NB*NB] block-reflector factors of the panels (kept for form Q)
};

size_t resident_smem(int n)
{
    const size_t ld = n + 8;
    return ((size_t)n * ld + NB * ld + NB * TLD + 2 * n + 2 * RW * NB + 2 * NB + NB * NB + (size_t)RGW * NB * NB + NB + 2) *
               sizeof(double) +
           ((size_t)n + 3 * NB + 4) * sizeof(int);
}

__global__ void __launch_bounds__(RTH, 1) udt_resident_kernel(ResArgs a)
{
    extern __shared__ __align__(16) double sm[];
    const int n = a.n, ld = n + 8, chain = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = lane & 3, rl = lane >> 2;
    double* A = sm;                                  // [n][ld] column-major
    double* Vs = A + (size_t)n * ld;                 // [NB][ld] reflector block, swizzled rows (vs_index)
    double* Ts = Vs + (size_t)NB * ld;               // [NB][TLD]
    double* nrm = Ts + NB * TLD;                     // [n] squared norms of the active rows of every column
    double* dvec = nrm + n;                          // [n] D
    double* part = dvec + n;                         // [2][RW][NB]
    double* rowk = part + 2 * RW * NB;               // [2][NB]
    double* Ssm = rowk + 2 * NB;                     // [NB][NB]
    double* Sred = Ssm + NB * NB;                    // [RGW][NB*NB]
    double* taus = Sred + (size_t)RGW * NB * NB;     // [NB]
    uint64_t* mbar = reinterpret_cast<uint64_t*>(taus + NB);
    int* perm = reinterpret_cast<int*>(mbar + 2);    // [n]
    int* sel = perm + n;                             // [NB]
    int* disp = sel + NB;                            // [NB]
    int* hole = disp + NB;                           // [NB]
    int* ndisp = hole + NB;
    const double* Xg = a.X + (size_t)chain * n * n;
    double* Tfac = a.Tfac + (size_t)chain * (n / NB) * NB * NB;

    // ---- load
    UDTR_T(0);
    if (tid == 0) {
        mbar_init(mbar, 1);
        mbar_expect_tx(mbar, (uint32_t)n * (uint32_t)n * sizeof(double));
    }
    for (int c = tid; c < n; c += RTH) perm[c] = c;
    __syncthreads();
    // one column per thread: the copies are issued in parallel (the barrier counts bytes, whoever issues them)
    if (tid < n) bulk_g2s(A + (size_t)tid * ld, Xg + (size_t)tid * n, (uint32_t)n * sizeof(double), mbar);
    mbar_wait(mbar, 0);
    for (int c = warp; c < n; c += RW) {
        const double* col = A + (size_t)c * ld;
        double s = 0.0;
        for (int i = lane; i < n; i += 32) s = fma(col[i], col[i], s);
        s = warp_sum(s);
        if (lane == 0) nrm[c] = (s == s) ? s : INFINITY;  // NaN columns sort first: the ranking needs a total order
    }

    const int npan = n / NB;
    UDTR_T(1);
    for (int p = 0; p < npan; ++p) {
        const int j = p * NB, m = n - j, mt = m / 8;
        __syncthreads();
        if (p == 1) UDTR_T(10);
        // ---- the NB largest remaining columns, by decreasing norm (ties: lower position first)
        qrb::select_by_rank<RTH>(nrm, j, n, tid, Sred, sel);   // Sred is free until the Gram product
        if (p == 1) UDTR_T(11);
        // ---- gather the selected columns (all rows: rows < j carry R of earlier panels) into the reflector buffer
        int ps = 0;
        if (tid < NB) ps = perm[sel[tid]];
        {
            const int k = warp;                         // RW == NB: one column per warp
            const double* col = A + (size_t)sel[k] * ld;
            for (int i = lane; i < n; i += 32) Vs[vs_index(k, i, ld)] = col[i];
        }
        if (warp == RW - 1) {
            // displaced panel positions (by increasing position) and holes = selected positions beyond the panel (by
            // increasing slot), by ballots: lane = slot / panel position (one thread walking the 16 x 16 comparisons took
            // about 1.5 k cycles while the other warps waited at the barrier)
            const int sv = lane < NB ? sel[lane] : 0x7fffffff;
            const bool inside = lane < NB && sv < j + NB;
            const unsigned chosen = __reduce_or_sync(0xffffffffu, inside ? 1u << (sv - j) : 0u);
            const unsigned outside = __ballot_sync(0xffffffffu, lane < NB && !inside);
            const unsigned freed = ~chosen & 0xffffu;
            if (lane < NB && ((freed >> lane) & 1u)) disp[__popc(freed & ((1u << lane) - 1u))] = j + lane;
            if (lane < NB && !inside) hole[__popc(outside & ((1u << lane) - 1u))] = sv;
            if (lane == 0) *ndisp = __popc(freed);
        }
        __syncthreads();
        // ---- displaced columns fill the holes
        const int nd = *ndisp;
        int moved = 0;
        if (tid < nd) moved = perm[disp[tid]];
        if (warp < nd) {
            const double* src = A + (size_t)disp[warp] * ld;
            double* dst = A + (size_t)hole[warp] * ld;
            for (int i = lane; i < n; i += 32) dst[i] = src[i];
        }
        __syncthreads();
        if (tid < nd) perm[hole[tid]] = moved;
        if (tid < NB) perm[j + tid] = ps;
        if (p == 1) UDTR_T(12);
        // ---- the panel in registers of PW warps, NB Householder steps (named barrier of these warps per step)
        __syncthreads();   // `part` was the selection's scratch
        if (p == 1) UDTR_T(12);
        if (warp < PW) {
            double x[RTPW][4];
#pragma unroll
            for (int u = 0; u < RTPW; ++u) {
                const int t = warp + PW * u;
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) x[u][cc] = (t < mt) ? Vs[vs_index(qrb::panel_col(q, cc), j + 8 * t + rl, ld)] : 0.0;
            }
            qrb::panel_steps<PW, RTPW, 1>(x, mt, warp, lane, part, rowk, taus, dvec + j);
#pragma unroll
            for (int u = 0; u < RTPW; ++u) {
                const int t = warp + PW * u;
                if (t < mt) {
#pragma unroll
                    for (int cc = 0; cc < 4; ++cc) Vs[vs_index(qrb::panel_col(q, cc), j + 8 * t + rl, ld)] = x[u][cc];
                }
            }
        }
        if (p == 1) UDTR_T(13);
        __syncthreads();
        // ---- R and V into the matrix; the buffer keeps the unit lower trapezoid
        {
            const int k = warp;
            double* col = A + (size_t)(j + k) * ld;
            for (int i = lane; i < n; i += 32) col[i] = Vs[vs_index(k, i, ld)];
        }
        __syncthreads();
        for (int e = tid; e < NB * NB; e += RTH) {
            const int k = e / NB, r = e % NB;
            if (r <= k) Vs[vs_index(k, j + r, ld)] = (r == k) ? 1.0 : 0.0;
        }
        __syncthreads();
        if (p == 1) UDTR_T(14);
        qrb::panel_gram_sw<RGW>(Vs, ld, j, mt, warp, lane, RTH, tid, Sred, Ssm);
        if (warp == 0) qrb::panel_tfactor(Ssm, taus, lane, Sred);   // Sred[0 .. 255] is free again: T, column-major
        __syncthreads();
        for (int e = tid; e < NB * NB; e += RTH) {
            const int k = e % NB, k2 = e / NB;                       // Sred[k + k2 * NB] = T[k][k2]
            const double v = Sred[e];
            Ts[k2 * TLD + k] = v;                                    // factorisation: Y' = W' T
            Tfac[(size_t)p * NB * NB + e] = v;
        }
        __syncthreads();
        if (p == 1) UDTR_T(15);
        // ---- trailing update: A[j:, j+NB:] <- (I - V T' V') A[j:, j+NB:], one 8-column strip per warp and round
        for (int strip = warp; strip < (m - NB) / 8; strip += RW) {
            const int c0 = j + NB + 8 * strip;
            double* Ccol = A + (size_t)(c0 + rl) * ld + j + 2 * q;
            double c[RRT][2];
#pragma unroll
            for (int t = 0; t < RRT; ++t) {
                if (t < mt) {
                    const double2 v = *reinterpret_cast<const double2*>(Ccol + 8 * t);
                    c[t][0] = v.x; c[t][1] = v.y;
                } else { c[t][0] = 0.0; c[t][1] = 0.0; }
            }
            double s = qrb::apply_strip_sw<RRT>(c, mt, NB / 8, Vs, ld, j, Ts, lane);
#pragma unroll
            for (int t = 0; t < RRT; ++t)
                if (t < mt) *reinterpret_cast<double2*>(Ccol + 8 * t) = make_double2(c[t][0], c[t][1]);
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            if (q == 0) nrm[c0 + rl] = (s == s) ? s : INFINITY;
        }
        if (p == 1) { __syncthreads(); UDTR_T(16); }
    }
    __syncthreads();
    UDTR_T(2);
    // ---- emit T = D^-1 R P' (apply_pivot) or D^-1 R in position order with a zero lower triangle; pivot; D
    {
        double* Tg = a.T + (size_t)chain * n * n;
        for (int pos = warp; pos < n; pos += RW) {
            const int dst = a.apply_pivot ? perm[pos] : pos;
            const double* col = A + (size_t)pos * ld;
            for (int i = lane; i < n; i += 32) Tg[i + (size_t)dst * n] = (i <= pos) ? col[i] / dvec[i] : 0.0;
        }
        for (int i = tid; i < n; i += RTH) {
            a.D[(size_t)chain * a.sD + i] = dvec[i];
            a.pivot[(size_t)chain * a.sP + i] = perm[i];
        }
    }
    __syncthreads();
    UDTR_T(3);
    // ---- form Q in place: Q = H_1 ... H_P I, block reflectors in reverse order
    for (int p = npan - 1; p >= 0; --p) {
        const int j = p * NB, m = n - j, mt = m / 8;
        {
            const int k = warp;
            const double* col = A + (size_t)(j + k) * ld;
            for (int i = j + lane; i < n; i += 32) {
                const int r = i - j;
                Vs[vs_index(k, i, ld)] = r > k ? col[i] : (r == k ? 1.0 : 0.0);
            }
        }
        for (int e = tid; e < NB * NB; e += RTH) {
            const int k = e % NB, k2 = e / NB;                       // Tfac[k + k2 * NB] = T[k][k2]
            Ts[k * TLD + k2] = Tfac[(size_t)p * NB * NB + e];         // form Q: Y' = W' T'
        }
        __syncthreads();
        {
            const int k = warp;
            double* col = A + (size_t)(j + k) * ld;
            for (int i = lane; i < n; i += 32) col[i] = (i == j + k) ? 1.0 : 0.0;
        }
        __syncthreads();
        for (int strip = warp; strip < m / 8; strip += RW) {
            const int c0 = j + 8 * strip;
            double* Ccol = A + (size_t)(c0 + rl) * ld + j + 2 * q;
            double c[RRT][2];
#pragma unroll
            for (int t = 0; t < RRT; ++t) {
                if (t < mt) {
                    const double2 v = *reinterpret_cast<const double2*>(Ccol + 8 * t);
                    c[t][0] = v.x; c[t][1] = v.y;
                } else { c[t][0] = 0.0; c[t][1] = 0.0; }
            }
            qrb::apply_strip_sw<RRT>(c, mt, RRT, Vs, ld, j, Ts, lane);
#pragma unroll
            for (int t = 0; t < RRT; ++t)
                if (t < mt) *reinterpret_cast<double2*>(Ccol + 8 * t) = make_double2(c[t][0], c[t][1]);
        }
        __syncthreads();
    }
    UDTR_T(4);
    // ---- U leaves by bulk copies
    fence_async_smem();
    __syncthreads();
    if (tid < n) {   // one column per thread, every issuing thread waits for its own bulk group
        double* Ug = a.U + (size_t)chain * n * n;
        bulk_s2g(Ug + (size_t)tid * n, A + (size_t)tid * ld, (uint32_t)n * sizeof(double));
        bulk_commit();
        bulk_wait_all();
    }
    UDTR_T(5);
}

} // namespace

bool udt_resident_supported(int n) { return n % NB == 0 && n >= 128 && n <= RMAXN; }

int init_udt_resident_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(udt_resident_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)resident_smem(RMAXN)));
    return 0;
}

// same contract as launch_udt (udt.cuh); X is only read
int launch_udt_resident(int n, int batch, const double* X, double* U, double* D, long long sD, double* T, int* pivot,
                        long long sP, int apply_pivot, cudaStream_t st)
{
    ResArgs a;
    a.n = n; a.X = X; a.U = U; a.D = D; a.sD = sD; a.T = T; a.pivot = pivot; a.sP = sP; a.apply_pivot = apply_pivot;
    if (scratch_alloc((void**)&a.Tfac, (size_t)batch * (n / NB) * NB * NB * sizeof(double), st)) return -2;
    {
        LaunchScope scope(LC_UDT, st, 8.0 / 3.0 * n * (double)n * n * batch);
        udt_resident_kernel<<<batch, RTH, resident_smem(n), st>>>(a);
    }
    MCB_CUDA(cudaGetLastError());
#ifdef UDTR_TIMING
    {
        long long h[64];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(h, udtr_clk, sizeof(h));
        fprintf(stderr, "udt_resident cycles: load+norms %lld | all panels %lld | emit %lld | form Q %lld | store %lld || panel 1: select %lld gather+move %lld steps %lld writeback+patch %lld gram+T %lld apply %lld\n",
                h[1] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3], h[5] - h[4], h[11] - h[10], h[12] - h[11], h[13] - h[12],
                h[14] - h[13], h[15] - h[14], h[16] - h[15]);
    }
#endif
    MCB_CUDA(cudaFreeAsync(a.Tfac, st));
    return 0;
}

} // namespace mcb
