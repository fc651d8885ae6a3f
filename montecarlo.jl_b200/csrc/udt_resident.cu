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

constexpr int RW = 16;              // warps per CTA
constexpr int RTH = RW * 32;
constexpr int RMAXN = 144;
constexpr int RRT = RMAXN / 8;      // row tiles of a strip
constexpr int RTPW = (RRT + RW - 1) / RW;
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
    if (tid == 0) {
        mbar_init(mbar, 1);
        const uint32_t colbytes = (uint32_t)n * sizeof(double);
        mbar_expect_tx(mbar, (uint32_t)n * colbytes);
        for (int c = 0; c < n; ++c) bulk_g2s(A + (size_t)c * ld, Xg + (size_t)c * n, colbytes, mbar);
    }
    for (int c = tid; c < n; c += RTH) perm[c] = c;
    __syncthreads();
    mbar_wait(mbar, 0);
    for (int c = warp; c < n; c += RW) {
        const double* col = A + (size_t)c * ld;
        double s = 0.0;
        for (int i = lane; i < n; i += 32) s = fma(col[i], col[i], s);
        s = warp_sum(s);
        if (lane == 0) nrm[c] = (s == s) ? s : INFINITY;  // NaN columns sort first: the ranking needs a total order
    }

    const int npan = n / NB;
    for (int p = 0; p < npan; ++p) {
        const int j = p * NB, m = n - j, mt = m / 8;
        __syncthreads();
        // ---- the NB largest remaining columns, by decreasing norm (ties: lower position first)
        qrb::select_largest<RW, 1>(nrm, j, n, tid, reinterpret_cast<unsigned long long*>(part),
                                   reinterpret_cast<int*>(part + 2 * RW), sel);
        // ---- gather the selected columns (all rows: rows < j carry R of earlier panels) into the reflector buffer
        int ps = 0;
        if (tid < NB) ps = perm[sel[tid]];
        {
            const int k = warp;                         // RW == NB: one column per warp
            const double* col = A + (size_t)sel[k] * ld;
            for (int i = lane; i < n; i += 32) Vs[vs_index(k, i, ld)] = col[i];
        }
        if (tid == 0) {
            int nd = 0, nh = 0;
            for (int pp = j; pp < j + NB; ++pp) {
                bool chosen = false;
                for (int s = 0; s < NB; ++s) chosen |= (sel[s] == pp);
                if (!chosen) disp[nd++] = pp;
            }
            for (int s = 0; s < NB; ++s)               // holes: selected positions beyond the panel
                if (sel[s] >= j + NB) hole[nh++] = sel[s];
            *ndisp = nd;
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
        // ---- the panel in registers, NB Householder steps
        double x[RTPW][4];
#pragma unroll
        for (int u = 0; u < RTPW; ++u) {
            const int t = warp + RW * u;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) x[u][cc] = (t < mt) ? Vs[vs_index(4 * q + cc, j + 8 * t + rl, ld)] : 0.0;
        }
        qrb::panel_steps<RW, RTPW>(x, mt, warp, lane, part, rowk, taus, dvec + j);
#pragma unroll
        for (int u = 0; u < RTPW; ++u) {
            const int t = warp + RW * u;
            if (t < mt) {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) Vs[vs_index(4 * q + cc, j + 8 * t + rl, ld)] = x[u][cc];
            }
        }
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
    }
    __syncthreads();
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
    // ---- U leaves by bulk copies
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
        double* Ug = a.U + (size_t)chain * n * n;
        for (int c = 0; c < n; ++c) bulk_s2g(Ug + (size_t)c * n, A + (size_t)c * ld, (uint32_t)n * sizeof(double));
        bulk_commit();
        bulk_wait_all();
    }
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
    MCB_CUDA(cudaMallocAsync(&a.Tfac, (size_t)batch * (n / NB) * NB * NB * sizeof(double), st));
    {
        LaunchScope scope(LC_UDT, st, 8.0 / 3.0 * n * (double)n * n * batch);
        udt_resident_kernel<<<batch, RTH, resident_smem(n), st>>>(a);
    }
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaFreeAsync(a.Tfac, st));
    return 0;
}

} // namespace mcb
