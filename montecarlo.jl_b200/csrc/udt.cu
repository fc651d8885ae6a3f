// One CTA per chain. The matrix lives in shared memory when it fits (B200: 227 KB per CTA),
// otherwise it is worked on in place in global memory (L2 resident). The code is the same for
// both: it works through generic pointers.
#include "udt.cuh"
#include "stab.cuh"
#include <stdlib.h>

namespace mcb {

bool force_general_udt();

namespace {

constexpr int SMEM_LIMIT = 227 * 1024;

struct UdtArgs {
    int n;
    double* X; long long sX;
    double* U; long long sU;
    double* D; long long sD;
    double* T; long long sT;
    int* pivot; long long sP;
    int apply_pivot;
    int mode; // 2: X and U in smem, 1: X in smem, 0: all global
};

__global__ void udt_kernel(UdtArgs a)
{
    extern __shared__ double sm[];
    const int n = a.n;
    const int chain = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    double* tau = sm;
    double* norms = tau + n;
    double* xs = norms + n;
    double* misc = xs + n;             // [0]=maxval [1]=xi
    int* ipiv = (int*)(misc + 2);      // n ints (+1 for jm)
    int* sh_jm = ipiv + n;
    double* mats = misc + 2 + (n + 2 + 1) / 2 + 1;
    double* Xg = a.X + (size_t)chain * a.sX;
    double* Ug = a.U + (size_t)chain * a.sU;
    double* A = a.mode >= 1 ? mats : Xg;
    double* U = a.mode == 2 ? mats + (size_t)n * n : Ug;
    const size_t nn = (size_t)n * n;

    if (a.mode >= 1)
        for (size_t e = tid; e < nn; e += blockDim.x) A[e] = Xg[e];
    for (int i = tid; i < n; i += blockDim.x) ipiv[i] = i;
    __syncthreads();
    // initial column norms
    for (int c = warp; c < n; c += nwarps) {
        double s = 0.0;
        for (int i = lane; i < n; i += 32) { const double v = A[i + (size_t)c * n]; s += v * v; }
        s = warp_sum(s);
        if (lane == 0) norms[c] = s;
    }

    for (int j = 0; j < n; ++j) {
        __syncthreads();
        if (warp == 0) {
            // first strict maximum of norms[j..n)  (indmaxcolumn, UDT.jl:151-168)
            double best = -1.0; int bi = n;
            for (int c = j + lane; c < n; c += 32) {
                const double v = norms[c];
                if (v > best) { best = v; bi = c; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
            }
            if (lane == 0) {
                *sh_jm = bi;
                misc[0] = best;
                misc[1] = A[j + (size_t)bi * n];
            }
        }
        __syncthreads();
        const int jm = *sh_jm;
        const double maxval = misc[0];
        double xi1 = misc[1];
        // reflector! (UDT.jl:133-148)
        double nu = 0.0, tj = 0.0;
        const bool degenerate = (maxval == 0.0);
        if (!degenerate) {
            nu = copysign(sqrt(maxval), xi1);
            xi1 += nu;
            tj = xi1 / nu;
        }
        for (int i = tid; i < n; i += blockDim.x) {
            double t = A[i + (size_t)jm * n];
            if (jm != j) A[i + (size_t)jm * n] = A[i + (size_t)j * n];
            if (!degenerate) {
                if (i == j) t = -nu;
                else if (i > j) t = t / xi1;
            }
            A[i + (size_t)j * n] = t;
            xs[i] = t;
        }
        if (tid == 0) {
            tau[j] = tj;
            if (jm != j) { const int tp = ipiv[jm]; ipiv[jm] = ipiv[j]; ipiv[j] = tp; }
        }
        __syncthreads();
        // reflectorApply! on trailing columns (UDT.jl:32-50), fused with the next step's norms
        for (int c = j + 1 + warp; c < n; c += nwarps) {
            double* col = A + (size_t)c * n;
            double dot = 0.0;
            for (int i = j + 1 + lane; i < n; i += 32) dot += xs[i] * col[i];
            dot = warp_sum(dot);
            const double v = tj * (col[j] + dot);
            __syncwarp();
            if (lane == 0) col[j] -= v;
            double nrm = 0.0;
            for (int i = j + 1 + lane; i < n; i += 32) {
                const double w = col[i] - xs[i] * v;
                col[i] = w;
                nrm += w * w;
            }
            nrm = warp_sum(nrm);
            if (lane == 0) norms[c] = nrm;
        }
    }
    __syncthreads();
    // explicit Q by backward accumulation (UDT.jl:250-266)
    for (size_t e = tid; e < nn; e += blockDim.x) U[e] = 0.0;
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) U[i + (size_t)i * n] = (i == n - 1) ? 1.0 - tau[n - 1] : 1.0;
    for (int k = n - 2; k >= 0; --k) {
        __syncthreads();
        const double* x = A + (size_t)k * n;
        const double tk = tau[k];
        for (int c = k + warp; c < n; c += nwarps) {
            double* col = U + (size_t)c * n;
            double dot = 0.0;
            for (int i = k + 1 + lane; i < n; i += 32) dot += x[i] * col[i];
            dot = warp_sum(dot);
            const double v = tk * (col[k] + dot);
            __syncwarp();
            if (lane == 0) col[k] -= v;
            for (int i = k + 1 + lane; i < n; i += 32) col[i] -= x[i] * v;
        }
    }
    __syncthreads();
    double* Dg = a.D + (size_t)chain * a.sD;
    for (int i = tid; i < n; i += blockDim.x) {
        const double d = fabs(A[i + (size_t)i * n]);
        Dg[i] = d;
        norms[i] = 1.0 / d;
        a.pivot[(size_t)chain * a.sP + i] = ipiv[i];
    }
    __syncthreads();
    double* Tg = a.T + (size_t)chain * a.sT;
    if (a.apply_pivot) { // _apply_pivot! Val(true)  (UDT.jl:283-297)
        for (size_t e = tid; e < nn; e += blockDim.x) {
            const int i = (int)(e % n), j = (int)(e / n);
            Tg[i + (size_t)ipiv[j] * n] = (j >= i) ? norms[i] * A[e] : 0.0;
        }
    } else { // Val(false): D^-1 R in the upper triangle, rest unspecified (UDT.jl:298-306)
        for (size_t e = tid; e < nn; e += blockDim.x) {
            const int i = (int)(e % n), j = (int)(e / n);
            Tg[e] = (j >= i) ? norms[i] * A[e] : A[e];
        }
    }
    if (a.mode == 2)
        for (size_t e = tid; e < nn; e += blockDim.x) Ug[e] = U[e];
}

struct RdivpArgs {
    int n;
    const double* in; long long sIn;
    double* out; long long sOut;
    const double* T; long long sT;
    const int* pivot; long long sP;
    int mode; // 2: O and T in smem, 1: O in smem, 0: global
};

__global__ void rdivp_kernel(RdivpArgs a)
{
    extern __shared__ double sm[];
    const int n = a.n, chain = blockIdx.x, tid = threadIdx.x;
    const size_t nn = (size_t)n * n;
    double* colj = sm;
    double* mats = sm + n;
    const double* in = a.in + (size_t)chain * a.sIn;
    double* outg = a.out + (size_t)chain * a.sOut;
    const double* Tg = a.T + (size_t)chain * a.sT;
    const int* piv = a.pivot + (size_t)chain * a.sP;
    double* O = a.mode >= 1 ? mats : outg;
    const double* T = Tg;
    if (a.mode == 2) {
        double* Ts = mats + nn;
        for (size_t e = tid; e < nn; e += blockDim.x) Ts[e] = Tg[e];
        T = Ts;
    }
    for (size_t e = tid; e < nn; e += blockDim.x) {
        const int i = (int)(e % n), j = (int)(e / n);
        O[e] = in[i + (size_t)piv[j] * n];
    }
    for (int j = 0; j < n; ++j) {
        __syncthreads();
        const double tjj = T[j + (size_t)j * n];
        for (int i = tid; i < n; i += blockDim.x) {
            const double v = O[i + (size_t)j * n] / tjj;
            colj[i] = v;
            if (a.mode >= 1) outg[i + (size_t)j * n] = v; else O[i + (size_t)j * n] = v;
        }
        __syncthreads();
        const int rem = n - j - 1;
        for (size_t e = tid; e < (size_t)rem * n; e += blockDim.x) {
            const int i = (int)(e % n), c = j + 1 + (int)(e / n);
            O[i + (size_t)c * n] -= colj[i] * T[j + (size_t)c * n];
        }
    }
}

} // namespace

static bool g_force_general_udt = false;
bool force_general_udt() { return g_force_general_udt; }
void set_force_general_udt(bool on) { g_force_general_udt = on; }

int init_udt_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(udt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    MCB_CUDA(cudaFuncSetAttribute(rdivp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    if (init_udt_blocked_kernels()) return -2;
    if (init_udt_resident_kernels()) return -2;
    if (init_rdivp_blocked_kernels()) return -2;
    return init_udt_cluster_kernels();
}

int launch_udt(int n, int batch, double* X, long long sX, double* U, long long sU, double* D,
               long long sD, double* T, long long sT, int* pivot, long long sP, int apply_pivot,
               cudaStream_t st)
{
    if (X == T || X == U || U == T) { set_error("udt: buffers alias"); return -1; }
    // blocked panel-pivoted QR with DMMA trailing updates (udt_blocked.cu); MCB_UDT_BLOCKED=0 restores the
    // step-by-step kernels below for comparisons
    static const bool blocked_on = !(getenv("MCB_UDT_BLOCKED") && atoi(getenv("MCB_UDT_BLOCKED")) == 0);
    // MCB_UDT_RESIDENT=0: the multi-kernel blocked path also where the matrix fits one SM's shared memory
    static const bool resident_on = !(getenv("MCB_UDT_RESIDENT") && atoi(getenv("MCB_UDT_RESIDENT")) == 0);
    if (blocked_on && resident_on && udt_resident_supported(n) && !force_general_udt() && sX == (long long)n * n &&
        sU == sX && sT == sX)
        return launch_udt_resident(n, batch, X, U, D, sD, T, pivot, sP, apply_pivot, st);
    if (blocked_on && udt_blocked_supported(n) && !force_general_udt() && sX == (long long)n * n && sU == sX && sT == sX)
        return launch_udt_blocked(n, batch, X, U, D, sD, T, pivot, sP, apply_pivot, st);
    static const bool prefer_smem = getenv("MCB_UDT_SMEM") && atoi(getenv("MCB_UDT_SMEM")) != 0;
    if (prefer_smem && n <= 200 && udt_cluster_supported(n) && !force_general_udt() && sX == (long long)n * n &&
        sU == sX && sT == sX && sD == n && sP == n)
        return launch_udt_cluster(n, batch, X, U, D, T, pivot, apply_pivot, st);
    // sizes beyond the chain-resident path that still fit one SM's registers: register-resident QR
    if (udt_panel_supported(n) && !force_general_udt() && sX == (long long)n * n && sU == sX && sT == sX &&
        sD == n && sP == n)
        return launch_udt_panel(n, batch, X, U, D, T, pivot, apply_pivot, st);
    // sizes beyond one SM: a cluster of 8 CTAs per chain holds the matrix in distributed shared memory
    if (udt_cluster_supported(n) && !force_general_udt() && sX == (long long)n * n && sU == sX && sT == sX &&
        sD == n && sP == n)
        return launch_udt_cluster(n, batch, X, U, D, T, pivot, apply_pivot, st);
    UdtArgs a{n, X, sX, U, sU, D, sD, T, sT, pivot, sP, apply_pivot, 0};
    const size_t hdr = (size_t)(3 * n + 2 + (n + 3) / 2 + 2) * sizeof(double);
    const size_t mat = (size_t)n * n * sizeof(double);
    size_t smem = hdr;
    if (hdr + 2 * mat <= (size_t)SMEM_LIMIT) { a.mode = 2; smem = hdr + 2 * mat; }
    else if (hdr + mat <= (size_t)SMEM_LIMIT) { a.mode = 1; smem = hdr + mat; }
    const int threads = n <= 32 ? 128 : (n <= 96 ? 256 : (n <= 200 ? 512 : 1024));
    LaunchScope scope(LC_UDT, st, 8.0 / 3.0 * n * (double)n * n * batch);
    udt_kernel<<<batch, threads, smem, st>>>(a);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

int launch_rdivp(int n, int batch, const double* in, long long sIn, double* out, long long sOut,
                 const double* T, long long sT, const int* pivot, long long sP, cudaStream_t st)
{
    if (in == out) { set_error("rdivp: in-place call"); return -1; }
    // MCB_RDIVP_BLOCKED=0: the row-by-row kernels below (comparisons)
    static const bool blocked_on = !(getenv("MCB_RDIVP_BLOCKED") && atoi(getenv("MCB_RDIVP_BLOCKED")) == 0);
    if (blocked_on && rdivp_blocked_supported(n) && !force_general_udt())
        return launch_rdivp_blocked(n, batch, in, sIn, out, sOut, T, sT, pivot, sP, st);
    if (rdivp_panel_supported(n) && !force_general_udt() && sIn == (long long)n * n && sOut == sIn &&
        sT == sIn && sP == n)
        return launch_rdivp_panel(n, batch, in, out, T, pivot, st);
    RdivpArgs a{n, in, sIn, out, sOut, T, sT, pivot, sP, 0};
    const size_t hdr = (size_t)n * sizeof(double);
    const size_t mat = (size_t)n * n * sizeof(double);
    size_t smem = hdr;
    if (hdr + 2 * mat <= (size_t)SMEM_LIMIT) { a.mode = 2; smem = hdr + 2 * mat; }
    else if (hdr + mat <= (size_t)SMEM_LIMIT) { a.mode = 1; smem = hdr + mat; }
    const int threads = n <= 32 ? 128 : (n <= 96 ? 256 : (n <= 200 ? 512 : 1024));
    LaunchScope scope(LC_RDIVP, st, 1.0 * n * (double)n * n * batch);
    rdivp_kernel<<<batch, threads, smem, st>>>(a);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mcb
