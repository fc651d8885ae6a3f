// Column-pivoted Householder QR (udt_AVX_pivot!, src/linalg/UDT.jl:192-306) for matrices that do not fit
// one SM (200 < N <= 448, N % 8 == 0; config D: N = 400): one thread-block CLUSTER of 8 CTAs per chain.
//
//   factor kernel  Every CTA keeps N/8 columns of the matrix in its shared memory for the whole
//                  factorisation; columns never move (the permutation is a position <-> column map that
//                  every CTA replicates). Per Householder step there is ONE cluster barrier:
//                    1. owners have pushed the fresh squared norm (rows >= j, as indmaxcolumn of
//                       UDT.jl:151-168 recomputes it) and the row-j element of each of their trailing
//                       columns into the tables of all 8 CTAs (st.shared::cluster),
//                    2. barrier.cluster; every warp of every CTA finds the pivot from its replica of the
//                       table (first strict maximum in position order) and derives the reflector scalars,
//                    3. every CTA pulls rows j+1.. of the pivot column out of the owner's shared memory
//                       (ld.shared::cluster), scaled into its own copy of the Householder vector,
//                    4. every warp applies the reflector to its two columns and computes their next norms.
//                  The owner also writes the scaled vector and tau into X (global), which is workspace.
//   form-Q kernel  U = H_0 ... H_{N-1} by backward accumulation (UDT.jl:250-266). Columns of U are
//                  independent: one warp owns the column pair (c, N-1-c) in registers (equal work per
//                  warp), the reflectors stream through shared memory in chunks of 8 (cp.async, double
//                  buffered). No cluster needed.
#include "udt.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace mcb {

namespace {

constexpr int CL = 8;          // CTAs per chain (portable cluster size)
constexpr int FW = 25;         // warps per CTA, factor kernel
constexpr int FT = FW * 32;
constexpr int QW = 8;          // warps per CTA, form-Q kernel (one column pair each)
constexpr int QT = QW * 32;
constexpr int QNR = 15;        // 32-row blocks per column held in registers (N <= 480)
constexpr int KC = 8;          // reflectors per shared-memory chunk
constexpr int NMAX = 448;       // n*n + 57 n bytes of shared memory per CTA must stay below 227 KB

struct ClusterUdtArgs {
    int n, nc;                 // nc = n / CL columns per CTA
    double* X;                 // in: the matrix; out: scaled Householder vectors below, tau on the diagonal
    double* D;
    double* T;
    int* pivot;
    int apply_pivot;
};

size_t factor_smem_bytes(int n)
{
    const int nc = n / CL;
    // A (nc*n) + vbuf (n) + norms (2n) + xrow (2n) + dtab (n) doubles, colpos + colat ints, elim bytes
    return (size_t)nc * n * 8 + (size_t)6 * n * 8 + (size_t)2 * n * 4 + (size_t)n + 16;
}

template <int K>
__device__ __forceinline__ void apply_reflector(double* const (&col)[K], const int (&cidx)[K], const double* vb,
                                                int j, int n, double tj, int lane, double* rnorm, double* rxrow)
{
    double dot[K];
#pragma unroll
    for (int k = 0; k < K; ++k) dot[k] = 0.0;
#pragma unroll 4
    for (int i = j + 1 + lane; i < n; i += 32) {
        const double v = vb[i];
#pragma unroll
        for (int k = 0; k < K; ++k) dot[k] = fma(v, col[k][i], dot[k]);
    }
    double w[K], nrm[K], first[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        dot[k] = warp_sum(dot[k]);
        w[k] = tj * (col[k][j] + dot[k]);
        nrm[k] = 0.0;
        first[k] = 0.0;
    }
#pragma unroll 4
    for (int i = j + 1 + lane; i < n; i += 32) {
        const double v = vb[i];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double x = fma(-v, w[k], col[k][i]);
            col[k][i] = x;
            nrm[k] = fma(x, x, nrm[k]);
            if (i == j + 1) first[k] = x;
        }
    }
    __syncwarp();
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) col[k][j] -= w[k];
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
        nrm[k] = warp_sum(nrm[k]);
        first[k] = __shfl_sync(0xffffffffu, first[k], 0);
    }
    if (lane < CL) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            rnorm[cidx[k]] = nrm[k];
            rxrow[cidx[k]] = first[k];
        }
    }
}

__global__ void __launch_bounds__(FT, 1) udt_cluster_factor_kernel(ClusterUdtArgs a)
{
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char smraw[];
    const int n = a.n, nc = a.nc;
    const int rank = (int)cluster.block_rank();
    const int chain = blockIdx.x / CL;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* A = reinterpret_cast<double*>(smraw);     // my nc columns, column-major, never permuted
    double* vbuf = A + (size_t)nc * n;                // Householder vector of the current step
    double* norms = vbuf + n;                         // [2][n] squared trailing norms by ORIGINAL column, -1 = eliminated
    double* xrow = norms + 2 * n;                     // [2][n] element of the current row of each column
    double* dtab = xrow + 2 * n;                      // |R_jj|
    int* colpos = reinterpret_cast<int*>(dtab + n);   // position of original column c
    int* colat = colpos + n;                          // original column at position p
    unsigned char* elim = reinterpret_cast<unsigned char*>(colat + n);
    const int c0 = rank * nc;
    double* Xg = a.X + (size_t)chain * n * n;

    for (int e = tid; e < nc * n; e += FT) A[e] = Xg[(size_t)c0 * n + e];
    for (int c = tid; c < n; c += FT) { colpos[c] = c; colat[c] = c; elim[c] = 0; }
    // remote views of the tables: lane r (< CL) writes into CTA r
    double* rnorms = norms;
    double* rxrows = xrow;
    if (lane < CL) {
        rnorms = cluster.map_shared_rank(norms, lane);
        rxrows = cluster.map_shared_rank(xrow, lane);
    }
    // every CTA of the cluster is running and has read its part of X (X is workspace from here on)
    cluster.sync();
    for (int lc = warp; lc < nc; lc += FW) {
        const double* col = A + (size_t)lc * n;
        double s = 0.0;
        for (int i = lane; i < n; i += 32) s = fma(col[i], col[i], s);
        s = warp_sum(s);
        if (lane < CL) { rnorms[c0 + lc] = s; rxrows[c0 + lc] = col[0]; }
    }

    int prev_jm = -1;
    for (int j = 0; j < n; ++j) {
        const int par = j & 1;
        cluster.sync();   // tables of parity `par` complete in every CTA; previous step fully applied
        // ---- pivot: first strict maximum in position order (indmaxcolumn), every warp redundantly
        const double* nt = norms + par * n;
        double best = -1.0;
        int bpos = 0x7fffffff, bc = 0;
        for (int c = lane; c < n; c += 32) {
            const double v = nt[c];
            const int p = colpos[c];
            if (v > best || (v == best && p < bpos)) { best = v; bpos = p; bc = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int op = __shfl_xor_sync(0xffffffffu, bpos, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (ov > best || (ov == best && op < bpos)) { best = ov; bpos = op; bc = oc; }
        }
        const int jm = bc;
        const double maxval = best;
        double xi1 = xrow[par * n + jm];
        const int owner = jm / nc, lcol = jm - owner * nc;
        // ---- reflector scalars (reflector!, UDT.jl:133-148), identical in every CTA
        const bool degenerate = !(maxval > 0.0);
        double nu = 0.0, tj = 0.0, rinv = 1.0;
        if (!degenerate) {
            nu = copysign(sqrt(maxval), xi1);
            xi1 += nu;
            tj = xi1 / nu;
            rinv = 1.0 / xi1;
        }
        // ---- pull the pivot column out of its owner's shared memory
        const double* rcol = cluster.map_shared_rank(A + (size_t)lcol * n, owner);
        double* Xj = Xg + (size_t)j * n;
        for (int i = j + 1 + tid; i < n; i += FT) {
            const double t = rcol[i] * rinv;
            vbuf[i] = t;
            if (rank == owner) Xj[i] = t;
        }
        __syncthreads();
        if (tid == 0) {
            const int cj = colat[j], p = colpos[jm];
            colat[j] = jm; colat[p] = cj;
            colpos[cj] = p; colpos[jm] = j;
            elim[jm] = 1;
            norms[(par ^ 1) * n + jm] = -1.0;
            if (prev_jm >= 0) norms[(par ^ 1) * n + prev_jm] = -1.0;
            dtab[j] = fabs(nu);
            if (rank == owner) {
                if (!degenerate) A[(size_t)lcol * n + j] = -nu;
                Xj[j] = tj;
            }
        }
        prev_jm = jm;
        // ---- reflectorApply! on my trailing columns (UDT.jl:32-50), fused with the next step's norms
        double* rn = rnorms + (par ^ 1) * n;
        double* rx = rxrows + (par ^ 1) * n;
        for (int base = warp; base < nc; base += 2 * FW) {
            const int l0 = base, l1 = base + FW;
            const int g0 = c0 + l0, g1 = c0 + l1;
            const bool act0 = !(g0 == jm || elim[g0]);
            const bool act1 = l1 < nc && !(g1 == jm || elim[g1]);
            if (act0 && act1) {
                double* const cols[2] = {A + (size_t)l0 * n, A + (size_t)l1 * n};
                const int idx[2] = {g0, g1};
                apply_reflector<2>(cols, idx, vbuf, j, n, tj, lane, rn, rx);
            } else if (act0 || act1) {
                double* const cols[1] = {A + (size_t)(act0 ? l0 : l1) * n};
                const int idx[1] = {act0 ? g0 : g1};
                apply_reflector<1>(cols, idx, vbuf, j, n, tj, lane, rn, rx);
            }
        }
    }
    // nobody leaves (or reuses vbuf) while a peer may still be reading its shared memory
    cluster.sync();
    for (int i = tid; i < n; i += FT) {
        const double d = dtab[i];
        vbuf[i] = 1.0 / d;
        if (rank == 0) {
            a.D[(size_t)chain * n + i] = d;
            a.pivot[(size_t)chain * n + i] = colat[i];
        }
    }
    __syncthreads();
    double* Tg = a.T + (size_t)chain * n * n;
    if (a.apply_pivot) {   // _apply_pivot! Val(true) (UDT.jl:283-297): column of T = original column
        for (int e = tid; e < nc * n; e += FT) {
            const int lc = e / n, i = e - lc * n;
            const int p = colpos[c0 + lc];
            Tg[(size_t)(c0 + lc) * n + i] = (i <= p) ? vbuf[i] * A[e] : 0.0;
        }
    } else {               // Val(false): D^-1 R in position order, below the diagonal unspecified
        for (int e = tid; e < nc * n; e += FT) {
            const int lc = e / n, i = e - lc * n;
            const int p = colpos[c0 + lc];
            Tg[(size_t)p * n + i] = (i <= p) ? vbuf[i] * A[e] : A[e];
        }
    }
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}

__global__ void __launch_bounds__(QT, 2) udt_formq_kernel(int n, const double* __restrict__ X, double* __restrict__ U)
{
    extern __shared__ __align__(16) unsigned char smraw[];
    double* buf = reinterpret_cast<double*>(smraw);   // [2][KC][n]
    const int chain = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int pair0 = blockIdx.x * QW;
    const int ca = pair0 + warp, cb = n - 1 - ca;     // ca < n/2 <= cb
    const bool valid = ca < n / 2;
    const double* Xc = X + (size_t)chain * n * n;
    double ua[QNR], ub[QNR];
#pragma unroll
    for (int r = 0; r < QNR; ++r) {
        const int i = 32 * r + lane;
        ua[r] = (i == ca) ? 1.0 : 0.0;
        ub[r] = (i == cb) ? 1.0 : 0.0;
    }
    // first chunk: the one that holds reflector n-1-pair0 (the longest column of this CTA)
    const int kstart = ((n - 1 - pair0) / KC) * KC;
    auto load_chunk = [&](int k0, int b) {
        const int half = (n - k0) >> 1;   // 16-byte pieces per column, rows k0..n-1 (k0, n even)
        double* dst = buf + (size_t)b * KC * n;
        for (int idx = tid; idx < KC * half; idx += QT) {
            const int col = idx / half, pc = idx - col * half;
            cp_async16(dst + (size_t)col * n + k0 + 2 * pc, Xc + (size_t)(k0 + col) * n + k0 + 2 * pc);
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };
    load_chunk(kstart, 0);
    asm volatile("cp.async.wait_all;\n" ::);
    __syncthreads();
    int b = 0;
    for (int k0 = kstart; k0 >= 0; k0 -= KC) {
        if (k0 >= KC) load_chunk(k0 - KC, b ^ 1);
        if (valid) {
            const double* chunk = buf + (size_t)b * KC * n;
            for (int kk = KC - 1; kk >= 0; --kk) {
                const int k = k0 + kk;
                if (k > cb) continue;
                const double* vk = chunk + (size_t)kk * n;
                const double tk = vk[k];
                const int rb = k >> 5;
                const bool doa = k <= ca;
                double v[QNR];
                double da = 0.0, db = 0.0;
#pragma unroll
                for (int r = 0; r < QNR; ++r) {
                    if (r >= rb) {
                        const int i = 32 * r + lane;
                        double t = 0.0;
                        if (i > k && i < n) t = vk[i];
                        if (i == k) t = 1.0;
                        v[r] = t;
                        db = fma(t, ub[r], db);
                        if (doa) da = fma(t, ua[r], da);
                    }
                }
                db = warp_sum(db);
                da = warp_sum(da);
                const double wb = tk * db, wa = tk * da;
#pragma unroll
                for (int r = 0; r < QNR; ++r) {
                    if (r >= rb) {
                        ub[r] = fma(-v[r], wb, ub[r]);
                        if (doa) ua[r] = fma(-v[r], wa, ua[r]);
                    }
                }
            }
        }
        asm volatile("cp.async.wait_all;\n" ::);
        __syncthreads();
        b ^= 1;
    }
    if (valid) {
        double* Uc = U + (size_t)chain * n * n;
#pragma unroll
        for (int r = 0; r < QNR; ++r) {
            const int i = 32 * r + lane;
            if (i < n) {
                Uc[(size_t)ca * n + i] = ua[r];
                Uc[(size_t)cb * n + i] = ub[r];
            }
        }
    }
}

bool g_cluster_ok = false;

} // namespace

bool udt_cluster_supported(int n) { return g_cluster_ok && n > 200 && n <= NMAX && n % 8 == 0; }

int init_udt_cluster_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(udt_cluster_factor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)factor_smem_bytes(NMAX)));
    MCB_CUDA(cudaFuncSetAttribute(udt_formq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  2 * KC * NMAX * (int)sizeof(double)));
    // can the device co-schedule one cluster of this shape at all?
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CL, 1, 1);
    cfg.blockDim = dim3(FT, 1, 1);
    cfg.dynamicSmemBytes = factor_smem_bytes(NMAX);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int nclusters = 0;
    if (cudaOccupancyMaxActiveClusters(&nclusters, udt_cluster_factor_kernel, &cfg) != cudaSuccess) {
        cudaGetLastError();
        nclusters = 0;
    }
    g_cluster_ok = nclusters > 0;
    return 0;
}

int launch_udt_cluster(int n, int batch, double* X, double* U, double* D, double* T, int* pivot, int apply_pivot,
                       cudaStream_t st)
{
    if (!udt_cluster_supported(n)) { set_error("udt_cluster: unsupported size"); return -1; }
    ClusterUdtArgs a{n, n / CL, X, D, T, pivot, apply_pivot};
    const double n3 = (double)n * n * n * batch;
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(batch * CL, 1, 1);
        cfg.blockDim = dim3(FT, 1, 1);
        cfg.dynamicSmemBytes = factor_smem_bytes(n);
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        LaunchScope scope(LC_UDT, st, 4.0 / 3.0 * n3);
        MCB_CUDA(cudaLaunchKernelEx(&cfg, udt_cluster_factor_kernel, a));
    }
    {
        LaunchScope scope(LC_UDT, st, 4.0 / 3.0 * n3);
        const dim3 grid((n / 2 + QW - 1) / QW, batch, 1);
        udt_formq_kernel<<<grid, QT, 2 * KC * n * sizeof(double), st>>>(n, X, U);
        MCB_CUDA(cudaGetLastError());
    }
    return 0;
}

} // namespace mcb
