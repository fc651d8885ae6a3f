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
#include <stdlib.h>

namespace cg = cooperative_groups;

namespace mcb {

namespace {

constexpr int CL = 8;          // CTAs per chain for 200 < N (portable cluster size); 1 CTA when the matrix fits one SM
constexpr int FW = 25;         // warps per CTA, factor kernel
constexpr int FT = FW * 32;
constexpr int QW = 8;          // warps per CTA, form-Q kernel (one column pair each)
constexpr int QT = QW * 32;
constexpr int KC = 8;          // reflectors per shared-memory chunk
constexpr int NMAX = 448;       // n*n + 57 n bytes of shared memory per CTA must stay below 227 KB

struct ClusterUdtArgs {
    int n, nc, cl;             // cl CTAs per chain, nc = n / cl columns per CTA
    double* X;                 // in: the matrix; out: scaled Householder vectors below, tau on the diagonal
    double* D;
    double* T;
    int* pivot;
    int apply_pivot;
    int exp;                   // timing experiments (MCB_UDT_EXP): phases switched off, results invalid
};

size_t factor_smem_bytes(int n, int cl)
{
    const int nc = n / cl;
    // A (nc*n) + vbuf (n) + norms (2n) + xrow (2n) + dtab (n) + rinv, tau (nc each) doubles,
    // colpos + colat ints, elim bytes, level-1 pivot candidates
    return (size_t)nc * n * 8 + (size_t)6 * n * 8 + (size_t)2 * nc * 8 + (size_t)2 * n * 4 + (size_t)n + 16 + 3 * 16 * 4 + 16;
}

// One 64-row block of the two passes of reflectorApply!: lane l owns rows 64 r + 2 l, + 1 (LDS.128).
template <int K>
__device__ __forceinline__ void dot_block(int r, int lane, int n, const double* vb, double* const (&col)[K],
                                          double (&dot)[2 * K])
{
    const int i0 = 64 * r + 2 * lane;
    if (i0 < n) {
        const double2 v = *reinterpret_cast<const double2*>(vb + i0);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const double2 c = *reinterpret_cast<const double2*>(col[k] + i0);
            dot[k] = fma(v.x, c.x, dot[k]);
            dot[K + k] = fma(v.y, c.y, dot[K + k]);
        }
    }
}

template <int K>
__device__ __forceinline__ void update_block(int r, int lane, int n, int j, const double* vb, double* const (&col)[K],
                                             const double (&w)[K], double (&nrm)[K])
{
    const int i0 = 64 * r + 2 * lane;
    if (i0 < n) {
        const double2 v = *reinterpret_cast<const double2*>(vb + i0);
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double2 c = *reinterpret_cast<double2*>(col[k] + i0);
            c.x = fma(-v.x, w[k], c.x);
            c.y = fma(-v.y, w[k], c.y);
            *reinterpret_cast<double2*>(col[k] + i0) = c;
            if (i0 > j) nrm[k] = fma(c.x, c.x, nrm[k]);       // rows <= j belong to R, not to the next norm
            if (i0 + 1 > j) nrm[k] = fma(c.y, c.y, nrm[k]);
        }
    }
}

// fall-through switch: only the blocks that hold rows > j are executed (real branches, no predicated-off work)
#define MCB_BLOCKS_FROM(rb0, CALL)            \
    switch (rb0) {                            \
    case 0: CALL(0); [[fallthrough]];         \
    case 1: CALL(1); [[fallthrough]];         \
    case 2: CALL(2); [[fallthrough]];         \
    case 3: CALL(3); [[fallthrough]];         \
    case 4: CALL(4); [[fallthrough]];         \
    case 5: CALL(5); [[fallthrough]];         \
    default: CALL(6);                         \
    }

// sums over the warp of K per-lane values, result in every lane. Two values share one butterfly: after the first
// exchange the lower half-warp carries the first sum and the upper half the second (7 instead of 10 shuffles;
// the shuffle / shared-memory pipe is what bounds this kernel).
template <int K>
__device__ __forceinline__ void warp_sums(double (&a)[K], int lane)
{
    if (K == 2) {
        const unsigned F = 0xffffffffu;
        const bool hi = lane & 16;
        double k = hi ? a[K - 1] : a[0];
        k += __shfl_xor_sync(F, hi ? a[0] : a[K - 1], 16);
        k += __shfl_xor_sync(F, k, 8);
        k += __shfl_xor_sync(F, k, 4);
        k += __shfl_xor_sync(F, k, 2);
        k += __shfl_xor_sync(F, k, 1);
        a[0] = __shfl_sync(F, k, 0);
        a[K - 1] = __shfl_sync(F, k, 16);
    } else {
#pragma unroll
        for (int i = 0; i < K; ++i) a[i] = warp_sum(a[i]);
    }
}

template <int K>
__device__ __forceinline__ void apply_reflector(double* const (&col)[K], const int (&cidx)[K], const double* vb,
                                                int j, int n, double tj, double rinv, int lane, int cl,
                                                double* rnorm, double* rxrow, int exp)
{
    const int rb0 = (j + 1) >> 6;   // vb is zero on the rows <= j of this block
    double dot[2 * K], w[K], wj[K], nrm[K];   // even / odd rows accumulate separately (shorter dependency chains)
#pragma unroll
    for (int k = 0; k < 2 * K; ++k) dot[k] = 0.0;
#define MCB_DOT(r) dot_block<K>(r, lane, n, vb, col, dot)
    if (!(exp & 64)) {
    MCB_BLOCKS_FROM(rb0, MCB_DOT)
    }
#undef MCB_DOT
    double dsum[K];
#pragma unroll
    for (int k = 0; k < K; ++k) dsum[k] = dot[k] + dot[K + k];
    warp_sums<K>(dsum, lane);
#pragma unroll
    for (int k = 0; k < K; ++k) {
        // vb holds the RAW pivot column: v = rinv * vb, so v.c = rinv * (vb.c) and c -= v w = vb (rinv w)
        const double d = dsum[k];
        wj[k] = tj * fma(rinv, d, col[k][j]);
        w[k] = rinv * wj[k];
        nrm[k] = 0.0;
    }
#define MCB_UPD(r) update_block<K>(r, lane, n, j, vb, col, w, nrm)
    if (!(exp & 32)) {
    MCB_BLOCKS_FROM(rb0, MCB_UPD)
    }
#undef MCB_UPD
    __syncwarp();
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) col[k][j] -= wj[k];
    }
    double xr[K];
    warp_sums<K>(nrm, lane);
#pragma unroll
    for (int k = 0; k < K; ++k) xr[k] = col[k][j + 1];
    if (lane < cl && !(exp & 16)) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            rnorm[cidx[k]] = nrm[k];
            rxrow[cidx[k]] = xr[k];
        }
    }
}

// (max norm, ties to the lowest position) over the warp on the integer pipe: squared norms are
// non-negative doubles, so their (hi, lo) words order like the values; eliminated columns carry -1.0.
__device__ __forceinline__ void warp_argmax(int hi, unsigned lo, int key, int& mh, unsigned& ml, int& bk)
{
    mh = __reduce_max_sync(0xffffffffu, hi);
    ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    bk = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? key : 0x7fffffff);
}

__global__ void __launch_bounds__(FT, 1) udt_cluster_factor_kernel(ClusterUdtArgs a)
{
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char smraw[];
    const int n = a.n, nc = a.nc, cl = a.cl;
    const int rank = (int)cluster.block_rank();
    const int chain = blockIdx.x / cl;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* A = reinterpret_cast<double*>(smraw);     // my nc columns, column-major, never permuted
    double* vbuf = A + (size_t)nc * n;                // Householder vector of the current step
    double* norms = vbuf + n;                         // [2][n] squared trailing norms by ORIGINAL column, -1 = eliminated
    double* xrow = norms + 2 * n;                     // [2][n] element of the current row of each column
    double* dtab = xrow + 2 * n;                      // |R_jj|
    double* rinvl = dtab + n;                         // per local column: 1 / xi1 of its step
    double* taul = rinvl + nc;                        // per local column: tau of its step
    int* colpos = reinterpret_cast<int*>(taul + nc);  // position of original column c
    int* colat = colpos + n;                          // original column at position p
    int* cand = colat + n;                            // [3][16] level-1 candidates (hi, lo, key)
    double* scal = reinterpret_cast<double*>(cand + 48);   // tau, 1 / xi1 of the current step
    unsigned char* elim = reinterpret_cast<unsigned char*>(scal + 2);
    const int c0 = rank * nc;
    const int nw1 = (n + 31) >> 5;                    // warps of the first argmax level (<= 14 <= FW)
    double* Xg = a.X + (size_t)chain * n * n;

    for (int e = tid; e < nc * n; e += FT) A[e] = Xg[(size_t)c0 * n + e];
    for (int c = tid; c < n; c += FT) { colpos[c] = c; colat[c] = c; elim[c] = 0; }
    // remote views of the tables: lane r (< cl) writes into CTA r
    double* rnorms = norms;
    double* rxrows = xrow;
    if (lane < cl) {
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
        if (lane < cl) { rnorms[c0 + lc] = s; rxrows[c0 + lc] = col[0]; }
    }

    int prev_jm = -1;
    for (int j = 0; j < n; ++j) {
        const int par = j & 1;
        if (a.exp & 8) __syncthreads(); else
        cluster.sync();   // tables of parity `par` complete in every CTA; previous step fully applied
        // ---- pivot: first strict maximum in position order (indmaxcolumn). Level 1: one column per thread
        if (!(a.exp & 4) && warp < nw1) {
            int hi = (int)0x80000000, key = 0x7fffffff;
            unsigned lo = 0u;
            if (tid < n) {
                const double v = norms[par * n + tid];
                hi = __double2hiint(v);
                lo = (unsigned)__double2loint(v);
                key = (colpos[tid] << 16) | tid;
            }
            int mh, bk;
            unsigned ml;
            warp_argmax(hi, lo, key, mh, ml, bk);
            if (lane == 0) { cand[warp] = mh; cand[16 + warp] = (int)ml; cand[32 + warp] = bk; }
        }
        __syncthreads();
        // level 2: every warp redundantly over the <= 14 candidates
        int jm;
        double maxval;
        {
            int hi = (int)0x80000000, key = 0x7fffffff;
            unsigned lo = 0u;
            if (lane < nw1) { hi = cand[lane]; lo = (unsigned)cand[16 + lane]; key = cand[32 + lane]; }
            int mh, bk;
            unsigned ml;
            warp_argmax(hi, lo, key, mh, ml, bk);
            jm = bk & 0xffff;
            if (a.exp & 4) jm = j;
            maxval = __hiloint2double(mh, (int)ml);
        }
        const int owner = jm / nc, lcol = jm - owner * nc;
        // ---- reflector scalars (reflector!, UDT.jl:133-148): warp 0 only, while the others fetch the column
        if (warp == 0) {
            double xi1 = xrow[par * n + jm];
            const bool degenerate = !(maxval > 0.0);
            double nu = 0.0, tj = 0.0, rinv = 1.0;
            if (!degenerate) {
                nu = copysign(sqrt(maxval), xi1);
                xi1 += nu;
                tj = xi1 / nu;
                rinv = 1.0 / xi1;
            }
            if (lane == 0) {
                scal[0] = tj; scal[1] = rinv;
                dtab[j] = fabs(nu);
                if (rank == owner) {
                    if (!degenerate) A[(size_t)lcol * n + j] = -nu;   // R_jj (the peers fetch rows > j only)
                    rinvl[lcol] = rinv;
                    taul[lcol] = tj;
                }
            }
        }
        // ---- pull the RAW pivot column out of its owner's shared memory (zero on the rows <= j of the first block)
        const double* rcol = cluster.map_shared_rank(A + (size_t)lcol * n, owner);
        if (!(a.exp & 2))
        for (int i = (((j + 1) >> 6) << 6) + tid; i < n; i += FT) vbuf[i] = (i > j) ? rcol[i] : 0.0;
        __syncthreads();
        const double tj = scal[0], rinv = scal[1];
        if (tid == 0) {
            const int cj = colat[j], p = colpos[jm];
            colat[j] = jm; colat[p] = cj;
            colpos[cj] = p; colpos[jm] = j;
            elim[jm] = 1;
            norms[(par ^ 1) * n + jm] = -1.0;
            if (prev_jm >= 0) norms[(par ^ 1) * n + prev_jm] = -1.0;
        }
        prev_jm = jm;
        // ---- reflectorApply! on my trailing columns (UDT.jl:32-50), fused with the next step's norms
        if (j + 1 < n && !(a.exp & 1)) {
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
                    apply_reflector<2>(cols, idx, vbuf, j, n, tj, rinv, lane, cl, rn, rx, a.exp);
                } else if (act0 || act1) {
                    double* const cols[1] = {A + (size_t)(act0 ? l0 : l1) * n};
                    const int idx[1] = {act0 ? g0 : g1};
                    apply_reflector<1>(cols, idx, vbuf, j, n, tj, rinv, lane, cl, rn, rx, a.exp);
                }
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
    // outputs of my columns: T = D^-1 R (P'), and for the form-Q kernel the scaled Householder vectors
    // below the diagonal of X in position order with tau on the diagonal
    double* Tg = a.T + (size_t)chain * n * n;
    for (int e = tid; e < nc * n; e += FT) {
        const int lc = e / n, i = e - lc * n;
        const int p = colpos[c0 + lc];
        const double v = A[e];
        if (a.apply_pivot)   // _apply_pivot! Val(true) (UDT.jl:283-297): column of T = original column
            Tg[(size_t)(c0 + lc) * n + i] = (i <= p) ? vbuf[i] * v : 0.0;
        else                 // Val(false): D^-1 R in position order, below the diagonal unspecified
            Tg[(size_t)p * n + i] = (i <= p) ? vbuf[i] * v : v;
        if (i >= p) Xg[(size_t)p * n + i] = (i == p) ? taul[lc] : v * rinvl[lc];
    }
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}


// reflector k applied to the warp's columns; RB = k / 32 is static so that only blocks r >= RB exist in the code
template <int QNR, int RB, bool BOTH>
__device__ __forceinline__ void formq_step(int k, int lane, const double* vk, double (&ua)[QNR], double (&ub)[QNR])
{
    const double tk = vk[k];
    double v0;
    {
        const int i = 32 * RB + lane;
        v0 = vk[i];
        if (i < k) v0 = 0.0;
        if (i == k) v0 = 1.0;
    }
    double db = v0 * ub[RB], da = BOTH ? v0 * ua[RB] : 0.0;
#pragma unroll
    for (int r = RB + 1; r < QNR; ++r) {
        const double t = vk[32 * r + lane];
        db = fma(t, ub[r], db);
        if (BOTH) da = fma(t, ua[r], da);
    }
    db = warp_sum(db);
    if (BOTH) da = warp_sum(da);
    const double wb = tk * db, wa = tk * da;
    ub[RB] = fma(-v0, wb, ub[RB]);
    if (BOTH) ua[RB] = fma(-v0, wa, ua[RB]);
#pragma unroll
    for (int r = RB + 1; r < QNR; ++r) {
        const double t = vk[32 * r + lane];
        ub[r] = fma(-t, wb, ub[r]);
        if (BOTH) ua[r] = fma(-t, wa, ua[r]);
    }
}

template <int QNR>
__device__ __forceinline__ void formq_load_chunk(int n, int k0, int tid, double* dst, const double* Xc)
{
    constexpr int QNP = 32 * QNR;
    const int half = (n - k0) >> 1;   // 16-byte pieces per column, rows k0..n-1 (k0, n even)
    for (int idx = tid; idx < KC * half; idx += QT) {
        const int col = idx / half, pc = idx - col * half;
        cp_async16(dst + (size_t)col * QNP + k0 + 2 * pc, Xc + (size_t)(k0 + col) * n + k0 + 2 * pc);
    }
    asm volatile("cp.async.commit_group;\n" ::);
}

// all reflectors k with k / 32 == RB (chunks of KC, next chunk in flight), then the levels below
template <int QNR, int RB>
__device__ __forceinline__ void formq_levels(int n, int kstart, bool valid, int ca, int cb, int lane, int tid,
                                             double* buf, const double* Xc, int& b, double (&ua)[QNR],
                                             double (&ub)[QNR])
{
    constexpr int QNP = 32 * QNR;   // padded column length of the reflector chunks (rows >= n stay zero)
    if (32 * RB <= kstart) {   // CTA-uniform
        for (int k0 = min(kstart, 32 * RB + 32 - KC); k0 >= 32 * RB; k0 -= KC) {
            if (k0 >= KC) formq_load_chunk<QNR>(n, k0 - KC, tid, buf + (size_t)(b ^ 1) * KC * QNP, Xc);
            if (valid) {
                const double* chunk = buf + (size_t)b * KC * QNP;
                for (int kk = KC - 1; kk >= 0; --kk) {
                    const int k = k0 + kk;
                    if (k > cb) continue;
                    if (k <= ca) formq_step<QNR, RB, true>(k, lane, chunk + (size_t)kk * QNP, ua, ub);
                    else formq_step<QNR, RB, false>(k, lane, chunk + (size_t)kk * QNP, ua, ub);
                }
            }
            asm volatile("cp.async.wait_all;\n" ::);
            __syncthreads();
            b ^= 1;
        }
    }
    if constexpr (RB > 0) formq_levels<QNR, RB - 1>(n, kstart, valid, ca, cb, lane, tid, buf, Xc, b, ua, ub);
}

// QNR = 32-row blocks per column held in registers (n <= 32 QNR)
template <int QNR>
__global__ void __launch_bounds__(QT, 2) udt_formq_kernel(int n, const double* __restrict__ X, double* __restrict__ U)
{
    constexpr int QNP = 32 * QNR;
    extern __shared__ __align__(16) unsigned char smraw[];
    double* buf = reinterpret_cast<double*>(smraw);   // [2][KC][QNP]
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
    for (int e = tid; e < 2 * KC * (QNP - n); e += QT) {
        const int c = e / (QNP - n), i = e - c * (QNP - n);
        buf[(size_t)c * QNP + n + i] = 0.0;
    }
    // first chunk: the one that holds reflector n-1-pair0 (the longest column of this CTA)
    const int kstart = ((n - 1 - pair0) / KC) * KC;
    formq_load_chunk<QNR>(n, kstart, tid, buf, Xc);
    asm volatile("cp.async.wait_all;\n" ::);
    __syncthreads();
    int b = 0;
    formq_levels<QNR, QNR - 1>(n, kstart, valid, ca, cb, lane, tid, buf, Xc, b, ua, ub);
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

int g_max_clusters = 0;   // clusters of CL CTAs the device runs at once (0: shape unavailable)
bool g_single_ok = false; // the one-CTA shape can be launched

int set_factor_smem(int n, int cl, cudaLaunchConfig_t* cfg, cudaLaunchAttribute* at, int batch, cudaStream_t st)
{
    *cfg = cudaLaunchConfig_t{};
    cfg->gridDim = dim3(batch * cl, 1, 1);
    cfg->blockDim = dim3(FT, 1, 1);
    cfg->dynamicSmemBytes = factor_smem_bytes(n, cl);
    cfg->stream = st;
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg->attrs = at;
    cfg->numAttrs = 1;
    return 0;
}

constexpr int N1MAX = 160;      // one CTA: n*n*8 + tables in 227 KB, and 5 row blocks in form-Q

} // namespace

bool udt_cluster_supported(int n)
{
    if (n % 8 != 0) return false;
    if (n > 200) return g_max_clusters > 0 && n <= NMAX;
    // the matrix fits one SM: same kernel with a "cluster" of one CTA (columns in shared memory, 25 warps)
    return g_single_ok && n > 128 && n <= N1MAX;
}

int init_udt_cluster_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(udt_cluster_factor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)factor_smem_bytes(NMAX, CL)));
    MCB_CUDA(cudaFuncSetAttribute(udt_formq_kernel<15>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  2 * KC * 32 * 15 * (int)sizeof(double)));
    MCB_CUDA(cudaFuncSetAttribute(udt_formq_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  2 * KC * 32 * 5 * (int)sizeof(double)));
    // can the device co-schedule clusters of this shape at all?
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute at[1];
    set_factor_smem(NMAX, CL, &cfg, at, 1, 0);
    if (cudaOccupancyMaxActiveClusters(&g_max_clusters, udt_cluster_factor_kernel, &cfg) != cudaSuccess) {
        cudaGetLastError();
        g_max_clusters = 0;
    }
    g_single_ok = factor_smem_bytes(N1MAX, 1) <= (size_t)factor_smem_bytes(NMAX, CL);
    return 0;
}

int launch_udt_cluster(int n, int batch, double* X, double* U, double* D, double* T, int* pivot, int apply_pivot,
                       cudaStream_t st)
{
    if (!udt_cluster_supported(n)) { set_error("udt_cluster: unsupported size"); return -1; }
    static const int exp_flags = experiment_flags("MCB_UDT_EXP");
    const int cl = n > 200 ? CL : 1;
    ClusterUdtArgs a{n, n / cl, cl, X, D, T, pivot, apply_pivot, exp_flags};
    const double n3 = (double)n * n * n * batch;
    {
        cudaLaunchConfig_t cfg;
        cudaLaunchAttribute at[1];
        set_factor_smem(n, cl, &cfg, at, batch, st);
        LaunchScope scope(LC_UDT, st, 4.0 / 3.0 * n3);
        MCB_CUDA(cudaLaunchKernelEx(&cfg, udt_cluster_factor_kernel, a));
    }
    {
        LaunchScope scope(LC_UDT, st, 4.0 / 3.0 * n3);
        const dim3 grid((n / 2 + QW - 1) / QW, batch, 1);
        if (n <= 160) udt_formq_kernel<5><<<grid, QT, 2 * KC * 32 * 5 * sizeof(double), st>>>(n, X, U);
        else udt_formq_kernel<15><<<grid, QT, 2 * KC * 32 * 15 * sizeof(double), st>>>(n, X, U);
        MCB_CUDA(cudaGetLastError());
    }
    return 0;
}

} // namespace mcb
