// Device-side building blocks of the blocked Householder QR (udt_blocked.cu):
//   panel_steps   NB unpivoted Householder steps on a register-held panel, one block barrier per step
//   panel_gram    S = V'V of the finished panel on the tensor pipe
//   panel_tfactor T of H_1 .. H_NB = I - V T V' from S and tau
//   apply_strip   C <- (I - V op(T) V') C for an 8-column strip held in registers, three DMMA products
// Layout conventions are documented at each function; NB = 16 throughout.
#pragma once
#include "common.cuh"

namespace mcb {
namespace qrb {

constexpr int NB = 16;       // panel width
constexpr int TLD = NB + 8;  // leading dimension of k-contiguous operands: 16-byte loads at (l/4) * TLD + 2 * (l%4)
                             // are bank-conflict free when TLD % 16 == 8

// ---------------------------------------------------------------------------------------------------------------------
// The NB largest of nrm[j .. n) (squared norms: non-negative, NaN already mapped to +inf), by decreasing value, ties
// to the lower position: sel[0 .. NB). NB rounds of a block-wide argmax on order-preserving integer keys (the bit
// pattern of a non-negative double is monotone), one barrier per round; wkey [2][NWARPS], wpos [2][NWARPS] are
// double-buffered by round parity. Every thread holds CPT candidates (position j + tid + u * NWARPS * 32).
// Requires n - j >= NB. Ends with a barrier: sel is visible to the block.
template <int NWARPS, int CPT>
__device__ __forceinline__ void select_largest(const double* nrm, int j, int n, int tid, unsigned long long* wkey,
                                               int* wpos, int* sel)
{
    const int lane = tid & 31, warp = tid >> 5;
    unsigned long long key[CPT];
#pragma unroll
    for (int u = 0; u < CPT; ++u) {
        const int c = j + tid + u * NWARPS * 32;
        key[u] = c < n ? (unsigned long long)__double_as_longlong(nrm[c]) + 1ull : 0ull;   // 0 = no candidate
    }
    for (int r = 0; r < NB; ++r) {
        unsigned long long bk = 0ull;
        int bp = 0x7fffffff;
#pragma unroll
        for (int u = 0; u < CPT; ++u) {   // ascending positions: strict > keeps the lower position on ties
            if (key[u] > bk) { bk = key[u]; bp = j + tid + u * NWARPS * 32; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ok = __shfl_xor_sync(0xffffffffu, bk, o);
            const int op = __shfl_xor_sync(0xffffffffu, bp, o);
            if (ok > bk || (ok == bk && op < bp)) { bk = ok; bp = op; }
        }
        const int buf = r & 1;
        if (lane == 0) { wkey[buf * NWARPS + warp] = bk; wpos[buf * NWARPS + warp] = bp; }
        __syncthreads();
        bk = wkey[buf * NWARPS];
        bp = wpos[buf * NWARPS];
#pragma unroll
        for (int w = 1; w < NWARPS; ++w) {
            const unsigned long long ok = wkey[buf * NWARPS + w];
            const int op = wpos[buf * NWARPS + w];
            if (ok > bk || (ok == bk && op < bp)) { bk = ok; bp = op; }
        }
#pragma unroll
        for (int u = 0; u < CPT; ++u)
            if (bp == j + tid + u * NWARPS * 32) key[u] = 0ull;   // the winner leaves the pool
        if (tid == 0) sel[r] = bp;
    }
    __syncthreads();
}

// The same selection by ranking (rank of a column = number of columns that precede it in the order above), in two
// stages. The norms are >= +0 or +inf, so their bit patterns, read as 64-bit integers, order like the values.
//   (1) A threshold: the candidates are dealt into NB groups (position e goes to group e % NB: interleaved, because
//       the columns that reach a factorisation are often sorted by norm already); one warp finds the maximum of every
//       group and tau = the smallest of these maxima. NB columns are >= tau, hence so are the NB largest.
//   (2) Only the columns >= tau (typically NB .. 3 NB of them) are ranked against each other, the comparisons of a
//       column split over NT / 128 groups of threads and combined with integer atomics (exact, hence independent of
//       the order in which the candidates were listed). A comparison is a 64-bit subtraction whose sign bit is the
//       result: w precedes v  <=>  v - [pos_w < pos_v] - w < 0.
// If every column passes the threshold (all norms equal) stage 2 is the full m x m ranking that used to run always
// (11 k cycles at m = 320 on 512 threads against 2 k for the two stages on DQMC matrices).
// scratch: 16 m + 16 bytes of shared memory, 8-byte aligned. Starts and ends with a barrier.
template <int NT>
__device__ __forceinline__ void select_by_rank(const double* nrm, int j, int n, int tid, void* scratch, int* sel)
{
    constexpr int G = NT / 128;
    const int m = n - j, lane = tid & 31;
    const long long* key = reinterpret_cast<const long long*>(nrm) + j;
    long long* ckey = reinterpret_cast<long long*>(scratch);          // [m] keys of the candidates
    long long* thr = ckey + m;                                        // threshold
    int* cnt = reinterpret_cast<int*>(thr + 1);                       // number of candidates (+ 1 int of padding)
    int* cpos = cnt + 2;                                              // [m] their positions
    int* crank = cpos + m;                                            // [m] their ranks
    if (tid < 32) {
        // lanes l and l + 16 share group l % 16 and take alternate members
        long long gm = -1;
#pragma unroll 4
        for (int e = lane; e < m; e += 32) gm = max(gm, key[e]);
        gm = max(gm, __shfl_xor_sync(0xffffffffu, gm, 16));
        long long t = gm;   // m >= NB: no group is empty
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) t = min(t, __shfl_xor_sync(0xffffffffu, t, o));
        if (lane == 0) { *thr = t; *cnt = 0; }
    }
    __syncthreads();
    const long long tau = *thr;
    for (int e = tid; e < m; e += NT) {
        const long long k = key[e];
        if (k >= tau) {
            const int idx = atomicAdd(cnt, 1);
            ckey[idx] = k; cpos[idx] = j + e; crank[idx] = 0;
        }
    }
    __syncthreads();
    const int nc = *cnt;
    {
        const int grp = tid >> 7, t = tid & 127;
        const int mq = (nc + G - 1) / G, o0 = grp * mq, o1 = min(nc, o0 + mq);
        for (int c = t; c < nc; c += 128) {
            const long long vk = ckey[c];
            const int vp = cpos[c];
            int rank = 0;
#pragma unroll 4
            for (int o = o0; o < o1; ++o)
                rank += (int)((unsigned long long)(vk - (cpos[o] < vp ? 1 : 0) - ckey[o]) >> 63);
            if (rank) atomicAdd(&crank[c], rank);
        }
    }
    __syncthreads();
    for (int c = tid; c < nc; c += NT) {
        const int r = crank[c];
        if (r < NB) sel[r] = cpos[c];
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------------------
// Panel layout: NW warps; for each of its row tiles t = warp + NW * u (8 rows of the active part) a thread owns row
// l/4 and the four columns 4 * (l%4) + {0..3}: x[u][cc].
//
// Step k (reflector! and reflectorApply!, src/linalg/UDT.jl:32-50, 133-148): with a = current panel,
//   s_c = sum_{i>k} a_ik a_ic  (c >= k; c = k gives the squared norm below the diagonal),
//   normu = sqrt(a_kk^2 + s_k), nu = copysign(normu, a_kk), xi1 = a_kk + nu, tau = xi1 / nu, R_kk = -nu,
//   v = [1; a_{k+1:,k} / xi1],  w_c = a_kc + s_c / xi1,  a_c -= tau w_c v.
// The s_c of all columns travel in ONE reduction: 3 shuffle levels inside a warp, per-warp partials through shared
// memory, one barrier. part: [2][NW][NB], rowk: [2][NB] doubles of shared memory; taus: [NB]; Dout: D + j (global).
// BAR_ID == 0: the NW warps are the whole CTA (__syncthreads); otherwise they are warps 0 .. NW-1 of a larger CTA
// and meet at the named barrier BAR_ID.
template <int NW, int BAR_ID>
__device__ __forceinline__ void panel_barrier()
{
    if constexpr (BAR_ID == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"n"(BAR_ID), "n"(NW * 32) : "memory");
}

// column of the panel held in register cc of a thread with q = lane % 4. Register-major: at step k every column of the
// registers cc < k / 4 is finished in ALL lanes, so their dot products, reductions and updates drop out at compile time
// (on average 3/8 of the FP64 and shuffle work of a panel).
__device__ __forceinline__ constexpr int panel_col(int q, int cc) { return 4 * cc + q; }

template <int NW, int TPW, int BAR_ID = 0>
__device__ __forceinline__ void panel_steps(double (&x)[TPW][4], int mt, int warp, int lane, double* part, double* rowk,
                                            double* taus, double* Dout)
{
    const int q = lane & 3, rl = lane >> 2;
    const int i0 = 8 * warp + rl; // row of tile u = 0; tiles u >= 1 start at row 8 * NW >= NB: always below the diagonal
    static_assert(8 * NW >= NB, "tile 0 of every warp must cover the rows on and above the panel's diagonal");
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        const int buf = k & 1;
        const int kc = k >> 2, kq = k & 3;   // column k = register kc of the lanes with q == kq
        double ak[TPW];
#pragma unroll
        for (int u = 0; u < TPW; ++u) ak[u] = __shfl_sync(0xffffffffu, x[u][kc], (lane & ~3) | kq);
        double p[4], p2[4];
        {
            const double a0 = i0 > k ? ak[0] : 0.0;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
                if (cc >= kc) { p[cc] = a0 * x[0][cc]; p2[cc] = 0.0; }
        }
#pragma unroll
        for (int u = 1; u < TPW; ++u) {
            if (warp + NW * u < mt) {
#pragma unroll
                for (int cc = 0; cc < 4; ++cc)
                    if (cc >= kc) {
                        if (u & 1) p2[cc] = fma(ak[u], x[u][cc], p2[cc]);
                        else p[cc] = fma(ak[u], x[u][cc], p[cc]);
                    }
            }
        }
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
            if (cc >= kc) {
                p[cc] += p2[cc];
                p[cc] += __shfl_xor_sync(0xffffffffu, p[cc], 4);
                p[cc] += __shfl_xor_sync(0xffffffffu, p[cc], 8);
                p[cc] += __shfl_xor_sync(0xffffffffu, p[cc], 16);
            }
        if (rl == 0) {
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
                if (cc >= kc) part[(buf * NW + warp) * NB + panel_col(q, cc)] = p[cc];
        }
        if (warp == (k >> 3) && rl == (k & 7)) { // the owner of row k publishes it
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
                if (cc >= kc) rowk[buf * NB + panel_col(q, cc)] = x[0][cc];
        }
        panel_barrier<NW, BAR_ID>();
        double s = 0.0;
        if (lane < NB) {   // lane = column; the entries of finished columns are stale and never used
            double s2 = 0.0;
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                if (w & 1) s2 += part[(buf * NW + w) * NB + lane];
                else s += part[(buf * NW + w) * NB + lane];
            }
            s += s2;
        }
        const double sk = __shfl_sync(0xffffffffu, s, k);
        const double akk = rowk[buf * NB + k];
        const double normu2 = fma(akk, akk, sk);
        const bool live = normu2 != 0.0;
        double nu = 0.0, rxi = 0.0, tau = 0.0;
        if (live) {
            const double rn = fast_rsqrt(normu2);
            nu = copysign(normu2 * rn, akk);
            const double xi1 = akk + nu;
            rxi = fast_rcp(xi1);
            tau = xi1 * copysign(rn, akk);
        }
        if (warp == 0 && lane == 0) { taus[k] = tau; Dout[k] = fabs(nu); }
        double f[4], g[4];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
            if (cc >= kc) {
                const int c = panel_col(q, cc);
                const double sc = __shfl_sync(0xffffffffu, s, c);
                const double wc = fma(sc, rxi, rowk[buf * NB + c]);     // v' a_c
                f[cc] = c > k ? tau * wc : 0.0;
                g[cc] = f[cc] * rxi;
            }
        // column k itself becomes v: scaled by 1 / xi1 below the diagonal (register kc of the lanes with q == kq)
        const double m1 = (q == kq && live) ? rxi : 1.0;
        // tile 0 holds the rows on and above the diagonal of this step
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
            if (cc >= kc) {
                const int c = panel_col(q, cc);
                const double old = x[0][cc];
                const double scaled = cc == kc ? old * m1 : old;
                const double below = fma(-g[cc], ak[0], scaled);
                const double onrow = (c == k && live) ? -nu : old - f[cc];
                x[0][cc] = i0 > k ? below : (i0 == k ? onrow : old);
            }
#pragma unroll
        for (int u = 1; u < TPW; ++u)
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
                if (cc >= kc) x[u][cc] = fma(-g[cc], ak[u], cc == kc ? x[u][cc] * m1 : x[u][cc]);
    }
}

// S = V'V (NB x NB, only the strict upper part is used) of the panel stored column-major in P (column c at
// P + c * ldp, active rows start at P[.. + row0], ldp % 16 == 8, row0 even), V = unit lower trapezoid: the caller has
// patched the top NB x NB block (ones on the diagonal, zeros above). Warps 0 .. GW-1 split the row tiles; their
// partial tiles are summed in fixed order through Sred [GW][NB*NB]. Result in Ssm[c * NB + k]. Two barriers inside.
// DMMA with the k index standing for row pairs: one 16-byte load feeds the A operand of row tile c and the B operand
// of column tile k.
template <int GW>
__device__ __forceinline__ void panel_gram(const double* P, int ldp, int row0, int mt, int warp, int lane, int nthreads,
                                           int tid, double* Sred, double* Ssm)
{
    const int q = lane & 3, rl = lane >> 2;
    if (warp < GW) {
        double acc[2][2][2] = {};
        for (int t = warp; t < mt; t += GW) {
            double2 a[2];
#pragma unroll
            for (int ct = 0; ct < 2; ++ct)
                a[ct] = *reinterpret_cast<const double2*>(P + (size_t)(ct * 8 + rl) * ldp + row0 + 8 * t + 2 * q);
#pragma unroll
            for (int ct = 0; ct < 2; ++ct)
#pragma unroll
                for (int kt = 0; kt < 2; ++kt) {
                    dmma884(acc[ct][kt][0], acc[ct][kt][1], a[ct].x, a[kt].x);
                    dmma884(acc[ct][kt][0], acc[ct][kt][1], a[ct].y, a[kt].y);
                }
        }
#pragma unroll
        for (int ct = 0; ct < 2; ++ct)
#pragma unroll
            for (int kt = 0; kt < 2; ++kt)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    Sred[(size_t)warp * NB * NB + (ct * 8 + rl) * NB + kt * 8 + 2 * q + e] = acc[ct][kt][e];
    }
    __syncthreads();
    for (int e = tid; e < NB * NB; e += nthreads) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < GW; ++w) s += Sred[(size_t)w * NB * NB + e];
        Ssm[e] = s;
    }
    __syncthreads();
}

// T of H_1 .. H_NB = I - V T V' (dlarft, forward / columnwise): T_kk = tau_k, T[0:k, k] = -tau_k T[0:k, 0:k] S[0:k, k].
// Row i lives in lane i of the calling warp (lanes >= NB idle). Tout: column-major NB x NB, Tout[i + k * NB].
__device__ __forceinline__ void panel_tfactor(const double* Ssm, const double* taus, int lane, double* Tout)
{
    if (lane >= NB) return;
    double Trow[NB];
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        double acc = 0.0;
#pragma unroll
        for (int mm = 0; mm < k; ++mm) {
            const double sv = Ssm[mm * NB + k];
            if (mm >= lane) acc = fma(Trow[mm], sv, acc);
        }
        const double tk = taus[k];
        Trow[k] = lane < k ? -tk * acc : (lane == k ? tk : 0.0);
    }
#pragma unroll
    for (int k = 0; k < NB; ++k) Tout[lane + k * NB] = Trow[k];
}

// ---------------------------------------------------------------------------------------------------------------------
// C <- (I - V op(T) V') C on an 8-column strip held in registers: c[t] = C[rows 8t + 2q + {0,1}][column l/4] for the mt
// active row tiles, i.e. the DMMA accumulator layout of the TRANSPOSED tile, which is also a valid A-operand layout if
// k slot q of two consecutive DMMA steps stands for rows 2q and 2q+1. Operands in shared memory:
//   Vs [NB][ldv]   V[r][k] at k * ldv + r      (ldv % 16 == 8)
//   Vt [m][TLD]    V[r][k] at r * TLD + k
//   Ts [NB][TLD]   op(T)[k][k2] at k2 * TLD + k   (op(T) = T' for Q' C, T for Q C; the caller fills it)
// Returns the squared norm of the rows >= 8 * t_norm of the thread's column (partial: sum over the 4 lanes of a column).
// BOTTOM: the tiles are counted from the last row upwards (c[t] = rows m - 8 (t + 1) + 2q + {0,1} of the m active
// ones), so that a strip kept in registers across panels of growing height always uses a prefix of its tiles (form Q).
template <int RT, bool BOTTOM = false>
__device__ __forceinline__ double apply_strip(double (&c)[RT][2], int mt, int t_norm, const double* Vs, int ldv,
                                              const double* Vt, const double* Ts, int lane, int m = 0)
{
    const int q = lane & 3, rl = lane >> 2;
    double w[2][2];
    {
        double wh[2][2][2] = {}; // even / odd row tiles: four independent DMMA chains
        // BOTTOM walks its tiles from the highest index down, i.e. through the rows from the top like the other mode, and
        // mt is even there (m % 16 == 0), so tile t is the (t + 1) & 1 -th chain: the same sums in the same order, bit for bit
#pragma unroll
        for (int tt = 0; tt < RT; ++tt) {
            const int t = BOTTOM ? RT - 1 - tt : tt;
            const int h = BOTTOM ? (t + 1) & 1 : t & 1;
            if (t < mt) {
#pragma unroll
                for (int kt = 0; kt < 2; ++kt) {
                    const int r0 = BOTTOM ? m - 8 * (t + 1) : 8 * t;
                    const double2 b = *reinterpret_cast<const double2*>(Vs + (size_t)(kt * 8 + rl) * ldv + r0 + 2 * q);
                    dmma884(wh[h][kt][0], wh[h][kt][1], c[t][0], b.x);
                    dmma884(wh[h][kt][0], wh[h][kt][1], c[t][1], b.y);
                }
            }
        }
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) { w[kt][0] = wh[0][kt][0] + wh[1][kt][0]; w[kt][1] = wh[0][kt][1] + wh[1][kt][1]; }
    }
    double y[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
    for (int ko = 0; ko < 2; ++ko)
#pragma unroll
        for (int ki = 0; ki < 2; ++ki) {
            const double2 b = *reinterpret_cast<const double2*>(Ts + (size_t)(ko * 8 + rl) * TLD + ki * 8 + 2 * q);
            dmma884(y[ko][0], y[ko][1], w[ki][0], b.x);
            dmma884(y[ko][0], y[ko][1], w[ki][1], b.y);
        }
#pragma unroll
    for (int ko = 0; ko < 2; ++ko) { y[ko][0] = -y[ko][0]; y[ko][1] = -y[ko][1]; }
    double nrm = 0.0;
#pragma unroll
    for (int t = 0; t < RT; ++t) {
        if (t < mt) {
#pragma unroll
            for (int kt = 0; kt < 2; ++kt) {
                const int r0 = BOTTOM ? m - 8 * (t + 1) : 8 * t;
                const double2 b = *reinterpret_cast<const double2*>(Vt + (size_t)(r0 + rl) * TLD + kt * 8 + 2 * q);
                dmma884(c[t][0], c[t][1], y[kt][0], b.x);
                dmma884(c[t][0], c[t][1], y[kt][1], b.y);
            }
            if (t >= t_norm) nrm += c[t][0] * c[t][0] + c[t][1] * c[t][1];
        }
    }
    return nrm;
}

// ---------------------------------------------------------------------------------------------------------------------
// Variants for the shared-memory-resident factorisation (udt_resident.cu). The reflector block lives ONCE in shared
// memory, column-major with absolute row indices, V[r][k] at vs_index(k, r, ld): the row is XOR-ed with 8 for the
// columns k with bit 1 set. That makes both access patterns of apply_strip conflict-free on one copy (ld % 16 == 8):
//   first product  (16-byte loads: column kt*8 + l/4, row pair 8t + 2q)            8 lanes -> 8 distinct 16-byte slots
//   third product  ( 8-byte loads: columns kt*8 + 2q (+1), row 8t + l/4)           2 lanes per bank pair = 2 wavefronts
// Rows come in aligned groups of 16 (the active part starts at a multiple of 16 and has an even number of row tiles).
__device__ __forceinline__ int vs_index(int k, int r, int ld) { return k * ld + (r ^ ((k & 2) << 2)); }

// panel_gram on the swizzled buffer: the panel's active rows are row0 .. row0 + 8 mt - 1 (absolute).
template <int GW>
__device__ __forceinline__ void panel_gram_sw(const double* Vs, int ld, int row0, int mt, int warp, int lane, int nthreads,
                                              int tid, double* Sred, double* Ssm)
{
    const int q = lane & 3, rl = lane >> 2;
    if (warp < GW) {
        double acc[2][2][2] = {};
        for (int t = warp; t < mt; t += GW) {
            double2 a[2];
#pragma unroll
            for (int ct = 0; ct < 2; ++ct)
                a[ct] = *reinterpret_cast<const double2*>(Vs + vs_index(ct * 8 + rl, row0 + 8 * t + 2 * q, ld));
#pragma unroll
            for (int ct = 0; ct < 2; ++ct)
#pragma unroll
                for (int kt = 0; kt < 2; ++kt) {
                    dmma884(acc[ct][kt][0], acc[ct][kt][1], a[ct].x, a[kt].x);
                    dmma884(acc[ct][kt][0], acc[ct][kt][1], a[ct].y, a[kt].y);
                }
        }
#pragma unroll
        for (int ct = 0; ct < 2; ++ct)
#pragma unroll
            for (int kt = 0; kt < 2; ++kt)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    Sred[(size_t)warp * NB * NB + (ct * 8 + rl) * NB + kt * 8 + 2 * q + e] = acc[ct][kt][e];
    }
    __syncthreads();
    for (int e = tid; e < NB * NB; e += nthreads) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < GW; ++w) s += Sred[(size_t)w * NB * NB + e];
        Ssm[e] = s;
    }
    __syncthreads();
}

// apply_strip on the swizzled buffer; the strip's tiles are rows row0 + 8t .. of its 8 columns.
template <int RT>
__device__ __forceinline__ double apply_strip_sw(double (&c)[RT][2], int mt, int t_norm, const double* Vs, int ld, int row0,
                                                 const double* Ts, int lane)
{
    const int q = lane & 3, rl = lane >> 2;
    double w[2][2];
    {
        double wh[2][2][2] = {};
#pragma unroll
        for (int t = 0; t < RT; ++t) {
            if (t < mt) {
#pragma unroll
                for (int kt = 0; kt < 2; ++kt) {
                    const double2 b = *reinterpret_cast<const double2*>(Vs + vs_index(kt * 8 + rl, row0 + 8 * t + 2 * q, ld));
                    dmma884(wh[t & 1][kt][0], wh[t & 1][kt][1], c[t][0], b.x);
                    dmma884(wh[t & 1][kt][0], wh[t & 1][kt][1], c[t][1], b.y);
                }
            }
        }
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) { w[kt][0] = wh[0][kt][0] + wh[1][kt][0]; w[kt][1] = wh[0][kt][1] + wh[1][kt][1]; }
    }
    double y[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
    for (int ko = 0; ko < 2; ++ko)
#pragma unroll
        for (int ki = 0; ki < 2; ++ki) {
            const double2 b = *reinterpret_cast<const double2*>(Ts + (size_t)(ko * 8 + rl) * TLD + ki * 8 + 2 * q);
            dmma884(y[ko][0], y[ko][1], w[ki][0], b.x);
            dmma884(y[ko][0], y[ko][1], w[ki][1], b.y);
        }
#pragma unroll
    for (int ko = 0; ko < 2; ++ko) { y[ko][0] = -y[ko][0]; y[ko][1] = -y[ko][1]; }
    double nrm = 0.0;
#pragma unroll
    for (int t = 0; t < RT; ++t) {
        if (t < mt) {
#pragma unroll
            for (int kt = 0; kt < 2; ++kt) {
                const int k0 = kt * 8 + 2 * q, r = (row0 + 8 * t + rl) ^ ((q & 1) << 3); // same swizzle for k0 and k0 + 1
                dmma884(c[t][0], c[t][1], y[kt][0], Vs[k0 * ld + r]);
                dmma884(c[t][0], c[t][1], y[kt][1], Vs[(k0 + 1) * ld + r]);
            }
            if (t >= t_norm) nrm += c[t][0] * c[t][0] + c[t][1] * c[t][1];
        }
    }
    return nrm;
}

} // namespace qrb
} // namespace mcb
