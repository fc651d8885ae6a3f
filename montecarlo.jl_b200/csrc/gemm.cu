// Batched FP64 DMMA GEMM, tile 64x64x16, 4 warps (2x2) of 32x32, register-staged double
// buffering (global loads of tile k+1 are in flight while tile k is multiplied).
#include "gemm.cuh"

namespace mcb {

namespace {

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int THREADS = 128;
constexpr int LDX = BM + 4;  // [k][x] layout: conflict-free fragment reads need ld % 16 == 4
constexpr int LDK = BK + 4;  // [x][k] layout
constexpr int TILE_ELEMS = BM * BK; // 1024, 8 per thread
constexpr int PER_THREAD = TILE_ELEMS / THREADS;
constexpr int SMEM_TILE = (BM * LDK > BK * LDX) ? BM * LDK : BK * LDX; // 1280 doubles

// XC = the tile is contiguous in its non-K index in global memory (A: !transA, B: transB)
template <bool XC>
__device__ __forceinline__ void load_tile(double (&r)[PER_THREAD], const double* __restrict__ P, int n,
                                          int x0, int k0, int tid)
{
#pragma unroll
    for (int e = 0; e < PER_THREAD; ++e) {
        const int idx = tid + e * THREADS;
        int x, k;
        if (XC) { x = idx % BM; k = idx / BM; }
        else { k = idx % BK; x = idx / BK; }
        const int gx = x0 + x, gk = k0 + k;
        double v = 0.0;
        if (gx < n && gk < n) v = XC ? P[(size_t)gx + (size_t)gk * n] : P[(size_t)gk + (size_t)gx * n];
        r[e] = v;
    }
}

template <bool XC, bool SCALE>
__device__ __forceinline__ void store_tile(const double (&r)[PER_THREAD], double* S, int tid,
                                           const Scale& mid, int chain, int k0, int n)
{
#pragma unroll
    for (int e = 0; e < PER_THREAD; ++e) {
        const int idx = tid + e * THREADS;
        int x, k;
        if (XC) { x = idx % BM; k = idx / BM; }
        else { k = idx % BK; x = idx / BK; }
        double v = r[e];
        if (SCALE) { const int gk = k0 + k; if (gk < n) v *= mid.at(chain, gk); }
        if (XC) S[k * LDX + x] = v; else S[x * LDK + k] = v;
    }
}

template <bool XC>
__device__ __forceinline__ double frag(const double* S, int x, int k)
{
    return XC ? S[k * LDX + x] : S[x * LDK + k];
}

template <bool TA, bool TB>
__global__ void __launch_bounds__(THREADS) gemm_kernel(GemmArgs g)
{
    __shared__ double As[2][SMEM_TILE];
    __shared__ double Bs[2][SMEM_TILE];
    const int n = g.n;
    const int tiles = (n + BM - 1) / BM;
    const int tm = blockIdx.x % tiles, tn = blockIdx.x / tiles;
    const int chain = blockIdx.y;
    const double* A = g.A + (size_t)chain * g.strideA;
    const double* B = g.B + (size_t)chain * g.strideB;
    double* C = g.C + (size_t)chain * g.strideC;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = (warp & 1) * 32, wn = (warp >> 1) * 32;
    const int r = lane >> 2, q = lane & 3;
    const int m0 = tm * BM, n0 = tn * BN;
    const bool has_mid = g.mid.active();

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double ra[PER_THREAD], rb[PER_THREAD];
    const int nk = (n + BK - 1) / BK;
    load_tile<!TA>(ra, A, n, m0, 0, tid);
    load_tile<TB>(rb, B, n, n0, 0, tid);
    if (has_mid) store_tile<!TA, true>(ra, As[0], tid, g.mid, chain, 0, n);
    else store_tile<!TA, false>(ra, As[0], tid, g.mid, chain, 0, n);
    store_tile<TB, false>(rb, Bs[0], tid, g.mid, chain, 0, n);
    __syncthreads();

    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) {
            load_tile<!TA>(ra, A, n, m0, (kt + 1) * BK, tid);
            load_tile<TB>(rb, B, n, n0, (kt + 1) * BK, tid);
        }
        const double* as = As[cur];
        const double* bs = Bs[cur];
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = frag<!TA>(as, wm + i * 8 + r, kk + q);
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = frag<TB>(bs, wn + j * 8 + r, kk + q);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        if (kt + 1 < nk) {
            const int nxt = cur ^ 1;
            if (has_mid) store_tile<!TA, true>(ra, As[nxt], tid, g.mid, chain, (kt + 1) * BK, n);
            else store_tile<!TA, false>(ra, As[nxt], tid, g.mid, chain, (kt + 1) * BK, n);
            store_tile<TB, false>(rb, Bs[nxt], tid, g.mid, chain, (kt + 1) * BK, n);
        }
        __syncthreads();
    }

    const bool has_row = g.row.active(), has_col = g.col.active();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + wm + i * 8 + r;
        if (gm >= n) continue;
        const double rs = has_row ? g.row.at(chain, gm) : 1.0;
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int gn = n0 + wn + j * 8 + 2 * q + c;
                if (gn >= n) continue;
                double v = acc[i][j][c];
                if (has_row) v *= rs;
                if (has_col) v *= g.col.at(chain, gn);
                if (g.add && gm == gn) v += g.add[(size_t)chain * g.stride_add + gm];
                C[(size_t)gm + (size_t)gn * n] = v;
            }
    }
}

} // namespace

int launch_gemm(const GemmArgs& g, cudaStream_t st)
{
    if (g.n <= 0 || g.batch <= 0) return 0;
    if (g.C == g.A || g.C == g.B) { set_error("gemm: output aliases an input"); return -1; }
    const int tiles = (g.n + BM - 1) / BM;
    dim3 grid(tiles * tiles, g.batch);
    LaunchScope scope(LC_GEMM, st, 2.0 * g.n * (double)g.n * g.n * g.batch);
    if (!g.transA && !g.transB) gemm_kernel<false, false><<<grid, THREADS, 0, st>>>(g);
    else if (!g.transA && g.transB) gemm_kernel<false, true><<<grid, THREADS, 0, st>>>(g);
    else if (g.transA && !g.transB) gemm_kernel<true, false><<<grid, THREADS, 0, st>>>(g);
    else gemm_kernel<true, true><<<grid, THREADS, 0, st>>>(g);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mcb
