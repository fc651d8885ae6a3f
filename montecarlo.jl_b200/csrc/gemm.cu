// Batched FP64 DMMA GEMM, square tile T x T x BK with T in {48, 64, 80} chosen so that it divides the
// matrix size (12x12 lattice -> 48, 8x8 -> 64, 20x20 -> 80; no padded tiles; BK = 16, or 8 for T = 80 so that
// three CTAs fit the register file of an SM: 253 -> 200 us per batched 400^3 product), 4 warps (2x2) of
// T/2 x T/2, register-staged double buffering (global loads of tile k+1 are in flight while tile k
// is multiplied).
#include "gemm.cuh"
#include <stdlib.h>

namespace mcb {

namespace {

constexpr int THREADS = 128;

template <int T>
struct Tile {
    // K depth of a shared-memory tile: 8 for the 80-wide tile (its 100 accumulator registers leave room for
    // 3 CTAs per SM only with half the staging registers), 16 otherwise
    static constexpr int BK = T >= 80 ? 8 : 16;
    static constexpr int LDK = BK + 4;  // [x][k] layout
    static constexpr int LDX = T + 4;  // [k][x] layout: conflict-free fragment reads need ld % 16 == 4
    static constexpr int PER_THREAD = T * BK / THREADS;
    static constexpr int SMEM_TILE = (T * LDK > BK * LDX) ? T * LDK : BK * LDX;
    static constexpr int WT = T / 16;  // DMMA tiles per warp in each direction (warp tile T/2 x T/2)
    static_assert((T * BK) % THREADS == 0 && T % 16 == 0 && LDX % 16 == 4, "unsupported tile");
};

// XC = the tile is contiguous in its non-K index in global memory (A: !transA, B: transB)
template <int T, bool XC>
__device__ __forceinline__ void load_tile(double (&r)[Tile<T>::PER_THREAD], const double* __restrict__ P, int n,
                                          int x0, int k0, int tid)
{
    constexpr int BM = T, BK = Tile<T>::BK;
#pragma unroll
    for (int e = 0; e < Tile<T>::PER_THREAD; ++e) {
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

template <int T, bool XC, bool SCALE>
__device__ __forceinline__ void store_tile(const double (&r)[Tile<T>::PER_THREAD], double* S, int tid,
                                           const Scale& mid, int chain, int k0, int n)
{
    constexpr int BM = T, LDX = Tile<T>::LDX, BK = Tile<T>::BK, LDK = Tile<T>::LDK;
#pragma unroll
    for (int e = 0; e < Tile<T>::PER_THREAD; ++e) {
        const int idx = tid + e * THREADS;
        int x, k;
        if (XC) { x = idx % BM; k = idx / BM; }
        else { k = idx % BK; x = idx / BK; }
        double v = r[e];
        if (SCALE) { const int gk = k0 + k; if (gk < n) v *= mid.at(chain, gk); }
        if (XC) S[k * LDX + x] = v; else S[x * LDK + k] = v;
    }
}

template <int T, bool XC>
__device__ __forceinline__ double frag(const double* S, int x, int k)
{
    return XC ? S[k * Tile<T>::LDX + x] : S[x * Tile<T>::LDK + k];
}

template <int T, bool TA, bool TB>
__global__ void __launch_bounds__(THREADS) gemm_kernel(GemmArgs g)
{
    constexpr int BM = T, BN = T, WT = Tile<T>::WT, PER_THREAD = Tile<T>::PER_THREAD, BK = Tile<T>::BK;
    extern __shared__ __align__(16) double gemm_smem[];
    double (*As)[Tile<T>::SMEM_TILE] = reinterpret_cast<double (*)[Tile<T>::SMEM_TILE]>(gemm_smem);
    double (*Bs)[Tile<T>::SMEM_TILE] = As + 2;
    const int n = g.n;
    const int tiles = (n + BM - 1) / BM;
    const int tm = blockIdx.x % tiles, tn = blockIdx.x / tiles;
    const int chain = blockIdx.y;
    const double* A = g.A + (size_t)chain * g.strideA;
    const double* B = g.B + (size_t)chain * g.strideB;
    double* C = g.C + (size_t)chain * g.strideC;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = (warp & 1) * (T / 2), wn = (warp >> 1) * (T / 2);
    const int r = lane >> 2, q = lane & 3;
    const int m0 = tm * BM, n0 = tn * BN;
    const bool has_mid = g.mid.active();

    double acc[WT][WT][2];
#pragma unroll
    for (int i = 0; i < WT; ++i)
#pragma unroll
        for (int j = 0; j < WT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double ra[PER_THREAD], rb[PER_THREAD];
    const int nk = (n + BK - 1) / BK;
    load_tile<T, !TA>(ra, A, n, m0, 0, tid);
    load_tile<T, TB>(rb, B, n, n0, 0, tid);
    if (has_mid) store_tile<T, !TA, true>(ra, As[0], tid, g.mid, chain, 0, n);
    else store_tile<T, !TA, false>(ra, As[0], tid, g.mid, chain, 0, n);
    store_tile<T, TB, false>(rb, Bs[0], tid, g.mid, chain, 0, n);
    __syncthreads();

    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) {
            load_tile<T, !TA>(ra, A, n, m0, (kt + 1) * BK, tid);
            load_tile<T, TB>(rb, B, n, n0, (kt + 1) * BK, tid);
        }
        const double* as = As[cur];
        const double* bs = Bs[cur];
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a[WT], b[WT];
#pragma unroll
            for (int i = 0; i < WT; ++i) a[i] = frag<T, !TA>(as, wm + i * 8 + r, kk + q);
#pragma unroll
            for (int j = 0; j < WT; ++j) b[j] = frag<T, TB>(bs, wn + j * 8 + r, kk + q);
#pragma unroll
            // the product is formed transposed (B fragments as the A operand): a thread's two accumulator elements are then
            // C(row, col), C(row + 1, col), adjacent in the column-major result: one 16-byte store instead of two of 8
            for (int i = 0; i < WT; ++i)
#pragma unroll
                for (int j = 0; j < WT; ++j) dmma884(acc[i][j][0], acc[i][j][1], b[j], a[i]);
        }
        if (kt + 1 < nk) {
            const int nxt = cur ^ 1;
            if (has_mid) store_tile<T, !TA, true>(ra, As[nxt], tid, g.mid, chain, (kt + 1) * BK, n);
            else store_tile<T, !TA, false>(ra, As[nxt], tid, g.mid, chain, (kt + 1) * BK, n);
            store_tile<T, TB, false>(rb, Bs[nxt], tid, g.mid, chain, (kt + 1) * BK, n);
        }
        __syncthreads();
    }

    // acc[i][j][c] = C(m0 + wm + 8 i + 2 q + c, n0 + wn + 8 j + r)
    const bool has_row = g.row.active(), has_col = g.col.active();
    const bool vec_ok = (n & 1) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0;
    double cs[WT];
#pragma unroll
    for (int j = 0; j < WT; ++j) {
        const int gn = n0 + wn + j * 8 + r;
        cs[j] = (has_col && gn < n) ? g.col.at(chain, gn) : 1.0;
    }
#pragma unroll
    for (int i = 0; i < WT; ++i) {
        const int gm = m0 + wm + i * 8 + 2 * q;
        if (gm >= n) continue;
        const bool two = gm + 1 < n;
        const double rs0 = has_row ? g.row.at(chain, gm) : 1.0;
        const double rs1 = (has_row && two) ? g.row.at(chain, gm + 1) : 1.0;
#pragma unroll
        for (int j = 0; j < WT; ++j) {
            const int gn = n0 + wn + j * 8 + r;
            if (gn >= n) continue;
            double v0 = acc[i][j][0], v1 = acc[i][j][1];
            if (has_row) { v0 *= rs0; v1 *= rs1; }
            if (has_col) { v0 *= cs[j]; v1 *= cs[j]; }
            if (g.add) {
                if (gm == gn) v0 += g.add[(size_t)chain * g.stride_add + gm];
                if (gm + 1 == gn) v1 += g.add[(size_t)chain * g.stride_add + gn];
            }
            double* dst = C + (size_t)gm + (size_t)gn * n;
            if (vec_ok) *reinterpret_cast<double2*>(dst) = make_double2(v0, v1);   // (n even: gm + 1 < n)
            else { dst[0] = v0; if (two) dst[1] = v1; }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// One CTA per chain, the whole NN x NN product in the registers of 12 warps (6 x 2 warps of 24 x 72,
// 3 x 9 DMMA tiles each): for 128 chains of N = 144 (config C) the tiled kernel above spends half its time in
// the prologue / epilogue of 1152 short-lived CTAs; here every SM streams one matrix product from start to end.
// K depth 8 per shared-memory stage, register-staged double buffering as above.
template <int NN>
struct Full {
    static constexpr int THREADS = 384, BK = 8;
    static constexpr int WM = NN / 6, WN = NN / 2;       // warp tile
    static constexpr int IT = WM / 8, JT = WN / 8;       // DMMA tiles per warp
    static constexpr int LDX = NN + 4, LDK = BK + 4;     // [k][x] and [x][k] layouts (conflict-free fragment reads)
    static constexpr int PER_THREAD = NN * BK / THREADS;
    static constexpr int SMEM_TILE = (NN * LDK > BK * LDX) ? NN * LDK : BK * LDX;
    static_assert(NN % 48 == 0 && (NN * BK) % THREADS == 0 && LDX % 16 == 4, "unsupported size");
};

template <int NN, bool XC>
__device__ __forceinline__ void full_load(double (&r)[Full<NN>::PER_THREAD], const double* __restrict__ P, int k0, int tid)
{
    constexpr int BK = Full<NN>::BK;
#pragma unroll
    for (int e = 0; e < Full<NN>::PER_THREAD; ++e) {
        const int idx = tid + e * Full<NN>::THREADS;
        int x, k;
        if (XC) { x = idx % NN; k = idx / NN; }
        else { k = idx % BK; x = idx / BK; }
        r[e] = XC ? P[(size_t)x + (size_t)(k0 + k) * NN] : P[(size_t)(k0 + k) + (size_t)x * NN];
    }
}

template <int NN, bool XC, bool SCALE>
__device__ __forceinline__ void full_store(const double (&r)[Full<NN>::PER_THREAD], double* S, int tid, const Scale& mid,
                                           int chain, int k0)
{
    constexpr int BK = Full<NN>::BK;
#pragma unroll
    for (int e = 0; e < Full<NN>::PER_THREAD; ++e) {
        const int idx = tid + e * Full<NN>::THREADS;
        int x, k;
        if (XC) { x = idx % NN; k = idx / NN; }
        else { k = idx % BK; x = idx / BK; }
        double v = r[e];
        if (SCALE) v *= mid.at(chain, k0 + k);
        if (XC) S[k * Full<NN>::LDX + x] = v; else S[x * Full<NN>::LDK + k] = v;
    }
}

template <int NN, bool TA, bool TB>
__global__ void __launch_bounds__(Full<NN>::THREADS, 1) gemm_full_kernel(GemmArgs g)
{
    using F = Full<NN>;
    constexpr int BK = F::BK, IT = F::IT, JT = F::JT;
    extern __shared__ __align__(16) double gemm_smem[];
    double (*As)[F::SMEM_TILE] = reinterpret_cast<double (*)[F::SMEM_TILE]>(gemm_smem);
    double (*Bs)[F::SMEM_TILE] = As + 2;
    const int chain = blockIdx.x;
    const double* A = g.A + (size_t)chain * g.strideA;
    const double* B = g.B + (size_t)chain * g.strideB;
    double* C = g.C + (size_t)chain * g.strideC;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = (warp % 6) * F::WM, wn = (warp / 6) * F::WN;
    const int r = lane >> 2, q = lane & 3;
    const bool has_mid = g.mid.active();

    double acc[IT][JT][2];
#pragma unroll
    for (int i = 0; i < IT; ++i)
#pragma unroll
        for (int j = 0; j < JT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    double ra[F::PER_THREAD], rb[F::PER_THREAD];
    constexpr int nk = NN / BK;
    full_load<NN, !TA>(ra, A, 0, tid);
    full_load<NN, TB>(rb, B, 0, tid);
    if (has_mid) full_store<NN, !TA, true>(ra, As[0], tid, g.mid, chain, 0);
    else full_store<NN, !TA, false>(ra, As[0], tid, g.mid, chain, 0);
    full_store<NN, TB, false>(rb, Bs[0], tid, g.mid, chain, 0);
    __syncthreads();
    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) {
            full_load<NN, !TA>(ra, A, (kt + 1) * BK, tid);
            full_load<NN, TB>(rb, B, (kt + 1) * BK, tid);
        }
        const double* as = As[cur];
        const double* bs = Bs[cur];
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a[IT], b[JT];
#pragma unroll
            for (int i = 0; i < IT; ++i) {
                const int x = wm + i * 8 + r, k = kk + q;
                a[i] = !TA ? as[k * F::LDX + x] : as[x * F::LDK + k];
            }
#pragma unroll
            for (int j = 0; j < JT; ++j) {
                const int x = wn + j * 8 + r, k = kk + q;
                b[j] = TB ? bs[k * F::LDX + x] : bs[x * F::LDK + k];
            }
#pragma unroll
            for (int i = 0; i < IT; ++i)   // (transposed product, as in gemm_kernel: 16-byte stores in the epilogue)
#pragma unroll
                for (int j = 0; j < JT; ++j) dmma884(acc[i][j][0], acc[i][j][1], b[j], a[i]);
        }
        if (kt + 1 < nk) {
            const int nxt = cur ^ 1;
            if (has_mid) full_store<NN, !TA, true>(ra, As[nxt], tid, g.mid, chain, (kt + 1) * BK);
            else full_store<NN, !TA, false>(ra, As[nxt], tid, g.mid, chain, (kt + 1) * BK);
            full_store<NN, TB, false>(rb, Bs[nxt], tid, g.mid, chain, (kt + 1) * BK);
        }
        __syncthreads();
    }
    // acc[i][j][c] = C(wm + 8 i + 2 q + c, wn + 8 j + r); NN is even and C is 16-byte aligned whenever the stride is even
    const bool has_row = g.row.active(), has_col = g.col.active();
    const bool vec_ok = (reinterpret_cast<uintptr_t>(C) & 15) == 0;
#pragma unroll
    for (int i = 0; i < IT; ++i) {
        const int gm = wm + i * 8 + 2 * q;
        const double rs0 = has_row ? g.row.at(chain, gm) : 1.0;
        const double rs1 = has_row ? g.row.at(chain, gm + 1) : 1.0;
#pragma unroll
        for (int j = 0; j < JT; ++j) {
            const int gn = wn + j * 8 + r;
            double v0 = acc[i][j][0], v1 = acc[i][j][1];
            if (has_row) { v0 *= rs0; v1 *= rs1; }
            if (has_col) { const double cv = g.col.at(chain, gn); v0 *= cv; v1 *= cv; }
            if (g.add) {
                if (gm == gn) v0 += g.add[(size_t)chain * g.stride_add + gm];
                if (gm + 1 == gn) v1 += g.add[(size_t)chain * g.stride_add + gn];
            }
            double* dst = C + (size_t)gm + (size_t)gn * NN;
            if (vec_ok) *reinterpret_cast<double2*>(dst) = make_double2(v0, v1);
            else { dst[0] = v0; dst[1] = v1; }
        }
    }
}

template <int NN>
int launch_gemm_full(const GemmArgs& g, cudaStream_t st)
{
    constexpr size_t smem = 4 * Full<NN>::SMEM_TILE * sizeof(double);
    if (!g.transA && !g.transB) gemm_full_kernel<NN, false, false><<<g.batch, Full<NN>::THREADS, smem, st>>>(g);
    else if (!g.transA && g.transB) gemm_full_kernel<NN, false, true><<<g.batch, Full<NN>::THREADS, smem, st>>>(g);
    else if (g.transA && !g.transB) gemm_full_kernel<NN, true, false><<<g.batch, Full<NN>::THREADS, smem, st>>>(g);
    else gemm_full_kernel<NN, true, true><<<g.batch, Full<NN>::THREADS, smem, st>>>(g);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

template <int T>
int launch_gemm_t(const GemmArgs& g, cudaStream_t st)
{
    const int tiles = (g.n + T - 1) / T;
    dim3 grid(tiles * tiles, g.batch);
    constexpr size_t smem = 4 * Tile<T>::SMEM_TILE * sizeof(double);
    if (!g.transA && !g.transB) gemm_kernel<T, false, false><<<grid, THREADS, smem, st>>>(g);
    else if (!g.transA && g.transB) gemm_kernel<T, false, true><<<grid, THREADS, smem, st>>>(g);
    else if (g.transA && !g.transB) gemm_kernel<T, true, false><<<grid, THREADS, smem, st>>>(g);
    else gemm_kernel<T, true, true><<<grid, THREADS, smem, st>>>(g);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

template <int T>
int set_gemm_attrs()
{
    constexpr int smem = 4 * Tile<T>::SMEM_TILE * (int)sizeof(double);
    MCB_CUDA(cudaFuncSetAttribute(gemm_kernel<T, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCB_CUDA(cudaFuncSetAttribute(gemm_kernel<T, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCB_CUDA(cudaFuncSetAttribute(gemm_kernel<T, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCB_CUDA(cudaFuncSetAttribute(gemm_kernel<T, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return 0;
}

} // namespace

static int g_sms = 148;

bool gemm_tma_routes_abt(int n, int batch)
{
    static const int tma_on = getenv("MCB_GEMM_TMA") ? atoi(getenv("MCB_GEMM_TMA")) : 1;
    return tma_on && n >= 144 && n % 16 == 0 && (n % 80 == 0 || n % 64 == 0 || n % 48 == 0) && !(n == 144 && batch >= g_sms / 2);
}

int init_gemm_kernels()
{
    {
        int dev = 0;
        MCB_CUDA(cudaGetDevice(&dev));
        MCB_CUDA(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
        constexpr int smem = 4 * Full<144>::SMEM_TILE * (int)sizeof(double);
        MCB_CUDA(cudaFuncSetAttribute(gemm_full_kernel<144, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        MCB_CUDA(cudaFuncSetAttribute(gemm_full_kernel<144, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        MCB_CUDA(cudaFuncSetAttribute(gemm_full_kernel<144, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        MCB_CUDA(cudaFuncSetAttribute(gemm_full_kernel<144, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    }
    if (set_gemm_attrs<48>() || set_gemm_attrs<64>() || set_gemm_attrs<80>()) return -2;
    if (init_gemm_tma_kernels()) return -2;
    return 0;
}

int launch_gemm(const GemmArgs& g, cudaStream_t st)
{
    if (g.n <= 0 || g.batch <= 0) return 0;
    if (g.C == g.A || g.C == g.B) { set_error("gemm: output aliases an input"); return -1; }
    LaunchScope scope(LC_GEMM, st, 2.0 * g.n * (double)g.n * g.n * g.batch);
    // a batch that fills the device with one CTA per matrix: the whole product stays in one CTA's registers
    static const int full_on = getenv("MCB_GEMM_FULL") ? atoi(getenv("MCB_GEMM_FULL")) : 1;
    // TMA-staged tiles where a tile divides the matrix (MCB_GEMM_TMA=0: the register-staged kernels below)
    static const int tma_on = getenv("MCB_GEMM_TMA") ? atoi(getenv("MCB_GEMM_TMA")) : 1;
    static const int full_first = getenv("MCB_GEMM_FULL_FIRST") ? atoi(getenv("MCB_GEMM_FULL_FIRST")) : 0;
    // (only where both tiles are contiguous along their long side, i.e. A B'; MCB_GEMM_TMA=2 forces every product)
    const bool tma_wins = !g.transA && g.transB && g.n >= 144 && !(g.n == 144 && g.batch >= g_sms / 2);
    if (tma_on && gemm_tma_supported(g) && (tma_wins || tma_on == 2) && !full_first) return launch_gemm_tma(g, st);
    if (g.n == 144 && full_on && g.batch >= g_sms / 2) return launch_gemm_full<144>(g, st);
    // the tile that wastes the least padded work
    auto padded = [&](int t) { const int k = (g.n + t - 1) / t * t; return (long long)k * k; };
    int best = 64;
    if (padded(80) < padded(best)) best = 80;
    if (padded(48) < padded(best)) best = 48;
    if (best == 48) return launch_gemm_t<48>(g, st);
    if (best == 80) return launch_gemm_t<80>(g, st);
    return launch_gemm_t<64>(g, st);
}

} // namespace mcb
