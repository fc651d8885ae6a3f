// Batched FP64 DMMA GEMM with TMA-staged tiles (same contract as gemm.cu, see gemm.cuh):
//   C[b] = diag(row) * opA(A[b]) * diag(mid) * opB(B[b]) * diag(col) + diag(add)
// for sizes the tile divides (N % T == 0, N % BK == 0; configs A / C / D: N = 64, 144, 400).
// Measured (batched 400^3 x 32, ncu, profiles/r02l): A B' 165 us against 184 us of the register-staged kernel, A B 208
// against 175, A' B 276 against 172: a tile whose K index is the contiguous one in memory needs T copies of 64 bytes
// per stage, which costs more than it saves (one 2-D tensor-map copy with a padded box would do; not built). launch_gemm
// therefore routes only the A B' products here.
//
// The operand tiles travel global -> shared memory as TMA bulk copies (cp.async.bulk, one copy per contiguous
// column segment of a tile) into a ring of STAGES buffers, each guarded by an mbarrier that counts the bytes; the
// threads that compute issue nothing but LDS + DMMA (the register-staged kernel of gemm.cu spends 9 instructions per
// DMMA on loads, stores and address arithmetic and sits at 52 % of the tensor pipe). The diagonal `mid` is applied to
// the A fragments, `row` / `col` / `add` to the accumulators; the result tile goes through shared memory and leaves as
// bulk stores of whole column segments.
#include "gemm.cuh"
#include "tma.cuh"

namespace mcb {

namespace {

constexpr int THREADS = 128;
constexpr int STAGES = 4;

template <int T>
struct TT {
    static constexpr int BK = T >= 80 ? 8 : 16;  // K depth of a stage (80: four stages of 15 KB leave room for 3 CTAs per SM)
    static constexpr int LDX = T + 4;             // [k][x] layout, x contiguous (LDX % 16 == 4: conflict-free fragment reads)
    static constexpr int LDK = BK + 4;            // [x][k] layout, k contiguous
    static constexpr int TILE = (BK * LDX > T * LDK) ? BK * LDX : T * LDK;   // doubles per operand per stage
    static constexpr int WT = T / 16;             // DMMA tiles per warp in each direction (2 x 2 warps of T/2 x T/2)
    static constexpr int LDC = T + 4;             // result tile [col][row]
    static constexpr size_t SMEM = ((size_t)2 * STAGES * TILE + (size_t)STAGES * BK + 2 * STAGES) * sizeof(double);   // + full / empty barriers
    static_assert(T % 16 == 0 && LDX % 16 == 4 && (size_t)T * LDC <= (size_t)2 * STAGES * TILE, "unsupported tile");
};

// one stage of one operand: XC = the tile is contiguous in its non-K index in global memory (A: !transA, B: transB)
// issued by one warp (lane = its lane index)
template <int T, bool XC>
__device__ __forceinline__ void issue_tile(double* S, const double* __restrict__ P, int n, int x0, int k0, int lane,
                                           uint64_t* bar)
{
    constexpr int BK = TT<T>::BK;
    if (XC) {   // BK column segments of T doubles
        if (lane < BK) bulk_g2s(S + lane * TT<T>::LDX, P + (size_t)x0 + (size_t)(k0 + lane) * n, T * sizeof(double), bar);
    } else {    // T row segments of BK doubles
        for (int x = lane; x < T; x += 32)
            bulk_g2s(S + x * TT<T>::LDK, P + (size_t)k0 + (size_t)(x0 + x) * n, BK * sizeof(double), bar);
    }
}

template <int T, bool XC>
__device__ __forceinline__ double frag(const double* S, int x, int k)
{
    return XC ? S[k * TT<T>::LDX + x] : S[x * TT<T>::LDK + k];
}

template <int T, bool TA, bool TB>
__global__ void __launch_bounds__(THREADS) gemm_tma_kernel(GemmArgs g)
{
    using P = TT<T>;
    constexpr int BK = P::BK, WT = P::WT, TILE = P::TILE;
    extern __shared__ __align__(128) double smem[];
    double* As = smem;                              // [STAGES][TILE]
    double* Bs = As + (size_t)STAGES * TILE;        // [STAGES][TILE]
    double* midv = Bs + (size_t)STAGES * TILE;      // [STAGES][BK]
    uint64_t* full = reinterpret_cast<uint64_t*>(midv + STAGES * BK);   // [STAGES]
    const int n = g.n, tiles = n / T;
    const int tm = blockIdx.x % tiles, tn = blockIdx.x / tiles, chain = blockIdx.y;
    const double* A = g.A + (size_t)chain * g.strideA;
    const double* B = g.B + (size_t)chain * g.strideB;
    double* C = g.C + (size_t)chain * g.strideC;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = (warp & 1) * (T / 2), wn = (warp >> 1) * (T / 2);
    const int r = lane >> 2, q = lane & 3;
    const int m0 = tm * T, n0 = tn * T;
    const bool has_mid = g.mid.active();
    const int nk = n / BK;

    uint64_t* empty = full + STAGES;   // one arrival per warp: the stage may be refilled
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, THREADS / 32); }
    }
    __syncthreads();
    // one warp issues a stage: the diagonal `mid` first (ordered before the consumers by the barrier's arrival), then the
    // byte count, then the copies. No CTA barrier inside the K loop: stage s is refilled for tile kt + STAGES by warp
    // kt % 4 once all four warps have released it.
    auto issue = [&](int kt) {
        const int s = kt % STAGES, k0 = kt * BK;
        if (has_mid && lane < BK) midv[s * BK + lane] = g.mid.at(chain, k0 + lane);
        __syncwarp();
        if (lane == 0) mbar_expect_tx(full + s, 2u * T * BK * sizeof(double));
        __syncwarp();
        issue_tile<T, !TA>(As + (size_t)s * TILE, A, n, m0, k0, lane, full + s);
        issue_tile<T, TB>(Bs + (size_t)s * TILE, B, n, n0, k0, lane, full + s);
    };
    if (warp == 0)
        for (int kt = 0; kt < STAGES && kt < nk; ++kt) issue(kt);

    double acc[WT][WT][2];
#pragma unroll
    for (int i = 0; i < WT; ++i)
#pragma unroll
        for (int j = 0; j < WT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int kt = 0; kt < nk; ++kt) {
        const int s = kt % STAGES;
        mbar_wait(full + s, (kt / STAGES) & 1);
        const double* as = As + (size_t)s * TILE;
        const double* bs = Bs + (size_t)s * TILE;
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a[WT], b[WT];
#pragma unroll
            for (int i = 0; i < WT; ++i) a[i] = frag<T, !TA>(as, wm + i * 8 + r, kk + q);
#pragma unroll
            for (int j = 0; j < WT; ++j) b[j] = frag<T, TB>(bs, wn + j * 8 + r, kk + q);
            if (has_mid) {
                const double mv = midv[s * BK + kk + q];
#pragma unroll
                for (int i = 0; i < WT; ++i) a[i] *= mv;
            }
#pragma unroll
            for (int i = 0; i < WT; ++i)
#pragma unroll
                for (int j = 0; j < WT; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + s);
        if (kt + STAGES < nk && warp == kt % (THREADS / 32)) {
            mbar_wait(empty + s, (kt / STAGES) & 1);
            issue(kt + STAGES);
        }
    }
    __syncthreads();   // every warp is done with the operand ring: it becomes the result tile

    // ---- epilogue: scalings in registers, the tile through shared memory, bulk stores of its column segments
    double* Cs = smem;                        // [T][LDC], the operand ring is free (barrier above)
    const bool has_row = g.row.active(), has_col = g.col.active();
#pragma unroll
    for (int i = 0; i < WT; ++i) {
        const int lm = wm + i * 8 + r, gm = m0 + lm;
        const double rs = has_row ? g.row.at(chain, gm) : 1.0;
#pragma unroll
        for (int j = 0; j < WT; ++j)
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int ln = wn + j * 8 + 2 * q + c, gn = n0 + ln;
                double v = acc[i][j][c];
                if (has_row) v *= rs;
                if (has_col) v *= g.col.at(chain, gn);
                if (g.add && gm == gn) v += g.add[(size_t)chain * g.stride_add + gm];
                Cs[ln * P::LDC + lm] = v;
            }
    }
    fence_async_smem();
    __syncthreads();
    if (tid < T) {
        bulk_s2g(C + (size_t)m0 + (size_t)(n0 + tid) * n, Cs + tid * P::LDC, T * sizeof(double));
        bulk_commit();
        bulk_wait_read();   // the shared source must outlive the copy; global visibility comes with kernel completion
    }
}

template <int T>
int launch_t(const GemmArgs& g, cudaStream_t st)
{
    const int tiles = g.n / T;
    dim3 grid(tiles * tiles, g.batch);
    constexpr size_t smem = TT<T>::SMEM;
    if (!g.transA && !g.transB) gemm_tma_kernel<T, false, false><<<grid, THREADS, smem, st>>>(g);
    else if (!g.transA && g.transB) gemm_tma_kernel<T, false, true><<<grid, THREADS, smem, st>>>(g);
    else if (g.transA && !g.transB) gemm_tma_kernel<T, true, false><<<grid, THREADS, smem, st>>>(g);
    else gemm_tma_kernel<T, true, true><<<grid, THREADS, smem, st>>>(g);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

template <int T>
int set_attrs()
{
    constexpr int smem = (int)TT<T>::SMEM;
    MCB_CUDA(cudaFuncSetAttribute(gemm_tma_kernel<T, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCB_CUDA(cudaFuncSetAttribute(gemm_tma_kernel<T, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCB_CUDA(cudaFuncSetAttribute(gemm_tma_kernel<T, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCB_CUDA(cudaFuncSetAttribute(gemm_tma_kernel<T, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return 0;
}

int tile_for(int n)
{
    if (n % 80 == 0) return 80;
    if (n % 64 == 0) return 64;
    if (n % 48 == 0) return 48;
    return 0;
}

} // namespace

int init_gemm_tma_kernels()
{
    if (set_attrs<48>() || set_attrs<64>() || set_attrs<80>()) return -2;
    return 0;
}

// bulk copies need 16-byte aligned segments: even strides and 16-byte aligned base pointers
bool gemm_tma_supported(const GemmArgs& g)
{
    const int t = tile_for(g.n);
    if (!t || g.n % 16 != 0) return false;
    auto ok = [](const void* p, long long stride) { return ((uintptr_t)p & 15) == 0 && (stride & 1) == 0; };
    return ok(g.A, g.strideA) && ok(g.B, g.strideB) && ok(g.C, g.strideC);
}

int launch_gemm_tma(const GemmArgs& g, cudaStream_t st)
{
    switch (tile_for(g.n)) {
    case 80: return launch_t<80>(g, st);
    case 64: return launch_t<64>(g, st);
    case 48: return launch_t<48>(g, st);
    }
    set_error("gemm_tma: unsupported size");
    return -1;
}

} // namespace mcb
