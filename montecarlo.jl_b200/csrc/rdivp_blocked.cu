// rdivp! (src/linalg/general.jl:138-166) for large matrices as a blocked forward substitution on the FP64 tensor pipe:
//   out = in[:, pivot] * triu(T)^-1 ,   column j:  x_j = (a_pivot[j] - sum_{i<j} x_i T[i, j]) / T[j, j]
// The rows of `in` are independent: a CTA owns 32 rows, each of its four computing warps 8 of them, and walks through
// the columns in blocks of 16:
//   W   = in[rows, pivot[J]] - X[rows, 0:j0] * T[0:j0, J]     8 x j0 x 16 per warp on DMMA: A = the warp's own rows of the X
//                                                              computed so far (shared memory, negated on the way into the
//                                                              fragment), B = the block column of T, brought in one block
//                                                              ahead by TMA bulk copies (one per column, two buffers);
//                                                              accumulators start as the permuted tile
//   X_J = W * inv(triu(T[J, J]))                              8 x 16 x 16 on DMMA. A fifth warp inverts the 16 x 16 diagonal
//                                                              blocks one block ahead (one column of the inverse per lane)
// so no warp waits for another inside a block; one CTA barrier per block hands over the buffers.
// The row-by-row kernel it replaces at N = 400 (rdivp_rows_kernel, stab.cu) does the N^3 flops on the FMA path: 656 us per
// batch of 32 matrices. Earlier forms of this kernel: B fragments straight from L2 (four warps per SM cannot hide that
// latency), 640 us; TMA-staged T with the diagonal block solved by substitution on one warp while three wait, 322 us.
#include "udt.cuh"
#include "prof.cuh"
#include "tma.cuh"

namespace mcb {

namespace {

constexpr int MR = 32;          // rows per CTA
constexpr int NB = 16;          // columns per block
constexpr int LDX = MR + 4;     // column stride of X in shared memory (% 16 == 4: conflict-free A fragments)

__host__ __device__ inline int ldt_of(int n) { return n + ((20 - n % 16) % 16 == 0 ? 16 : (20 - n % 16) % 16); }   // % 16 == 4, > n

constexpr int LDI = 20;         // row stride of an inverted diagonal block (% 16 == 4: conflict-free B fragments)

__global__ void __launch_bounds__(160) rdivp_blocked_kernel(int n, const double* __restrict__ in, long long sIn,
                                                            double* __restrict__ out, long long sOut,
                                                            const double* __restrict__ T, long long sT,
                                                            const int* __restrict__ pivot, long long sP)
{
    extern __shared__ __align__(128) double sm[];
    const int LDT = ldt_of(n);
    double* Xs = sm;                              // [n][LDX]
    double* Ts = Xs + (size_t)n * LDX;            // [2][NB][LDT] block columns of T (rows 0 .. j0 + NB of 16 columns)
    double* Ri = Ts + (size_t)2 * NB * LDT;       // [2][NB][LDI] inverse of triu(T[J, J]): element (k, c) at Ri[k * LDI + c]
    double* Rt = Ri + 2 * NB * LDI;               // [NB][NB + 1] staging of a diagonal block for the inverting warp
    uint64_t* bar = reinterpret_cast<uint64_t*>(Rt + NB * (NB + 1));
    int* piv_s = reinterpret_cast<int*>(bar + 2);   // [n] the pivot vector
    const int chain = blockIdx.y, row0 = blockIdx.x * MR;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, r = lane >> 2, q = lane & 3;
    const double* A = in + (size_t)chain * sIn;
    double* O = out + (size_t)chain * sOut;
    const double* Tg = T + (size_t)chain * sT;
    const int* piv = pivot + (size_t)chain * sP;
    const int myrow = row0 + 8 * warp + r;   // the row of this lane's accumulator elements (computing warps)
    const bool rowok = myrow < n;
    const int nblk = n / NB;

    if (tid == 0) { mbar_init(bar, 1); mbar_init(bar + 1, 1); }
    for (int i = tid; i < n; i += blockDim.x) piv_s[i] = piv[i];
    __syncthreads();
    auto issue = [&](int jb) {   // block column jb into buffer jb & 1
        const int j0 = jb * NB;
        const uint32_t bytes = (uint32_t)((j0 + NB) * sizeof(double));
        if (lane == 0) mbar_expect_tx(bar + (jb & 1), NB * bytes);
        __syncwarp();
        if (lane < NB) bulk_g2s(Ts + (size_t)((jb & 1) * NB + lane) * LDT, Tg + (size_t)(j0 + lane) * n, bytes, bar + (jb & 1));
    };
    // inverse of the upper triangular diagonal block jb: lane c < 16 solves R y = e_c backwards. The block itself was
    // fetched from global memory into registers one call earlier (dpre); this call fetches the next one.
    double dpre[8];
    auto fetch_diag = [&](int jb) {
        const int j0 = jb * NB;
#pragma unroll
        for (int z = 0; z < 8; ++z) {
            const int idx = lane + 32 * z, i = idx & 15, c = idx >> 4;
            dpre[z] = Tg[(size_t)(j0 + i) + (size_t)(j0 + c) * n];
        }
    };
    auto invert = [&](int jb) {
#pragma unroll
        for (int z = 0; z < 8; ++z) {
            const int idx = lane + 32 * z, i = idx & 15, c = idx >> 4;
            Rt[i * (NB + 1) + c] = dpre[z];
        }
        if (jb + 1 < nblk) fetch_diag(jb + 1);
        __syncwarp();
        const double rd = 1.0 / Rt[(lane & 15) * (NB + 1) + (lane & 15)];
        double y[NB];
#pragma unroll
        for (int i = NB - 1; i >= 0; --i) {
            double s0 = 0.0, s1 = 0.0;   // two partial sums halve the dependent chain
#pragma unroll
            for (int k = i + 1; k < NB; k += 2) {
                s0 = fma(Rt[i * (NB + 1) + k], y[k], s0);
                if (k + 1 < NB) s1 = fma(Rt[i * (NB + 1) + k + 1], y[k + 1], s1);
            }
            const double rii = __shfl_sync(0xffffffffu, rd, i);
            y[i] = i > (lane & 15) ? 0.0 : ((i == (lane & 15) ? 1.0 : 0.0) - (s0 + s1)) * rii;
        }
        double* dst = Ri + (size_t)(jb & 1) * NB * LDI;
        if (lane < NB) {
#pragma unroll
            for (int i = 0; i < NB; ++i) dst[i * LDI + lane] = y[i];
        }
        __syncwarp();
    };
    auto load_tile = [&](int jb, double (&t)[2][2]) {   // the permuted tile of `in` in accumulator layout
#pragma unroll
        for (int y = 0; y < 2; ++y)
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int col = piv_s[jb * NB + 8 * y + 2 * q + cc];
                t[y][cc] = rowok ? A[(size_t)col * n + myrow] : 0.0;
            }
    };
    double cur[2][2] = {{0.0, 0.0}, {0.0, 0.0}}, nxt[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    if (warp == 4) { issue(0); fetch_diag(0); invert(0); }
    else load_tile(0, cur);

    for (int jb = 0; jb < nblk; ++jb) {
        const int j0 = jb * NB, buf = jb & 1;
        __syncthreads();   // block jb - 1 is complete everywhere: its T buffer and inverse are free, the inverse of block jb is there
        if (warp == 4) {
            if (jb + 1 < nblk) { issue(jb + 1); invert(jb + 1); }
            continue;
        }
        if (jb + 1 < nblk) load_tile(jb + 1, nxt);
        mbar_wait(bar + buf, (jb >> 1) & 1);
        const double* tb0 = Ts + (size_t)(buf * NB + r) * LDT + q;
        const double* tb1 = tb0 + (size_t)8 * LDT;
        const double* xa = Xs + q * LDX + 8 * warp + r;
        // W -= X[:, 0:j0] T[0:j0, J]. One computing warp per scheduler: four accumulator sets (eight independent DMMA
        // chains) and operands fetched one iteration ahead keep its tensor pipe issuing.
        double c2[2][2] = {{0.0, 0.0}, {0.0, 0.0}}, c3[2][2] = {{0.0, 0.0}, {0.0, 0.0}}, c4[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
        double a[4], b0[4], b1[4];
        if (j0 > 0) {
#pragma unroll
            for (int z = 0; z < 4; ++z) { a[z] = xa[(size_t)(4 * z) * LDX]; b0[z] = tb0[4 * z]; b1[z] = tb1[4 * z]; }
        }
        for (int k = 0; k < j0; k += 16) {
            double an[4], b0n[4], b1n[4];
            const int kn = k + 16 < j0 ? k + 16 : k;   // (the last iteration re-reads its own operands)
#pragma unroll
            for (int z = 0; z < 4; ++z) {
                an[z] = xa[(size_t)(kn + 4 * z) * LDX]; b0n[z] = tb0[kn + 4 * z]; b1n[z] = tb1[kn + 4 * z];
            }
            dmma884(cur[0][0], cur[0][1], -a[0], b0[0]);
            dmma884(cur[1][0], cur[1][1], -a[0], b1[0]);
            dmma884(c2[0][0], c2[0][1], -a[1], b0[1]);
            dmma884(c2[1][0], c2[1][1], -a[1], b1[1]);
            dmma884(c3[0][0], c3[0][1], -a[2], b0[2]);
            dmma884(c3[1][0], c3[1][1], -a[2], b1[2]);
            dmma884(c4[0][0], c4[0][1], -a[3], b0[3]);
            dmma884(c4[1][0], c4[1][1], -a[3], b1[3]);
#pragma unroll
            for (int z = 0; z < 4; ++z) { a[z] = an[z]; b0[z] = b0n[z]; b1[z] = b1n[z]; }
        }
        // W through shared memory into A-fragment layout (the warp's own rows: a warp-level exchange), X_J = W inv(T_JJ)
        double* xw = Xs + (size_t)j0 * LDX + 8 * warp + r;
#pragma unroll
        for (int y = 0; y < 2; ++y)
#pragma unroll
            for (int cc = 0; cc < 2; ++cc)
                xw[(size_t)(8 * y + 2 * q + cc) * LDX] = (cur[y][cc] + c2[y][cc]) + (c3[y][cc] + c4[y][cc]);
        __syncwarp();
        const double* ri = Ri + (size_t)buf * NB * LDI;
        double x[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
        for (int k0 = 0; k0 < NB; k0 += 4) {
            const double av = xw[(size_t)(k0 + q) * LDX];
            dmma884(x[0][0], x[0][1], av, ri[(k0 + q) * LDI + r]);
            dmma884(x[1][0], x[1][1], av, ri[(k0 + q) * LDI + 8 + r]);
        }
        __syncwarp();   // every lane has read W before X overwrites it
#pragma unroll
        for (int y = 0; y < 2; ++y)
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int col = j0 + 8 * y + 2 * q + cc;
                xw[(size_t)(8 * y + 2 * q + cc) * LDX] = x[y][cc];
                if (rowok) O[(size_t)col * n + myrow] = x[y][cc];
            }
        __syncwarp();
#pragma unroll
        for (int y = 0; y < 2; ++y)
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) cur[y][cc] = nxt[y][cc];
    }
}

} // namespace

// n a multiple of 16 whose X block and two block columns of T fit one SM's shared memory
bool rdivp_blocked_supported(int n) { return n % NB == 0 && n >= 256 && n <= 400; }

static size_t rdivp_blocked_smem(int n)
{
    return ((size_t)n * LDX + (size_t)2 * NB * ldt_of(n) + 2 * NB * LDI + NB * (NB + 1) + 2) * sizeof(double) + (size_t)n * sizeof(int);
}

int init_rdivp_blocked_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(rdivp_blocked_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rdivp_blocked_smem(400)));
    return 0;
}

int launch_rdivp_blocked(int n, int batch, const double* in, long long sIn, double* out, long long sOut, const double* T,
                         long long sT, const int* pivot, long long sP, cudaStream_t st)
{
    if (!rdivp_blocked_supported(n)) { set_error("rdivp_blocked: unsupported size"); return -1; }
    LaunchScope scope(LC_RDIVP, st, 1.0 * n * (double)n * n * batch);
    dim3 grid((n + MR - 1) / MR, batch);
    rdivp_blocked_kernel<<<grid, 160, rdivp_blocked_smem(n), st>>>(n, in, sIn, out, sOut, T, sT, pivot, sP);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mcb
