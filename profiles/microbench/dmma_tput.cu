// FP64 tensor-pipe throughput by mma.sync shape, warps per SM and independent accumulators per warp.
// One CTA per SM; every warp issues REP rounds of ILP independent MMAs. Prints flop / clk / SM.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma_tput dmma_tput.cu
#include <cstdio>
#include <cuda_runtime.h>

#define REP 256

template <int SHAPE> struct Frag;
template <> struct Frag<0> { static constexpr int NA = 1, NB = 1, NC = 2, FLOP = 2 * 8 * 8 * 4; };    // m8n8k4
template <> struct Frag<1> { static constexpr int NA = 2, NB = 1, NC = 4, FLOP = 2 * 16 * 8 * 4; };   // m16n8k4
template <> struct Frag<2> { static constexpr int NA = 4, NB = 2, NC = 4, FLOP = 2 * 16 * 8 * 8; };   // m16n8k8
template <> struct Frag<3> { static constexpr int NA = 8, NB = 4, NC = 4, FLOP = 2 * 16 * 8 * 16; };  // m16n8k16

template <int SHAPE>
__device__ __forceinline__ void mma(double* c, const double* a, const double* b)
{
    if constexpr (SHAPE == 0)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[0]), "+d"(c[1]) : "d"(a[0]), "d"(b[0]));
    else if constexpr (SHAPE == 1)
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
    else if constexpr (SHAPE == 2)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, "
                     "{%12,%13,%14,%15}, {%0,%1,%2,%3};"
                     : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]),
                       "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

// DISTINCT: every MMA of a round reads its own A / B registers (as a GEMM inner loop does) instead of one shared pair
template <int SHAPE, int ILP, bool DISTINCT>
__global__ void k(double* out, long long* cyc, double seed)
{
    using F = Frag<SHAPE>;
    double c[ILP][F::NC], a[DISTINCT ? ILP : 1][F::NA], b[DISTINCT ? ILP : 1][F::NB];
    for (int i = 0; i < ILP; ++i)
        for (int e = 0; e < F::NC; ++e) c[i][e] = seed * (threadIdx.x + i + e);
    for (int i = 0; i < (DISTINCT ? ILP : 1); ++i) {
        for (int e = 0; e < F::NA; ++e) a[i][e] = seed + 1e-3 * (threadIdx.x + 3 * i + e);
        for (int e = 0; e < F::NB; ++e) b[i][e] = seed - 1e-3 * (threadIdx.x + 5 * i + e);
    }
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 4
    for (int r = 0; r < REP; ++r) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) mma<SHAPE>(c[i], a[DISTINCT ? i : 0], b[DISTINCT ? i : 0]);
    }
    __syncthreads();
    const long long t1 = clock64();
    double s = 0;
    for (int i = 0; i < ILP; ++i)
        for (int e = 0; e < F::NC; ++e) s += c[i][e];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int SHAPE, int ILP, bool DISTINCT>
void run(const char* name, double* out, long long* cyc)
{
    for (int warps : {1, 2, 4, 8, 12, 16}) {
        k<SHAPE, ILP, DISTINCT><<<148, warps * 32>>>(out, cyc, 1.0);
        cudaDeviceSynchronize();
        k<SHAPE, ILP, DISTINCT><<<148, warps * 32>>>(out, cyc, 1.0);
        cudaDeviceSynchronize();
        long long h = 0;
        cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        const double flop = (double)Frag<SHAPE>::FLOP * ILP * REP * warps;
        printf("%-10s ilp %2d %s warps/SM %2d : %8lld cycles  %7.1f flop/clk/SM  (%5.1f cycles per MMA per SMSP)\n", name, ILP,
               DISTINCT ? "distinct-ab" : "shared-ab  ", warps, h, flop / h, (double)h / (ILP * REP * (warps + 3) / 4));
    }
}

// GEMM-style register tile: c[i][j] += a[i] * b[j], IT x JT accumulator tiles per warp (what a warp tile of
// 8 IT x 8 JT does per k4 step), operands refreshed from shared memory every round like the real inner loop.
template <int IT, int JT, bool LDS_OPERANDS>
__global__ void kt(double* out, long long* cyc, double seed)
{
    __shared__ double sa[64 * 40], sb[64 * 40];
    for (int e = threadIdx.x; e < 64 * 40; e += blockDim.x) { sa[e] = seed + 1e-4 * e; sb[e] = seed - 1e-4 * e; }
    double c[IT][JT][2], a[IT], b[JT];
    for (int i = 0; i < IT; ++i)
        for (int j = 0; j < JT; ++j) c[i][j][0] = c[i][j][1] = seed * (threadIdx.x + i + j);
    for (int i = 0; i < IT; ++i) a[i] = seed + 1e-3 * (threadIdx.x + 3 * i);
    for (int j = 0; j < JT; ++j) b[j] = seed - 1e-3 * (threadIdx.x + 5 * j);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long t0 = clock64();
#pragma unroll 2
    for (int r = 0; r < REP; ++r) {
        if (LDS_OPERANDS) {
#pragma unroll
            for (int i = 0; i < IT; ++i) a[i] = sa[((r & 7) * 4 + (lane & 3)) * 40 + i * 8 + (lane >> 2)];
#pragma unroll
            for (int j = 0; j < JT; ++j) b[j] = sb[((r & 7) * 4 + (lane & 3)) * 40 + j * 8 + (lane >> 2)];
        }
#pragma unroll
        for (int i = 0; i < IT; ++i)
#pragma unroll
            for (int j = 0; j < JT; ++j) mma<0>(c[i][j], &a[i], &b[j]);
    }
    __syncthreads();
    const long long t1 = clock64();
    double s = 0;
    for (int i = 0; i < IT; ++i)
        for (int j = 0; j < JT; ++j) s += c[i][j][0] + c[i][j][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int IT, int JT, bool L>
void run_tile(double* out, long long* cyc)
{
    for (int warps : {4, 8, 12, 16}) {
        cudaFuncSetAttribute(kt<IT, JT, L>, cudaFuncAttributePreferredSharedMemoryCarveout, 50);
        kt<IT, JT, L><<<148, warps * 32>>>(out, cyc, 1.0);
        if (cudaDeviceSynchronize() != cudaSuccess || cudaGetLastError() != cudaSuccess) { printf("tile %dx%d warps %d: launch failed\n", IT, JT, warps); cudaGetLastError(); continue; }
        kt<IT, JT, L><<<148, warps * 32>>>(out, cyc, 1.0);
        cudaDeviceSynchronize();
        long long h = 0;
        cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        const double flop = 512.0 * IT * JT * REP * warps;
        printf("tile %dx%d %s warps/SM %2d : %8lld cycles  %7.1f flop/clk/SM\n", IT, JT, L ? "lds-operands" : "reg-operands", warps, h, flop / h);
    }
}

int main()
{
    double* out; long long* cyc;
    cudaMalloc(&out, 148 * 512 * sizeof(double));
    cudaMalloc(&cyc, sizeof(long long));
    run_tile<3, 9, false>(out, cyc);
    run_tile<3, 9, true>(out, cyc);
    run_tile<2, 4, true>(out, cyc);
    run_tile<4, 4, true>(out, cyc);
    run_tile<2, 8, true>(out, cyc);
    run_tile<4, 8, true>(out, cyc);
    run_tile<3, 6, true>(out, cyc);
    run_tile<6, 6, true>(out, cyc);
    run<0, 8, false>("m8n8k4", out, cyc);
    run<0, 8, true>("m8n8k4", out, cyc);
    run<0, 27, true>("m8n8k4", out, cyc);
    run<1, 8, true>("m16n8k4", out, cyc);
    run<2, 8, true>("m16n8k8", out, cyc);
    run<3, 8, true>("m16n8k16", out, cyc);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
    return 0;
}
