// Dependent-issue latencies of the FP64 instructions the QR / site kernels chain together (one warp,
// clock64 around a fully dependent sequence). Build: nvcc -gencode arch=compute_100a,code=sm_100a -o lat lat.cu
#include <cstdio>
#include <cuda_runtime.h>
#define REP 512
__global__ void k_dfma(double* out, long long* cyc, double a, double b)
{
    double x = out[threadIdx.x];
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) x = fma(x, a, b);
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_dfma_ilp4(double* out, long long* cyc, double a, double b)
{
    double x0 = out[threadIdx.x], x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3;
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) { x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b); }
    long long t1 = clock64();
    out[threadIdx.x] = x0 + x1 + x2 + x3;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_shfl(double* out, long long* cyc)
{
    double x = out[threadIdx.x];
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) x = __shfl_xor_sync(0xffffffffu, x, 4);
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_shfl_add(double* out, long long* cyc)
{
    double x = out[threadIdx.x];
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) x += __shfl_xor_sync(0xffffffffu, x, 4);
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_dmma(double* out, long long* cyc, double a, double b)
{
    double c0 = out[threadIdx.x], c1 = c0;
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
    long long t1 = clock64();
    out[threadIdx.x] = c0 + c1;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_dmma_ilp8(double* out, long long* cyc, double a, double b)
{
    double c[8][2];
    for (int k = 0; k < 8; ++k) c[k][0] = c[k][1] = out[threadIdx.x] + k;
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP / 8; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[k][0]), "+d"(c[k][1]) : "d"(a), "d"(b));
    long long t1 = clock64();
    double s = 0; for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_rcp(double* out, long long* cyc)
{
    double x = out[threadIdx.x];
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) { double y; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); x = y; }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_lds(double* out, long long* cyc)
{
    __shared__ int idx[256];
    idx[threadIdx.x] = (threadIdx.x * 7 + 1) & 255;
    __syncthreads();
    int p = threadIdx.x;
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) p = idx[p];
    long long t1 = clock64();
    out[threadIdx.x] = p;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_bar(double* out, long long* cyc)
{
    long long t0 = clock64();
#pragma unroll
    for (int i = 0; i < REP; ++i) asm volatile("bar.sync 1, 128;" ::: "memory");
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_bar_reg(double* out, long long* cyc)
{
    const int grp = threadIdx.x >> 7;
    const int id = 1 + grp;
    __shared__ double buf[512];
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < REP; ++i) {
        buf[threadIdx.x] = i;
        asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(128) : "memory");
    }
    long long t1 = clock64();
    out[threadIdx.x] = buf[(threadIdx.x + 1) & 511];
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_bar_one_works(double* out, long long* cyc)
{
    // warp 0 of each group does ~200 cycles of dependent work, the other 3 wait
    const int grp = threadIdx.x >> 7, w = (threadIdx.x >> 5) & 3;
    const int id = 1 + grp;
    double x = out[threadIdx.x];
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < REP; ++i) {
        if (w == 0) {
#pragma unroll
            for (int k = 0; k < 24; ++k) x = fma(x, 1.0000001, 1e-9);
        }
        asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(128) : "memory");
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <class F> void run(const char* name, F f, int ops = REP)
{
    long long* cyc; double* out;
    cudaMalloc(&cyc, 8); cudaMalloc(&out, 1024 * 8); cudaMemset(out, 0, 1024 * 8);
    f(out, cyc); f(out, cyc);
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-28s %8.2f cycles/op  (%s)\n", name, (double)h / ops, cudaGetErrorString(cudaGetLastError()));
    cudaFree(cyc); cudaFree(out);
}
int main()
{
    run("DFMA dependent (1 warp)", [](double* o, long long* c) { k_dfma<<<1, 32>>>(o, c, 1.0000001, 1e-9); });
    run("DFMA dependent (4 warps/SMSP)", [](double* o, long long* c) { k_dfma<<<1, 512>>>(o, c, 1.0000001, 1e-9); });
    run("DFMA 4 indep chains (per 4)", [](double* o, long long* c) { k_dfma_ilp4<<<1, 32>>>(o, c, 1.0000001, 1e-9); });
    run("SHFL.64 dependent", [](double* o, long long* c) { k_shfl<<<1, 32>>>(o, c); });
    run("SHFL.64 + DADD dependent", [](double* o, long long* c) { k_shfl_add<<<1, 32>>>(o, c); });
    run("DMMA dependent", [](double* o, long long* c) { k_dmma<<<1, 32>>>(o, c, 1.0, 1e-9); });
    run("DMMA 8 indep (per DMMA)", [](double* o, long long* c) { k_dmma_ilp8<<<1, 32>>>(o, c, 1.0, 1e-9); });
    run("DMMA 8 indep, 4 warps/SMSP", [](double* o, long long* c) { k_dmma_ilp8<<<1, 512>>>(o, c, 1.0, 1e-9); });
    run("MUFU.RCP64H dependent", [](double* o, long long* c) { k_rcp<<<1, 32>>>(o, c); });
    run("LDS dependent", [](double* o, long long* c) { k_lds<<<1, 256>>>(o, c); });
    run("bar.sync 128 thr", [](double* o, long long* c) { k_bar<<<1, 128>>>(o, c); });
    run("bar.sync reg-id, 4 groups x128 + STS", [](double* o, long long* c) { k_bar_reg<<<1, 512>>>(o, c); });
    run("bar.sync reg-id, 1 group  x128 + STS", [](double* o, long long* c) { k_bar_reg<<<1, 128>>>(o, c); });
    run("24 dep DFMA in 1 of 4 warps + bar, 4 grp", [](double* o, long long* c) { k_bar_one_works<<<1, 512>>>(o, c); });
    run("24 dep DFMA in 1 of 4 warps + bar, 1 grp", [](double* o, long long* c) { k_bar_one_works<<<1, 128>>>(o, c); });
    return 0;
}
