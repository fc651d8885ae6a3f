// The sequential 8-site decision chain of segment.cu (phase 2a) and the vector recurrences (phase 2b)
// in isolation: one warp (or W warps on one SM), clock64 around REP blocks of 8 sites.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../montecarlo.jl_b200/csrc/common.cuh"
using namespace mcb;
#define REP 64
__global__ void k_chain(double* out, long long* cyc, double gp, double gm, double ep, double em)
{
    __shared__ double unif[64], nuS[8][64], rowS[8][64], xS[8][8];
    __shared__ int8_t s_in[64], s_out[64];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, r8 = l & 7;
    if (threadIdx.x < 64) { unif[threadIdx.x] = 0.37 + 0.001 * threadIdx.x; s_in[threadIdx.x] = (threadIdx.x & 1) ? 1 : -1; }
    __syncthreads();
    double Dr[8];
    for (int c = 0; c < 8; ++c) Dr[c] = out[r8 * 8 + c];
    int nacc = 0;
    long long t0 = clock64();
    for (int rep = 0; rep < REP; ++rep) {
        int sgn = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) sgn |= (s_in[j] > 0 ? 1 : 0) << j;
        int flips = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const double gii = __shfl_sync(0xffffffffu, Dr[j], j);
            double rowj[8];
#pragma unroll
            for (int c = j + 1; c < 8; ++c) rowj[c] = __shfl_sync(0xffffffffu, Dr[c], j);
            const bool up = (sgn >> j) & 1;
            const double gam = up ? gp : gm;
            const double efc = up ? ep : em;
            const double dm = Dr[j] - (r8 == j ? 1.0 : 0.0);
            const double rt = 1.0 + gam * (1.0 - gii);
            const double p = efc * (rt * rt);
            const bool accept = unif[j] < p;
            const double xs = gam * fast_rcp(rt);
            const double x = accept ? xs : 0.0;
            const double nu = accept ? xs * dm : 0.0;
#pragma unroll
            for (int c = j + 1; c < 8; ++c) Dr[c] += nu * rowj[c];
            flips |= (accept ? 1 : 0) << j;
            if (l < 8) nuS[w][j * 8 + l] = nu;
            if (l == 0) {
                xS[w][j] = x;
#pragma unroll
                for (int c = j + 1; c < 8; ++c) rowS[w][j * 8 + c] = rowj[c];
            }
        }
        nacc += __popc(flips);
        if (l < 8) s_out[l] = ((flips >> l) & 1) ? -s_in[l] : s_in[l];
        // keep the values bounded
#pragma unroll
        for (int c = 0; c < 8; ++c) Dr[c] = Dr[c] * 0.5 + 0.01 * (c + r8);
    }
    long long t1 = clock64();
    double s = nacc;
    for (int c = 0; c < 8; ++c) s += Dr[c];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

// the same chain with G_ii of the next site taken from shuffles issued before 1/R is known (segment.cu as built)
__global__ void k_chain2(double* out, long long* cyc, double gp, double gm, double ep, double em)
{
    __shared__ double unif[64], nuS[8][64], rowS[8][64], xS[8][8];
    __shared__ int8_t s_in[64];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, r8 = l & 7;
    if (threadIdx.x < 64) { unif[threadIdx.x] = 0.37 + 0.001 * threadIdx.x; s_in[threadIdx.x] = (threadIdx.x & 1) ? 1 : -1; }
    __syncthreads();
    double Dr[8];
    for (int c = 0; c < 8; ++c) Dr[c] = out[r8 * 8 + c];
    int nacc = 0;
    long long t0 = clock64();
    for (int rep = 0; rep < REP; ++rep) {
        int sgn = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) sgn |= (s_in[j] > 0 ? 1 : 0) << j;
        int flips = 0;
        double gii = __shfl_sync(0xffffffffu, Dr[0], 0);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            double rowj[8];
#pragma unroll
            for (int c = j + 1; c < 8; ++c) rowj[c] = __shfl_sync(0xffffffffu, Dr[c], j);
            double a1 = 0.0, q1 = 0.0;
            if (j < 7) {
                a1 = __shfl_sync(0xffffffffu, Dr[j], j + 1);
                q1 = __shfl_sync(0xffffffffu, Dr[j < 7 ? j + 1 : 7], j + 1);
            }
            const bool up = (sgn >> j) & 1;
            const double gam = up ? gp : gm;
            const double efc = up ? ep : em;
            const double dm = Dr[j] - (r8 == j ? 1.0 : 0.0);
            const double rt = 1.0 + gam * (1.0 - gii);
            const double p = efc * (rt * rt);
            const bool accept = unif[j] < p;
            const double xs = gam * fast_rcp(rt);
            const double x = accept ? xs : 0.0;
            const double nu = accept ? xs * dm : 0.0;
            if (j < 7) {
                const double nu1 = accept ? xs * a1 : 0.0;
                gii = fma(nu1, rowj[j < 7 ? j + 1 : 7], q1);
            }
#pragma unroll
            for (int c = j + 1; c < 8; ++c) Dr[c] += nu * rowj[c];
            flips |= (accept ? 1 : 0) << j;
            if (l < 8) nuS[w][j * 8 + l] = nu;
            if (l == 0) {
                xS[w][j] = x;
#pragma unroll
                for (int c = j + 1; c < 8; ++c) rowS[w][j * 8 + c] = rowj[c];
            }
        }
        nacc += __popc(flips);
#pragma unroll
        for (int c = 0; c < 8; ++c) Dr[c] = Dr[c] * 0.5 + 0.01 * (c + r8);
    }
    long long t1 = clock64();
    double s = nacc;
    for (int c = 0; c < 8; ++c) s += Dr[c];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

// the decision chain in warp 0 of every group of 4 warps; the other 3 wait at the group's named barrier
__global__ void k_chain_wait(double* out, long long* cyc, double gp, double gm, double ep, double em, int nregs_dummy)
{
    __shared__ double unif[64], nuS[4][64], rowS[4][64], xS[4][8];
    __shared__ int8_t s_in[64];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, r8 = l & 7, grp = w >> 2;
    if (threadIdx.x < 64) { unif[threadIdx.x] = 0.37 + 0.001 * threadIdx.x; s_in[threadIdx.x] = (threadIdx.x & 1) ? 1 : -1; }
    __syncthreads();
    double Dr[8];
    for (int c = 0; c < 8; ++c) Dr[c] = out[r8 * 8 + c];
    int nacc = 0;
    long long t0 = clock64();
    for (int rep = 0; rep < REP; ++rep) {
        if ((w & 3) == 0) {
            int sgn = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) sgn |= (s_in[j] > 0 ? 1 : 0) << j;
            int flips = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const double gii = __shfl_sync(0xffffffffu, Dr[j], j);
                double rowj[8];
#pragma unroll
                for (int c = j + 1; c < 8; ++c) rowj[c] = __shfl_sync(0xffffffffu, Dr[c], j);
                const bool up = (sgn >> j) & 1;
                const double gam = up ? gp : gm;
                const double efc = up ? ep : em;
                const double dm = Dr[j] - (r8 == j ? 1.0 : 0.0);
                const double rt = 1.0 + gam * (1.0 - gii);
                const double p = efc * (rt * rt);
                const bool accept = unif[j] < p;
                const double xs = gam * fast_rcp(rt);
                const double x = accept ? xs : 0.0;
                const double nu = accept ? xs * dm : 0.0;
#pragma unroll
                for (int c = j + 1; c < 8; ++c) Dr[c] += nu * rowj[c];
                flips |= (accept ? 1 : 0) << j;
                if (l < 8) nuS[grp][j * 8 + l] = nu;
                if (l == 0) {
                    xS[grp][j] = x;
#pragma unroll
                    for (int c = j + 1; c < 8; ++c) rowS[grp][j * 8 + c] = rowj[c];
                }
            }
            nacc += __popc(flips);
#pragma unroll
            for (int c = 0; c < 8; ++c) Dr[c] = Dr[c] * 0.5 + 0.01 * (c + r8);
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
    }
    long long t1 = clock64();
    double s = nacc;
    for (int c = 0; c < 8; ++c) s += Dr[c];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

// phase 2b: one row per thread
__global__ void k_vec(double* out, long long* cyc)
{
    constexpr int LD = 68;
    __shared__ double NUX[8 * LD], V[8 * LD], Rb[8 * LD], nuS[64], rowS[64], xS[8];
    for (int i = threadIdx.x; i < 8 * LD; i += blockDim.x) { NUX[i] = 0.01 * i; Rb[i] = 0.02 * i; }
    if (threadIdx.x < 64) { nuS[threadIdx.x] = 0.1; rowS[threadIdx.x] = 0.2; }
    if (threadIdx.x < 8) xS[threadIdx.x] = 0.3;
    __syncthreads();
    long long t0 = clock64();
    for (int rep = 0; rep < REP; ++rep) {
        if (threadIdx.x < 64) {
            const int rho = threadIdx.x;
            double nux[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                double v = NUX[k * LD + rho] - (rho == k ? 1.0 : 0.0);
#pragma unroll
                for (int m = 0; m < k; ++m) v += nux[m] * rowS[m * 8 + k];
                nux[k] = xS[k] * v;
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) NUX[k * LD + rho] = nux[k];
        } else {
            const int kap = threadIdx.x - 64;
            double vv[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                double v = Rb[k * LD + kap];
#pragma unroll
                for (int m = 0; m < k; ++m) v += nuS[m * 8 + k] * vv[m];
                vv[k] = v;
                V[k * LD + kap] = v;
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    long long t1 = clock64();
    out[threadIdx.x] = NUX[threadIdx.x] + V[threadIdx.x];
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

int main()
{
    long long* cyc; double* out;
    cudaMalloc(&cyc, 8); cudaMalloc(&out, 1024 * 8);
    double h[1024]; for (int i = 0; i < 1024; ++i) h[i] = 0.3 + 0.001 * i;
    for (int warps = 1; warps <= 4; warps *= 2) {
        cudaMemcpy(out, h, sizeof(h), cudaMemcpyHostToDevice);
        k_chain<<<1, 32 * warps>>>(out, cyc, 0.7, -0.4, 1.2, 0.8);
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("decision chain, %d warp(s): %.1f cycles per site (%s)\n", warps, (double)c / (REP * 8), cudaGetErrorString(cudaGetLastError()));
    }
    {
        cudaMemcpy(out, h, sizeof(h), cudaMemcpyHostToDevice);
        k_chain2<<<1, 32>>>(out, cyc, 0.7, -0.4, 1.2, 0.8);
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("decision chain with early next-G_ii shuffles, 1 warp: %.1f cycles per site (%s)\n", (double)c / (REP * 8), cudaGetErrorString(cudaGetLastError()));
    }
    for (int groups = 1; groups <= 4; groups *= 2) {
        cudaMemcpy(out, h, sizeof(h), cudaMemcpyHostToDevice);
        k_chain_wait<<<1, 128 * groups>>>(out, cyc, 0.7, -0.4, 1.2, 0.8, 0);
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("decision chain + 3 waiting warps, %d group(s): %.1f cycles per site (%s)\n", groups, (double)c / (REP * 8), cudaGetErrorString(cudaGetLastError()));
    }
    cudaMemcpy(out, h, sizeof(h), cudaMemcpyHostToDevice);
    k_vec<<<1, 128>>>(out, cyc);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("vector recurrences (phase 2b), 4 warps: %.1f cycles per block (%s)\n", (double)c / REP, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
