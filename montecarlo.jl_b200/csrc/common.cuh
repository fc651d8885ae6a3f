// Shared device/host helpers of libmcb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <utility>
#include "prof.cuh"

namespace mcb {

// Switches that turn phases of a kernel off for timing runs (results invalid) are read from the environment only
// in -DMCB_EXPERIMENTS builds (profiles/); the shipped library ignores them.
inline int experiment_flags(const char* name)
{
#ifdef MCB_EXPERIMENTS
    const char* e = getenv(name);
    return e ? atoi(e) : 0;
#else
    (void)name;
    return 0;
#endif
}

void set_error(const std::string& msg);

#define MCB_CUDA(call)                                                                          \
    do {                                                                                        \
        cudaError_t err__ = (call);                                                             \
        if (err__ != cudaSuccess) {                                                             \
            char buf__[512];                                                                    \
            snprintf(buf__, sizeof(buf__), "%s:%d: %s -> %s", __FILE__, __LINE__, #call,        \
                     cudaGetErrorString(err__));                                                \
            mcb::set_error(buf__);                                                              \
            return -2;                                                                          \
        }                                                                                       \
    } while (0)

// A diagonal scaling applied inside a kernel. Either a double vector (optionally inverted), or
// the Hubbard-Stratonovich field of one time slice: value(i) = conf[i] > 0 ? ep : em, i.e.
// exp(+-lambda * s_i) of interaction_matrix_exp! (HubbardModelAttractive.jl:100-110) as a
// sign-select between two host-computed constants.
struct Scale {
    const double* d = nullptr;  // [batch][n] (stride_d per chain) or null
    const int8_t* s = nullptr;  // conf row of the slice: [batch] stride_s, or null
    long long stride_d = 0;
    long long stride_s = 0;
    double ep = 1.0, em = 1.0;
    int inv = 0;

    __host__ __device__ bool active() const { return d != nullptr || s != nullptr; }
    __device__ __forceinline__ double at(int chain, int i) const
    {
        if (d) {
            const double v = d[(long long)chain * stride_d + i];
            return inv ? 1.0 / v : v;
        }
        if (s) return s[(long long)chain * stride_s + i] > 0 ? ep : em;
        return 1.0;
    }
};

// ---------------------------------------------------------------- Philox4x32-10
// Stream specification: SURVEY.md §8(d) / Appendix D.
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                       uint32_t c3, uint32_t k0, uint32_t k1,
                                                       uint32_t out[4])
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
#else
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__host__ __device__ __forceinline__ uint32_t seed_key0(uint64_t seed)
{
    return (uint32_t)seed ^ (uint32_t)(seed >> 32);
}

// acceptance uniform of (sweep, visit, site) for chain key k1: 53-bit, [0,1)
__device__ __forceinline__ double dqmc_uniform(uint32_t k0, uint32_t k1, uint32_t sweep,
                                               uint32_t visit, uint32_t site)
{
    uint32_t o[4];
    philox4x32_10(sweep, visit, site, 0u, k0, k1, o);
    const uint64_t x = ((uint64_t)o[1] << 32) | o[0];
    return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

// ---------------------------------------------------------------- FP64 tensor core
// mma.sync m8n8k4 f64 (SASS: DMMA.8x8x4). Fragment ownership for lane l:
//   a = A[l/4][l%4]   b = B[l%4][l/4]   c0,c1 = C[l/4][2*(l%4) + {0,1}]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// 1/x and 1/sqrt(x) to ~1 ulp for normal, positive-magnitude x: hardware seed (20 mantissa bits) and
// two Newton steps; no special-case branches (callers guard x == 0 themselves).
__device__ __forceinline__ double fast_rcp(double x)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    r = fma(fma(-x, r, 1.0), r, r);
    r = fma(fma(-x, r, 1.0), r, r);
    return r;
}
__device__ __forceinline__ double fast_rsqrt(double x)
{
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double hx = 0.5 * x;
    r = fma(fma(-hx * r, r, 0.5), r, r);
    r = fma(fma(-hx * r, r, 0.5), r, r);
    return r;
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization attribute (launch_pdl) may
// be scheduled while its predecessor on the stream still runs (once every CTA of the predecessor has executed
// griddep_launch or exited). Nothing before griddep_wait() may read or write global memory that the predecessor
// touches; the wait returns when the predecessor has completed and its writes are visible. Both instructions are no-ops
// in a kernel launched the plain way. Used where it was measured to pay: the chain-resident kernels of N <= 64 (segment.cu,
// stab.cu: wait and trigger first thing, 16.38 -> 16.20 ms per sweep of dqmc_A512) and the panel / apply / form-Q chain of
// the blocked QR (udt_blocked.cu: trigger late, so that a waiting grid does not hold SMs the other stream group needs;
// 257 -> 241 ms per sweep of dqmc_D32). The N = 144 / 400 GEMM, sweep and rdivp kernels gained nothing (one CTA per SM
// and kernels of 40 - 500 us) and stay plain launches.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// host side: the launch; MCB_PDL=0 makes it a plain launch (comparisons). Only for kernels that call griddep_wait().
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args)
{
    static const bool pdl = !(getenv("MCB_PDL") && atoi(getenv("MCB_PDL")) == 0);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

} // namespace mcb
