// sweep_spatial (src/flavors/DQMC/DQMC.jl:546-582) with propose_local / accept_local!
// (src/models/HubbardModel/HubbardModelAttractive.jl:113-155), one CTA per chain.
// G is held in shared memory for the whole site loop when it fits (N <= ~165 on B200),
// otherwise the rank-1 updates run on the L2-resident copy.
#include "sweep.cuh"

namespace mcb {

namespace {

constexpr int SMEM_LIMIT = 227 * 1024;

__global__ void sweep_kernel(SweepArgs a, int in_smem)
{
    extern __shared__ double sm[];
    const int n = a.n, chain = blockIdx.x, tid = threadIdx.x;
    const size_t nn = (size_t)n * n;
    const int nb = a.G2 ? 2 : 1; // HubbardModelRepulsive: second sector on the negated field (sweep.cuh)
    double* u = sm;
    double* v = u + n;
    double* unif = v + n;
    int8_t* sconf = (int8_t*)(unif + n);
    double* Gs = unif + n + (n + 7) / 8;
    double* Gg[2] = {a.G + (size_t)chain * a.sG, a.G2 ? a.G2 + (size_t)chain * a.sG : nullptr};
    double* G[2] = {in_smem ? Gs : Gg[0], in_smem ? Gs + nn : Gg[1]};
    int8_t* cg = a.conf + (size_t)chain * a.sConf;

    if (in_smem)
        for (int b = 0; b < nb; ++b)
            for (size_t e = tid; e < nn; e += blockDim.x) G[b][e] = Gg[b][e];
    for (int i = tid; i < n; i += blockDim.x) {
        sconf[i] = cg[i];
        unif[i] = dqmc_uniform(a.key0, a.chain_offset + (uint32_t)chain, a.sweep, a.visit, (uint32_t)i);
    }
    __syncthreads();

    int nacc = 0;
    for (int i = 0; i < n; ++i) {
        const int s = sconf[i];
        // sector 0 sees the field s, sector 1 (if any) the field -s
        const double gam[2] = {s > 0 ? a.gamma_p : a.gamma_m, s > 0 ? a.gamma_m : a.gamma_p};
        const double efac = s > 0 ? a.efac_p : a.efac_m;
        const double r0 = 1.0 + gam[0] * (1.0 - G[0][i + (size_t)i * n]);
        const double r1 = nb == 2 ? 1.0 + gam[1] * (1.0 - G[1][i + (size_t)i * n]) : r0;
        const double p = efac * (r0 * r1);
        if (p > 1.0 || unif[i] < p) { // uniform over the CTA: every thread sees the same G_ii
            for (int b = 0; b < nb; ++b) {
                double* Gb = G[b];
                for (int k = tid; k < n; k += blockDim.x) {
                    double uk = -Gb[k + (size_t)i * n];
                    if (k == i) uk += 1.0;
                    u[k] = uk;
                    v[k] = Gb[i + (size_t)k * n];
                }
                __syncthreads();
                const double x = gam[b] / (1.0 + gam[b] * u[i]);
                for (size_t e = tid; e < nn; e += blockDim.x) {
                    const int k = (int)(e % n), l = (int)(e / n);
                    Gb[e] -= u[k] * (x * v[l]);
                }
                __syncthreads();
            }
            if (tid == 0) sconf[i] = (int8_t)-s;
            ++nacc;
            __syncthreads();
        }
    }
    bool bad = false;
    if (in_smem) {
        for (int b = 0; b < nb; ++b)
            for (size_t e = tid; e < nn; e += blockDim.x) { const double g = G[b][e]; Gg[b][e] = g; bad |= !isfinite(g); }
    }
    for (int i = tid; i < n; i += blockDim.x) {
        cg[i] = sconf[i];
        if (nb == 2) a.conf2[(size_t)chain * a.sConf + i] = (int8_t)-sconf[i];
    }
    if (__syncthreads_or(bad) && tid == 0) a.nonfinite[chain] += 1;
    if (tid == 0) { a.acc[chain] += nacc; a.prop[chain] += n; }
}

} // namespace

int init_sweep_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    return 0;
}

int launch_sweep_spatial(const SweepArgs& a, cudaStream_t st)
{
    const int n = a.n;
    const size_t hdr = (size_t)(3 * n + (n + 7) / 8) * sizeof(double);
    const size_t mat = (size_t)n * n * sizeof(double) * (a.G2 ? 2 : 1);
    const int in_smem = hdr + mat <= (size_t)SMEM_LIMIT;
    const int threads = n <= 32 ? 128 : (n <= 96 ? 256 : (n <= 200 ? 512 : 1024));
    LaunchScope scope(LC_SWEEP, st, 0.0);
    sweep_kernel<<<a.batch, threads, in_smem ? hdr + mat : hdr, st>>>(a, in_smem);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mcb
