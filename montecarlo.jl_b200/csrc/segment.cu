// wrap_greens! (src/flavors/DQMC/stack.jl:491-500) + sweep_spatial (src/flavors/DQMC/DQMC.jl:546-582)
// with propose_local / accept_local! (src/models/HubbardModel/HubbardModelAttractive.jl:113-155),
// fused over a run of consecutive time slices. One CTA per chain; G stays in registers (panel.cuh).
//
// Wrap: the two N^3 products run on the FP64 tensor pipe from shared-memory operands; the slice
// matrices B_l = e^{-dtau T} diag(e^{lambda s}) are never materialised (slice_matrices.jl:23-39):
// the diagonal factor is applied while G is staged to shared memory.
// Local updates: one __syncthreads per site. The owners of row i and column i publish them
// (double-buffered), every thread then evaluates the Metropolis decision redundantly from G_ii and
// applies the rank-1 update G -= (e_i - G[:,i]) x G[i,:] to its 16 register elements.
#include "panel.cuh"
#include "segment.cuh"

namespace mcb {

namespace {

using namespace panel;

template <int N>
struct SegSmem {
    static constexpr int LD = Geo<N>::LD;
    double eT2[N * LD];
    double eTinv2[N * LD];
    double X[N * LD];
    double colbuf[2][N];
    double rowbuf[2][N];
    double unif[N];
    double sc_row[N];
    double sc_col[N];
    int8_t s_in[N];
    int8_t s_out[N];
};

template <int N>
__global__ void __launch_bounds__(Geo<N>::THREADS) segment_kernel(SegmentArgs a)
{
    constexpr int MT = Geo<N>::MT, TH = Geo<N>::THREADS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SegSmem<N>& sm = *reinterpret_cast<SegSmem<N>*>(smem_raw);
    const Lane t;
    const int chain = blockIdx.x;
    double* Gg = a.G + (size_t)chain * a.sG;
    int8_t* cg = a.conf + (size_t)chain * a.sConf;
    const uint32_t k1 = a.chain_offset + (uint32_t)chain;

    double acc[MT][2];
    g2r<N>(acc, Gg, t);
    g2s<N, false>(sm.eT2, a.eT2, t.tid);
    g2s<N, false>(sm.eTinv2, a.eTinv2, t.tid);

    int nacc = 0;
    for (int v = 0; v < a.n_visits; ++v) {
        const int slice = a.start_slice + a.direction * v;
        if (v > 0 || a.first_wrap) {
            // ---- wrap_greens!
            const int l = a.direction > 0 ? slice - 1 : slice; // slice matrix B_l
            __syncthreads();
            for (int i = t.tid; i < N; i += TH) {
                const int s = cg[(size_t)(l - 1) * N + i];
                if (a.direction > 0) { sm.sc_row[i] = s > 0 ? a.eVp[0] : a.eVp[1]; sm.sc_col[i] = s > 0 ? a.eVm[0] : a.eVm[1]; }
                else { sm.sc_row[i] = s > 0 ? a.eVm[0] : a.eVm[1]; sm.sc_col[i] = s > 0 ? a.eVp[0] : a.eVp[1]; }
            }
            __syncthreads();
            if (a.direction > 0) {
                // G <- e^{-dtau T} diag(e^{+lambda s}) G diag(e^{-lambda s}) e^{+dtau T}
                scale_rows<N>(acc, sm.sc_row, t);
                r2s<N, false>(acc, sm.X, t);
                __syncthreads();
                gemm<N, false, false>(acc, sm.eT2, sm.X, t);
                __syncthreads();
                scale_cols<N>(acc, sm.sc_col, t);
                r2s<N, false>(acc, sm.X, t);
                __syncthreads();
                gemm<N, false, false>(acc, sm.X, sm.eTinv2, t);
            } else {
                // G <- diag(e^{-lambda s}) e^{+dtau T} G e^{-dtau T} diag(e^{+lambda s})
                r2s<N, false>(acc, sm.X, t);
                __syncthreads();
                gemm<N, false, false>(acc, sm.eTinv2, sm.X, t);
                __syncthreads();
                scale_rows<N>(acc, sm.sc_row, t);
                r2s<N, false>(acc, sm.X, t);
                __syncthreads();
                gemm<N, false, false>(acc, sm.X, sm.eT2, t);
                scale_cols<N>(acc, sm.sc_col, t);
            }
        }
        // ---- sweep_spatial of `slice`
        const long long upd = a.n_updates0 + v;
        const uint32_t sweep = (uint32_t)(upd / (2LL * a.n_slices));
        const uint32_t visit = (uint32_t)(upd % (2LL * a.n_slices));
        __syncthreads();
        for (int i = t.tid; i < N; i += TH) {
            sm.s_in[i] = cg[(size_t)(slice - 1) * N + i];
            sm.unif[i] = dqmc_uniform(a.key0, k1, sweep, visit, (uint32_t)i);
        }
        // (the first __syncthreads of the site loop orders these writes before their reads)
#pragma unroll
        for (int it = 0; it < MT; ++it) {
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
                const int i = 8 * it + rr;
                double* colb = sm.colbuf[rr & 1];
                double* rowb = sm.rowbuf[rr & 1];
                if (t.w == it && t.q == (rr >> 1)) {
#pragma unroll
                    for (int m = 0; m < MT; ++m) {
                        const double g = (rr & 1) ? acc[m][1] : acc[m][0];
                        colb[8 * m + t.r] = ((8 * m + t.r) == i ? 1.0 : 0.0) - g;
                    }
                }
                if (t.r == rr) {
                    rowb[8 * t.w + 2 * t.q] = acc[it][0];
                    rowb[8 * t.w + 2 * t.q + 1] = acc[it][1];
                }
                __syncthreads();
                const double gii = rowb[i];
                const int s = sm.s_in[i];
                const double gamma = s > 0 ? a.gamma_p : a.gamma_m;
                const double efac = s > 0 ? a.efac_p : a.efac_m;
                const double rt = 1.0 + gamma * (1.0 - gii);
                const double p = efac * (rt * rt);
                const bool accept = (p > 1.0 || sm.unif[i] < p);
                if (accept) {
                    const double x = gamma / (1.0 + gamma * colb[i]);
                    const double v0 = x * rowb[8 * t.w + 2 * t.q], v1 = x * rowb[8 * t.w + 2 * t.q + 1];
#pragma unroll
                    for (int m = 0; m < MT; ++m) {
                        const double u = colb[8 * m + t.r];
                        acc[m][0] -= u * v0;
                        acc[m][1] -= u * v1;
                    }
                    ++nacc;
                }
                if (t.tid == 0) sm.s_out[i] = (int8_t)(accept ? -s : s);
            }
        }
        __syncthreads();
        for (int i = t.tid; i < N; i += TH) cg[(size_t)(slice - 1) * N + i] = sm.s_out[i];
    }
    bool bad = false;
#pragma unroll
    for (int m = 0; m < MT; ++m) bad |= !isfinite(acc[m][0]) || !isfinite(acc[m][1]);
    r2g<N>(acc, Gg, t);
    const int anybad = __syncthreads_or(bad);
    if (t.tid == 0) {
        a.acc[chain] += nacc;
        a.prop[chain] += (long long)N * a.n_visits;
        if (anybad) a.nonfinite[chain] += 1;
    }
}

template <int N>
int launch_t(const SegmentArgs& a, cudaStream_t st)
{
    // algorithmic flop: 4 N^3 per wrap; the rank-1 updates are added by the caller from the
    // acceptance counters
    const int wraps = a.n_visits - 1 + (a.first_wrap ? 1 : 0);
    LaunchScope scope(LC_FUSED, st, 4.0 * N * (double)N * N * wraps * a.batch);
    segment_kernel<N><<<a.batch, Geo<N>::THREADS, sizeof(SegSmem<N>), st>>>(a);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace

bool segment_supported(int n) { return n == 16 || n == 32 || n == 64; }

int init_segment_kernels()
{
    MCB_CUDA(cudaFuncSetAttribute(segment_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem<16>)));
    MCB_CUDA(cudaFuncSetAttribute(segment_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem<32>)));
    MCB_CUDA(cudaFuncSetAttribute(segment_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SegSmem<64>)));
    return 0;
}

int launch_segment(const SegmentArgs& a, cudaStream_t st)
{
    if (a.n_visits <= 0 || a.batch <= 0) return 0;
    switch (a.n) {
    case 16: return launch_t<16>(a, st);
    case 32: return launch_t<32>(a, st);
    case 64: return launch_t<64>(a, st);
    default: set_error("segment kernel: unsupported matrix size"); return -1;
    }
}

} // namespace mcb
