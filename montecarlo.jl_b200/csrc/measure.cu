// Measurement kernels. One launch per measurement for the whole batch.
//
//   greens_kernel      G                                     measurements/measurements.jl:31-37
//   occupation_kernel  1 - G_ii over EachSiteAndFlavor        :41-44, generic.jl:299-304
//   cdc_kernel         over EachSitePairByDistance            :58-72 (two flavors),
//                                                             HubbardModelAttractive.jl:221-224 (one flavor)
//   mz_kernel          G_dn,ii - G_up,ii                      :122-125 (mx = my = 0 for a block-diagonal G)
//   sdc_{x,y,z}_kernel over EachSitePairByDistance            :148-186, HubbardModelAttractive.jl:235-237
//   pc_kernel          over EachLocalQuadByDistance{K}        :197-207, HubbardModelAttractive.jl:243-245
//   prepare! / apply! / finish! (the 1/N)                     generic.jl:279-347
//   double occupancy, kinetic / interaction energy            SURVEY.md 8a O5 (no reference kernel)
//
// The lattice iterators arrive as ONE table trg_of[dir][src] (lattice_iterators.jl:158-189: a source has at most
// one target per direction, and the reference lists the pairs of a direction by ascending source). A thread owns
// one output bin and adds its terms in the reference's order, so there are no atomics and the result does not
// depend on the launch; scalars go through a fixed-shape block reduction. The two-flavor formulas are the
// reference's general kernels with the off-diagonal blocks of the BlockDiagonal G dropped; with Gu == Gd they
// are the HubbardModelAttractive overloads.
#include "measure.cuh"
#include "common.cuh"
#include "prof.cuh"

namespace mcb {
namespace {

constexpr int MT = 256; // threads per block

__device__ __forceinline__ double block_sum(double v, double* red)
{
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < MT / 32; ++w) t += red[w];
    return t;
}

template <bool TWO>
__global__ void __launch_bounds__(MT) measure_kernel(const double* __restrict__ Gu_all, const double* __restrict__ Gd_all,
                                                     const int32_t* __restrict__ trg_of, const double* __restrict__ hopT,
                                                     int n, int n_dirs, int K, double U_signed, int obs_len,
                                                     double* __restrict__ obs_all, double* __restrict__ Gsum_all)
{
    __shared__ double red[MT / 32];
    const int chain = blockIdx.x, tid = threadIdx.x;
    const size_t nn = (size_t)n * n;
    const double* gu = Gu_all + (size_t)chain * nn;
    const double* gd = TWO ? Gd_all + (size_t)chain * nn : gu;
    double* obs = obs_all + (size_t)chain * obs_len;
    const double invN = 1.0 / n;
    const int gtid = blockIdx.y * MT + tid, gthreads = gridDim.y * MT;

    // <G>: element-wise sum
    {
        double* su = Gsum_all + (size_t)chain * (TWO ? 2 : 1) * nn;
        for (size_t e = gtid; e < nn; e += gthreads) {
            su[e] += gu[e];
            if (TWO) su[nn + e] += gd[e];
        }
    }
    // direction-resolved bins: [cdc | sdc_x | sdc_z | pc]
    const int n_pair_bins = 3 * n_dirs, n_bins = n_pair_bins + n_dirs * K * K;
    for (int b = gtid; b < n_bins; b += gthreads) {
        double acc = 0.0;
        if (b < n_pair_bins) {
            const int kind = b / n_dirs, d = b - kind * n_dirs;
            const int32_t* trg = trg_of + (size_t)d * n;
            for (int i = 0; i < n; ++i) {           // i = src, j = trg
                const int j = trg[i];
                if (j < 0) continue;
                const double dji = (i == j) ? 1.0 : 0.0;
                const double uij = gu[i + (size_t)j * n], uji = gu[j + (size_t)i * n];
                if (!TWO) {
                    if (kind == 0) acc += 4 * (1 - gu[i + (size_t)i * n]) * (1 - gu[j + (size_t)j * n]) + 2 * (dji - uji) * uij;
                    else acc += 2 * (dji - uji) * uij;
                } else {
                    const double dij_ = gd[i + (size_t)j * n], dji_ = gd[j + (size_t)i * n];
                    const double nui = 1 - gu[i + (size_t)i * n], nuj = 1 - gu[j + (size_t)j * n];
                    const double ndi = 1 - gd[i + (size_t)i * n], ndj = 1 - gd[j + (size_t)j * n];
                    if (kind == 0) acc += nui * nuj + (dji - uji) * uij + nui * ndj + ndi * nuj + ndi * ndj + (dji - dji_) * dij_;
                    else if (kind == 1) acc += (dji - uji) * dij_ + (dji - dji_) * uij;
                    else acc += nui * nuj + (dji - uji) * uij - nui * ndj - ndi * nuj + ndi * ndj + (dji - dji_) * dij_;
                }
            }
        } else {
            const int q = b - n_pair_bins;          // q = dir12 + n_dirs * (dir1 + K * dir2)
            const int d12 = q % n_dirs, k1 = (q / n_dirs) % K, k2 = q / (n_dirs * K);
            const int32_t* t12 = trg_of + (size_t)d12 * n;
            const int32_t* t1 = trg_of + (size_t)k1 * n;
            const int32_t* t2 = trg_of + (size_t)k2 * n;
            for (int s1 = 0; s1 < n; ++s1) {
                const int s2 = t12[s1];
                if (s2 < 0) continue;
                const int r1 = t1[s1], r2 = t2[s2];
                if (r1 < 0 || r2 < 0) continue;
                acc += gu[s1 + (size_t)s2 * n] * gd[r1 + (size_t)r2 * n];
            }
        }
        obs[4 + b] += acc * invN;
    }
    if (blockIdx.y != 0) return;
    // per-site outputs and scalars
    const int off_occ = 4 + n_bins;
    double occ = 0.0, docc = 0.0, ekin = 0.0;
    for (int i = tid; i < n; i += MT) {
        const double nu = 1.0 - gu[i + (size_t)i * n], nd = 1.0 - gd[i + (size_t)i * n];
        obs[off_occ + i] += nu;
        if (TWO) {
            obs[off_occ + n + i] += nd;
            obs[off_occ + 2 * n + i] += nu - nd;    // mz = G_dn,ii - G_up,ii
        }
        occ += nu + nd;
        docc += nu * nd;
    }
    for (size_t e = tid; e < nn; e += MT) {
        const int i = (int)(e % n), j = (int)(e / n);
        const double dij = (i == j) ? 1.0 : 0.0;
        const size_t f = j + (size_t)i * n;
        ekin += hopT[e] * ((dij - gu[f]) + (dij - gd[f]));
    }
    occ = block_sum(occ, red);
    docc = block_sum(docc, red);
    ekin = block_sum(ekin, red);
    if (tid == 0) {
        obs[0] += occ * invN;
        obs[1] += docc * invN;
        obs[2] += ekin * invN;
        obs[3] += U_signed * (docc - 0.5 * occ + 0.25 * n) * invN;
    }
}

// One slice of apply!(::CombinedGreensIterator, ...) (measurements/generic.jl:227-243) for charge_density_susceptibility,
// spin_density_susceptibility and pairing_susceptibility (measurements.jl:56-58, 144-146, 207). Kernels for packed Green's
// functions (G00, G0l, Gl0, Gll):
//   one flavor (HubbardModelAttractive.jl:225-229, 239-241, 246-248)
//     cdc  4 (1 - Gll[i,i]) (1 - G00[j,j]) - 2 G0l[j,i] Gl0[i,j]     sdc  -2 G0l[j,i] Gl0[i,j]  (x = y = z)
//     pc   Gl0[src1,src2] Gl0[trg1,trg2]
//   two flavors (measurements.jl:75-90, 157-196, 215-219 with the off-diagonal blocks of the BlockDiagonal matrices dropped;
//   u / d = spin-up / spin-down block)
//     cdc    sum over the four spin pairs of (1 - Gll_s[i,i]) (1 - G00_s'[j,j]), minus G0l_u[j,i] Gl0_u[i,j] + G0l_d[j,i] Gl0_d[i,j]
//     sdc_x  -G0l_u[j,i] Gl0_d[i,j] - G0l_d[j,i] Gl0_u[i,j]           (= sdc_y)
//     sdc_z  the cdc terms with a minus sign on the up-down and down-up products
//     pc     Gl0_u[src1,src2] Gl0_d[trg1,trg2]
// and current_current_susceptibility (measurements.jl:254-262) over EachLocalQuadBySyncedDistance{K} (both targets in the
// same direction, lattice_iterators.jl:390-464) with cc_kernel (HubbardModelAttractive.jl:249-266; general form
// measurements.jl:268-300), T = mc.s.hopping_matrix: with A_s = T[s1,t1] Gll_s[t1,s1] - T[t1,s1] Gll_s[s1,t1],
// B_s = T[s2,t2] G00_s[t2,s2] - T[t2,s2] G00_s[s2,t2] the kernel is (A_u + A_d)(B_u + B_d) + X_u + X_d,
//     X_s = - T[t1,s1] T[t2,s2] G0l_s[s2,t1] Gl0_s[s1,t2] + T[s1,t1] T[t2,s2] G0l_s[s2,s1] Gl0_s[t1,t2]
//           + T[t1,s1] T[s2,t2] G0l_s[t2,t1] Gl0_s[s1,s2] - T[s1,t1] T[s2,t2] G0l_s[t2,s1] Gl0_s[t1,s2]
// (one flavor: d = u, i.e. 4 A B + 2 X).
// over the same iterator tables as the equal-time kernel; factor = delta_tau / N (finish!, generic.jl:272-286) is applied
// per slice (the sum over l is linear). A thread owns one bin: fixed order, no atomics.
struct TdMats { const double* m[4]; };   // G00, G0l, Gl0, Gll of one spin block, [C][N*N]

template <bool TWO>
__global__ void __launch_bounds__(MT) measure_td_kernel(TdMats up, TdMats dn, const int32_t* __restrict__ trg_of,
                                                        const double* __restrict__ hop, int n, int n_dirs, int K, double factor, int td_len,
                                                        double* __restrict__ td_all)
{
    const int chain = blockIdx.x;
    const size_t nn = (size_t)n * n, off = (size_t)chain * nn;
    const double *u00 = up.m[0] + off, *u0l = up.m[1] + off, *ul0 = up.m[2] + off, *ull = up.m[3] + off;
    const double *d00 = dn.m[0] + off, *d0l = dn.m[1] + off, *dl0 = dn.m[2] + off, *dll = dn.m[3] + off;
    double* td = td_all + (size_t)chain * td_len;
    const int gtid = blockIdx.y * MT + threadIdx.x, gthreads = gridDim.y * MT;
    const int n_pair_bins = (TWO ? 3 : 2) * n_dirs, n_quad_bins = n_dirs * K * K, n_bins = n_pair_bins + n_quad_bins + n_dirs * K;
    for (int b = gtid; b < n_bins; b += gthreads) {
        double acc = 0.0;
        if (b >= n_pair_bins + n_quad_bins) {
            const int q = b - n_pair_bins - n_quad_bins;   // q = dir12 + n_dirs * dir1
            const int d12 = q % n_dirs, k1 = q / n_dirs;
            const int32_t* t12 = trg_of + (size_t)d12 * n;
            const int32_t* tk = trg_of + (size_t)k1 * n;
            auto at = [n](const double* g, int a, int c) { return g[a + (size_t)c * n]; };   // g[a, c]
            for (int s1 = 0; s1 < n; ++s1) {
                const int s2 = t12[s1];
                if (s2 < 0) continue;
                const int t1 = tk[s1], t2 = tk[s2];
                if (t1 < 0 || t2 < 0) continue;
                const double T11 = at(hop, s1, t1), T11r = at(hop, t1, s1), T22 = at(hop, s2, t2), T22r = at(hop, t2, s2);
                double A = T11 * at(ull, t1, s1) - T11r * at(ull, s1, t1);
                double B = T22 * at(u00, t2, s2) - T22r * at(u00, s2, t2);
                double X = -T11r * T22r * at(u0l, s2, t1) * at(ul0, s1, t2) + T11 * T22r * at(u0l, s2, s1) * at(ul0, t1, t2)
                           + T11r * T22 * at(u0l, t2, t1) * at(ul0, s1, s2) - T11 * T22 * at(u0l, t2, s1) * at(ul0, t1, s2);
                if (TWO) {
                    A += T11 * at(dll, t1, s1) - T11r * at(dll, s1, t1);
                    B += T22 * at(d00, t2, s2) - T22r * at(d00, s2, t2);
                    X += -T11r * T22r * at(d0l, s2, t1) * at(dl0, s1, t2) + T11 * T22r * at(d0l, s2, s1) * at(dl0, t1, t2)
                         + T11r * T22 * at(d0l, t2, t1) * at(dl0, s1, s2) - T11 * T22 * at(d0l, t2, s1) * at(dl0, t1, s2);
                    acc += A * B + X;
                } else {
                    acc += 4.0 * A * B + 2.0 * X;
                }
            }
        } else if (b < n_pair_bins) {
            const int kind = b / n_dirs, d = b - kind * n_dirs;
            const int32_t* trg = trg_of + (size_t)d * n;
            for (int i = 0; i < n; ++i) {           // i = src, j = trg; G[i, j] at i + j * n
                const int j = trg[i];
                if (j < 0) continue;
                const size_t ji = j + (size_t)i * n, ij = i + (size_t)j * n, ii = i + (size_t)i * n, jj = j + (size_t)j * n;
                if (!TWO) {
                    const double ex = -2.0 * u0l[ji] * ul0[ij];
                    acc += kind == 0 ? 4.0 * (1.0 - ull[ii]) * (1.0 - u00[jj]) + ex : ex;
                } else if (kind == 1) {
                    acc += -u0l[ji] * dl0[ij] - d0l[ji] * ul0[ij];
                } else {
                    const double nui = 1.0 - ull[ii], ndi = 1.0 - dll[ii], nuj = 1.0 - u00[jj], ndj = 1.0 - d00[jj];
                    const double same = nui * nuj - u0l[ji] * ul0[ij] + ndi * ndj - d0l[ji] * dl0[ij];
                    const double mixed = nui * ndj + ndi * nuj;
                    acc += kind == 0 ? same + mixed : same - mixed;
                }
            }
        } else {
            const int q = b - n_pair_bins;          // q = dir12 + n_dirs * (dir1 + K * dir2)
            const int d12 = q % n_dirs, k1 = (q / n_dirs) % K, k2 = q / (n_dirs * K);
            const int32_t* t12 = trg_of + (size_t)d12 * n;
            const int32_t* t1 = trg_of + (size_t)k1 * n;
            const int32_t* t2 = trg_of + (size_t)k2 * n;
            for (int s1 = 0; s1 < n; ++s1) {
                const int s2 = t12[s1];
                if (s2 < 0) continue;
                const int r1 = t1[s1], r2 = t2[s2];
                if (r1 < 0 || r2 < 0) continue;
                acc += ul0[s1 + (size_t)s2 * n] * dl0[r1 + (size_t)r2 * n];
            }
        }
        td[b] += acc * factor;
    }
}

} // namespace

int measure_setup(MeasureState& m, int N, int C, int flavors, const int32_t* trg_of_host, int n_dirs, int K,
                  const double* hop_host, double U_signed, cudaStream_t st)
{
    measure_free(m);
    m.N = N; m.C = C; m.flavors = flavors; m.n_dirs = n_dirs; m.K = K; m.U_signed = U_signed;
    const size_t nn = (size_t)N * N;
    MCB_CUDA(cudaMalloc(&m.trg_of, (size_t)n_dirs * N * sizeof(int32_t)));
    MCB_CUDA(cudaMalloc(&m.hopT, nn * sizeof(double)));
    MCB_CUDA(cudaMalloc(&m.obs, (size_t)C * m.obs_len() * sizeof(double)));
    MCB_CUDA(cudaMalloc(&m.Gsum, (size_t)C * flavors * nn * sizeof(double)));
    MCB_CUDA(cudaMalloc(&m.td, (size_t)C * (m.td_len() > 0 ? m.td_len() : 1) * sizeof(double)));
    MCB_CUDA(cudaMemcpyAsync(m.trg_of, trg_of_host, (size_t)n_dirs * N * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    MCB_CUDA(cudaMemcpyAsync(m.hopT, hop_host, nn * sizeof(double), cudaMemcpyHostToDevice, st));
    MCB_CUDA(cudaStreamSynchronize(st));
    return measure_reset(m, st);
}

void measure_free(MeasureState& m)
{
    cudaFree(m.trg_of); cudaFree(m.hopT); cudaFree(m.obs); cudaFree(m.Gsum); cudaFree(m.td); cudaFree(m.G00);
    m.trg_of = nullptr; m.hopT = nullptr; m.obs = nullptr; m.Gsum = nullptr; m.td = nullptr; m.G00 = nullptr;
    m.n_meas = 0; m.n_td = 0;
}

int measure_reset(MeasureState& m, cudaStream_t st)
{
    MCB_CUDA(cudaMemsetAsync(m.obs, 0, (size_t)m.C * m.obs_len() * sizeof(double), st));
    MCB_CUDA(cudaMemsetAsync(m.Gsum, 0, (size_t)m.C * m.flavors * m.N * m.N * sizeof(double), st));
    if (m.td) MCB_CUDA(cudaMemsetAsync(m.td, 0, (size_t)m.C * (m.td_len() > 0 ? m.td_len() : 1) * sizeof(double), st));
    MCB_CUDA(cudaStreamSynchronize(st));
    m.n_meas = 0; m.n_td = 0;
    return 0;
}

int measure_accumulate(MeasureState& m, const double* Gu, const double* Gd, cudaStream_t st)
{
    const int n_bins = 3 * m.n_dirs + m.n_dirs * m.K * m.K;
    int gy = (n_bins + MT - 1) / MT;
    if (gy < 1) gy = 1;
    if (gy > 64) gy = 64;
    const dim3 grid(m.C, gy);
    LaunchScope scope(LC_MEASURE, st, 0.0);
    if (m.flavors == 2)
        measure_kernel<true><<<grid, MT, 0, st>>>(Gu, Gd, m.trg_of, m.hopT, m.N, m.n_dirs, m.K, m.U_signed, m.obs_len(), m.obs, m.Gsum);
    else
        measure_kernel<false><<<grid, MT, 0, st>>>(Gu, Gu, m.trg_of, m.hopT, m.N, m.n_dirs, m.K, m.U_signed, m.obs_len(), m.obs, m.Gsum);
    MCB_CUDA(cudaGetLastError());
    m.n_meas += 1;
    return 0;
}

int measure_td_g00(MeasureState& m, double* g00[2])
{
    const size_t per = (size_t)m.C * m.N * m.N;
    if (!m.G00) MCB_CUDA(cudaMalloc(&m.G00, (size_t)m.flavors * per * sizeof(double)));
    g00[0] = m.G00;
    g00[1] = m.flavors == 2 ? m.G00 + per : m.G00;
    return 0;
}

int measure_td_accumulate(MeasureState& m, const double* const up[4], const double* const dn[4], double factor, cudaStream_t st)
{
    const int n_bins = m.td_len();
    if (n_bins <= 0) return 0;
    int gy = (n_bins + MT - 1) / MT;
    if (gy > 64) gy = 64;
    TdMats u, d;
    for (int w = 0; w < 4; ++w) { u.m[w] = up[w]; d.m[w] = dn[w]; }
    LaunchScope scope(LC_MEASURE, st, 0.0);
    if (m.flavors == 2)
        measure_td_kernel<true><<<dim3(m.C, gy), MT, 0, st>>>(u, d, m.trg_of, m.hopT, m.N, m.n_dirs, m.K, factor, n_bins, m.td);
    else
        measure_td_kernel<false><<<dim3(m.C, gy), MT, 0, st>>>(u, u, m.trg_of, m.hopT, m.N, m.n_dirs, m.K, factor, n_bins, m.td);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mcb
