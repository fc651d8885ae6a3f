// Host-side schedule of the DQMC sweep, batched over chains.
//
// The slice schedule of the reference (src/flavors/DQMC/stack.jl:502-631) depends
// only on (current_slice, direction), never on a chain's data, so one host state machine drives all
// chains of the handle with batched kernels. Stack slots are addressed through a logical->physical
// table with one spare buffer, so the reference's slot<->work-matrix copies (stack.jl:507-528,
// 570-594) become pointer swaps, and calculate_greens reads the slots without clobbering them.
#include "dqmc_engine.cuh"
#include <math.h>
#include <stdlib.h>
#include <utility>

namespace mcb {

namespace {

#define MCB_TRY(expr)                        \
    do {                                     \
        const int rc__ = (expr);             \
        if (rc__ != 0) return rc__;          \
    } while (0)

__global__ void set_identity_kernel(double* U, double* T, double* D, int n, size_t total)
{
    const size_t nn = (size_t)n * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total;
         e += (size_t)gridDim.x * blockDim.x) {
        const size_t r = e % nn;
        const double v = (r % n == r / n) ? 1.0 : 0.0;
        U[e] = v;
        if (T) T[e] = v;
        if (D && r < (size_t)n) D[(e / nn) * n + r] = 1.0;
    }
}

__global__ void max_abs_diff_kernel(const double* A, const double* B, double* out, size_t nn)
{
    __shared__ double red[32];
    const size_t base = (size_t)blockIdx.x * nn;
    double m = 0.0;
    for (size_t e = threadIdx.x; e < nn; e += blockDim.x) m = fmax(m, fabs(A[base + e] - B[base + e]));
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = fmax(m, red[w]);
        out[blockIdx.x] = fmax(out[blockIdx.x], m);
    }
}

int set_identity(mcb_dqmc* h, double* U, double* T, double* D)
{
    LaunchScope scope(LC_MISC, h->st, 0.0);
    set_identity_kernel<<<296, 256, 0, h->st>>>(U, T, D, h->N, h->mat);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

// multiply_slice_matrix_left!: Y = e^{-dtau T} diag(e^{+lambda s_l}) X   (slice_matrices.jl:42-48)
int mul_slice_left(mcb_dqmc* h, int slice, const double* X, double* Y, const Scale* col = nullptr)
{
    GemmArgs g = base_gemm(h, h->hop[2], true, X, false, Y);
    g.mid = conf_scale(h, slice, true);
    if (col) g.col = *col;
    return launch_gemm(g, h->st);
}
// multiply_daggered_slice_matrix_left!: Y = diag(e^{+lambda s_l}) (e^{-dtau T})' X  (slice_matrices.jl:70-76)
int mul_daggered_slice_left(mcb_dqmc* h, int slice, const double* X, double* Y, const Scale* col = nullptr)
{
    GemmArgs g = base_gemm(h, h->hop[2], true, X, false, Y);
    g.transA = true;
    g.row = conf_scale(h, slice, true);
    if (col) g.col = *col;
    return launch_gemm(g, h->st);
}

// wrap_greens! (stack.jl:491-500). In place on Gx through the scratch Gt.
int wrap_greens(mcb_dqmc* h, double* Gx, double* Gt, int curr_slice, int direction)
{
    if (direction == -1) {
        const int l = curr_slice - 1;
        // multiply_slice_matrix_inv_left!: Gt = diag(e^{-lambda s}) e^{+dtau T} G
        GemmArgs a = base_gemm(h, h->hop[3], true, Gx, false, Gt);
        a.row = conf_scale(h, l, false);
        MCB_TRY(launch_gemm(a, h->st));
        // multiply_slice_matrix_right!: G = Gt e^{-dtau T} diag(e^{+lambda s})
        // (as Gt (e^{-dtau T})'' with the stored transpose where that form goes to the TMA-staged kernel: both tiles are then
        // contiguous along their long side)
        const bool abt = !h->use_resident_segment && h->hopT[0] && gemm_tma_routes_abt(h->N, h->C);
        GemmArgs b = base_gemm(h, Gt, false, abt ? h->hopT[0] : h->hop[2], true, Gx);
        b.transB = abt;
        b.col = conf_scale(h, l, true);
        MCB_TRY(launch_gemm(b, h->st));
    } else {
        const int l = curr_slice;
        // multiply_slice_matrix_left!: Gt = e^{-dtau T} diag(e^{+lambda s}) G
        MCB_TRY(mul_slice_left(h, l, Gx, Gt));
        // multiply_slice_matrix_inv_right!: G = Gt diag(e^{-lambda s}) e^{+dtau T}
        const bool abt = !h->use_resident_segment && h->hopT[1] && gemm_tma_routes_abt(h->N, h->C);
        GemmArgs b = base_gemm(h, Gt, false, abt ? h->hopT[1] : h->hop[3], true, Gx);
        b.transB = abt;
        b.mid = conf_scale(h, l, false);
        MCB_TRY(launch_gemm(b, h->st));
    }
    return 0;
}

} // namespace

// ------------------------------------------------------------------ calculate_greens
// calculate_greens_AVX! (stack.jl:337-393) without clobbering its inputs.
int calc_greens_raw(int n, int batch, const double* Ul, const double* Dl, const double* Tl,
                    const double* Ur, const double* Dr, const double* Tr, double* G, double* const W[6],
                    double* const Dw[2], int* pivot, cudaStream_t st, bool resident)
{
    const long long nn = (long long)n * n;
    auto gemm = [&](const double* A, bool ta, const double* B, bool tb, double* C) {
        GemmArgs g;
        g.n = n; g.batch = batch;
        g.A = A; g.strideA = nn; g.transA = ta;
        g.B = B; g.strideB = nn; g.transB = tb;
        g.C = C; g.strideC = nn;
        return g;
    };
    if (resident && stab_supported(n)) {
        // one CTA per chain: the scaled product inside the first factorisation, everything between and
        // after the factorisations in the two glue kernels (stab.cuh)
        MCB_TRY(launch_udt_of_product(n, batch, Tl, Tr, Dl, Dr, W[1], Dw[0], W[2], pivot, st));
        MCB_TRY(launch_greens_mid(n, batch, Ul, W[1], W[2], pivot, Dw[0], Ur, W[3], W[4], W[0], st));
        MCB_TRY(launch_udt(n, batch, W[0], nn, W[1], nn, Dw[1], n, W[2], nn, pivot, n, 0, st));
        MCB_TRY(launch_greens_end(n, batch, W[3], W[1], W[2], pivot, Dw[1], W[4], G, st));
        return 0;
    }
    // B1: W0 = diag(Dl) (Tl Tr') diag(Dr) ; udt -> Q1 = W1, D1 = Dw0, dirty T = W2
    {
        GemmArgs g = gemm(Tl, false, Tr, true, W[0]);
        g.row = vec_scale(Dl, n);
        g.col = vec_scale(Dr, n);
        MCB_TRY(launch_gemm(g, st));
    }
    MCB_TRY(launch_udt(n, batch, W[0], nn, W[1], nn, Dw[0], n, W[2], nn, pivot, n, 0, st));
    // B2: W3 = Ul Q1 ; W4 = Ur[:,p] triu(W2)^-1 ; W0 = W3' W4 + diag(D1)   (B3 fused)
    MCB_TRY(launch_gemm(gemm(Ul, false, W[1], false, W[3]), st));
    MCB_TRY(launch_rdivp(n, batch, Ur, nn, W[4], nn, W[2], nn, pivot, n, st));
    {
        GemmArgs g = gemm(W[3], true, W[4], false, W[0]);
        g.add = Dw[0]; g.stride_add = n;
        MCB_TRY(launch_gemm(g, st));
    }
    // B4: udt -> Q2 = W1, D2 = Dw1, dirty T = W2 ; W5 = W4[:,p] triu(W2)^-1 ; W0 = W3 Q2
    MCB_TRY(launch_udt(n, batch, W[0], nn, W[1], nn, Dw[1], n, W[2], nn, pivot, n, 0, st));
    MCB_TRY(launch_rdivp(n, batch, W[4], nn, W[5], nn, W[2], nn, pivot, n, st));
    MCB_TRY(launch_gemm(gemm(W[3], false, W[1], false, W[0]), st));
    // B5/B6: G = (W5 diag(1/D2)) W0'
    {
        GemmArgs g = gemm(W[5], false, W[0], true, G);
        g.mid = vec_scale(Dw[1], n, true);
        MCB_TRY(launch_gemm(g, st));
    }
    return 0;
}

namespace {

int calc_greens(mcb_dqmc* h, int physL, int physR, double* out)
{
    return calc_greens_raw(h->N, h->C, h->bufU[physL], h->bufD[physL], h->bufT[physL], h->bufU[physR],
                           h->bufD[physR], h->bufT[physR], out, h->W, h->Dw, h->pivot, h->st, h->use_panel);
}

// add_slice_sequence_left (stack.jl:272-287): phys buffer `dst` <- udt(B_hi ... B_lo * U_src D_src) , T_dst = T T_src
// add_slice_sequence_right (stack.jl:296-311) for left == false (daggered, descending slices)
int advance(mcb_dqmc* h, bool left, int src, int dst, int lo, int hi)
{
    return advance_raw(h, left, h->bufU[src], h->bufD[src], h->bufT[src], h->bufU[dst], h->bufD[dst], h->bufT[dst], lo,
                       hi);
}

} // namespace

int advance_raw(mcb_dqmc* h, bool left, const double* Usrc, const double* Dsrc, const double* Tsrc, double* Udst,
                double* Ddst, double* Tdst, int lo, int hi)
{
    if (h->use_panel) {
        AdvanceArgs a;
        a.n = h->N; a.batch = h->C;
        a.Usrc = Usrc; a.Dsrc = Dsrc; a.Tsrc = Tsrc;
        a.Udst = Udst; a.Ddst = Ddst; a.Tdst = Tdst;
        a.eT2 = h->hop[2];
        a.conf = h->conf; a.sConf = (long long)h->M * h->N;
        a.lo = lo; a.hi = hi; a.left = left ? 1 : 0;
        a.ep = h->eVp[0]; a.em = h->eVp[1];
        MCB_TRY(launch_slice_products(a, h->W[0], h->st));
        if (!force_general_udt())
            return launch_udt_times(h->N, h->C, h->W[0], Udst, Ddst, Tsrc, Tdst,
                                    h->pivot, h->st);
        MCB_TRY(launch_udt(h->N, h->C, h->W[0], h->nn, Udst, h->nn, Ddst, h->N, h->W[2], h->nn,
                           h->pivot, h->N, 1, h->st));
        return launch_gemm(base_gemm(h, h->W[2], false, Tsrc, false, Tdst), h->st);
    }
    if (h->use_resident_segment && hi >= lo) {
        // the whole chain of slice products in one launch (segment_resident.cu)
        SliceChainArgs c;
        c.n = h->N; c.batch = h->C;
        c.Usrc = Usrc; c.Dsrc = Dsrc; c.dst = h->W[0];
        c.conf = h->conf; c.sConf = (long long)h->M * h->N;
        c.eT2 = h->hop[2]; c.eT2T = h->hopT[0];
        c.lo = lo; c.hi = hi; c.left = left ? 1 : 0;
        c.ep = h->eVp[0]; c.em = h->eVp[1];
        MCB_TRY(launch_slice_chain_resident(c, h->st));
        MCB_TRY(launch_udt(h->N, h->C, h->W[0], h->nn, Udst, h->nn, Ddst, h->N, h->W[2], h->nn,
                           h->pivot, h->N, 1, h->st));
        return launch_gemm(base_gemm(h, h->W[2], false, Tsrc, false, Tdst), h->st);
    }
    const double* cur = Usrc;
    const Scale dcol = vec_scale(Dsrc, h->N);
    int pp = 0;
    const int cnt = hi - lo + 1;
    for (int t = 0; t < cnt; ++t) {
        const int slice = left ? lo + t : hi - t;
        double* nxt = h->W[pp];
        const Scale* col = (t == cnt - 1) ? &dcol : nullptr; // tmp1 = curr_U * Diagonal(d) fused
        if (left) MCB_TRY(mul_slice_left(h, slice, cur, nxt, col));
        else MCB_TRY(mul_daggered_slice_left(h, slice, cur, nxt, col));
        cur = nxt;
        pp ^= 1;
    }
    double* X = const_cast<double*>(cur);
    MCB_TRY(launch_udt(h->N, h->C, X, h->nn, Udst, h->nn, Ddst, h->N, h->W[2], h->nn,
                       h->pivot, h->N, 1, h->st));
    GemmArgs g = base_gemm(h, h->W[2], false, Tsrc, false, Tdst);
    return launch_gemm(g, h->st);
}

namespace {

int record_prop_err(mcb_dqmc* h)
{
    LaunchScope scope(LC_MISC, h->st, 0.0);
    max_abs_diff_kernel<<<h->C, 256, 0, h->st>>>(h->Gtmp, h->G, h->prop_err, h->nn);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

} // namespace

// ------------------------------------------------------------------ allocation
int engine_alloc(mcb_dqmc* h)
{
    const int nbuf = h->ne + 1;
    for (int i = 0; i < 4; ++i) MCB_CUDA(cudaMalloc(&h->hop[i], h->nn * sizeof(double)));
    MCB_CUDA(cudaMalloc(&h->conf, (size_t)h->C * h->M * h->N));
    MCB_CUDA(cudaMalloc(&h->G, h->mat * sizeof(double)));
    MCB_CUDA(cudaMalloc(&h->Gtmp, h->mat * sizeof(double)));
    for (int i = 0; i < 8; ++i) MCB_CUDA(cudaMalloc(&h->W[i], h->mat * sizeof(double)));
    for (int i = 0; i < 2; ++i) MCB_CUDA(cudaMalloc(&h->Dw[i], h->vec * sizeof(double)));
    MCB_CUDA(cudaMalloc(&h->Dx, h->vec * sizeof(double)));
    MCB_CUDA(cudaMalloc(&h->pivot, h->vec * sizeof(int)));
    h->bufU.assign(nbuf, nullptr); h->bufT.assign(nbuf, nullptr); h->bufD.assign(nbuf, nullptr);
    for (int i = 0; i < nbuf; ++i) {
        MCB_CUDA(cudaMalloc(&h->bufU[i], h->mat * sizeof(double)));
        MCB_CUDA(cudaMalloc(&h->bufT[i], h->mat * sizeof(double)));
        MCB_CUDA(cudaMalloc(&h->bufD[i], h->vec * sizeof(double)));
    }
    h->slot.resize(h->ne);
    for (int i = 0; i < h->ne; ++i) h->slot[i] = i;
    h->spare = h->ne;
    MCB_CUDA(cudaMalloc(&h->acc, h->C * sizeof(long long)));
    MCB_CUDA(cudaMalloc(&h->prop, h->C * sizeof(long long)));
    MCB_CUDA(cudaMalloc(&h->nonfinite, h->C * sizeof(long long)));
    MCB_CUDA(cudaMalloc(&h->prop_err, h->C * sizeof(double)));
    MCB_CUDA(cudaMemsetAsync(h->acc, 0, h->C * sizeof(long long), h->st));
    MCB_CUDA(cudaMemsetAsync(h->prop, 0, h->C * sizeof(long long), h->st));
    MCB_CUDA(cudaMemsetAsync(h->nonfinite, 0, h->C * sizeof(long long), h->st));
    MCB_CUDA(cudaMemsetAsync(h->prop_err, 0, h->C * sizeof(double), h->st));
    MCB_CUDA(cudaMemsetAsync(h->G, 0, h->mat * sizeof(double), h->st));
    return 0;
}

void engine_free(mcb_dqmc* h)
{
    unequal_free(h);
    for (int i = 0; i < 4; ++i) cudaFree(h->hop[i]);
    for (int i = 0; i < 2; ++i) cudaFree(h->hopT[i]);
    cudaFree(h->conf); cudaFree(h->conf_bits); cudaFree(h->G); cudaFree(h->Gtmp);
    for (int i = 0; i < 8; ++i) cudaFree(h->W[i]);
    for (int i = 0; i < 2; ++i) cudaFree(h->Dw[i]);
    cudaFree(h->pivot); cudaFree(h->Dx);
    for (auto p : h->bufU) cudaFree(p);
    for (auto p : h->bufT) cudaFree(p);
    for (auto p : h->bufD) cudaFree(p);
    cudaFree(h->acc); cudaFree(h->prop); cudaFree(h->nonfinite); cudaFree(h->prop_err);
    measure_free(h->ms);
}

// ------------------------------------------------------------------ stack schedule
// build_stack (stack.jl:242-255)
int engine_build_stack(mcb_dqmc* h)
{
    for (int i = 0; i < h->ne; ++i) h->slot[i] = i;
    h->spare = h->ne;
    MCB_TRY(set_identity(h, h->bufU[h->slot[0]], h->bufT[h->slot[0]], h->bufD[h->slot[0]]));
    for (int k = 1; k <= h->ne - 1; ++k)
        MCB_TRY(advance(h, true, h->slot[k - 1], h->slot[k], (k - 1) * h->s + 1, k * h->s));
    h->current_slice = h->M + 1;
    h->direction = -1;
    return 0;
}

// propagate (stack.jl:502-631); slots are 0-based here: slot k = udt(B_{ks}...B_1) while moving up,
// udt(B'_{ks+1}...B'_M) while moving down.
int engine_propagate(mcb_dqmc* h, bool defer_wrap, bool* wrap_pending)
{
    if (wrap_pending) *wrap_pending = false;
    const int s = h->s, M = h->M;
    double* wrap_tmp = h->W[6];
    if (h->direction == 1) {
        if (h->current_slice % s == 0) {
            h->current_slice += 1;
            if (h->current_slice == 1) {
                // Ur <- stack[1]; stack[1] <- I; Ul <- I
                MCB_TRY(set_identity(h, h->bufU[h->spare], h->bufT[h->spare], h->bufD[h->spare]));
                std::swap(h->slot[0], h->spare);
                MCB_TRY(calc_greens(h, h->slot[0], h->spare, h->G));
            } else if (h->current_slice <= M) {
                const int k = (h->current_slice - 1) / s;
                MCB_TRY(advance(h, true, h->slot[k - 1], h->spare, (k - 1) * s + 1, k * s));
                std::swap(h->slot[k], h->spare); // spare now holds the old (right) decomposition
                if (h->p.check_propagation) {
                    MCB_CUDA(cudaMemcpyAsync(h->Gtmp, h->G, h->mat * sizeof(double),
                                             cudaMemcpyDeviceToDevice, h->st));
                    MCB_TRY(wrap_greens(h, h->Gtmp, wrap_tmp, h->current_slice - 1, 1));
                }
                MCB_TRY(calc_greens(h, h->slot[k], h->spare, h->G));
                if (h->p.check_propagation) MCB_TRY(record_prop_err(h));
            } else {
                const int k = h->ne - 1;
                MCB_TRY(advance(h, true, h->slot[k - 1], h->spare, (k - 1) * s + 1, k * s));
                std::swap(h->slot[k], h->spare);
                h->direction = -1;
                h->current_slice = M + 1;
                return engine_propagate(h, defer_wrap, wrap_pending);
            }
        } else {
            MCB_TRY(wrap_greens(h, h->G, wrap_tmp, h->current_slice, 1));
            h->current_slice += 1;
        }
    } else {
        if ((h->current_slice - 1) % s == 0) {
            h->current_slice -= 1;
            if (h->current_slice == M) {
                const int last = h->ne - 1;
                MCB_TRY(set_identity(h, h->bufU[h->spare], h->bufT[h->spare], h->bufD[h->spare]));
                std::swap(h->slot[last], h->spare); // spare = old left decomposition, slot = identity
                MCB_TRY(calc_greens(h, h->spare, h->slot[last], h->G));
                if (defer_wrap) *wrap_pending = true;
                else MCB_TRY(wrap_greens(h, h->G, wrap_tmp, h->current_slice + 1, -1));
            } else if (h->current_slice > 0) {
                const int k = h->current_slice / s;
                MCB_TRY(advance(h, false, h->slot[k + 1], h->spare, k * s + 1, (k + 1) * s));
                std::swap(h->slot[k], h->spare); // spare = old left decomposition
                if (h->p.check_propagation)
                    MCB_CUDA(cudaMemcpyAsync(h->Gtmp, h->G, h->mat * sizeof(double),
                                             cudaMemcpyDeviceToDevice, h->st));
                MCB_TRY(calc_greens(h, h->spare, h->slot[k], h->G));
                if (h->p.check_propagation) MCB_TRY(record_prop_err(h));
                if (defer_wrap) *wrap_pending = true;
                else MCB_TRY(wrap_greens(h, h->G, wrap_tmp, h->current_slice + 1, -1));
            } else {
                MCB_TRY(advance(h, false, h->slot[1], h->spare, 1, s));
                std::swap(h->slot[0], h->spare);
                h->direction = 1;
                h->current_slice = 0;
                return engine_propagate(h, defer_wrap, wrap_pending);
            }
        } else {
            MCB_TRY(wrap_greens(h, h->G, wrap_tmp, h->current_slice, -1));
            h->current_slice -= 1;
        }
    }
    return 0;
}

// update (DQMC.jl:523-538) = propagate + sweep_spatial, general path (any N): separate launches
int engine_update(mcb_dqmc* h)
{
    h->conf_version++;
    MCB_TRY(engine_propagate(h, false, nullptr));
    SweepArgs a;
    a.n = h->N; a.batch = h->C;
    a.G = h->G; a.sG = (long long)h->nn;
    a.conf = h->conf + (size_t)(h->current_slice - 1) * h->N;
    a.sConf = (long long)h->M * h->N;
    const double dEp = -2.0 * h->lambda * 1.0, dEm = -2.0 * h->lambda * -1.0;
    a.gamma_p = exp(dEp) - 1.0; a.efac_p = exp(-dEp);
    a.gamma_m = exp(dEm) - 1.0; a.efac_m = exp(-dEm);
    a.key0 = seed_key0(h->p.seed);
    a.chain_offset = h->p.chain_offset;
    a.sweep = (uint32_t)(h->n_updates / (2 * (int64_t)h->M));
    a.visit = (uint32_t)(h->n_updates % (2 * (int64_t)h->M));
    a.acc = h->acc; a.prop = h->prop; a.nonfinite = h->nonfinite;
    // immediate rank-1 updates only where the whole update is a handful of elements per thread
    if (h->N >= 16) MCB_TRY(launch_sweep_delayed(a, h->st));
    else MCB_TRY(launch_sweep_spatial(a, h->st));
    h->n_updates += 1;
    return 0;
}

// A run of update() calls on the chain-resident path (segment.cu): the stabilisation part of
// propagate (if the next update needs one) runs through the stack kernels, then ONE launch does
// [wrap +] sweep_spatial for this update and for every following update that is a plain wrap in the
// same direction (at most safe_mult - 1 of them), up to max_updates.
int engine_update_batch(mcb_dqmc* h, int64_t max_updates, bool stop_at_measure_point, int64_t* done)
{
    h->conf_version++;
    *done = 0;
    if (max_updates <= 0) return 0;
    if (!h->use_panel && !h->use_resident_segment) {
        MCB_TRY(engine_update(h));
        *done = 1;
        return 0;
    }
    const int s = h->s;
    const bool plain = h->direction == 1 ? (h->current_slice % s != 0) : ((h->current_slice - 1) % s != 0);
    bool first_wrap = true;
    if (plain) h->current_slice += h->direction;
    else MCB_TRY(engine_propagate(h, true, &first_wrap));
    const int start = h->current_slice, d = h->direction;
    int64_t further = d == 1 ? (start % s == 0 ? 0 : s - start % s) : (start - 1) % s;
    if (stop_at_measure_point && start == 1 && d == 1) further = 0;
    const int64_t visits = 1 + further < max_updates ? 1 + further : max_updates;

    SegmentArgs a;
    a.n = h->N; a.batch = h->C;
    a.G = h->G; a.sG = (long long)h->nn;
    a.conf = h->conf; a.sConf = (long long)h->M * h->N;
    a.eT2 = h->hop[2]; a.eTinv2 = h->hop[3];
    a.n_slices = h->M;
    a.start_slice = start; a.direction = d; a.n_visits = (int)visits; a.first_wrap = first_wrap ? 1 : 0;
    a.eVp[0] = h->eVp[0]; a.eVp[1] = h->eVp[1]; a.eVm[0] = h->eVm[0]; a.eVm[1] = h->eVm[1];
    const double dEp = -2.0 * h->lambda * 1.0, dEm = -2.0 * h->lambda * -1.0;
    a.gamma_p = exp(dEp) - 1.0; a.efac_p = exp(-dEp);
    a.gamma_m = exp(dEm) - 1.0; a.efac_m = exp(-dEm);
    a.key0 = seed_key0(h->p.seed);
    a.chain_offset = h->p.chain_offset;
    a.n_updates0 = h->n_updates;
    a.acc = h->acc; a.prop = h->prop; a.nonfinite = h->nonfinite;
    {
        // all chains of a CTA wrap in the same window (segment.cuh); MCB_PHASE_SYNC=0 turns it off
        const char* ev = getenv("MCB_PHASE_SYNC");
        a.phase_sync = ((!ev || atoi(ev)) && h->C % 4 == 0) ? 1 : 0;
    }
    if (h->use_panel) {
        MCB_TRY(launch_segment(a, h->st));
    } else {
        a.eT2T = h->hopT[0]; a.eTinv2T = h->hopT[1];
        MCB_TRY(launch_segment_resident(a, h->st));
    }
    h->current_slice = start + d * (int)(visits - 1);
    h->n_updates += visits;
    *done = visits;
    return 0;
}

// ------------------------------------------------------------------ from-scratch G
namespace {

// UDT of a slice-matrix chain with a stabilisation every safe_mult slices
// (the two loops of calculate_greens(mc, slice), stack.jl:429-476).
int scratch_side(mcb_dqmc* h, bool left, int lo, int hi, double* dstU, double* dstD, double* dstT)
{
    MCB_TRY(set_identity(h, dstU, dstT, dstD));
    const int cnt = hi - lo + 1;
    if (cnt <= 0) return 0;
    MCB_TRY(set_identity(h, h->W[3], nullptr, nullptr));
    const double* cur = h->W[3];
    double* Tacc = dstT;
    double* Tother = h->W[7];
    int pp = 0;
    const Scale dcol = vec_scale(dstD, h->N);
    for (int t = 0; t < cnt; ++t) {
        const int slice = left ? lo + t : hi - t;
        const bool last = (t == cnt - 1);
        const bool stab = last || (slice % h->s == 0);
        double* nxt = h->W[pp];
        if (left) MCB_TRY(mul_slice_left(h, slice, cur, nxt, stab ? &dcol : nullptr));
        else MCB_TRY(mul_daggered_slice_left(h, slice, cur, nxt, stab ? &dcol : nullptr));
        if (stab) {
            double* Uout = last ? dstU : h->W[3];
            MCB_TRY(launch_udt(h->N, h->C, nxt, h->nn, Uout, h->nn, dstD, h->N, h->W[2], h->nn, h->pivot,
                               h->N, 1, h->st));
            MCB_TRY(launch_gemm(base_gemm(h, h->W[2], false, Tacc, false, Tother), h->st));
            std::swap(Tacc, Tother);
            cur = Uout;
        } else {
            cur = nxt;
            pp ^= 1;
        }
    }
    if (Tacc != dstT)
        MCB_CUDA(cudaMemcpyAsync(dstT, Tacc, h->mat * sizeof(double), cudaMemcpyDeviceToDevice, h->st));
    return 0;
}

} // namespace

// calculate_greens(mc, slice) (stack.jl:422-480): G = [1 + B_slice..B_1 B_M..B_{slice+1}]^-1.
// The stack is untouched: the two temporary decompositions live in the spare slot buffer and in
// (Gtmp, Dx, W[6]).
int engine_calculate_greens_at(mcb_dqmc* h, int slice)
{
    // right triple -> spare slot buffer ; left triple -> (Gtmp, Dx, W[6]).
    // calc_greens_raw uses W[0..5] and Dw[0..1] as scratch, so the left D lives in its own buffer.
    double* leftD = h->Dx;
    const int sp = h->spare;
    MCB_TRY(scratch_side(h, false, slice + 1, h->M, h->bufU[sp], h->bufD[sp], h->bufT[sp]));
    MCB_TRY(scratch_side(h, true, 1, slice, h->Gtmp, leftD, h->W[6]));
    return calc_greens_raw(h->N, h->C, h->Gtmp, leftD, h->W[6], h->bufU[sp], h->bufD[sp], h->bufT[sp],
                           h->G, h->W, h->Dw, h->pivot, h->st, h->use_panel);
}

// _greens! (DQMC.jl:721-730): out = e^{+dtau T/2} (G e^{-dtau T/2})
int engine_physical_greens(mcb_dqmc* h, double* out)
{
    MCB_TRY(launch_gemm(base_gemm(h, h->G, false, h->hop[0], true, h->W[0]), h->st));
    MCB_TRY(launch_gemm(base_gemm(h, h->hop[1], true, h->W[0], false, out), h->st));
    return 0;
}

// ------------------------------------------------------------------ measurements
// apply!(::Greens, ...) (measurements/generic.jl:207-215): G = greens!(mc), then every measurement of the group
int engine_measure(mcb_dqmc* h)
{
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    if (h->ms.flavors != 1) { set_error("two-flavor tables on a one-flavor handle"); return MCB_ERR_STATE; }
    MCB_TRY(engine_physical_greens(h, h->W[1]));
    return measure_accumulate(h->ms, h->W[1], h->W[1], h->st);
}

// apply!(::CombinedGreensIterator, ...) (measurements/generic.jl:227-243): G00 = greens!(mc), then for every slice of the
// iterator the time-displaced kernels of the group on (G00, G0l, Gl0, Gll); finish! scales by delta_tau / N. The iterator
// runs on the device (unequal.cu) and leaves the sweep state as it found it.
int engine_measure_td(mcb_dqmc* h, int recalculate, bool from_g0)
{
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    if (h->ms.flavors != 1) { set_error("two-flavor tables on a one-flavor handle"); return MCB_ERR_STATE; }
    if (from_g0) MCB_TRY(engine_cgi_begin_from_g0(h, recalculate));   // replay!: G = G(0) of the recorded configuration
    else MCB_TRY(engine_cgi_begin(h, recalculate));                   // checks the measurement point
    double* g00[2];
    MCB_TRY(measure_td_g00(h->ms, g00));
    MCB_TRY(engine_physical_greens(h, g00[0]));
    const double factor = h->p.delta_tau / (double)h->N;
    for (;;) {
        int l = 0;
        const double* res[3] = {nullptr, nullptr, nullptr};
        MCB_TRY(engine_cgi_next(h, 1, &l, res));
        if (l == 0) break;
        const double* const mats[4] = {g00[0], res[0], res[1], res[2]};
        MCB_TRY(measure_td_accumulate(h->ms, mats, mats, factor, h->st));
    }
    h->ms.n_td += 1;
    return 0;
}

} // namespace mcb
