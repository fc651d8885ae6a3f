// Unequal-time Green's function G(k <- l) = <c_i(k dtau) c_j^+(l dtau)> of the current configuration,
// batched over the chains of a handle. Follows src/flavors/DQMC/unequal_time_stack.jl:
//   UnequalTimeStack / build_stack           :1-166   -> ut_build      (forward, backward, inverse stacks)
//   compute_inverse_udt_block!               :326-378 -> inverse_block
//   compute_forward_/backward_udt_block!     :391-439 -> forward_block / backward_block
//   calculate_greens_full1! / _full2!        :443-622 -> full1 / full2
// Every matrix product / factorisation / solve is one of the batched kernels of the sweep (gemm.cu, udt.cu,
// stab.cu); the diagonal scalings of the reference are fused into the products as Scale operands, the
// min(1, d) / 1/max(1, d) splittings of the scales are small vector kernels. The stacks are rebuilt when
// the configuration has changed since the last call (h->conf_version).
#include "dqmc_engine.cuh"
#include <utility>

struct UtStack {
    std::vector<double*> fu, fd, ft, bu, bd, bt, iu, id, it;
    double* U[2] = {nullptr, nullptr};
    double* T[2] = {nullptr, nullptr};
    double* D = nullptr;
    double *Ul = nullptr, *Dl = nullptr, *Tl = nullptr, *Ur = nullptr, *Dr = nullptr, *Tr = nullptr;
    double *t1 = nullptr, *t2 = nullptr, *t3 = nullptr, *dv = nullptr, *dw = nullptr;
    double* G = nullptr; // result (ut_stack.greens): the sweep's own G (h->G) is never touched
    std::vector<void*> all;
    int64_t built_version = -1;
    // CombinedGreensIterator state: (Ul Dl Tl) -> G(l <- 0), (Ur Dr Tr) -> G(0 <- l), (U D T) -> G(l <- l); the triples of the
    // iterator are its own (the stabilised evaluations above clobber Ul .. Tr). out[0..2] = G0l, Gl0, Gll (effective),
    // phys[0..2] the same after _greens!.
    double *cUl = nullptr, *cDl = nullptr, *cTl = nullptr, *cUr = nullptr, *cDr = nullptr, *cTr = nullptr;
    double *cU = nullptr, *cD = nullptr, *cT = nullptr, *cX = nullptr;
    double* out[3] = {nullptr, nullptr, nullptr};
    double* phys[3] = {nullptr, nullptr, nullptr};
    const double* last[3] = {nullptr, nullptr, nullptr};   // results of the last step
    int cgi_l = 0, cgi_recalc = 0;   // next slice to produce (0: no iteration in progress)
    int64_t cgi_version = -1;
};

namespace mcb {

namespace {

#define MCB_TRY(expr)                        \
    do {                                     \
        const int rc__ = (expr);             \
        if (rc__ != 0) return rc__;          \
    } while (0)

enum VecOp { V_MAXINV = 0, V_MIN = 1, V_INV = 2 };

__global__ void vec_op_kernel(double* out, const double* in, size_t total, int mode)
{
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const double w = in[e];
        out[e] = mode == V_MAXINV ? 1.0 / fmax(1.0, w) : (mode == V_MIN ? fmin(1.0, w) : 1.0 / w);
    }
}

// mode 0: C = A + B ; 1: C = -A ; 2: C = diag(d) (d per chain) ; 3: C = A * diag(d) ; 4: C = identity, d = 1 ;
// 5: C = diag(B) * A * diag(d) (B, d vectors per chain) ; 6: C = A - I ; 7: C = diag(d) * A
__global__ void mat_op_kernel(double* C, const double* A, const double* B, double* d, int n, size_t total, int mode)
{
    const size_t nn = (size_t)n * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t chain = e / nn, r = e % nn;
        const int i = (int)(r % n), j = (int)(r / n);
        double v;
        if (mode == 0) v = A[e] + B[e];
        else if (mode == 1) v = -A[e];
        else if (mode == 2) v = i == j ? d[chain * n + i] : 0.0;
        else if (mode == 3) v = A[e] * d[chain * n + j];
        else if (mode == 5) v = B[chain * n + i] * A[e] * d[chain * n + j];
        else if (mode == 6) v = A[e] - (i == j ? 1.0 : 0.0);
        else if (mode == 7) v = d[chain * n + i] * A[e];
        else {
            v = i == j ? 1.0 : 0.0;
            if (d && j == 0) d[chain * n + i] = 1.0;
        }
        C[e] = v;
    }
}

int vec_op(mcb_dqmc* h, double* out, const double* in, int mode)
{
    LaunchScope scope(LC_MISC, h->st, 0.0);
    vec_op_kernel<<<64, 256, 0, h->st>>>(out, in, h->vec, mode);
    MCB_CUDA(cudaGetLastError());
    return 0;
}
int mat_op(mcb_dqmc* h, double* C, const double* A, const double* B, double* d, int mode)
{
    LaunchScope scope(LC_MISC, h->st, 0.0);
    mat_op_kernel<<<296, 256, 0, h->st>>>(C, A, B, d, h->N, h->mat, mode);
    MCB_CUDA(cudaGetLastError());
    return 0;
}

// C = diag(row) opA(A) opB(B) diag(col); null scale pointers = no scaling
int mm(mcb_dqmc* h, const double* A, bool ta, const double* B, bool tb, double* C, const double* row = nullptr,
       const double* col = nullptr)
{
    GemmArgs g = base_gemm(h, A, false, B, false, C);
    g.transA = ta; g.transB = tb;
    if (row) g.row = vec_scale(row, h->N);
    if (col) g.col = vec_scale(col, h->N);
    return launch_gemm(g, h->st);
}

int udt(mcb_dqmc* h, double* X, double* U, double* D, double* T, int apply_pivot)
{
    return launch_udt(h->N, h->C, X, h->nn, U, h->nn, D, h->N, T, h->nn, h->pivot, h->N, apply_pivot, h->st);
}
int rdivp(mcb_dqmc* h, const double* in, double* out, const double* T)
{
    return launch_rdivp(h->N, h->C, in, h->nn, out, h->nn, T, h->nn, h->pivot, h->N, h->st);
}

int ut_alloc(mcb_dqmc* h)
{
    if (h->ut) return 0;
    UtStack* u = new UtStack();
    h->ut = u;
    auto mat = [&](double*& p) {
        MCB_CUDA(cudaMalloc(&p, h->mat * sizeof(double)));
        u->all.push_back(p);
        return 0;
    };
    auto vec = [&](double*& p) {
        MCB_CUDA(cudaMalloc(&p, h->vec * sizeof(double)));
        u->all.push_back(p);
        return 0;
    };
    const int ne = h->ne, R = h->ne - 1;
    for (auto* v : {&u->fu, &u->ft, &u->bu, &u->bt}) {
        v->assign(ne, nullptr);
        for (auto& p : *v) MCB_TRY(mat(p));
    }
    for (auto* v : {&u->fd, &u->bd}) {
        v->assign(ne, nullptr);
        for (auto& p : *v) MCB_TRY(vec(p));
    }
    for (auto* v : {&u->iu, &u->it}) {
        v->assign(R, nullptr);
        for (auto& p : *v) MCB_TRY(mat(p));
    }
    u->id.assign(R, nullptr);
    for (auto& p : u->id) MCB_TRY(vec(p));
    for (double** p : {&u->U[0], &u->U[1], &u->T[0], &u->T[1], &u->Ul, &u->Tl, &u->Ur, &u->Tr, &u->t1, &u->t2, &u->t3, &u->G})
        MCB_TRY(mat(*p));
    for (double** p : {&u->D, &u->Dl, &u->Dr, &u->dv, &u->dw}) MCB_TRY(vec(*p));
    return 0;
}

// initialize_stack + build_stack(mc, s::UnequalTimeStack)   unequal_time_stack.jl:54-166
int ut_build(mcb_dqmc* h)
{
    UtStack* u = h->ut;
    if (u->built_version == h->conf_version) return 0;
    const int s = h->s, R = h->ne - 1;
    MCB_TRY(mat_op(h, u->fu[0], nullptr, nullptr, u->fd[0], 4));
    MCB_TRY(mat_op(h, u->ft[0], nullptr, nullptr, nullptr, 4));
    MCB_TRY(mat_op(h, u->bu[R], nullptr, nullptr, u->bd[R], 4));
    MCB_TRY(mat_op(h, u->bt[R], nullptr, nullptr, nullptr, 4));
    for (int idx = 1; idx <= R; ++idx) // forward: B_{idx s} ... B_1                                  (:113-126)
        MCB_TRY(advance_raw(h, true, u->fu[idx - 1], u->fd[idx - 1], u->ft[idx - 1], u->fu[idx], u->fd[idx], u->ft[idx],
                            (idx - 1) * s + 1, idx * s));
    for (int idx = R; idx >= 1; --idx) // backward: B'_{(idx-1)s+1} ... B'_M                          (:129-142)
        MCB_TRY(advance_raw(h, false, u->bu[idx], u->bd[idx], u->bt[idx], u->bu[idx - 1], u->bd[idx - 1], u->bt[idx - 1],
                            (idx - 1) * s + 1, idx * s));
    for (int idx = 1; idx <= R; ++idx) { // inverse blocks B^-1_{(idx-1)s+1} ... B^-1_{idx s}           (:146-157)
        MCB_TRY(mat_op(h, u->t1, nullptr, nullptr, nullptr, 4));
        double *cur = u->t1, *nxt = u->t2;
        for (int slice = idx * s; slice >= (idx - 1) * s + 1; --slice) {
            GemmArgs g = base_gemm(h, h->hop[3], true, cur, false, nxt); // multiply_slice_matrix_inv_left!
            g.row = conf_scale(h, slice, false);
            MCB_TRY(launch_gemm(g, h->st));
            std::swap(cur, nxt);
        }
        MCB_TRY(udt(h, cur, u->iu[idx - 1], u->id[idx - 1], u->it[idx - 1], 1));
    }
    u->built_version = h->conf_version;
    return 0;
}

// compute_inverse_udt_block!  (:326-378): U D T = B_{low+1}^-1 ... B_high^-1 -> (u->U[0], u->D, u->T[0])
int inverse_block(mcb_dqmc* h, int low, int high)
{
    UtStack* u = h->ut;
    const int s = h->s;
    double *U = u->U[0], *U2 = u->U[1], *T = u->T[0], *T2 = u->T[1];
    const int lower = (low + 1 + s - 2) / s + 1;
    const int upper = high / s;
    MCB_TRY(mat_op(h, U, nullptr, nullptr, u->D, 4));
    MCB_TRY(mat_op(h, T, nullptr, nullptr, nullptr, 4));
    for (int idx = lower; idx <= upper; ++idx) {
        // tmp1 = diag(D) (T inv_u) diag(inv_d) ; udt -> (tmp2, D, tmp1) ; T = tmp1 inv_t ; U = U tmp2
        MCB_TRY(mm(h, T, false, u->iu[idx - 1], false, u->t1, u->D, u->id[idx - 1]));
        MCB_TRY(udt(h, u->t1, u->t2, u->D, u->t3, 1));
        MCB_TRY(mm(h, u->t3, false, u->it[idx - 1], false, T));
        MCB_TRY(mm(h, U, false, u->t2, false, U2));
        std::swap(U, U2);
    }
    const int lower_slice = (lower - 1) * s + 1, upper_slice = upper * s;
    const int top = lower_slice - 1 < high ? lower_slice - 1 : high;
    if (top >= low + 1) {
        for (int slice = top; slice >= low + 1; --slice) {
            GemmArgs g = base_gemm(h, h->hop[3], true, U, false, U2);
            g.row = conf_scale(h, slice, false);
            if (slice == low + 1) g.col = vec_scale(u->D, h->N); // tmp1 = U * Diagonal(D) fused
            MCB_TRY(launch_gemm(g, h->st));
            std::swap(U, U2);
        }
        MCB_TRY(udt(h, U, U2, u->D, u->t3, 1));
        std::swap(U, U2);
        MCB_TRY(mm(h, u->t3, false, T, false, T2));
        std::swap(T, T2);
    }
    const int from = upper_slice + 1 > top + 1 ? upper_slice + 1 : top + 1;
    for (int slice = from; slice <= high; ++slice) {
        GemmArgs g = base_gemm(h, T, false, h->hop[3], true, T2); // multiply_slice_matrix_inv_right!
        g.mid = conf_scale(h, slice, false);
        MCB_TRY(launch_gemm(g, h->st));
        std::swap(T, T2);
    }
    // results where the callers expect them
    if (U != u->U[0]) std::swap(u->U[0], u->U[1]);
    if (T != u->T[0]) std::swap(u->T[0], u->T[1]);
    return 0;
}

// one side of the product with a partial range: (U, D, T) = udt(prod * Usrc * diag(Dsrc)), T <- T * Tsrc
int side_block(mcb_dqmc* h, bool left, const double* Usrc, const double* Dsrc, const double* Tsrc, double* U, double* D,
               double* T, int lo, int hi)
{
    UtStack* u = h->ut;
    if (hi >= lo) return advance_raw(h, left, Usrc, Dsrc, Tsrc, U, D, T, lo, hi);
    MCB_TRY(mat_op(h, u->t1, Usrc, nullptr, const_cast<double*>(Dsrc), 3));
    MCB_TRY(udt(h, u->t1, U, D, u->t2, 1));
    return mm(h, u->t2, false, Tsrc, false, T);
}

// compute_forward_udt_block!  (:391-408): Ul Dl Tl = B_slice ... B_1
int forward_block(mcb_dqmc* h, int slice)
{
    UtStack* u = h->ut;
    const int idx = (slice - 1) / h->s; // Julia's div truncates: slice = 0 -> 0
    return side_block(h, true, u->fu[idx], u->fd[idx], u->ft[idx], u->Ul, u->Dl, u->Tl, h->s * idx + 1, slice);
}
// compute_backward_udt_block! (:422-439): (Ur Dr Tr)' = B_M ... B_{slice+1}
int backward_block(mcb_dqmc* h, int slice)
{
    UtStack* u = h->ut;
    const int idx = (slice + h->s - 1) / h->s;
    return side_block(h, false, u->bu[idx], u->bd[idx], u->bt[idx], u->Ur, u->Dr, u->Tr, slice + 1, h->s * idx);
}

// calculate_greens_full1!  (:443-525): G(slice1 <- slice2), slice1 >= slice2
int full1(mcb_dqmc* h, int slice1, int slice2)
{
    UtStack* u = h->ut;
    MCB_TRY(inverse_block(h, slice2, slice1));
    MCB_TRY(forward_block(h, slice2));
    MCB_TRY(backward_block(h, slice1));
    double *U = u->U[0], *D = u->D, *T = u->T[0], *G = u->G;
    // B1 + udt: [U D T + Ul (Dl Tl Tr' Dr) Ur']^-1 ; (Tr, Dr, t1) = udt(Dl Tl Tr' Dr), t1 "dirty"
    MCB_TRY(mm(h, u->Tl, false, u->Tr, true, G, u->Dl, u->Dr));
    MCB_TRY(udt(h, G, u->Tr, u->Dr, u->t1, 0));
    // B2: Tl = Ul Tr ; Ur <- Ur[:, p] triu(t1)^-1  (-> t2)
    MCB_TRY(mm(h, u->Ul, false, u->Tr, false, u->Tl));
    MCB_TRY(rdivp(h, u->Ur, u->t2, u->t1));
    // B3: Tr = 1/max(1, D) (U' Tl) min(1, Dr)
    MCB_TRY(vec_op(h, u->dv, D, V_MAXINV));
    MCB_TRY(vec_op(h, u->dw, u->Dr, V_MIN));
    MCB_TRY(mm(h, U, true, u->Tl, false, u->Tr, u->dv, u->dw));
    // B4: Tl = min(1, D) (T Ur) 1/max(1, Dr)
    MCB_TRY(vec_op(h, u->dv, D, V_MIN));
    MCB_TRY(vec_op(h, u->dw, u->Dr, V_MAXINV));
    MCB_TRY(mm(h, T, false, u->t2, false, u->Tl, u->dv, u->dw));
    // sum, UDT: (Tr, Dl, Tl) = udt(Tl + Tr), Tl "dirty"
    MCB_TRY(mat_op(h, u->t3, u->Tl, u->Tr, nullptr, 0));
    MCB_TRY(udt(h, u->t3, u->Tr, u->Dl, u->Tl, 0));
    // B5: G = [((1/Dr_max) / Tl) 1/Dl] Tr' 1/D_max
    MCB_TRY(vec_op(h, u->Dr, u->Dr, V_MAXINV));
    MCB_TRY(mat_op(h, u->Ul, nullptr, nullptr, u->Dr, 2));
    MCB_TRY(rdivp(h, u->Ul, u->t1, u->Tl));
    MCB_TRY(vec_op(h, u->dv, D, V_MAXINV));
    {
        GemmArgs g = base_gemm(h, u->t1, false, u->Tr, false, G);
        g.transB = true;
        g.mid = vec_scale(u->Dl, h->N, true);
        g.col = vec_scale(u->dv, h->N);
        MCB_TRY(launch_gemm(g, h->st));
    }
    // B6: G = Ur (G U')
    MCB_TRY(mm(h, G, false, U, true, u->t3));
    return mm(h, u->t2, false, u->t3, false, G);
}

// calculate_greens_full2!  (:529-622): G(slice1 <- slice2), slice1 <= slice2
int full2(mcb_dqmc* h, int slice1, int slice2)
{
    UtStack* u = h->ut;
    MCB_TRY(inverse_block(h, slice1, slice2));
    MCB_TRY(forward_block(h, slice1));
    MCB_TRY(backward_block(h, slice2));
    double *U = u->U[0], *D = u->D, *T = u->T[0], *G = u->G;
    // B1 + udt
    MCB_TRY(mm(h, u->Tl, false, u->Tr, true, G, u->Dl, u->Dr));
    MCB_TRY(udt(h, G, u->Tr, u->Dr, u->t1, 0));
    // B2: Tl = Ul Tr (kept) ; t2 = 1/max(1, D) (U' Tl) min(1, Dr)
    MCB_TRY(mm(h, u->Ul, false, u->Tr, false, u->Tl));
    MCB_TRY(vec_op(h, u->dv, D, V_MAXINV));
    MCB_TRY(vec_op(h, u->dw, u->Dr, V_MIN));
    MCB_TRY(mm(h, U, true, u->Tl, false, u->t2, u->dv, u->dw));
    // B3: Tr = min(1, D) ((T Ur)[:, p] triu(t1)^-1) 1/max(1, Dr)
    MCB_TRY(mm(h, T, false, u->Ur, false, u->t3));
    MCB_TRY(rdivp(h, u->t3, U, u->t1));
    MCB_TRY(vec_op(h, u->dv, D, V_MIN));
    MCB_TRY(vec_op(h, u->dw, u->Dr, V_MAXINV));
    MCB_TRY(mat_op(h, u->Tr, U, u->dv, u->dw, 5));
    // sum, udt: (Ul, Dl, Tr) = udt(Tr + t2), Tr "dirty"
    MCB_TRY(mat_op(h, u->t3, u->Tr, u->t2, nullptr, 0));
    MCB_TRY(udt(h, u->t3, u->Ul, u->Dl, u->Tr, 0));
    // B4: t3 = {[(Dr_min / Tr) 1/Dl] Ul'} D_min
    MCB_TRY(vec_op(h, u->Dr, u->Dr, V_MIN));
    MCB_TRY(mat_op(h, u->t2, nullptr, nullptr, u->Dr, 2));
    MCB_TRY(rdivp(h, u->t2, U, u->Tr));
    MCB_TRY(vec_op(h, u->dv, D, V_MIN));
    {
        GemmArgs g = base_gemm(h, U, false, u->Ul, false, u->t3);
        g.transB = true;
        g.mid = vec_scale(u->Dl, h->N, true);
        g.col = vec_scale(u->dv, h->N);
        MCB_TRY(launch_gemm(g, h->st));
    }
    // B6: G = -Tl (t3 T)
    MCB_TRY(mm(h, u->t3, false, T, false, u->t1));
    MCB_TRY(mm(h, u->Tl, false, u->t1, false, G));
    return mat_op(h, G, G, nullptr, nullptr, 1);
}

int cgi_alloc(mcb_dqmc* h)
{
    UtStack* u = h->ut;
    if (u->cUl) return 0;
    auto mat = [&](double*& p) {
        MCB_CUDA(cudaMalloc(&p, h->mat * sizeof(double)));
        u->all.push_back(p);
        return 0;
    };
    auto vec = [&](double*& p) {
        MCB_CUDA(cudaMalloc(&p, h->vec * sizeof(double)));
        u->all.push_back(p);
        return 0;
    };
    for (double** p : {&u->cUl, &u->cTl, &u->cUr, &u->cTr, &u->cU, &u->cT, &u->cX, &u->out[0], &u->out[1], &u->out[2],
                       &u->phys[0], &u->phys[1], &u->phys[2]})
        MCB_TRY(mat(*p));
    for (double** p : {&u->cDl, &u->cDr, &u->cD}) MCB_TRY(vec(*p));
    return 0;
}

int copy_mat(mcb_dqmc* h, double* dst, const double* src)
{
    MCB_CUDA(cudaMemcpyAsync(dst, src, h->mat * sizeof(double), cudaMemcpyDeviceToDevice, h->st));
    return 0;
}
int copy_vec(mcb_dqmc* h, double* dst, const double* src)
{
    MCB_CUDA(cudaMemcpyAsync(dst, src, h->vec * sizeof(double), cudaMemcpyDeviceToDevice, h->st));
    return 0;
}
// (U, D, T) = udt(X) with X kept
int udt_of(mcb_dqmc* h, const double* X, double* U, double* D, double* T)
{
    MCB_TRY(copy_mat(h, h->ut->cX, X));
    return udt(h, h->ut->cX, U, D, T, 1);
}
// X <- B_l X (multiply_slice_matrix_left!) and X <- X B_l^-1 (multiply_slice_matrix_inv_right!), through cX
int cgi_mul_left(mcb_dqmc* h, int l, double*& X)
{
    GemmArgs g = base_gemm(h, h->hop[2], true, X, false, h->ut->cX);
    g.mid = conf_scale(h, l, true);
    MCB_TRY(launch_gemm(g, h->st));
    std::swap(X, h->ut->cX);
    return 0;
}
int cgi_mul_inv_right(mcb_dqmc* h, int l, double*& X)
{
    GemmArgs g = base_gemm(h, X, false, h->hop[3], true, h->ut->cX);
    g.mid = conf_scale(h, l, false);
    MCB_TRY(launch_gemm(g, h->st));
    std::swap(X, h->ut->cX);
    return 0;
}
// C = A diag(d) B
int mdm(mcb_dqmc* h, const double* A, const double* d, const double* B, double* C)
{
    GemmArgs g = base_gemm(h, A, false, B, false, C);
    g.mid = vec_scale(d, h->N);
    return launch_gemm(g, h->st);
}

// iterate(it::CombinedGreensIterator) (:768-787): the three triples from G(0) = mc.s.greens
int cgi_begin(mcb_dqmc* h, int recalculate)
{
    UtStack* u = h->ut;
    MCB_TRY(cgi_alloc(h));
    MCB_TRY(udt_of(h, h->G, u->cUl, u->cDl, u->cTl));
    MCB_TRY(copy_mat(h, u->cU, u->cUl));
    MCB_TRY(copy_vec(h, u->cD, u->cDl));
    MCB_TRY(copy_mat(h, u->cT, u->cTl));
    MCB_TRY(mat_op(h, u->cX, h->G, nullptr, nullptr, 6));
    MCB_TRY(udt(h, u->cX, u->cUr, u->cDr, u->cTr, 1));
    u->cgi_l = 1;
    u->last[0] = u->last[1] = u->last[2] = nullptr;
    u->cgi_recalc = recalculate > 0 ? recalculate : 4 * h->s;
    u->cgi_version = h->conf_version;
    return 0;
}

// iterate(it, l) (:789-885): recalculation, stabilisation or quick advance
int cgi_step(mcb_dqmc* h)
{
    UtStack* u = h->ut;
    const int l = u->cgi_l;
    double *G0l = u->out[0], *Gl0 = u->out[1], *Gll = u->out[2];
    if (l % u->cgi_recalc == 0) {   // (:795-822)
        MCB_TRY(full1(h, l, 0));
        MCB_TRY(copy_mat(h, Gl0, u->G));
        MCB_TRY(full2(h, 0, l));
        MCB_TRY(copy_mat(h, G0l, u->G));
        MCB_TRY(full1(h, l, l));
        MCB_TRY(copy_mat(h, Gll, u->G));
        MCB_TRY(udt_of(h, Gll, u->cU, u->cD, u->cT));
        MCB_TRY(udt_of(h, Gl0, u->cUl, u->cDl, u->cTl));
        MCB_TRY(udt_of(h, G0l, u->cUr, u->cDr, u->cTr));
        return 0;
    }
    MCB_TRY(cgi_mul_left(h, l, u->cUl));
    MCB_TRY(cgi_mul_inv_right(h, l, u->cTr));
    MCB_TRY(cgi_mul_left(h, l, u->cU));
    MCB_TRY(cgi_mul_inv_right(h, l, u->cT));
    MCB_TRY(mdm(h, u->cUl, u->cDl, u->cTl, Gl0));
    MCB_TRY(mdm(h, u->cUr, u->cDr, u->cTr, G0l));
    MCB_TRY(mdm(h, u->cU, u->cD, u->cT, Gll));
    if (l % h->s != 0) return 0;    // quick advance (:862-885)
    // stabilisation (:824-861): re-factorise the large side of every triple
    // Gl0: (Ul, Dl, t) = udt(Ul Dl), Tl = t Tl
    MCB_TRY(mat_op(h, u->t1, u->cUl, nullptr, u->cDl, 3));
    MCB_TRY(udt(h, u->t1, u->cUl, u->cDl, u->t2, 1));
    MCB_TRY(mm(h, u->t2, false, u->cTl, false, u->cX));
    std::swap(u->cTl, u->cX);
    // G0l: (t, Dr, Tr) = udt(Dr Tr), Ur = Ur t
    MCB_TRY(mat_op(h, u->t1, u->cTr, nullptr, u->cDr, 7));
    MCB_TRY(udt(h, u->t1, u->t2, u->cDr, u->cTr, 1));
    MCB_TRY(mm(h, u->cUr, false, u->t2, false, u->cX));
    std::swap(u->cUr, u->cX);
    // Gll: (Q, D, t) = udt(U D) ; (t', D, T) = udt(D (t T)) ; U = Q t'
    MCB_TRY(mat_op(h, u->t1, u->cU, nullptr, u->cD, 3));
    MCB_TRY(udt(h, u->t1, u->t2, u->cD, u->t3, 1));
    MCB_TRY(mm(h, u->t3, false, u->cT, false, u->t1, u->cD));
    MCB_TRY(udt(h, u->t1, u->t3, u->cD, u->cT, 1));
    return mm(h, u->t2, false, u->t3, false, u->cU);
}

} // namespace

int engine_greens_unequal(mcb_dqmc* h, int slice1, int slice2)
{
    MCB_TRY(ut_alloc(h));
    MCB_TRY(ut_build(h));
    return slice1 >= slice2 ? full1(h, slice1, slice2) : full2(h, slice1, slice2);
}

double* unequal_result(mcb_dqmc* h) { return h->ut ? h->ut->G : nullptr; }

// CombinedGreensIterator(mc, recalculate) (unequal_time_stack.jl:746-787): needs G = G(tau = 0), i.e. the measurement
// point current_slice == 1 on the way up (the assertion of :770; DQMC.jl's run loop measures only there)
int engine_cgi_begin(mcb_dqmc* h, int recalculate)
{
    if (!h->stack_built || h->current_slice != 1 || h->direction != 1) {
        set_error("CombinedGreensIterator: the handle must be at current_slice == 1 on the way up (G = G(0))");
        return MCB_ERR_STATE;
    }
    MCB_TRY(ut_alloc(h));
    MCB_TRY(ut_build(h));
    return cgi_begin(h, recalculate);
}

// the same start for replay!, where calculate_greens(mc, 0) has just put G(0) of a recorded configuration into mc.s.greens
// and there is no sweep state to check (DQMC.jl:649-660)
int engine_cgi_begin_from_g0(mcb_dqmc* h, int recalculate)
{
    MCB_TRY(ut_alloc(h));
    MCB_TRY(ut_build(h));
    return cgi_begin(h, recalculate);
}

// one step: *l_out = the slice the results belong to (1 .. M), 0 when the iteration is over. results[0..2] = device
// pointers to G(0 <- l), G(l <- 0), G(l <- l), [chain][col][row], valid until the next step.
int engine_cgi_next(mcb_dqmc* h, int physical, int* l_out, const double* results[3])
{
    UtStack* u = h->ut;
    if (!u || u->cgi_l == 0 || u->cgi_version != h->conf_version) {
        set_error("CombinedGreensIterator: no iteration in progress for this configuration (call cgi_begin)");
        return MCB_ERR_STATE;
    }
    if (u->cgi_l > h->M) {
        u->cgi_l = 0;
        *l_out = 0;
        return 0;
    }
    MCB_TRY(cgi_step(h));
    for (int w = 0; w < 3; ++w) {
        if (physical) { // _greens! (DQMC.jl:721-730)
            MCB_TRY(launch_gemm(base_gemm(h, u->out[w], false, h->hop[0], true, u->cX), h->st));
            MCB_TRY(launch_gemm(base_gemm(h, h->hop[1], true, u->cX, false, u->phys[w]), h->st));
            results[w] = u->phys[w];
        } else {
            results[w] = u->out[w];
        }
    }
    for (int w = 0; w < 3; ++w) u->last[w] = results[w];
    *l_out = u->cgi_l++;
    return 0;
}

int engine_cgi_last(mcb_dqmc* h, const double* results[3])
{
    UtStack* u = h->ut;
    if (!u || !u->last[0] || u->cgi_version != h->conf_version) {
        set_error("CombinedGreensIterator: no step has been taken for this configuration");
        return MCB_ERR_STATE;
    }
    for (int w = 0; w < 3; ++w) results[w] = u->last[w];
    return 0;
}

void unequal_free(mcb_dqmc* h)
{
    if (!h->ut) return;
    for (void* p : h->ut->all) cudaFree(p);
    delete h->ut;
    h->ut = nullptr;
}

} // namespace mcb
