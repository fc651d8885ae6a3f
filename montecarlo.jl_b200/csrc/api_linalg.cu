// Kernel-level entry points on host buffers (include/mcb200.h, "batched linear algebra").
#include "../../include/mcb200.h"
#include "dqmc_engine.cuh"
#include <vector>

using namespace mcb;

namespace {

struct DevBufs {
    std::vector<void*> ptrs;
    ~DevBufs() { for (void* p : ptrs) cudaFree(p); }
    template <typename T> T* alloc(size_t count)
    {
        void* p = nullptr;
        if (cudaMalloc(&p, count * sizeof(T)) != cudaSuccess) return nullptr;
        ptrs.push_back(p);
        return (T*)p;
    }
};

int prepare(int dev)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        set_error("no CUDA device: libmcb200 has no CPU fallback");
        return MCB_ERR_CUDA;
    }
    MCB_CUDA(cudaSetDevice(dev));
    if (init_gemm_kernels() != 0) return MCB_ERR_CUDA;
    if (init_udt_kernels() != 0) return MCB_ERR_CUDA;
    if (init_stab_kernels() != 0) return MCB_ERR_CUDA;
    return 0;
}

bool g_force_general = false;
bool use_panel(int n) { return !g_force_general && stab_supported(n); }

#define CHECK_ALLOC(p) do { if (!(p)) { set_error("cudaMalloc failed"); return MCB_ERR_CUDA; } } while (0)
#define L_TRY(expr) do { const int rc__ = (expr); if (rc__ != 0) return rc__; } while (0)

} // namespace

extern "C" {

int mcb_linalg_force_general(int32_t on)
{
    g_force_general = on != 0;
    set_force_general_udt(on != 0);
    return MCB_OK;
}

int mcb_linalg_vmul(int32_t n, int32_t batch, const double* A, const double* B, double* C, int32_t mode,
                    int32_t device_id)
{
    if (!A || !B || !C || n <= 0 || batch <= 0 || mode < 0 || mode > 2) { set_error("bad argument"); return MCB_ERR_ARG; }
    L_TRY(prepare(device_id));
    const size_t cnt = (size_t)n * n * batch;
    DevBufs d;
    double* dA = d.alloc<double>(cnt); double* dB = d.alloc<double>(cnt); double* dC = d.alloc<double>(cnt);
    CHECK_ALLOC(dA && dB && dC);
    MCB_CUDA(cudaMemcpy(dA, A, cnt * sizeof(double), cudaMemcpyHostToDevice));
    MCB_CUDA(cudaMemcpy(dB, B, cnt * sizeof(double), cudaMemcpyHostToDevice));
    GemmArgs g;
    g.n = n; g.batch = batch;
    g.A = dA; g.strideA = (long long)n * n; g.transA = (mode == 2);
    g.B = dB; g.strideB = (long long)n * n; g.transB = (mode == 1);
    g.C = dC; g.strideC = (long long)n * n;
    L_TRY(launch_gemm(g, 0));
    MCB_CUDA(cudaDeviceSynchronize());
    MCB_CUDA(cudaMemcpy(C, dC, cnt * sizeof(double), cudaMemcpyDeviceToHost));
    return MCB_OK;
}

int mcb_linalg_udt(int32_t n, int32_t batch, const double* X, double* U, double* D, double* T,
                   int32_t* pivot, int32_t apply_pivot, int32_t device_id)
{
    if (!X || !U || !D || !T || !pivot || n <= 0 || batch <= 0) { set_error("bad argument"); return MCB_ERR_ARG; }
    L_TRY(prepare(device_id));
    const size_t cnt = (size_t)n * n * batch, vcnt = (size_t)n * batch;
    DevBufs d;
    double* dX = d.alloc<double>(cnt); double* dU = d.alloc<double>(cnt); double* dT = d.alloc<double>(cnt);
    double* dD = d.alloc<double>(vcnt); int* dP = d.alloc<int>(vcnt);
    CHECK_ALLOC(dX && dU && dT && dD && dP);
    MCB_CUDA(cudaMemcpy(dX, X, cnt * sizeof(double), cudaMemcpyHostToDevice));
    if (use_panel(n)) L_TRY(launch_udt_panel(n, batch, dX, dU, dD, dT, dP, apply_pivot, 0));
    else
        L_TRY(launch_udt(n, batch, dX, (long long)n * n, dU, (long long)n * n, dD, n, dT, (long long)n * n, dP, n,
                         apply_pivot, 0));
    MCB_CUDA(cudaDeviceSynchronize());
    MCB_CUDA(cudaMemcpy(U, dU, cnt * sizeof(double), cudaMemcpyDeviceToHost));
    MCB_CUDA(cudaMemcpy(T, dT, cnt * sizeof(double), cudaMemcpyDeviceToHost));
    MCB_CUDA(cudaMemcpy(D, dD, vcnt * sizeof(double), cudaMemcpyDeviceToHost));
    MCB_CUDA(cudaMemcpy(pivot, dP, vcnt * sizeof(int), cudaMemcpyDeviceToHost));
    return MCB_OK;
}

int mcb_linalg_rdivp(int32_t n, int32_t batch, double* A, const double* T, const int32_t* pivot,
                     int32_t device_id)
{
    if (!A || !T || !pivot || n <= 0 || batch <= 0) { set_error("bad argument"); return MCB_ERR_ARG; }
    L_TRY(prepare(device_id));
    const size_t cnt = (size_t)n * n * batch, vcnt = (size_t)n * batch;
    for (size_t i = 0; i < vcnt; ++i)
        if (pivot[i] < 0 || pivot[i] >= n) { set_error("pivot out of range"); return MCB_ERR_ARG; }
    DevBufs d;
    double* dA = d.alloc<double>(cnt); double* dO = d.alloc<double>(cnt); double* dT = d.alloc<double>(cnt);
    int* dP = d.alloc<int>(vcnt);
    CHECK_ALLOC(dA && dO && dT && dP);
    MCB_CUDA(cudaMemcpy(dA, A, cnt * sizeof(double), cudaMemcpyHostToDevice));
    MCB_CUDA(cudaMemcpy(dT, T, cnt * sizeof(double), cudaMemcpyHostToDevice));
    MCB_CUDA(cudaMemcpy(dP, pivot, vcnt * sizeof(int), cudaMemcpyHostToDevice));
    L_TRY(launch_rdivp(n, batch, dA, (long long)n * n, dO, (long long)n * n, dT, (long long)n * n, dP, n, 0));
    MCB_CUDA(cudaDeviceSynchronize());
    MCB_CUDA(cudaMemcpy(A, dO, cnt * sizeof(double), cudaMemcpyDeviceToHost));
    return MCB_OK;
}

int mcb_linalg_calculate_greens(int32_t n, int32_t batch, const double* Ul, const double* Dl, const double* Tl,
                                const double* Ur, const double* Dr, const double* Tr, double* G,
                                int32_t device_id)
{
    if (!Ul || !Dl || !Tl || !Ur || !Dr || !Tr || !G || n <= 0 || batch <= 0) { set_error("bad argument"); return MCB_ERR_ARG; }
    L_TRY(prepare(device_id));
    const size_t cnt = (size_t)n * n * batch, vcnt = (size_t)n * batch;
    DevBufs d;
    double* m[5]; const double* hm[4] = {Ul, Tl, Ur, Tr};
    for (int i = 0; i < 5; ++i) { m[i] = d.alloc<double>(cnt); CHECK_ALLOC(m[i]); }
    double* v[2]; const double* hv[2] = {Dl, Dr};
    for (int i = 0; i < 2; ++i) { v[i] = d.alloc<double>(vcnt); CHECK_ALLOC(v[i]); }
    double* W[6]; for (int i = 0; i < 6; ++i) { W[i] = d.alloc<double>(cnt); CHECK_ALLOC(W[i]); }
    double* Dw[2]; for (int i = 0; i < 2; ++i) { Dw[i] = d.alloc<double>(vcnt); CHECK_ALLOC(Dw[i]); }
    int* piv = d.alloc<int>(vcnt); CHECK_ALLOC(piv);
    for (int i = 0; i < 4; ++i) MCB_CUDA(cudaMemcpy(m[i], hm[i], cnt * sizeof(double), cudaMemcpyHostToDevice));
    for (int i = 0; i < 2; ++i) MCB_CUDA(cudaMemcpy(v[i], hv[i], vcnt * sizeof(double), cudaMemcpyHostToDevice));
    L_TRY(calc_greens_raw(n, batch, m[0], v[0], m[1], m[2], v[1], m[3], m[4], W, Dw, piv, 0, !force_general_udt()));
    MCB_CUDA(cudaDeviceSynchronize());
    MCB_CUDA(cudaMemcpy(G, m[4], cnt * sizeof(double), cudaMemcpyDeviceToHost));
    return MCB_OK;
}

} // extern "C"
