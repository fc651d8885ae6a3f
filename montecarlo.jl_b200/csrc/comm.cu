// The one collective of a multi-GPU run behind the C ABI (dqmc_allreduce_observables): chains are
// independent Markov chains (src/flavors/DQMC/DQMC.jl:369-515 touches only its own mc), so ranks exchange nothing until
// the end of a run, when the moments of the per-chain observable means (sum, sum of squares, chain count) are summed
// over all ranks. A Julia host calls this instead of reducing in Julia; the Python host keeps torch.distributed
// (montecarlo.jl_b200/sharding.py) and uses this entry point only in its tests.
//
// NCCL is not a link-time dependency: libnccl.so.2 (or the file named by MCB_NCCL_LIB) is opened with dlopen on first
// use and five functions are looked up. The communicator is created from a 128-byte unique id that rank 0 obtains
// from mcb_comm_unique_id and hands to the other ranks by whatever means the host has (a file, MPI, a socket).
#include "dqmc_engine.cuh"
#include <dlfcn.h>
#include <mutex>
#include <stdlib.h>
#include <string.h>

namespace {

struct NcclId { char bytes[128]; };                      // ncclUniqueId
typedef int (*fn_get_unique_id)(NcclId*);
typedef int (*fn_comm_init_rank)(void**, int, NcclId, int);   // the id travels by value
typedef int (*fn_all_reduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_comm_destroy)(void*);
typedef const char* (*fn_error_string)(int);
constexpr int NCCL_FLOAT64 = 8, NCCL_SUM = 0;           // ncclDataType_t / ncclRedOp_t (nccl.h)

struct Nccl {
    void* lib = nullptr;
    fn_get_unique_id get_unique_id = nullptr;
    fn_comm_init_rank comm_init_rank = nullptr;
    fn_all_reduce all_reduce = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_error_string error_string = nullptr;
};
Nccl g_nccl;
std::mutex g_nccl_mutex;

int load_nccl()
{
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    if (g_nccl.lib) return 0;
    // an NCCL that the process has loaded already (a host with torch) is reused: two different libnccl.so.2 in one process
    // do not mix (the loader resolves the later one's symbols against the first)
    const char* path = getenv("MCB_NCCL_LIB");
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
    if (!lib) lib = dlopen(path ? path : "libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!lib) {
        mcb::set_error(std::string("NCCL not found (dlopen: ") + dlerror() + "); set MCB_NCCL_LIB to the path of libnccl.so.2");
        return MCB_ERR_STATE;
    }
    Nccl n;
    n.lib = lib;
    n.get_unique_id = (fn_get_unique_id)dlsym(lib, "ncclGetUniqueId");
    n.comm_init_rank = (fn_comm_init_rank)dlsym(lib, "ncclCommInitRank");
    n.all_reduce = (fn_all_reduce)dlsym(lib, "ncclAllReduce");
    n.comm_destroy = (fn_comm_destroy)dlsym(lib, "ncclCommDestroy");
    n.error_string = (fn_error_string)dlsym(lib, "ncclGetErrorString");
    if (!n.get_unique_id || !n.comm_init_rank || !n.all_reduce || !n.comm_destroy || !n.error_string) {
        mcb::set_error("NCCL library lacks one of ncclGetUniqueId / ncclCommInitRank / ncclAllReduce / ncclCommDestroy");
        dlclose(lib);
        return MCB_ERR_STATE;
    }
    g_nccl = n;
    return 0;
}

int nccl_check(int rc, const char* what)
{
    if (rc == 0) return 0;
    mcb::set_error(std::string(what) + ": " + g_nccl.error_string(rc));
    return MCB_ERR_CUDA;
}

// per-chain means -> this rank's moments: out[e] = sum_c x_ce, out[len + e] = sum_c x_ce^2 with x_ce = sums[c][e] / n_meas;
// one thread per entry, chains in a fixed order (bitwise reproducible)
__global__ void moments_kernel(const double* __restrict__ sums, int n_chains, int len, double inv_meas, double* __restrict__ out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= len) return;
    double s = 0.0, s2 = 0.0;
    for (int c = 0; c < n_chains; ++c) {
        const double x = sums[(size_t)c * len + e] * inv_meas;
        s += x;
        s2 = fma(x, x, s2);
    }
    out[e] = s;
    out[len + e] = s2;
}

} // namespace

struct mcb_comm {
    void* comm = nullptr;
    int n_ranks = 1, rank = 0, device = 0;
    double* buf = nullptr;     // device staging of the payload
    size_t buf_len = 0;
};

extern "C" {

int mcb_comm_unique_id(void* id)
{
    if (!id) { mcb::set_error("null argument"); return MCB_ERR_ARG; }
    if (load_nccl()) return MCB_ERR_STATE;
    NcclId nid;
    if (nccl_check(g_nccl.get_unique_id(&nid), "ncclGetUniqueId")) return MCB_ERR_CUDA;
    memcpy(id, nid.bytes, sizeof(nid.bytes));
    return MCB_OK;
}

int mcb_comm_create(const void* id, int32_t n_ranks, int32_t rank, int32_t device_id, mcb_comm** out)
{
    if (!id || !out || n_ranks < 1 || rank < 0 || rank >= n_ranks) { mcb::set_error("bad argument"); return MCB_ERR_ARG; }
    if (load_nccl()) return MCB_ERR_STATE;
    MCB_CUDA(cudaSetDevice(device_id));
    NcclId nid;
    memcpy(nid.bytes, id, sizeof(nid.bytes));
    mcb_comm* c = new mcb_comm();
    c->n_ranks = n_ranks; c->rank = rank; c->device = device_id;
    if (nccl_check(g_nccl.comm_init_rank(&c->comm, n_ranks, nid, rank), "ncclCommInitRank")) { delete c; return MCB_ERR_CUDA; }
    *out = c;
    return MCB_OK;
}

int mcb_comm_destroy(mcb_comm* c)
{
    if (!c) return MCB_OK;
    cudaSetDevice(c->device);
    if (c->buf) cudaFree(c->buf);
    if (c->comm) g_nccl.comm_destroy(c->comm);
    delete c;
    return MCB_OK;
}

int32_t mcb_dqmc_moments_len(mcb_dqmc* h)
{
    if (!h || !h->ms.ready()) return 0;
    return 2 * (h->ms.obs_len() + h->ms.flavors * (int32_t)h->nn) + 1;
}

// moments, length mcb_dqmc_moments_len(h) = 2 (R + F N^2) + 1 with R = mcb_dqmc_observable_len, F = flavors:
//   [ sum_c x_c (R) | sum_c x_c^2 (R) | sum_c G_c (F N^2) | sum_c G_c^2 (F N^2) | number of chains ]
// summed over the chains of ALL ranks of the communicator; x_c, G_c = the per-chain means over this run's measurements.
// A rank that has measured nothing yet contributes zeros and a chain count of zero. comm == NULL: this rank alone.
int mcb_dqmc_allreduce_observables(mcb_dqmc* h, mcb_comm* c, double* moments)
{
    if (!h || !moments) { mcb::set_error("null argument"); return MCB_ERR_ARG; }
    if (!h->ms.ready()) { mcb::set_error("measurement tables not set"); return MCB_ERR_STATE; }
    if (c && c->device != h->p.device_id) { mcb::set_error("communicator and handle live on different devices"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    const int R = h->ms.obs_len(), GN = h->ms.flavors * (int)h->nn;
    const size_t len = (size_t)2 * (R + GN) + 1;
    double* buf = nullptr;
    double* own = nullptr;
    if (c) {
        if (c->buf_len < len) {
            if (c->buf) cudaFree(c->buf);
            c->buf = nullptr; c->buf_len = 0;
            MCB_CUDA(cudaMalloc(&c->buf, len * sizeof(double)));
            c->buf_len = len;
        }
        buf = c->buf;
    } else {
        MCB_CUDA(cudaMalloc(&own, len * sizeof(double)));
        buf = own;
    }
    const bool measured = h->ms.n_meas > 0;
    const double inv = measured ? 1.0 / (double)h->ms.n_meas : 0.0;
    MCB_CUDA(cudaMemsetAsync(buf, 0, len * sizeof(double), h->st));
    if (measured) {
        moments_kernel<<<(R + 127) / 128, 128, 0, h->st>>>(h->ms.obs, h->C, R, inv, buf);
        // Gsum is [C][F][N*N]: one record of F N^2 entries per chain
        moments_kernel<<<(GN + 127) / 128, 128, 0, h->st>>>(h->ms.Gsum, h->C, GN, inv, buf + 2 * R);
        MCB_CUDA(cudaGetLastError());
        const double count = (double)h->C;
        MCB_CUDA(cudaMemcpyAsync(buf + len - 1, &count, sizeof(double), cudaMemcpyHostToDevice, h->st));
    }
    int rc = MCB_OK;
    if (c && c->n_ranks > 1)
        rc = nccl_check(g_nccl.all_reduce(buf, buf, len, NCCL_FLOAT64, NCCL_SUM, c->comm, h->st), "ncclAllReduce");
    if (rc == MCB_OK) {
        cudaError_t e = cudaMemcpyAsync(moments, buf, len * sizeof(double), cudaMemcpyDeviceToHost, h->st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->st);
        if (e != cudaSuccess) { mcb::set_error(cudaGetErrorString(e)); rc = MCB_ERR_CUDA; }
    }
    if (own) cudaFree(own);
    return rc;
}

// The same reduction for the time-displaced records (mcb_dqmc_get_susceptibilities): moments, length 2 L + 1 with
// L = mcb_dqmc_susceptibility_len(h): [ sum_c x_c (L) | sum_c x_c^2 (L) | number of chains ], x_c = chain c's mean over the
// time-displaced measurements taken so far; a rank without such measurements contributes zeros and zero chains.
int mcb_dqmc_allreduce_susceptibilities(mcb_dqmc* h, mcb_comm* c, double* moments)
{
    if (!h || !moments) { mcb::set_error("null argument"); return MCB_ERR_ARG; }
    if (!h->ms.ready()) { mcb::set_error("measurement tables not set"); return MCB_ERR_STATE; }
    if (c && c->device != h->p.device_id) { mcb::set_error("communicator and handle live on different devices"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    const int L = h->ms.td_len();
    const size_t len = (size_t)2 * L + 1;
    double* buf = nullptr;
    double* own = nullptr;
    if (c) {
        if (c->buf_len < len) {
            if (c->buf) cudaFree(c->buf);
            c->buf = nullptr; c->buf_len = 0;
            MCB_CUDA(cudaMalloc(&c->buf, len * sizeof(double)));
            c->buf_len = len;
        }
        buf = c->buf;
    } else {
        MCB_CUDA(cudaMalloc(&own, len * sizeof(double)));
        buf = own;
    }
    const bool measured = h->ms.n_td > 0;
    MCB_CUDA(cudaMemsetAsync(buf, 0, len * sizeof(double), h->st));
    if (measured && L > 0) {
        moments_kernel<<<(L + 127) / 128, 128, 0, h->st>>>(h->ms.td, h->C, L, 1.0 / (double)h->ms.n_td, buf);
        MCB_CUDA(cudaGetLastError());
        const double count = (double)h->C;
        MCB_CUDA(cudaMemcpyAsync(buf + len - 1, &count, sizeof(double), cudaMemcpyHostToDevice, h->st));
    }
    int rc = MCB_OK;
    if (c && c->n_ranks > 1)
        rc = nccl_check(g_nccl.all_reduce(buf, buf, len, NCCL_FLOAT64, NCCL_SUM, c->comm, h->st), "ncclAllReduce");
    if (rc == MCB_OK) {
        cudaError_t e = cudaMemcpyAsync(moments, buf, len * sizeof(double), cudaMemcpyDeviceToHost, h->st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->st);
        if (e != cudaSuccess) { mcb::set_error(cudaGetErrorString(e)); rc = MCB_ERR_CUDA; }
    }
    if (own) cudaFree(own);
    return rc;
}

} // extern "C"
