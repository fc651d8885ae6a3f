// extern "C" entry points of the DQMC engine and of the batched linear algebra (include/mcb200.h).
#include "dqmc_engine.cuh"
#include <math.h>
#include <mutex>
#include <string.h>
#include <stdlib.h>

namespace mcb {
static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
} // namespace mcb

using namespace mcb;

#define API_TRY(expr)                        \
    do {                                     \
        const int rc__ = (expr);             \
        if (rc__ != 0) return rc__;          \
    } while (0)
#define REQUIRE(cond, msg)                   \
    do {                                     \
        if (!(cond)) { set_error(msg); return MCB_ERR_ARG; } \
    } while (0)

static int use_device(int dev)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        set_error("no CUDA device: libmcb200 has no CPU fallback");
        return MCB_ERR_CUDA;
    }
    MCB_CUDA(cudaSetDevice(dev));
    static bool inited[64] = {};
    static std::mutex init_mutex;   // handles may be created from several host threads
    std::lock_guard<std::mutex> lock(init_mutex);
    if (dev >= 0 && dev < 64 && !inited[dev]) {
        API_TRY(init_gemm_kernels());
        API_TRY(init_udt_kernels());
        API_TRY(init_sweep_kernels());
        API_TRY(init_sweep_delayed_kernels());
        API_TRY(init_segment_kernels());
        API_TRY(init_segment_resident_kernels());
        API_TRY(init_stab_kernels());
        inited[dev] = true;
    }
    return 0;
}

namespace {
struct ProfBind {
    explicit ProfBind(mcb_dqmc* h) { set_current_profiler(&h->prof); }
    ~ProfBind() { set_current_profiler(nullptr); }
};

// compress / decompress (HubbardModel.jl:56-59): a configuration travels as BitArray(conf .== 1), i.e. bit b of
// 64-bit chunk b >> 6 is spin b = i + N * l (Julia column-major), stored at the same linear index here.
__global__ void unpack_conf_kernel(int8_t* conf, const unsigned long long* bits, size_t per_chain, size_t chunks,
                                   size_t total)
{
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const size_t chain = e / per_chain, b = e % per_chain;
        conf[e] = ((bits[chain * chunks + (b >> 6)] >> (b & 63)) & 1ull) ? 1 : -1;
    }
}
__global__ void pack_conf_kernel(unsigned long long* bits, const int8_t* conf, size_t per_chain, size_t chunks,
                                 size_t total_chunks)
{
    for (size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x; c < total_chunks; c += (size_t)gridDim.x * blockDim.x) {
        const size_t chain = c / chunks, k = c % chunks;
        unsigned long long w = 0;
        for (int j = 0; j < 64; ++j) {
            const size_t b = k * 64 + j;
            if (b < per_chain && conf[chain * per_chain + b] == 1) w |= 1ull << j;
        }
        bits[c] = w;
    }
}
} // namespace

extern "C" {

const char* mcb_last_error(void) { return g_error.c_str(); }
const char* mcb_version(void) { return "0.1 sm_100a"; }

int mcb_dqmc_create(const mcb_dqmc_params* p, mcb_dqmc** out)
{
    REQUIRE(p && out, "null argument");
    REQUIRE(p->n_sites > 0 && p->n_slices > 0 && p->safe_mult > 0 && p->n_chains > 0, "non-positive size");
    REQUIRE(p->n_slices % p->safe_mult == 0,
            "n_slices must be a multiple of safe_mult (stack.jl:115)");
    REQUIRE(p->U >= 0.0, "U must be positive (HubbardModelAttractive.jl:31)");
    REQUIRE(p->n_chains <= 65535, "at most 65535 chains per handle");
    API_TRY(use_device(p->device_id));
    mcb_dqmc* h = new mcb_dqmc();
    h->p = *p;
    h->N = p->n_sites; h->M = p->n_slices; h->s = p->safe_mult; h->C = p->n_chains;
    h->ne = h->M / h->s + 1;
    h->use_panel = segment_supported(h->N) && stab_supported(h->N) && p->general_path == 0;
    {
        // MCB_SEGMENT_RESIDENT=0: the GEMM + delayed-sweep launches per slice instead (comparisons)
        const char* ev = getenv("MCB_SEGMENT_RESIDENT");
        h->use_resident_segment = !h->use_panel && segment_resident_supported(h->N) && p->general_path == 0 &&
                                  !(ev && atoi(ev) == 0);
    }
    h->nn = (size_t)h->N * h->N;
    h->mat = h->nn * h->C;
    h->vec = (size_t)h->N * h->C;
    h->lambda = acosh(exp(0.5 * p->U * p->delta_tau)); // HubbardModelAttractive.jl:103
    h->eVp[0] = exp(1.0 * h->lambda * 1.0);  h->eVp[1] = exp(1.0 * h->lambda * -1.0);
    h->eVm[0] = exp(-1.0 * h->lambda * 1.0); h->eVm[1] = exp(-1.0 * h->lambda * -1.0);
    if (cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking) != cudaSuccess) {
        set_error("cudaStreamCreate failed");
        delete h;
        return MCB_ERR_CUDA;
    }
    if (cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess) {
        set_error("cudaEventCreate failed");
        cudaStreamDestroy(h->st);
        delete h;
        return MCB_ERR_CUDA;
    }
    const int rc = engine_alloc(h);
    if (rc != 0) { engine_free(h); cudaStreamDestroy(h->st); delete h; return rc; }
    MCB_CUDA(cudaStreamSynchronize(h->st));
    *out = h;
    return MCB_OK;
}

int mcb_dqmc_destroy(mcb_dqmc* h)
{
    if (!h) return MCB_OK;
    cudaSetDevice(h->p.device_id);
    cudaStreamSynchronize(h->st);
    h->prof.collect();
    engine_free(h);
    cudaEventDestroy(h->ev0);
    cudaEventDestroy(h->ev1);
    cudaStreamDestroy(h->st);
    delete h;
    return MCB_OK;
}

__global__ void transpose_kernel(double* dst, const double* src, int n)
{
    const size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (e < (size_t)n * n) dst[e] = src[(e % n) * n + e / n];
}

int mcb_dqmc_set_hopping(mcb_dqmc* h, const double* eT_half_minus, const double* eT_half_plus,
                         const double* eT_minus, const double* eT_plus)
{
    REQUIRE(h && eT_half_minus && eT_half_plus && eT_minus && eT_plus, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    const double* src[4] = {eT_half_minus, eT_half_plus, eT_minus, eT_plus};
    for (int i = 0; i < 4; ++i)
        MCB_CUDA(cudaMemcpyAsync(h->hop[i], src[i], h->nn * sizeof(double), cudaMemcpyHostToDevice, h->st));
    if (h->use_resident_segment || gemm_tma_routes_abt(h->N, h->C)) {   // transposes of e^{-+dtau T} (streamed as columns)
        for (int i = 0; i < 2; ++i) {
            if (!h->hopT[i]) MCB_CUDA(cudaMalloc(&h->hopT[i], h->nn * sizeof(double)));
            transpose_kernel<<<(unsigned)((h->nn + 255) / 256), 256, 0, h->st>>>(h->hopT[i], h->hop[2 + i], h->N);
            MCB_CUDA(cudaGetLastError());
        }
    }
    MCB_CUDA(cudaStreamSynchronize(h->st));
    h->hopping_set = true;
    return MCB_OK;
}

int mcb_dqmc_set_conf(mcb_dqmc* h, const int8_t* conf)
{
    REQUIRE(h && conf, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(h->conf, conf, (size_t)h->C * h->M * h->N, cudaMemcpyHostToDevice, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    h->conf_set = true;
    h->stack_built = false;
    h->conf_version++;
    return MCB_OK;
}

int mcb_dqmc_get_conf(mcb_dqmc* h, int8_t* conf)
{
    REQUIRE(h && conf, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(conf, h->conf, (size_t)h->C * h->M * h->N, cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_build_stack(mcb_dqmc* h)
{
    REQUIRE(h, "null handle");
    if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    API_TRY(engine_build_stack(h));
    API_TRY(engine_propagate(h, false, nullptr));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    h->stack_built = true;
    return MCB_OK;
}

int mcb_dqmc_resume(mcb_dqmc* h, int64_t last_sweep)
{
    REQUIRE(h && last_sweep >= 0, "bad argument");
    API_TRY(mcb_dqmc_build_stack(h));
    h->n_updates = last_sweep * 2 * (int64_t)h->M; // Philox counters and the sweep index carry on
    return MCB_OK;
}

int mcb_dqmc_update(mcb_dqmc* h, int64_t n_updates, int32_t measure_every)
{
    REQUIRE(h && n_updates >= 0, "bad argument");
    if (!h->stack_built) { set_error("build_stack first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    MCB_CUDA(cudaEventRecord(h->ev0, h->st));
    for (int64_t left = n_updates; left > 0;) {
        int64_t done = 0;
        API_TRY(engine_update_batch(h, left, measure_every > 0, &done));
        left -= done;
        if (measure_every > 0 && h->current_slice == 1 && h->direction == 1) {
            const int64_t sweep = (h->n_updates - 1) / (2 * (int64_t)h->M) + 1; // 1-based like DQMC.jl:420
            if (sweep % measure_every == 0) {
                API_TRY(engine_measure(h));
                if (h->td_scheduled) API_TRY(engine_measure_td(h, h->td_recalculate));
            }
        }
    }
    MCB_CUDA(cudaEventRecord(h->ev1, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    MCB_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    return MCB_OK;
}

int mcb_dqmc_sweep(mcb_dqmc* h, int64_t n_sweeps, int32_t measure_every)
{
    REQUIRE(h && n_sweeps >= 0, "bad argument");
    return mcb_dqmc_update(h, n_sweeps * 2 * (int64_t)h->M, measure_every);
}

int mcb_dqmc_get_greens(mcb_dqmc* h, double* G, int32_t physical)
{
    REQUIRE(h && G, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    const double* src = h->G;
    ProfBind bind(h);
    if (physical) { API_TRY(engine_physical_greens(h, h->W[1])); src = h->W[1]; }
    MCB_CUDA(cudaMemcpyAsync(G, src, h->mat * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_calculate_greens_at(mcb_dqmc* h, int32_t slice, double* G)
{
    REQUIRE(h && G && slice >= 0 && slice <= h->M, "bad argument");
    if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    API_TRY(engine_calculate_greens_at(h, slice));
    MCB_CUDA(cudaMemcpyAsync(G, h->G, h->mat * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_greens_unequal(mcb_dqmc* h, int32_t slice1, int32_t slice2, double* G, int32_t physical)
{
    REQUIRE(h && G && slice1 >= 0 && slice1 <= h->M && slice2 >= 0 && slice2 <= h->M, "bad argument");
    if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    API_TRY(engine_greens_unequal(h, slice1, slice2));
    const double* src = unequal_result(h);
    if (physical) { // _greens! (DQMC.jl:721-730): e^{+dtau T/2} G e^{-dtau T/2}
        API_TRY(launch_gemm(base_gemm(h, src, false, h->hop[0], true, h->W[0]), h->st));
        API_TRY(launch_gemm(base_gemm(h, h->hop[1], true, h->W[0], false, h->W[1]), h->st));
        src = h->W[1];
    }
    MCB_CUDA(cudaMemcpyAsync(G, src, h->mat * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_cgi_begin(mcb_dqmc* h, int32_t recalculate)
{
    REQUIRE(h, "null handle");
    if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    API_TRY(engine_cgi_begin(h, recalculate));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_cgi_next(mcb_dqmc* h, int32_t physical, int32_t* slice, double* G0l, double* Gl0, double* Gll)
{
    REQUIRE(h && slice, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    int l = 0;
    const double* res[3] = {nullptr, nullptr, nullptr};
    API_TRY(engine_cgi_next(h, physical, &l, res));
    *slice = l;
    if (l > 0) {
        double* dst[3] = {G0l, Gl0, Gll};
        for (int w = 0; w < 3; ++w)
            if (dst[w]) MCB_CUDA(cudaMemcpyAsync(dst[w], res[w], h->mat * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    }
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_cgi_device_results(mcb_dqmc* h, const double** G0l, const double** Gl0, const double** Gll)
{
    REQUIRE(h && G0l && Gl0 && Gll, "null argument");
    const double* res[3];
    if (engine_cgi_last(h, res) != 0) return MCB_ERR_STATE;
    *G0l = res[0]; *Gl0 = res[1]; *Gll = res[2];
    return MCB_OK;
}

int mcb_dqmc_conf_chunks(mcb_dqmc* h) { return h ? (int)(((size_t)h->N * h->M + 63) / 64) : 0; }

static int upload_bits(mcb_dqmc* h, const uint64_t* chunks)
{
    const size_t per = (size_t)h->N * h->M, nch = (per + 63) / 64;
    if (!h->conf_bits) MCB_CUDA(cudaMalloc(&h->conf_bits, (size_t)h->C * nch * sizeof(uint64_t)));
    MCB_CUDA(cudaMemcpyAsync(h->conf_bits, chunks, (size_t)h->C * nch * sizeof(uint64_t), cudaMemcpyHostToDevice, h->st));
    unpack_conf_kernel<<<296, 256, 0, h->st>>>(h->conf, (const unsigned long long*)h->conf_bits, per, nch, (size_t)h->C * per);
    MCB_CUDA(cudaGetLastError());
    h->conf_set = true;
    h->stack_built = false;
    h->conf_version++;
    return 0;
}

int mcb_dqmc_set_conf_bits(mcb_dqmc* h, const uint64_t* chunks)
{
    REQUIRE(h && chunks, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    API_TRY(upload_bits(h, chunks));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_get_conf_bits(mcb_dqmc* h, uint64_t* chunks)
{
    REQUIRE(h && chunks, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    const size_t per = (size_t)h->N * h->M, nch = (per + 63) / 64;
    if (!h->conf_bits) MCB_CUDA(cudaMalloc(&h->conf_bits, (size_t)h->C * nch * sizeof(uint64_t)));
    pack_conf_kernel<<<296, 256, 0, h->st>>>((unsigned long long*)h->conf_bits, h->conf, per, nch, (size_t)h->C * nch);
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaMemcpyAsync(chunks, h->conf_bits, (size_t)h->C * nch * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_replay(mcb_dqmc* h, const uint64_t* chunks, int64_t n_confs)
{
    REQUIRE(h && chunks && n_confs > 0, "bad argument");
    REQUIRE(n_confs % h->C == 0, "replay: the number of configurations must be a multiple of n_chains");
    if (!h->hopping_set) { set_error("set_hopping first"); return MCB_ERR_STATE; }
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    const size_t nch = ((size_t)h->N * h->M + 63) / 64;
    for (int64_t b = 0; b < n_confs / h->C; ++b) {
        API_TRY(upload_bits(h, chunks + (size_t)b * h->C * nch));
        API_TRY(engine_calculate_greens_at(h, 0));
        API_TRY(engine_measure(h));
        if (h->td_scheduled) API_TRY(engine_measure_td(h, h->td_recalculate, true));
        MCB_CUDA(cudaStreamSynchronize(h->st)); // the staging buffer is reused by the next batch
    }
    return MCB_OK;
}

int mcb_dqmc_get_state(mcb_dqmc* h, mcb_dqmc_state* st)
{
    REQUIRE(h && st, "null argument");
    st->current_slice = h->current_slice;
    st->direction = h->direction;
    st->n_updates = h->n_updates;
    st->n_measurements = h->ms.n_meas;
    return MCB_OK;
}

int mcb_dqmc_get_stats(mcb_dqmc* h, mcb_dqmc_chain_stats* out)
{
    REQUIRE(h && out, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    std::vector<long long> acc(h->C), prop(h->C), nf(h->C);
    std::vector<double> pe(h->C);
    MCB_CUDA(cudaMemcpyAsync(acc.data(), h->acc, h->C * sizeof(long long), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaMemcpyAsync(prop.data(), h->prop, h->C * sizeof(long long), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaMemcpyAsync(nf.data(), h->nonfinite, h->C * sizeof(long long), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaMemcpyAsync(pe.data(), h->prop_err, h->C * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    for (int c = 0; c < h->C; ++c) {
        out[c].prop_local = prop[c];
        out[c].acc_local = acc[c];
        out[c].max_propagation_error = pe[c];
        out[c].n_negative = 0;
        out[c].n_nonfinite = nf[c];
    }
    return MCB_OK;
}

int mcb_dqmc_set_measurement_tables(mcb_dqmc* h, const int32_t* trg_of, int32_t n_dirs, int32_t K, const double* hop)
{
    REQUIRE(h && trg_of && hop && n_dirs > 0 && K >= 0 && K <= n_dirs, "bad argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    for (size_t e = 0; e < (size_t)n_dirs * h->N; ++e) REQUIRE(trg_of[e] >= -1 && trg_of[e] < h->N, "target site out of range");
    return measure_setup(h->ms, h->N, h->C, 1, trg_of, n_dirs, K, hop, -h->p.U, h->st);
}

int mcb_dqmc_reset_observables(mcb_dqmc* h)
{
    REQUIRE(h, "null handle");
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    return measure_reset(h->ms, h->st);
}

int mcb_dqmc_measure(mcb_dqmc* h)
{
    REQUIRE(h, "null handle");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    API_TRY(engine_measure(h));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_measure_susceptibilities(mcb_dqmc* h, int32_t recalculate)
{
    REQUIRE(h, "null handle");
    if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    ProfBind bind(h);
    API_TRY(engine_measure_td(h, recalculate));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_dqmc_schedule_susceptibilities(mcb_dqmc* h, int32_t enable, int32_t recalculate)
{
    REQUIRE(h, "null handle");
    h->td_scheduled = enable != 0;
    h->td_recalculate = recalculate;
    return MCB_OK;
}

int32_t mcb_dqmc_susceptibility_len(mcb_dqmc* h) { return (h && h->ms.ready()) ? h->ms.td_len() : 0; }

int mcb_dqmc_get_susceptibilities(mcb_dqmc* h, double* sums, int64_t* count)
{
    REQUIRE(h && sums && count, "null argument");
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(sums, h->ms.td, (size_t)h->C * h->ms.td_len() * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    *count = h->ms.n_td;
    return MCB_OK;
}

int mcb_dqmc_profile(mcb_dqmc* h, int32_t enable_timing, int32_t reset)
{
    REQUIRE(h, "null handle");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    if (reset) h->prof.reset();
    h->prof.timing = enable_timing != 0;
    return MCB_OK;
}

int mcb_dqmc_get_profile(mcb_dqmc* h, int64_t* launches, double* ms, int64_t* timed, double* flops)
{
    REQUIRE(h && launches && ms && timed && flops, "null argument");
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    h->prof.collect();
    for (int i = 0; i < LC_COUNT; ++i) {
        launches[i] = h->prof.launches[i];
        ms[i] = h->prof.ms[i];
        timed[i] = h->prof.timed[i];
        flops[i] = h->prof.flops[i];
    }
    return MCB_OK;
}

double mcb_dqmc_last_elapsed_ms(mcb_dqmc* h) { return h ? (double)h->last_ms : -1.0; }

int32_t mcb_dqmc_observable_len(mcb_dqmc* h) { return h ? h->ms.obs_len() : 0; }

int mcb_dqmc_get_observables(mcb_dqmc* h, double* sums, int64_t* count)
{
    REQUIRE(h && sums && count, "null argument");
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(sums, h->ms.obs, (size_t)h->C * h->ms.obs_len() * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    *count = h->ms.n_meas;
    return MCB_OK;
}

int mcb_dqmc_get_greens_sum(mcb_dqmc* h, double* Gsum)
{
    REQUIRE(h && Gsum, "null argument");
    if (!h->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(Gsum, h->ms.Gsum, h->mat * sizeof(double), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

} // extern "C"
