// extern "C" entry points of DQMC{HubbardModelRepulsive} (include/mcb200.h).
//
// The reference keeps the two spin sectors as the two blocks of BlockDiagonal matrices
// (src/models/HubbardModel/HubbardModelRepulsive.jl:72-124, src/linalg/blockdiagonal.jl:156-428): every stack,
// wrap and UDT operation acts on the blocks independently; the sectors meet only in propose_local
// (:126-154), where the determinant ratio is the product of the two diagonal 2x2 entries, and accept_local!
// (:156-232) updates each block with its own rank-1 term. The spin-down block sees exp(-sign(power) lambda s),
// i.e. it is the attractive model's slice matrix on the NEGATED field. Here a repulsive run therefore is a
// pair of ordinary engine handles on one stream - `up` on the field s, `dn` on -s - that propagate
// separately (all stabilisation kernels are reused as they are) and share one sweep_spatial launch
// (sweep_delayed.cu: a cluster of two CTAs per chain, one per sector).
#include "dqmc_engine.cuh"
#include <math.h>
#include <vector>

using namespace mcb;

struct mcb_dqmc_rep {
    mcb_dqmc* up = nullptr;
    mcb_dqmc* dn = nullptr;
    cudaStream_t dn_stream = nullptr; // the stream dn was created with (unused: dn runs on up's stream)
    std::vector<int8_t> neg;          // host staging for the negated field
};

#define API_TRY(expr)                        \
    do {                                     \
        const int rc__ = (expr);             \
        if (rc__ != 0) return rc__;          \
    } while (0)
#define REQUIRE(cond, msg)                   \
    do {                                     \
        if (!(cond)) { set_error(msg); return MCB_ERR_ARG; } \
    } while (0)

namespace {
struct ProfBind {
    explicit ProfBind(mcb_dqmc* h) { set_current_profiler(&h->prof); }
    ~ProfBind() { set_current_profiler(nullptr); }
};

// update (DQMC.jl:523-538) of the pair: propagate both sectors, one joint sweep_spatial
int rep_update(mcb_dqmc_rep* r)
{
    mcb_dqmc* u = r->up;
    mcb_dqmc* d = r->dn;
    u->conf_version++; d->conf_version++;
    API_TRY(engine_propagate(u, false, nullptr));
    API_TRY(engine_propagate(d, false, nullptr));
    SweepArgs a;
    a.n = u->N; a.batch = u->C;
    a.G = u->G; a.G2 = d->G; a.sG = (long long)u->nn;
    a.conf = u->conf + (size_t)(u->current_slice - 1) * u->N;
    a.conf2 = d->conf + (size_t)(d->current_slice - 1) * d->N;
    a.sConf = (long long)u->M * u->N;
    // Delta = exp(-+2 alpha s) - 1 (HubbardModelRepulsive.jl:136-140); no bosonic factor in the weight
    const double dEp = -2.0 * u->lambda * 1.0, dEm = -2.0 * u->lambda * -1.0;
    a.gamma_p = exp(dEp) - 1.0; a.efac_p = 1.0;
    a.gamma_m = exp(dEm) - 1.0; a.efac_m = 1.0;
    a.key0 = seed_key0(u->p.seed);
    a.chain_offset = u->p.chain_offset;
    a.sweep = (uint32_t)(u->n_updates / (2 * (int64_t)u->M));
    a.visit = (uint32_t)(u->n_updates % (2 * (int64_t)u->M));
    a.acc = u->acc; a.prop = u->prop; a.nonfinite = u->nonfinite;
    // N < 16: immediate rank-1 updates in shared memory (sweep.cu); otherwise the delayed-update cluster kernel
    if (u->N < 16) API_TRY(launch_sweep_spatial(a, u->st));
    else API_TRY(launch_sweep_delayed(a, u->st));
    u->n_updates += 1;
    d->n_updates += 1;
    return 0;
}
} // namespace

extern "C" {

int mcb_dqmc_rep_create(const mcb_dqmc_params* p, mcb_dqmc_rep** out)
{
    REQUIRE(p && out, "null argument");
    mcb_dqmc_rep* r = new mcb_dqmc_rep();
    int rc = mcb_dqmc_create(p, &r->up);
    if (rc == MCB_OK) rc = mcb_dqmc_create(p, &r->dn);
    if (rc != MCB_OK) {
        if (r->up) mcb_dqmc_destroy(r->up);
        delete r;
        return rc;
    }
    // one stream for the pair: the joint sweep needs both propagations
    r->dn_stream = r->dn->st;
    r->dn->st = r->up->st;
    r->neg.resize((size_t)p->n_chains * p->n_slices * p->n_sites);
    *out = r;
    return MCB_OK;
}

int mcb_dqmc_rep_destroy(mcb_dqmc_rep* r)
{
    if (!r) return MCB_OK;
    cudaSetDevice(r->up->p.device_id);
    cudaStreamSynchronize(r->up->st);
    r->dn->st = r->dn_stream;
    mcb_dqmc_destroy(r->dn);
    mcb_dqmc_destroy(r->up);
    delete r;
    return MCB_OK;
}

int mcb_dqmc_rep_set_hopping(mcb_dqmc_rep* r, const double* eT_half_minus, const double* eT_half_plus,
                             const double* eT_minus, const double* eT_plus)
{
    REQUIRE(r, "null handle");
    // hopping_matrix returns BlockDiagonal(T, copy(T)) (HubbardModelRepulsive.jl:84-100): same block twice
    API_TRY(mcb_dqmc_set_hopping(r->up, eT_half_minus, eT_half_plus, eT_minus, eT_plus));
    return mcb_dqmc_set_hopping(r->dn, eT_half_minus, eT_half_plus, eT_minus, eT_plus);
}

int mcb_dqmc_rep_set_conf(mcb_dqmc_rep* r, const int8_t* conf)
{
    REQUIRE(r && conf, "null argument");
    for (size_t e = 0; e < r->neg.size(); ++e) r->neg[e] = (int8_t)-conf[e];
    API_TRY(mcb_dqmc_set_conf(r->up, conf));
    return mcb_dqmc_set_conf(r->dn, r->neg.data());
}

int mcb_dqmc_rep_get_conf(mcb_dqmc_rep* r, int8_t* conf)
{
    REQUIRE(r && conf, "null argument");
    return mcb_dqmc_get_conf(r->up, conf);
}

int mcb_dqmc_rep_build_stack(mcb_dqmc_rep* r)
{
    REQUIRE(r, "null handle");
    API_TRY(mcb_dqmc_build_stack(r->up));
    return mcb_dqmc_build_stack(r->dn);
}

// apply!(::Greens, ...) for the two-block handle: physical G of both sectors, one measurement launch
static int rep_measure(mcb_dqmc_rep* r)
{
    mcb_dqmc* u = r->up;
    mcb_dqmc* d = r->dn;
    if (!u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    API_TRY(engine_physical_greens(u, u->W[1]));
    API_TRY(engine_physical_greens(d, d->W[1]));
    return measure_accumulate(u->ms, u->W[1], d->W[1], u->st);
}

// apply!(::CombinedGreensIterator, ...) (measurements/generic.jl:227-243) for the two-block handle: charge / spin density,
// pairing and current-current susceptibility from the iterators of both blocks, advanced in step (measure.cu:
// measure_td_kernel<true>). Enqueues only.
static int rep_measure_td(mcb_dqmc_rep* r, int recalculate, bool from_g0 = false)
{
    mcb_dqmc* u = r->up;
    mcb_dqmc* d = r->dn;
    if (!u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    if (u->ms.flavors != 2) { set_error("one-flavor tables on a two-flavor handle"); return MCB_ERR_STATE; }
    for (mcb_dqmc* h : {u, d}) {
        if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
        // replay!: G = G(0) of the recorded configuration, no sweep state to check
        if (from_g0) API_TRY(engine_cgi_begin_from_g0(h, recalculate));
        else API_TRY(engine_cgi_begin(h, recalculate));
    }
    double* g00[2];
    API_TRY(measure_td_g00(u->ms, g00));
    API_TRY(engine_physical_greens(u, g00[0]));
    API_TRY(engine_physical_greens(d, g00[1]));
    const double factor = u->p.delta_tau / (double)u->N;
    for (;;) {
        int lu = 0, ld = 0;
        const double* ru[3] = {nullptr, nullptr, nullptr};
        const double* rd[3] = {nullptr, nullptr, nullptr};
        API_TRY(engine_cgi_next(u, 1, &lu, ru));
        API_TRY(engine_cgi_next(d, 1, &ld, rd));
        if (lu != ld) { set_error("CombinedGreensIterator: the two blocks are out of step"); return MCB_ERR_STATE; }
        if (lu == 0) break;
        const double* const mu[4] = {g00[0], ru[0], ru[1], ru[2]};
        const double* const md[4] = {g00[1], rd[0], rd[1], rd[2]};
        API_TRY(measure_td_accumulate(u->ms, mu, md, factor, u->st));
    }
    u->ms.n_td += 1;
    return MCB_OK;
}

int mcb_dqmc_rep_update(mcb_dqmc_rep* r, int64_t n_updates, int32_t measure_every)
{
    REQUIRE(r && n_updates >= 0, "bad argument");
    mcb_dqmc* u = r->up;
    if (!u->stack_built || !r->dn->stack_built) { set_error("build_stack first"); return MCB_ERR_STATE; }
    if (measure_every > 0 && !u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(u->p.device_id));
    ProfBind bind(u);
    MCB_CUDA(cudaEventRecord(u->ev0, u->st));
    for (int64_t k = 0; k < n_updates; ++k) {
        API_TRY(rep_update(r));
        if (measure_every > 0 && u->current_slice == 1 && u->direction == 1) { // DQMC.jl:429-436
            const int64_t sweep = (u->n_updates - 1) / (2 * (int64_t)u->M) + 1;
            if (sweep % measure_every == 0) {
                API_TRY(rep_measure(r));
                if (u->td_scheduled) API_TRY(rep_measure_td(r, u->td_recalculate));
            }
        }
    }
    MCB_CUDA(cudaEventRecord(u->ev1, u->st));
    MCB_CUDA(cudaStreamSynchronize(u->st));
    MCB_CUDA(cudaEventElapsedTime(&u->last_ms, u->ev0, u->ev1));
    return MCB_OK;
}

int mcb_dqmc_rep_sweep(mcb_dqmc_rep* r, int64_t n_sweeps, int32_t measure_every)
{
    REQUIRE(r && n_sweeps >= 0, "bad argument");
    return mcb_dqmc_rep_update(r, n_sweeps * 2 * (int64_t)r->up->M, measure_every);
}

int mcb_dqmc_rep_set_measurement_tables(mcb_dqmc_rep* r, const int32_t* trg_of, int32_t n_dirs, int32_t K, const double* hop)
{
    REQUIRE(r && trg_of && hop && n_dirs > 0 && K >= 0 && K <= n_dirs, "bad argument");
    mcb_dqmc* u = r->up;
    MCB_CUDA(cudaSetDevice(u->p.device_id));
    for (size_t e = 0; e < (size_t)n_dirs * u->N; ++e) REQUIRE(trg_of[e] >= -1 && trg_of[e] < u->N, "target site out of range");
    return measure_setup(u->ms, u->N, u->C, 2, trg_of, n_dirs, K, hop, +u->p.U, u->st);
}

int mcb_dqmc_rep_measure(mcb_dqmc_rep* r)
{
    REQUIRE(r, "null handle");
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    API_TRY(rep_measure(r));
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

int mcb_dqmc_rep_reset_observables(mcb_dqmc_rep* r)
{
    REQUIRE(r, "null handle");
    if (!r->up->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    return measure_reset(r->up->ms, r->up->st);
}

int32_t mcb_dqmc_rep_observable_len(mcb_dqmc_rep* r) { return r ? r->up->ms.obs_len() : 0; }

int mcb_dqmc_rep_get_observables(mcb_dqmc_rep* r, double* sums, int64_t* count)
{
    REQUIRE(r && sums && count, "null argument");
    mcb_dqmc* u = r->up;
    if (!u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(u->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(sums, u->ms.obs, (size_t)u->C * u->ms.obs_len() * sizeof(double), cudaMemcpyDeviceToHost, u->st));
    MCB_CUDA(cudaStreamSynchronize(u->st));
    *count = u->ms.n_meas;
    return MCB_OK;
}

int mcb_dqmc_rep_get_greens_sum(mcb_dqmc_rep* r, double* Gsum)
{
    REQUIRE(r && Gsum, "null argument");
    mcb_dqmc* u = r->up;
    if (!u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(u->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(Gsum, u->ms.Gsum, 2 * u->mat * sizeof(double), cudaMemcpyDeviceToHost, u->st));
    MCB_CUDA(cudaStreamSynchronize(u->st));
    return MCB_OK;
}

// replay!(mc, configurations) measurement loop (DQMC.jl:649-660) on both blocks
int mcb_dqmc_rep_replay(mcb_dqmc_rep* r, const uint64_t* chunks, int64_t n_confs)
{
    REQUIRE(r && chunks && n_confs > 0, "bad argument");
    mcb_dqmc* u = r->up;
    REQUIRE(n_confs % u->C == 0, "replay: the number of configurations must be a multiple of n_chains");
    if (!u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    const size_t nch = (size_t)mcb_dqmc_conf_chunks(u);
    for (int64_t b = 0; b < n_confs / u->C; ++b) {
        API_TRY(mcb_dqmc_rep_set_conf_bits(r, chunks + (size_t)b * u->C * nch));
        MCB_CUDA(cudaSetDevice(u->p.device_id));
        ProfBind bind(u);
        API_TRY(engine_calculate_greens_at(u, 0));
        API_TRY(engine_calculate_greens_at(r->dn, 0));
        API_TRY(rep_measure(r));
        if (u->td_scheduled) API_TRY(rep_measure_td(r, u->td_recalculate, true));
        MCB_CUDA(cudaStreamSynchronize(u->st));
    }
    return MCB_OK;
}

int mcb_dqmc_rep_get_greens(mcb_dqmc_rep* r, double* G, int32_t physical)
{
    REQUIRE(r && G, "null argument");
    mcb_dqmc* hs[2] = {r->up, r->dn};
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    const size_t nn = r->up->nn;
    for (int f = 0; f < 2; ++f) {
        mcb_dqmc* h = hs[f];
        const double* src = h->G;
        if (physical) { API_TRY(engine_physical_greens(h, h->W[1])); src = h->W[1]; }
        // [chain][sector][col][row]: the two diagonal blocks of the reference's BlockDiagonal G
        MCB_CUDA(cudaMemcpy2DAsync(G + f * nn, 2 * nn * sizeof(double), src, nn * sizeof(double), nn * sizeof(double),
                                   h->C, cudaMemcpyDeviceToHost, h->st));
    }
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

int mcb_dqmc_rep_calculate_greens_at(mcb_dqmc_rep* r, int32_t slice, double* G)
{
    REQUIRE(r && G, "null argument");
    mcb_dqmc* hs[2] = {r->up, r->dn};
    const size_t nn = r->up->nn;
    REQUIRE(slice >= 0 && slice <= r->up->M, "bad slice");
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    for (int f = 0; f < 2; ++f) {
        mcb_dqmc* h = hs[f];
        if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
        API_TRY(engine_calculate_greens_at(h, slice));
        MCB_CUDA(cudaMemcpy2DAsync(G + f * nn, 2 * nn * sizeof(double), h->G, nn * sizeof(double), nn * sizeof(double),
                                   h->C, cudaMemcpyDeviceToHost, h->st));
    }
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

int mcb_dqmc_rep_get_state(mcb_dqmc_rep* r, mcb_dqmc_state* st)
{
    REQUIRE(r, "null handle");
    return mcb_dqmc_get_state(r->up, st);
}

int mcb_dqmc_rep_get_stats(mcb_dqmc_rep* r, mcb_dqmc_chain_stats* out)
{
    REQUIRE(r && out, "null argument");
    mcb_dqmc* u = r->up;
    API_TRY(mcb_dqmc_get_stats(u, out));
    // propagation error and non-finite flags: the worse of the two sectors
    std::vector<mcb_dqmc_chain_stats> d(u->C);
    API_TRY(mcb_dqmc_get_stats(r->dn, d.data()));
    for (int c = 0; c < u->C; ++c) {
        if (d[c].max_propagation_error > out[c].max_propagation_error)
            out[c].max_propagation_error = d[c].max_propagation_error;
        out[c].n_nonfinite += d[c].n_nonfinite;
    }
    return MCB_OK;
}

// greens(mc, k, l) = G(k <- l) of both blocks (unequal_time_stack.jl:275-305 on BlockDiagonal matrices)
int mcb_dqmc_rep_greens_unequal(mcb_dqmc_rep* r, int32_t slice1, int32_t slice2, double* G, int32_t physical)
{
    REQUIRE(r && G, "null argument");
    mcb_dqmc* hs[2] = {r->up, r->dn};
    const size_t nn = r->up->nn;
    REQUIRE(slice1 >= 0 && slice1 <= r->up->M && slice2 >= 0 && slice2 <= r->up->M, "bad slice");
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    for (int f = 0; f < 2; ++f) {
        mcb_dqmc* h = hs[f];
        if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
        API_TRY(engine_greens_unequal(h, slice1, slice2));
        const double* src = unequal_result(h);
        if (physical) { // _greens! (DQMC.jl:721-730): e^{+dtau T/2} G e^{-dtau T/2}
            API_TRY(launch_gemm(base_gemm(h, src, false, h->hop[0], true, h->W[0]), h->st));
            API_TRY(launch_gemm(base_gemm(h, h->hop[1], true, h->W[0], false, h->W[1]), h->st));
            src = h->W[1];
        }
        MCB_CUDA(cudaMemcpy2DAsync(G + f * nn, 2 * nn * sizeof(double), src, nn * sizeof(double), nn * sizeof(double),
                                   h->C, cudaMemcpyDeviceToHost, h->st));
    }
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

// CombinedGreensIterator on both blocks (unequal_time_stack.jl:746-885)
int mcb_dqmc_rep_cgi_begin(mcb_dqmc_rep* r, int32_t recalculate)
{
    REQUIRE(r, "null handle");
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    for (mcb_dqmc* h : {r->up, r->dn}) {
        if (!h->hopping_set || !h->conf_set) { set_error("set_hopping and set_conf first"); return MCB_ERR_STATE; }
        API_TRY(engine_cgi_begin(h, recalculate));
    }
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

int mcb_dqmc_rep_cgi_next(mcb_dqmc_rep* r, int32_t physical, int32_t* slice, double* G0l, double* Gl0, double* Gll)
{
    REQUIRE(r && slice, "null argument");
    mcb_dqmc* hs[2] = {r->up, r->dn};
    const size_t nn = r->up->nn;
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    double* dst[3] = {G0l, Gl0, Gll};
    int ls[2] = {0, 0};
    for (int f = 0; f < 2; ++f) {
        mcb_dqmc* h = hs[f];
        const double* res[3] = {nullptr, nullptr, nullptr};
        API_TRY(engine_cgi_next(h, physical, &ls[f], res));
        if (ls[f] == 0) continue;
        for (int w = 0; w < 3; ++w)
            if (dst[w])
                MCB_CUDA(cudaMemcpy2DAsync(dst[w] + f * nn, 2 * nn * sizeof(double), res[w], nn * sizeof(double),
                                           nn * sizeof(double), h->C, cudaMemcpyDeviceToHost, h->st));
    }
    if (ls[0] != ls[1]) { set_error("CombinedGreensIterator: the two blocks are out of step"); return MCB_ERR_STATE; }
    *slice = ls[0];
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

int mcb_dqmc_rep_measure_susceptibilities(mcb_dqmc_rep* r, int32_t recalculate)
{
    REQUIRE(r, "null handle");
    MCB_CUDA(cudaSetDevice(r->up->p.device_id));
    ProfBind bind(r->up);
    API_TRY(rep_measure_td(r, recalculate));
    MCB_CUDA(cudaStreamSynchronize(r->up->st));
    return MCB_OK;
}

int mcb_dqmc_rep_schedule_susceptibilities(mcb_dqmc_rep* r, int32_t enable, int32_t recalculate)
{
    REQUIRE(r, "null handle");
    r->up->td_scheduled = enable != 0;
    r->up->td_recalculate = recalculate;
    return MCB_OK;
}

int32_t mcb_dqmc_rep_susceptibility_len(mcb_dqmc_rep* r) { return (r && r->up->ms.ready()) ? r->up->ms.td_len() : 0; }

int mcb_dqmc_rep_get_susceptibilities(mcb_dqmc_rep* r, double* sums, int64_t* count)
{
    REQUIRE(r && sums && count, "null argument");
    mcb_dqmc* u = r->up;
    if (!u->ms.ready()) { set_error("measurement tables not set"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(u->p.device_id));
    MCB_CUDA(cudaMemcpyAsync(sums, u->ms.td, (size_t)u->C * u->ms.td_len() * sizeof(double), cudaMemcpyDeviceToHost, u->st));
    MCB_CUDA(cudaStreamSynchronize(u->st));
    *count = u->ms.n_td;
    return MCB_OK;
}

// compress / decompress (HubbardModel.jl:56-59): the field travels as BitArray(conf .== 1); the spin-down stack
// lives on the negated field, i.e. on the complemented bits
int mcb_dqmc_rep_conf_chunks(mcb_dqmc_rep* r) { return r ? mcb_dqmc_conf_chunks(r->up) : 0; }

int mcb_dqmc_rep_get_conf_bits(mcb_dqmc_rep* r, uint64_t* chunks)
{
    REQUIRE(r && chunks, "null argument");
    return mcb_dqmc_get_conf_bits(r->up, chunks);
}

int mcb_dqmc_rep_set_conf_bits(mcb_dqmc_rep* r, const uint64_t* chunks)
{
    REQUIRE(r && chunks, "null argument");
    const size_t n = (size_t)r->up->C * mcb_dqmc_conf_chunks(r->up);
    std::vector<uint64_t> inv(n);
    for (size_t e = 0; e < n; ++e) inv[e] = ~chunks[e];
    API_TRY(mcb_dqmc_set_conf_bits(r->up, chunks));
    return mcb_dqmc_set_conf_bits(r->dn, inv.data());
}

// resume!(mc) (DQMC.jl:517-521): rebuild both stacks on the uploaded field, continue the counters
int mcb_dqmc_rep_resume(mcb_dqmc_rep* r, int64_t last_sweep)
{
    REQUIRE(r && last_sweep >= 0, "bad argument");
    API_TRY(mcb_dqmc_resume(r->up, last_sweep));
    return mcb_dqmc_resume(r->dn, last_sweep);
}

double mcb_dqmc_rep_last_elapsed_ms(mcb_dqmc_rep* r) { return r ? r->up->last_ms : 0.0; }

} // extern "C"
