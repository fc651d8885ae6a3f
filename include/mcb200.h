/* mcb200.h — C ABI of the B200-native DQMC / Ising sweep engine (libmcb200.so, sm_100a).
 *
 * The reference (ffreyer/MonteCarlo.jl) is pure Julia and has no FFI of its own; its
 * "plugin API" is multiple dispatch on the flavor/model types. The entry points below are
 * what a Julia host binds with `ccall` to replace the hot path of
 *   DQMC{HubbardModelAttractive, CheckerboardFalse}  and  MC{IsingModel{SquareLattice}}
 * (see INTEGRATION.md for the Julia-side methods). Every function cites the reference
 * function(s) it replaces, paths relative to the reference repository root.
 *
 * Conventions
 *  - all pointers are HOST pointers owned by the caller and only valid during the call;
 *    the library owns all device memory (freed by *_destroy)
 *  - matrices are column-major (Julia layout): G[i,j] at i + j*N; batched arrays put the
 *    chain / replica index slowest
 *  - return value 0 = ok, negative = error; mcb_last_error() gives a thread-local message
 *  - a handle is bound to one device and one stream and is not thread-safe; calls are
 *    synchronous (they return after the stream has drained)
 *  - there is no CPU fallback: without a CUDA device *_create fails with MCB_ERR_CUDA
 */
#ifndef MCB200_H
#define MCB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCB_OK 0
#define MCB_ERR_ARG (-1)
#define MCB_ERR_CUDA (-2)
#define MCB_ERR_STATE (-3)

const char* mcb_last_error(void);
/* "major.minor sm_100a" */
const char* mcb_version(void);

/* ------------------------------------------------------------------ DQMC ---- */

typedef struct mcb_dqmc mcb_dqmc;

/* Mirrors DQMCParameters (src/flavors/DQMC/DQMC.jl:52-125) + model constants
 * (src/models/HubbardModel/HubbardModelAttractive.jl:27-40). */
typedef struct {
    int32_t n_sites;           /* N = length(lattice) */
    int32_t n_slices;          /* M = beta / delta_tau ; must be a multiple of safe_mult (stack.jl:115) */
    int32_t safe_mult;         /* DQMC.jl:79 */
    int32_t n_chains;          /* independent Markov chains in this handle */
    double delta_tau;          /* DQMC.jl:85 */
    double U;                  /* |U| of the attractive model */
    uint64_t seed;             /* Philox key word 0 */
    uint32_t chain_offset;     /* global id of chain 0 (Philox key word 1 = chain_offset + c) */
    int32_t device_id;
    int32_t check_propagation; /* DQMCParameters.check_propagation_error (stack.jl:540-550) */
    int32_t general_path;      /* 0 = library default (chain-resident kernels where N allows);
                                  1 = force the general any-N kernels (testing / comparison) */
} mcb_dqmc_params;

typedef struct {
    int64_t current_slice; /* mc.s.current_slice (stack.jl) */
    int64_t direction;     /* mc.s.direction */
    int64_t n_updates;     /* update() calls so far; sweep = n_updates / 2M */
    int64_t n_measurements;
} mcb_dqmc_state;

/* per chain: DQMCAnalysis (DQMC.jl:36-47) */
typedef struct {
    int64_t prop_local;
    int64_t acc_local;
    double max_propagation_error;
    int64_t n_negative; /* proposals with detratio < 0 (DQMC.jl:562-568); always 0 for this model */
    int64_t n_nonfinite; /* updates after which G contained NaN/Inf */
} mcb_dqmc_chain_stats;

/* DQMC(model; kwargs...) constructor + initialize_stack (DQMC.jl:250-289, stack.jl:108-159) */
int mcb_dqmc_create(const mcb_dqmc_params* p, mcb_dqmc** out);
int mcb_dqmc_destroy(mcb_dqmc* h);

/* init_hopping_matrix_exp (stack.jl:167-181): the four N x N matrices
 * exp(-dtau T/2), exp(+dtau T/2), exp(-dtau T/2)^2, exp(+dtau T/2)^2 computed by the host
 * exactly as the reference does, shared by all chains. */
int mcb_dqmc_set_hopping(mcb_dqmc* h, const double* eT_half_minus, const double* eT_half_plus,
                         const double* eT_minus, const double* eT_plus);

/* mc.conf (HubbardModel.jl:4): int8 +-1, layout [chain][slice][site] */
int mcb_dqmc_set_conf(mcb_dqmc* h, const int8_t* conf);
int mcb_dqmc_get_conf(mcb_dqmc* h, int8_t* conf);

/* run!'s prologue: build_stack(mc, mc.s); propagate(mc)  (DQMC.jl:412-414, stack.jl:242-255).
 * Resets the update counter's slice position to (M, -1); G = G_M afterwards. */
int mcb_dqmc_build_stack(mcb_dqmc* h);

/* n x update(mc, i) = propagate + sweep_spatial  (DQMC.jl:523-582, stack.jl:502-631).
 * If measure_every > 0 the built-in observables are accumulated after every update that ends at
 * current_slice == 1 && direction == +1 and whose sweep index is a multiple of measure_every
 * (DQMC.jl:429-436). */
/* resume!(mc) (src/flavors/DQMC/DQMC.jl:517-521, FileIO.jl:38-156): continue a run whose configuration was saved
 * after `last_sweep` complete sweeps: build_stack + propagate on the uploaded configuration (as run! does after
 * loading), with the update counter (= random-number counters, measurement schedule) set to last_sweep * 2M.
 * The checkpoint itself is {conf (mcb_dqmc_get_conf / _get_conf_bits), last_sweep, seed, chain_offset, parameters}. */
int mcb_dqmc_resume(mcb_dqmc* h, int64_t last_sweep);
int mcb_dqmc_update(mcb_dqmc* h, int64_t n_updates, int32_t measure_every);
/* n_sweeps * 2M updates (the body of run!'s sweep loop, DQMC.jl:420-438) */
int mcb_dqmc_sweep(mcb_dqmc* h, int64_t n_sweeps, int32_t measure_every);

/* mc.s.greens (physical == 0) or greens(mc) = e^{+dtau T/2} G e^{-dtau T/2} (DQMC.jl:711-730);
 * layout [chain][col][row] */
int mcb_dqmc_get_greens(mcb_dqmc* h, double* G, int32_t physical);
/* calculate_greens(mc, slice) from scratch, stack untouched (stack.jl:422-480);
 * overwrites the handle's current G like the reference does. */
int mcb_dqmc_calculate_greens_at(mcb_dqmc* h, int32_t slice, double* G);
/* greens(mc, k, l): unequal-time G(k <- l) = <c_i(k dtau) c_j^+(l dtau)> of the current configuration,
 * 0 <= k, l <= n_slices, through the UnequalTimeStack (src/flavors/DQMC/unequal_time_stack.jl:275-305, 443-622).
 * physical != 0 applies the same transform as mcb_dqmc_get_greens. Does not disturb the sweep state; the
 * stacks are rebuilt only if the configuration changed since the last call. Layout [chain][col][row]. */
int mcb_dqmc_greens_unequal(mcb_dqmc* h, int32_t slice1, int32_t slice2, double* G, int32_t physical);

/* The one collective of a multi-GPU run (dqmc_allreduce_observables): the reference runs one
 * independent Markov chain per process (src/flavors/DQMC/DQMC.jl:369-515) and combines results afterwards; here the ranks
 * of a run sum the moments of the per-chain observable means at the end. NCCL is loaded at run time (dlopen of
 * libnccl.so.2, or of the file named by the environment variable MCB_NCCL_LIB); the library does not link against it.
 *   mcb_comm_unique_id   rank 0 obtains the 128-byte id and hands it to the other ranks (file, MPI, socket ...)
 *   mcb_comm_create      every rank, after cudaSetDevice-able device_id is known; collective over the n_ranks ranks
 *   mcb_dqmc_allreduce_observables   moments[mcb_dqmc_moments_len(h)] =
 *        [ sum_c x_c (R) | sum_c x_c^2 (R) | sum_c G_c (F N^2) | sum_c G_c^2 (F N^2) | number of chains ]
 *        over the chains c of ALL ranks, x_c / G_c = chain c's means over the measurements taken so far (record layout R =
 *        mcb_dqmc_observable_len as in mcb_dqmc_get_observables, G as in mcb_dqmc_get_greens_sum, F = 1). A rank without
 *        measurements contributes zeros and zero chains. comm == NULL: the moments of this rank alone, no NCCL needed.
 *        mean = sum / chains, standard error^2 = (sumsq / chains - mean^2) / (chains - 1). */
typedef struct mcb_comm mcb_comm;
#define MCB_UNIQUE_ID_BYTES 128
int mcb_comm_unique_id(void* id);
int mcb_comm_create(const void* id, int32_t n_ranks, int32_t rank, int32_t device_id, mcb_comm** out);
int mcb_comm_destroy(mcb_comm* c);
int32_t mcb_dqmc_moments_len(mcb_dqmc* h);
int mcb_dqmc_allreduce_observables(mcb_dqmc* h, mcb_comm* c, double* moments);
/* the same for the time-displaced records: moments[2 L + 1] = [ sum_c x_c | sum_c x_c^2 | number of chains ],
 * L = mcb_dqmc_susceptibility_len(h), x_c = chain c's mean record of mcb_dqmc_get_susceptibilities */
int mcb_dqmc_allreduce_susceptibilities(mcb_dqmc* h, mcb_comm* c, double* moments);

/* CombinedGreensIterator(mc, recalculate) (src/flavors/DQMC/unequal_time_stack.jl:746-885) on the device: yields, for
 * l = 1 .. n_slices, G(0 <- l), G(l <- 0) and G(l <- l) of the current configuration by advancing three UDT triples one
 * slice at a time (quick advance :862-885), re-factorising them every safe_mult slices (:824-861) and recomputing them from
 * the UnequalTimeStack every `recalculate` slices (:795-822; recalculate <= 0 selects the reference's default 4 safe_mult).
 * mcb_dqmc_cgi_begin needs the handle at its measurement point (current_slice == 1 on the way up, where mc.s.greens = G(0);
 * the reference asserts the slice at :770) and fails with MCB_ERR_STATE otherwise. mcb_dqmc_cgi_next advances one slice:
 * *slice receives l, or 0 once all n_slices steps have been produced; each non-NULL host pointer receives the matrix
 * ([chain][col][row], after _greens! when physical != 0). The results also stay on the device until the next step:
 * mcb_dqmc_cgi_device_results returns the three device pointers (same order: G0l, Gl0, Gll) for on-device consumers.
 * Sweeping or changing the configuration invalidates an iteration in progress. The sweep state is not disturbed. */
int mcb_dqmc_cgi_begin(mcb_dqmc* h, int32_t recalculate);
int mcb_dqmc_cgi_next(mcb_dqmc* h, int32_t physical, int32_t* slice, double* G0l, double* Gl0, double* Gll);
int mcb_dqmc_cgi_device_results(mcb_dqmc* h, const double** G0l, const double** Gl0, const double** Gll);

/* Recorded configurations in the reference's wire format compress(mc, model, conf) = BitArray(conf .== 1)
 * (src/models/HubbardModel/HubbardModel.jl:56-59, ConfigRecorder src/configurations.jl:24-43): per chain
 * mcb_dqmc_conf_chunks(h) = ceil(N*M/64) little-endian 64-bit chunks, bit b = spin i + N*l. */
int mcb_dqmc_conf_chunks(mcb_dqmc* h);
int mcb_dqmc_set_conf_bits(mcb_dqmc* h, const uint64_t* chunks);
int mcb_dqmc_get_conf_bits(mcb_dqmc* h, uint64_t* chunks);
/* The measurement loop of replay!(mc, configurations) (src/flavors/DQMC/DQMC.jl:649-660): for every recorded
 * configuration decompress, calculate_greens(mc, 0), measure. n_confs must be a multiple of n_chains (one
 * configuration per chain and batch); results accumulate in the observables of the handle. The sweep state is
 * invalid afterwards (build_stack again before update). */
int mcb_dqmc_replay(mcb_dqmc* h, const uint64_t* chunks, int64_t n_confs);

int mcb_dqmc_get_state(mcb_dqmc* h, mcb_dqmc_state* st);
int mcb_dqmc_get_stats(mcb_dqmc* h, mcb_dqmc_chain_stats* per_chain);

/* Observables: the reference's equal-time measurement kernels over its lattice iterators
 * (greens_kernel, occupation_kernel src/flavors/DQMC/measurements/measurements.jl:31-44; cdc / sdc / pc overloads
 * src/models/HubbardModel/HubbardModelAttractive.jl:221-245; apply! / finish! src/flavors/DQMC/measurements/generic.jl:279-347;
 * SURVEY.md 8a O1-O5, 8f rank 1).
 * trg_of: n_dirs * N int32, trg_of[dir*N + src] = target site (0-based) of the pair (src, trg) that
 * EachSitePairByDistance files under direction `dir` (src/lattices/lattice_iterators.jl:131-221), -1 if src has none;
 * K: EachLocalQuadByDistance{K} (lattice_iterators.jl:258-355) pairs every source with its targets in the first K
 * directions (0 <= K <= n_dirs; 5 = on-site + 4 nearest neighbours on the square lattice, measurements.jl:197-201);
 * hop: N x N hopping matrix T (kinetic energy). */
int mcb_dqmc_set_measurement_tables(mcb_dqmc* h, const int32_t* trg_of, int32_t n_dirs, int32_t K,
                                    const double* hop);
/* accumulate the observables of the current G now (what apply!(groups) does, generic.jl:194-243) */
int mcb_dqmc_measure(mcb_dqmc* h);
int mcb_dqmc_reset_observables(mcb_dqmc* h);
/* Time-displaced measurements: charge_density_susceptibility, spin_density_susceptibility, pairing_susceptibility
 * (src/flavors/DQMC/measurements/measurements.jl:56-58, 144-146, 207) through apply!(::CombinedGreensIterator, ...)
 * (src/flavors/DQMC/measurements/generic.jl:227-243): with G00 = greens!(mc), for every slice l = 1..M of the iterator the
 * kernels of HubbardModelAttractive for packed Green's functions (HubbardModelAttractive.jl:225-229, 239-241, 246-248) are
 * summed over the lattice iterators of mcb_dqmc_set_measurement_tables, then scaled by delta_tau / N (finish!, :272-286).
 * The handle has to be at its measurement point (as for mcb_dqmc_cgi_begin; `recalculate` as there); the sweep state is
 * left untouched. One call = one measurement; per chain the record [cds[n_dirs] | sds[n_dirs] | ps[n_dirs*K*K] | ccs[n_dirs*K]]
 * (ps column-major (dir12, dir1, dir2); sds = the x, y and z components, equal for this model; ccs =
 * current_current_susceptibility, measurements.jl:254-262, cc_kernel HubbardModelAttractive.jl:249-266 over
 * EachLocalQuadBySyncedDistance{K}, column-major (dir12, dir1), with the hopping matrix of the tables) accumulates as a SUM over
 * calls, reset by mcb_dqmc_reset_observables. (Two-flavor handles: mcb_dqmc_rep_measure_susceptibilities.) */
int mcb_dqmc_measure_susceptibilities(mcb_dqmc* h, int32_t recalculate);
/* enable != 0: every scheduled measurement of mcb_dqmc_update / mcb_dqmc_sweep (measure_every > 0) also takes one
 * time-displaced measurement, as run! does for a measurement group on the CombinedGreensIterator (DQMC.jl:423-438), and so
 * does every configuration of mcb_dqmc_replay (DQMC.jl:649-660) */
int mcb_dqmc_schedule_susceptibilities(mcb_dqmc* h, int32_t enable, int32_t recalculate);
int32_t mcb_dqmc_susceptibility_len(mcb_dqmc* h);
int mcb_dqmc_get_susceptibilities(mcb_dqmc* h, double* sums /*[chain][len]*/, int64_t* count);
/* Length of one chain's observable record, and the records: per chain the SUM over measurements of
 *   [ n = mean_i 2(1-G_ii), double occupancy mean_i (1-G_ii)^2, E_kin/N, E_int/N          
 *   | cdc[n_dirs] | sdc_x[n_dirs] (= sdc_y) | sdc_z[n_dirs]                                (outputs of EachSitePairByDistance, /N)
 *   | pc[n_dirs*K*K], column-major (dir12, dir1, dir2) like the reference's output array   (EachLocalQuadByDistance{K}, /N)
 *   | occupation[N] = 1 - G_ii ]                                                           (EachSiteAndFlavor)
 * followed by the number of measurements. Sums are formed in a fixed order: repeated runs agree bit for bit. */
int32_t mcb_dqmc_observable_len(mcb_dqmc* h);
int mcb_dqmc_get_observables(mcb_dqmc* h, double* sums /*[chain][len]*/, int64_t* count);
/* per chain SUM over measurements of G_phys, [chain][col][row] */
int mcb_dqmc_get_greens_sum(mcb_dqmc* h, double* Gsum);

/* Launch accounting and device timing (the role of the reference's @bm / TimerOutputs hooks,
 * src/helpers.jl:160-236). Launch classes, in order: gemm, udt, rdivp, sweep_spatial, fused
 * wrap+sweep, measure, misc, ising, chain-resident add_slice_sequence, chain-resident
 * calculate_greens  (MCB_N_LAUNCH_CLASSES entries per output array).
 * mcb_dqmc_profile: enable_timing != 0 brackets every launch with CUDA events on the handle's
 * stream; reset != 0 clears all counters. get_profile: launches since reset, event-timed ms and
 * number of timed launches, algorithmic FP64 flop per class. */
#define MCB_N_LAUNCH_CLASSES 10
int mcb_dqmc_profile(mcb_dqmc* h, int32_t enable_timing, int32_t reset);
int mcb_dqmc_get_profile(mcb_dqmc* h, int64_t* launches, double* ms, int64_t* timed, double* flops);
/* device time (CUDA events on the handle's stream) of the last mcb_dqmc_update / _sweep call */
double mcb_dqmc_last_elapsed_ms(mcb_dqmc* h);

/* --------------------------------------------------- batched linear algebra ---- *
 * Kernel-level entry points (the methods a custom matrix type overloads in the
 * reference, cf. src/linalg/blockdiagonal.jl:156-428). Host buffers, [batch][col][row].   */

/* Testing aid: on != 0 makes the entry points below use the general any-N kernels even where a
 * chain-resident (register/shared-memory) kernel exists for n; default 0. Process-wide. */
int mcb_linalg_force_general(int32_t on);

/* vmul!: mode 0 C=A*B, 1 C=A*B', 2 C=A'*B  (src/linalg/general.jl:7-45) */
int mcb_linalg_vmul(int32_t n, int32_t batch, const double* A, const double* B, double* C,
                    int32_t mode, int32_t device_id);
/* udt_AVX_pivot!(U, D, T, pivot, temp, Val(apply_pivot)) (src/linalg/UDT.jl:192-306);
 * X is the input, T the output; pivot is 0-based int32 */
int mcb_linalg_udt(int32_t n, int32_t batch, const double* X, double* U, double* D, double* T,
                   int32_t* pivot, int32_t apply_pivot, int32_t device_id);
/* rdivp!(A, T, O, pivot): A <- A[:,pivot] * triu(T)^-1 (src/linalg/general.jl:138-166) */
int mcb_linalg_rdivp(int32_t n, int32_t batch, double* A, const double* T, const int32_t* pivot,
                     int32_t device_id);
/* Allocation-free variant of mcb_linalg_udt. It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it). Unlike every other entry point it takes DEVICE pointers (the caller owns all buffers):
 * X_dev [batch][n*n] input, used as workspace and overwritten; U_dev, T_dev [batch][n*n]; D_dev [batch][n];
 * pivot_dev [batch][n] 0-based int32. `stream` is a cudaStream_t (NULL = default stream); the call only enqueues
 * kernels and returns without synchronising. It runs the blocked panel-pivoted factorisation (csrc/udt_blocked.cu),
 * same contract as mcb_linalg_udt: U D T = X, pivot order panel-wise.
 * Limits: n a multiple of 16 with 128 <= n <= 400, batch <= 8; anything larger returns MCB_ERR_ARG. The scratch
 * (448 000 bytes) is a static device array that every GPU that loads the library holds until exit; all calls share
 * it, so calls that overlap on the GPU (two streams, two host threads) race. After the first call per device (which
 * sets kernel attributes and looks up the array's address) it performs no cudaMalloc, cudaFree or other allocation. */
int mcb_linalg_udt_noalloc(int32_t n, int32_t batch, double* X_dev, double* U_dev, double* D_dev, double* T_dev,
                           int32_t* pivot_dev, int32_t apply_pivot, int32_t device_id, void* stream);
/* calculate_greens_AVX!(Ul,Dl,Tl,Ur,Dr,Tr,G) (src/flavors/DQMC/stack.jl:337-393); inputs are NOT
 * clobbered here */
int mcb_linalg_calculate_greens(int32_t n, int32_t batch, const double* Ul, const double* Dl,
                                const double* Tl, const double* Ur, const double* Dr,
                                const double* Tr, double* G, int32_t device_id);

/* ------------------------------------------------- DQMC{HubbardModelRepulsive} ---- */
/* SURVEY.md 8(f) rank 4. The reference keeps the two spin sectors as the blocks of BlockDiagonal matrices
 * (src/models/HubbardModel/HubbardModelRepulsive.jl:72-232, src/linalg/blockdiagonal.jl:156-428); here they are
 * two batched stacks that share one field and one sweep_spatial launch. Same parameter struct, same call
 * sequence and error codes as the mcb_dqmc_* functions above.
 * Green's functions travel as [chain][sector (up, down)][col][row], i.e. the two diagonal blocks of mc.s.greens. */
typedef struct mcb_dqmc_rep mcb_dqmc_rep;
int mcb_dqmc_rep_create(const mcb_dqmc_params* p, mcb_dqmc_rep** out);
int mcb_dqmc_rep_destroy(mcb_dqmc_rep* h);
/* hopping_matrix = BlockDiagonal(T, copy(T)) (HubbardModelRepulsive.jl:84-100): one block is passed */
int mcb_dqmc_rep_set_hopping(mcb_dqmc_rep* h, const double* eT_half_minus, const double* eT_half_plus,
                             const double* eT_minus, const double* eT_plus);
int mcb_dqmc_rep_set_conf(mcb_dqmc_rep* h, const int8_t* conf);
int mcb_dqmc_rep_get_conf(mcb_dqmc_rep* h, int8_t* conf);
int mcb_dqmc_rep_build_stack(mcb_dqmc_rep* h);
/* n x update(mc, i): propagate on both blocks, then sweep_spatial with propose_local / accept_local! of
 * HubbardModelRepulsive.jl:126-232 (detratio = R_up R_down, no bosonic factor, one rank-1 update per block) */
int mcb_dqmc_rep_update(mcb_dqmc_rep* h, int64_t n_updates, int32_t measure_every);
int mcb_dqmc_rep_sweep(mcb_dqmc_rep* h, int64_t n_sweeps, int32_t measure_every);
int mcb_dqmc_rep_get_greens(mcb_dqmc_rep* h, double* G, int32_t physical);
int mcb_dqmc_rep_calculate_greens_at(mcb_dqmc_rep* h, int32_t slice, double* G);
int mcb_dqmc_rep_get_state(mcb_dqmc_rep* h, mcb_dqmc_state* st);
int mcb_dqmc_rep_get_stats(mcb_dqmc_rep* h, mcb_dqmc_chain_stats* out);
/* greens(mc, k, l), the recorded-configuration wire format and resume!(mc) for the two-block handle: as the
 * mcb_dqmc_* functions of the same names (the spin-down stack sees the complemented bits) */
int mcb_dqmc_rep_greens_unequal(mcb_dqmc_rep* h, int32_t slice1, int32_t slice2, double* G, int32_t physical);
/* the CombinedGreensIterator on both blocks; host layout [chain][flavor][col][row] like mcb_dqmc_rep_greens_unequal */
int mcb_dqmc_rep_cgi_begin(mcb_dqmc_rep* h, int32_t recalculate);
int mcb_dqmc_rep_cgi_next(mcb_dqmc_rep* h, int32_t physical, int32_t* slice, double* G0l, double* Gl0, double* Gll);
/* the time-displaced measurements of mcb_dqmc_measure_susceptibilities with the reference's general (two-flavor) kernels for
 * packed Green's functions (src/flavors/DQMC/measurements/measurements.jl:75-90 cdc, :157-196 sdc_x/y/z, :215-219 pc,
 * :268-300 cc) on the BlockDiagonal matrices. Record per chain:
 * [cds[n_dirs] | sds_x[n_dirs] (= sds_y) | sds_z[n_dirs] | ps[n_dirs*K*K] | ccs[n_dirs*K]]. */
int mcb_dqmc_rep_measure_susceptibilities(mcb_dqmc_rep* h, int32_t recalculate);
/* as mcb_dqmc_schedule_susceptibilities: scheduled measurements of mcb_dqmc_rep_update / _sweep and every configuration of
 * mcb_dqmc_rep_replay take the time-displaced group along */
int mcb_dqmc_rep_schedule_susceptibilities(mcb_dqmc_rep* h, int32_t enable, int32_t recalculate);
int32_t mcb_dqmc_rep_susceptibility_len(mcb_dqmc_rep* h);
int mcb_dqmc_rep_get_susceptibilities(mcb_dqmc_rep* h, double* sums /*[chain][len]*/, int64_t* count);
int mcb_dqmc_rep_conf_chunks(mcb_dqmc_rep* h);
int mcb_dqmc_rep_get_conf_bits(mcb_dqmc_rep* h, uint64_t* chunks);
int mcb_dqmc_rep_set_conf_bits(mcb_dqmc_rep* h, const uint64_t* chunks);
int mcb_dqmc_rep_resume(mcb_dqmc_rep* h, int64_t last_sweep);
double mcb_dqmc_rep_last_elapsed_ms(mcb_dqmc_rep* h);
/* Observables of the two-flavor model: the general kernels of src/flavors/DQMC/measurements/measurements.jl:41-207
 * (occupation, cdc, mz, sdc_x/y/z, pc) on the BlockDiagonal G (its off-diagonal blocks vanish, so mx = my = 0 and
 * sdc_y = sdc_x). Tables as for mcb_dqmc_set_measurement_tables. Record per chain:
 *   [ n, double occupancy <n_up n_dn>, E_kin/N, E_int/N | cdc[n_dirs] | sdc_x[n_dirs] | sdc_z[n_dirs]
 *   | pc[n_dirs*K*K] | occupation[2N] (EachSiteAndFlavor: up sites, then down sites) | mz[N] ]
 * mcb_dqmc_rep_get_greens_sum: [chain][sector][col][row]. replay: as mcb_dqmc_replay. */
int mcb_dqmc_rep_set_measurement_tables(mcb_dqmc_rep* h, const int32_t* trg_of, int32_t n_dirs, int32_t K,
                                        const double* hop);
int mcb_dqmc_rep_measure(mcb_dqmc_rep* h);
int mcb_dqmc_rep_reset_observables(mcb_dqmc_rep* h);
int32_t mcb_dqmc_rep_observable_len(mcb_dqmc_rep* h);
int mcb_dqmc_rep_get_observables(mcb_dqmc_rep* h, double* sums, int64_t* count);
int mcb_dqmc_rep_get_greens_sum(mcb_dqmc_rep* h, double* Gsum);
int mcb_dqmc_rep_replay(mcb_dqmc_rep* h, const uint64_t* chunks, int64_t n_confs);

/* ----------------------------------------------------------------- Ising ---- */

typedef struct mcb_ising mcb_ising;

/* MC(IsingModel(L=L, dims=2); beta) (src/flavors/MC/MC.jl:50-80). L even. Checkerboard schedule
 * and Philox stream as specified in SURVEY.md Appendix D. */
int mcb_ising_create(int32_t L, double beta, int32_t n_replicas, uint64_t seed,
                     uint32_t replica_offset, int32_t device_id, mcb_ising** out);
int mcb_ising_destroy(mcb_ising* h);
/* mc.conf (IsingModel.jl:3-5): int8 +-1, layout [replica][y][x] (site index x + y*L) */
int mcb_ising_set_spins(mcb_ising* h, const int8_t* spins);
int mcb_ising_get_spins(mcb_ising* h, int8_t* spins);
/* The same configuration in the recorder's wire format compress = BitArray(conf .== 1) (the scheme of
 * src/models/HubbardModel/HubbardModel.jl:56-59 and ConfigRecorder src/configurations.jl:24-43 applied to IsingConf): per
 * replica mcb_ising_spin_chunks(h) = ceil(L*L/64) little-endian 64-bit chunks, bit s = site x + y*L. 1 bit per spin
 * over PCIe instead of 8. */
int mcb_ising_spin_chunks(mcb_ising* h);
int mcb_ising_set_spins_bits(mcb_ising* h, const uint64_t* chunks);
int mcb_ising_get_spins_bits(mcb_ising* h, uint64_t* chunks);
/* n x sweep(mc) (MC.jl:316-332) */
int mcb_ising_sweep(mcb_ising* h, int64_t n_sweeps);
/* energy(mc, m, conf) (IsingModel.jl:177-185) and sum(conf) (Ising/measurements.jl:72-78) */
int mcb_ising_measure(mcb_ising* h, int64_t* E, int64_t* Msum);
/* sweeps done so far (Philox counter word 0) */
int64_t mcb_ising_sweep_count(mcb_ising* h);
/* device time (CUDA events on the handle's stream) of the last mcb_ising_sweep call, and the number
 * of kernel launches since create */
double mcb_ising_last_elapsed_ms(mcb_ising* h);
int64_t mcb_ising_launch_count(mcb_ising* h);

#ifdef __cplusplus
}
#endif
#endif /* MCB200_H */
