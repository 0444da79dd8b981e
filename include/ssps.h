/*
 * ssps.h -- C ABI of libssps_b200.so: the B200-native replacement for the ONE hot path of
 * gitter-lab/ssps: scoring proposed parent sets (per-node DBN marginal likelihood + edge prior)
 * inside the Metropolis-Hastings sweep, batched over chains x vertices x proposals.
 *
 * What each entry point replaces in the reference (paths relative to the reference checkout):
 *
 *   ssps_set_data / ssps_set_timeseries
 *        combine_X / standardize_X / vectorize_X        SSPS/src/dbn_preprocess.jl:70-108
 *        compute_B_col / construct_B                    SSPS/src/dbn_distributions.jl:168-194
 *        prob_param_i, lambda_min                       SSPS/src/SSPS.jl:164-169
 *   ssps_score
 *        log_marg_lik / compute_lml / safe_solve        SSPS/src/dbn_distributions.jl:201-264
 *        logpdf(::ConfParentPrior, ...)                 SSPS/src/dbn_distributions.jl:96-153
 *        logpdf(::CPDMarginal, ...)                     SSPS/src/dbn_distributions.jl:303-306
 *   ssps_init_chains
 *        Gen.generate(vertex_lambda_dbn_model, ...)     SSPS/src/mcmc_inference.jl:40,
 *                                                       SSPS/src/dbn_models.jl:66-99,
 *        random(::ConfParentPrior, ...)                 SSPS/src/dbn_distributions.jl:75-89
 *        recover_state! (resume)                        SSPS/src/SSPS.jl:24-43,187-196
 *   ssps_run / ssps_replay
 *        mcmc_inference loop                            SSPS/src/mcmc_inference.jl:58-79
 *        vertex_lambda_update_loop / multithread_...    SSPS/src/state_updates.jl:35-93
 *        smart_proposal / smart_involution              SSPS/src/dbn_proposals.jl:60-130
 *        lambda_vec_proposal                            SSPS/src/dbn_proposals.jl:244-248
 *        Gen.mh acceptance (Gen 0.4.5, not in tree)     call sites state_updates.jl:44-47
 *        config.proposal = 1: uniform_update_loop       SSPS/src/state_updates.jl:98-112
 *        unpaired_edges / uniform_proposal / _involution SSPS/src/dbn_proposals.jl:142-230
 *   ssps_export_draws
 *        the stream of Gen.normal / rand() / categorical / uniform_discrete draws of the sweep
 *        (Gen 0.4.5 samplers; call sites SSPS/src/dbn_proposals.jl:67-77,246)
 *   ssps_edge_counts / ssps_edge_posterior
 *        increment_counts! / update_results_summary     SSPS/src/state_updates.jl:120-140,240-249
 *        evaluate_summary                               scripts/postprocess_samples.jl:183-191
 *   ssps_changepoints_*
 *        update_results_storediff                       SSPS/src/state_updates.jl:161-227
 *   ssps_psrf / ssps_psrf_sums / ssps_psrf_finish
 *        evaluate_convergence / compute_psrf_neff       scripts/postprocess_samples.jl:99-180
 *   ssps_convergence
 *        correlation_sum / compute_psrf_neff (N_eff)    scripts/postprocess_samples.jl:58-138
 *        evaluate_convergence / evaluate_summary        scripts/postprocess_samples.jl:155-191
 *   ssps_comm_* / ssps_run_multi / ssps_*_global
 *        pooling of per-chain JSON files                scripts/postprocess_samples.jl:242-259
 *   ssps_score_static
 *        no counterpart in the reference: ssps_score on device
 *        pointers with a fixed in-library scratch array, for callers that cannot allocate
 *
 * Conventions: all indices 0-based (the 1-based <-> 0-based shift lives in the language shim);
 * matrices column-major; the caller owns every buffer passed in or out (inputs are copied to the
 * device inside the call); every function returns 0 or a negative ssps_status and leaves a
 * message for ssps_last_error() (thread-local).  No exceptions cross this boundary.  A handle
 * is bound to one CUDA device; calls on one handle are not re-entrant.
 * There is NO CPU fallback: without a CUDA device every call fails with SSPS_ERR_CUDA.
 */
#ifndef SSPS_H
#define SSPS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSPS_ABI_VERSION 4

typedef enum {
    SSPS_OK = 0,
    SSPS_ERR_INVALID = -1,      /* bad argument / call order */
    SSPS_ERR_CUDA = -2,         /* CUDA runtime error (message has the cudaError string) */
    SSPS_ERR_UNSUPPORTED = -3,  /* e.g. a set too large for the interaction-column path, --proposal uniform */
    SSPS_ERR_NOMEM = -4,
    SSPS_ERR_INDEGREE = -5,     /* an INITIAL parent set exceeds max_indegree (ssps_init_chains); proposals beyond the bound
                                   are rejected, not errors */
    SSPS_ERR_UNAVAILABLE = -6   /* e.g. PSRF requested before any snapshot exists */
} ssps_status;

typedef struct ssps_handle ssps_handle;

typedef struct {
    int32_t struct_size;      /* = sizeof(ssps_config) */
    int32_t device;           /* CUDA ordinal; -1 = current device */
    int32_t regression_deg;   /* --regression-deg (SSPS.jl:75-78).  1: the tabulated-Gram path.  >= 2: interaction columns
                                 (construct_B), every proposal scored from scratch: one thread per set up to 7 design columns,
                                 the whole warp per set up to n - 2 columns, closed form from n - 1 columns on (any number of
                                 parents); any degree >= max_indegree is Hill's "full" (all non-empty product subsets).
                                 <= 0 (the reference CLI's default -1): the reference builds an EMPTY design -- flat
                                 likelihood -- reproduced as is (quirk Q2) */
    int32_t max_indegree;     /* bound on |S| (0 = min(V, 128), the most the factor arena supports).  A proposal that would
                                 grow a set beyond the bound is REJECTED (counted in n_indegree_rejects): the chain samples
                                 the posterior truncated at |S| <= max_indegree, never an invalid run.  The reference has
                                 no bound; with the default it is not reached unless V > 128 */
    double  large_indeg;      /* --large-indeg   (SSPS.jl:64-67)  */
    double  lambda_max;       /* --lambda-max    (SSPS.jl:68-70)  */
    double  lambda_prop_std;  /* --lambda-prop-std (SSPS.jl:75-78) */
    int32_t psrf_grid;        /* snapshot the per-chain tallies every this many recorded samples
                                 (0 = default 250, <0 = never) */
    int32_t record_changepoints; /* 1: keep the reference's change-point log (small runs only) */
    int32_t collect_khist;    /* 1: histogram of proposed |S'| (diagnostics; slower) */
    int32_t verify;           /* 1: additionally (a) score EVERY proposal from scratch (O(k^3)) and count proposals whose
                                 incrementally maintained factor gives a different score, (b) redo EVERY acceptance test
                                 (parent-set and lambda moves) in the reference order with fp64 logarithms and count
                                 decisions that differ from the screened test (diagnostics; much slower) */
    double  min_burnin;       /* smallest burn-in fraction later passed to ssps_psrf* / ssps_edge_posterior (0 = 0.5,
                                 the reference's default, postprocess_samples.jl:272): tally snapshots older than
                                 min_burnin * samples are recycled first when the snapshot pool is full */
    int32_t snapshot_slots;   /* size of the tally-snapshot pool, allocated ONCE by ssps_init_chains (0 = 32, clamped to
                                 a quarter of the free device memory, at least 4); ssps_run never allocates */
    int32_t proposal;         /* --proposal (SSPS.jl:87-88): 0 = "sparse", the parent-set proposal (smart_proposal, the default and what
                                 --multithread forces); 1 = "uniform", the deprecated uniform graph proposal (uniform_update_loop:
                                 add / remove a uniformly drawn pair or reverse an unpaired edge) -- one thread per chain, every
                                 changed set scored from scratch; regression_deg <= 1 only */
} ssps_config;

typedef struct {
    int64_t sweeps_done;        /* per chain */
    int64_t samples_recorded;   /* per chain, the reference's results["n"] */
    int64_t n_proposals;        /* parent-set proposals scored, all chains (lambda moves not counted) */
    int64_t n_accepted;         /* parent-set proposals accepted */
    int64_t n_lambda_accepted;
    int64_t sum_k;              /* sum over proposals of |S'| */
    double  algorithmic_flops;  /* sum over proposals of F(|S'|) = k^3/3 + 2k^2 + 2k */
    double  algorithmic_bytes;  /* sum over proposals of 8 (k(k+1)/2 + k) */
    double  kernel_ms;          /* device time of the MH kernel launches (CUDA events) */
    int64_t kernel_launches;    /* all kernels launched by this handle */
    int64_t mh_launches;        /* MH sweep kernel launches */
    /* diagnostics of the MH kernel (csrc/ssps_mh.cuh): */
    int64_t n_tier1;            /* parent-set tests whose fp32 screening was too close: redone in reference order, fp64 logs */
    int64_t n_lambda_exact;     /* lambda-move tests redone in reference order */
    int64_t n_big_steps;        /* proposals of lanes whose factor lives in the global-memory tier (|S| > 6) */
    int64_t big_cycles;         /* warp cycles spent in the cooperative global-memory tier (clock64, summed over warps) */
    int64_t n_mismatch;         /* verify = 1 only: proposals whose cached-factor score differs from the from-scratch
                                   score of the same parent list by even one bit (must be 0) */
    int64_t n_decision_mismatch;        /* verify = 1 only: parent-set moves whose screened accept/reject differs from the
                                           reference-order fp64 test log(u) < alpha (must be 0) */
    int64_t n_lambda_decision_mismatch; /* verify = 1 only: the same for lambda moves (must be 0) */
    double  margin_use_max;     /* verify = 1 only: max over screened parent-set tests of |screening error| / margin;
                                   a decision can only flip when this exceeds 1 */
    double  lambda_margin_use_max;      /* the same for lambda moves */
    int64_t n_indegree_rejects; /* proposals rejected because |S'| would exceed max_indegree (0 unless V > 128 or the
                                   caller lowered max_indegree) */
    int64_t n_mid_steps;        /* proposals of lanes in the lane-owned MID tier (SSPS_KS < |S| < SSPS_KM) */
    int64_t snapshot_slots;     /* size of the tally-snapshot pool of this handle */
    int64_t snapshots_live;     /* snapshots currently held */
    int64_t snapshot_grid;      /* current spacing of the snapshots in recorded samples (doubles when the pool fills) */
    int64_t device_allocs_in_run;   /* device allocations made inside ssps_run / ssps_replay since ssps_init_chains
                                       (change-point logging and replay uploads only; 0 on the Philox path) */
} ssps_run_stats;

const char* ssps_last_error(void);
int  ssps_abi_version(void);

int  ssps_create(const ssps_config* cfg, ssps_handle** out);
void ssps_destroy(ssps_handle* h);
/* Launch all work on this CUDA stream (a cudaStream_t); NULL = the legacy default stream. */
int  ssps_set_stream(ssps_handle* h, void* cuda_stream);

/* K1: raw lagged design -> standardised Z, Y -> G = Z'Z, R = Z'Y (fp64) + prior tables.
 * Xminus, Xplus: n x V column-major, RAW (only Xplus is pre-standardised by the reference,
 * SSPS.jl:156; Xminus is standardised per column inside compute_B_col).
 * prior: V x V column-major, entry [parent j + V * child i], values <= 0 mean "no confidence". */
int  ssps_set_data(ssps_handle* h, const double* Xminus, const double* Xplus, int32_t n, int32_t V,
                   const double* prior);
/* Same, from M concatenated time series (rows of series m are contiguous, already sorted by
 * timestep): X is (sum series_len) x V column-major; lagging is done on the device. */
int  ssps_set_timeseries(ssps_handle* h, const double* X, const int32_t* series_len, int32_t n_series,
                         int32_t V, const double* prior);
/* test hooks: copy back what K1 built (any pointer may be NULL) */
int  ssps_get_design(ssps_handle* h, double* Z, double* Y);
int  ssps_get_gram(ssps_handle* h, double* G, double* R, double* yty);
int  ssps_get_derived(ssps_handle* h, double* lambda_min, double* prob_param /* V */);
int  ssps_get_dims(ssps_handle* h, int32_t* n, int32_t* V, int32_t* n_chains);

/* K2: score n_sets (child, S) pairs.  S sorted ascending, unique.  lambda may be NULL when
 * logprior_out is NULL.  min_pivot_out (nullable) = smallest accepted Schur pivot / G_jj, a
 * conditioning diagnostic (inf for k = 0). */
int  ssps_score(ssps_handle* h, int32_t n_sets, const int32_t* child, const int32_t* set_offsets,
                const int32_t* set_members, const double* lambda, double* lml_out, double* logprior_out,
                double* min_pivot_out);

/* Chains.  chain ids are chain_id_offset .. chain_id_offset + n_chains - 1 (they key the Philox
 * streams, so ranks of a multi-GPU job pass disjoint offsets).  init_offsets (n_chains*V + 1),
 * init_members, init_lambda (n_chains*V, [chain*V + vertex]) are all-or-nothing; NULL => draw
 * the initial state from the model (sweep-0 Philox blocks). */
int  ssps_init_chains(ssps_handle* h, int32_t n_chains, uint64_t seed, uint32_t chain_id_offset,
                      const int32_t* init_offsets, const int32_t* init_members, const double* init_lambda);

/* K3: run up to n_sweeps sweeps per chain (one sweep = V lambda moves + V parent-set moves),
 * recording every `thinning`-th state; stops early once timeout_s (<= 0: none) is exceeded,
 * checked between kernel launches. */
int  ssps_run(ssps_handle* h, int64_t n_sweeps, double timeout_s, int32_t thinning, int64_t* sweeps_done);
/* Same sweeps, but driven by a recorded stream: draws is n_sweeps x n_chains x V x 6 doubles
 * (z_lambda, u_lambda, u_action, u_add, u_rem, u_accept), host memory. */
int  ssps_replay(ssps_handle* h, int64_t n_sweeps, const double* draws, int32_t thinning);

/* The random numbers ssps_run will use (or used) for sweeps first_sweep+1 .. first_sweep+n_sweeps of the
 * local chains, in ssps_replay's layout: out[n_sweeps][n_chains][V][6], host memory.  Feeding them to
 * ssps_replay (or to the CPU oracle) from the same initial state reproduces ssps_run bit for bit. */
int  ssps_export_draws(ssps_handle* h, int64_t first_sweep, int64_t n_sweeps, double* out);

/* Current state of one chain: indeg[V], members (concatenated sorted parent lists, capacity cap,
 * *n_members = total), lambda[V], lml[V] (any may be NULL). */
int  ssps_get_state(ssps_handle* h, int32_t chain, int32_t* indeg, int32_t* members, int64_t cap,
                    int64_t* n_members, double* lambda, double* lml);

/* Tallies over recorded samples 1..n of all local chains: counts[child*V + parent]. */
int  ssps_edge_counts(ssps_handle* h, double* counts_child_major, int64_t* n_samples, int32_t* n_chains);
/* Per-chain tallies: counts[(chain*V + child)*V + parent], uint32. */
int  ssps_chain_edge_counts(ssps_handle* h, uint32_t* counts);
/* evaluate_summary: pooled post-burn-in edge frequencies probs[child*V + parent];
 * burnin_used/n_used report the (snapshot-aligned) indices actually used. */
int  ssps_edge_posterior(ssps_handle* h, double burnin, double* probs_child_major, double* lambda_mean,
                         int64_t* burnin_used, int64_t* n_used);

/* K4: split-chain PSRF of every edge indicator and every lambda_i (postprocess_samples.jl:99-180).
 * ssps_psrf_sums writes the additive sufficient statistics of the LOCAL chains:
 *   sums[0 .. 3*V*V)        per edge: sum_h m_h, sum_h m_h^2, sum_h var_h   (h = half-chains)
 *   sums[3*V*V .. +3*V)     per lambda_i: same three
 *   sums[.. +2]             number of half-chains H, and H * n (n = half length), so that n = tail[1] / tail[0]
 *                           survives a SUM all-reduce
 * so that ranks can add them (ncclAllReduce / torch.distributed.all_reduce) before
 * ssps_psrf_finish.  on_device != 0: `sums` is a DEVICE pointer on this handle's device. */
int64_t ssps_psrf_sums_len(int32_t V);
int  ssps_psrf_sums(ssps_handle* h, double burnin, double* sums, int32_t on_device);
int  ssps_psrf_finish(ssps_handle* h, const double* sums, int32_t on_device, double* max_edge_psrf,
                      double* edge_psrf_child_major /* V*V, nullable, NaN where W == 0 */,
                      double* lambda_psrf /* V, nullable */);
/* local convenience = sums + finish */
int  ssps_psrf(ssps_handle* h, double burnin, double* max_edge_psrf, double* edge_psrf_child_major);

/* Change-point log of one chain (needs record_changepoints = 1): two-call protocol.
 * Parent events: (sample n, child, parent, value 0|1), ordered by (child, parent, n).
 * Lambda events: (sample n, child, value), ordered by (child, n).  Index 0 = initial state. */
int  ssps_changepoints_size(ssps_handle* h, int32_t chain, int64_t* n_parent_events, int64_t* n_lambda_events);
int  ssps_changepoints_get(ssps_handle* h, int32_t chain,
                           int64_t* pe_n, int32_t* pe_child, int32_t* pe_parent, int32_t* pe_value,
                           int64_t* le_n, int32_t* le_child, double* le_value);

/* ---------------------------------------------------------------------------------------------------------------
 * Multi-GPU.  Chains are independent: every handle (one per GPU) runs its own block of chain ids on its own copy of the
 * tables; the only exchange is a SUM all-reduce of additive statistics over an NCCL communicator OWNED BY THE LIBRARY
 * (libnccl.so.2 is resolved with dlopen when first needed; single-GPU use needs no NCCL).  Replaces the reference's
 * offline pooling of per-chain JSON files, scripts/postprocess_samples.jl:242-259.  Two host models:
 *   one process per GPU : rank 0 calls ssps_comm_unique_id and hands the 128 bytes to the other ranks by any host
 *                         means (a file, MPI, torch.distributed); every rank calls ssps_comm_init_rank; the *_global
 *                         calls are collective (every rank calls them, n = 1);
 *   one process, n GPUs : ssps_comm_init_all on the n handles (distinct devices); ssps_run_multi keeps all GPUs busy
 *                         from one host thread; the *_global calls take all n handles.
 * Pass disjoint chain_id_offset blocks to ssps_init_chains: an N-GPU job then draws exactly the chains a 1-GPU job of
 * the same total would. */
#define SSPS_COMM_ID_BYTES 128
int  ssps_comm_unique_id(void* id_out /* SSPS_COMM_ID_BYTES */);
int  ssps_comm_init_rank(ssps_handle* h, int32_t n_ranks, int32_t rank, const void* id /* SSPS_COMM_ID_BYTES */);
int  ssps_comm_init_all(ssps_handle* const* handles, int32_t n);
int  ssps_comm_destroy(ssps_handle* h);
int  ssps_comm_info(ssps_handle* h, int32_t* n_ranks, int32_t* rank, int32_t* nccl_version /* any may be NULL */);
/* ssps_run on n handles at once (launches enqueued round-robin, one synchronisation at the end);
 * *sweeps_done = the fewest sweeps any handle completed. */
int  ssps_run_multi(ssps_handle* const* handles, int32_t n, int64_t n_sweeps, double timeout_s, int32_t thinning,
                    int64_t* sweeps_done);
/* ssps_psrf / ssps_edge_posterior over the chains of ALL ranks (outputs as ssps_psrf_finish / ssps_edge_posterior).
 * Without a communicator and n = 1 they equal the local calls. */
int  ssps_psrf_global(ssps_handle* const* handles, int32_t n, double burnin, double* max_edge_psrf,
                      double* edge_psrf_child_major /* nullable */, double* lambda_psrf /* nullable */);
int  ssps_edge_posterior_global(ssps_handle* const* handles, int32_t n, double burnin, double* probs_child_major,
                                double* lambda_mean, int64_t* burnin_used, int64_t* n_used, int64_t* n_chains_total);

/* K5: evaluate_convergence + evaluate_summary (scripts/postprocess_samples.jl:58-191) on the device, from the change-point
 * logs of the chains of the given handles (all in this process; needs record_changepoints = 1): for every edge indicator
 * and every lambda_i the pooled post-burn-in mean, the split-chain PSRF and N_eff (variogram sum with the reference's
 * stopping rule) at stop point `length` (0 = all recorded samples).  Burn-in index floor(burnin * length) and split index
 * are EXACT (no snapshot grid).  Edge outputs V*V child-major, lambda outputs V; any may be NULL.  PSRF is NaN where the
 * within-chain variance is 0, as in the reference (JSON null). */
int  ssps_convergence(ssps_handle* const* handles, int32_t n, double burnin, int64_t length,
                      double* edge_mean, double* edge_psrf, double* edge_neff,
                      double* lambda_mean, double* lambda_psrf, double* lambda_neff,
                      int64_t* burnin_index, int64_t* split_index);

int  ssps_get_stats(ssps_handle* h, ssps_run_stats* out);
int  ssps_get_khist(ssps_handle* h, int64_t* hist, int32_t len);
/* The MH kernel's raw counter block (csrc/ssps_mh.cuh, enum ST_*), for profiling builds. */
int  ssps_get_raw_stats(ssps_handle* h, int64_t* out, int32_t len);
/* Diagnostics of the task scheduler: per vertex, the warp cycles its tasks cost in the last MH launch, 1000 x the
 * mean in-degree over the local chains now, and the number of warps its slabs were spread over (any may be NULL). */
int  ssps_get_vertex_cost(ssps_handle* h, int64_t* cycles /* V */, int32_t* mean_indeg_x1000 /* V */, int32_t* split /* V */);
/* Device-side microbenchmark used by bench.py for the FP64 roofline denominator: dependent-free
 * DFMA streams on every SM; returns achieved TFLOP/s. */
int  ssps_measure_fp64_peak(ssps_handle* h, double* tflops);
/* The same for the gather roofline (achieved_gather / measured L2 or shared-memory bandwidth): random 8-byte
 * loads from a table of `table_bytes` (pass the size of G + R), 8 in flight per thread, all SMs.  Returns USEFUL GB/s
 * (8 bytes per load) through the read-only L1 path the kernel uses (l1_gbs), from L2 only (ld.global.cg, l2_gbs) and from
 * shared memory (smem_gbs); any pointer may be NULL. */
int  ssps_measure_gather_peak(ssps_handle* h, int64_t table_bytes, double* l1_gbs, double* l2_gbs, double* smem_gbs);
/* Device blocks released by destroyed handles are kept in a process-wide cache for the next handle of the same
 * shapes (cudaMalloc of the 100 MB-class buffers costs 2..100 ms).  This returns the cached blocks of `device` to the
 * driver; live handles are not affected.  No counterpart in the reference (Julia's GC owns its memory). */
int  ssps_release_cached_memory(int32_t device);
/* Debug builds only (-DSSPS_BOUNDS: index checks on every scratch / Stage / slot accessor of the kernels): number of failed
 * checks since the library was loaded, on the current device; -1 = the library was built without the checks. */
int64_t ssps_bounds_failures(void);

/* K2 without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same
 * scores as ssps_score for regression_deg <= 1 (same device routine, same operation order), with two differences:
 *   - every array argument is a DEVICE pointer into the caller's own buffers (ssps_score takes host pointers and stages
 *     them through device blocks it allocates); the results are written there;
 *   - the per-thread factor scratch comes from ONE file-scope `static __device__` array of
 *     SSPS_SCORE_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until exit) instead
 *     of an arena sized per call.  A fixed number of threads walks the sets grid-stride, so n_sets is not limited by it.
 * The call performs no cudaMalloc / cudaFree (nor any allocating library routine): it validates the sizes, enqueues one
 * kernel on the handle's stream and RETURNS WITHOUT SYNCHRONISING -- order your own work against that stream
 * (ssps_set_stream) or synchronise the device.  The first call on a device resolves the array's device address
 * (cudaGetSymbolAddress); make it once before a graph capture begins.
 * Limits (beyond them: SSPS_ERR_INVALID and a message, nothing is launched): 0 <= max_parents <=
 * SSPS_SCORE_STATIC_MAX_PARENTS, 0 <= n_sets <= SSPS_SCORE_STATIC_MAX_SETS.  max_parents is the caller's bound on the
 * set sizes (the host cannot read device offsets without a copy): a set with more members than that, a child or member
 * outside [0, V), or members not strictly ascending gets NaN in every requested output -- checked on the device.
 * regression_deg >= 2 handles: SSPS_ERR_UNSUPPORTED (their scratch is orders of magnitude larger).
 * All calls in a process share the array per device: two calls that overlap on the GPU (different streams, different
 * handles, different host threads) RACE on it -- serialise them on one stream.  The array keeps the last call's contents;
 * nothing in it has to start at zero (every factor row, w entry and id is written before it is read within one set), so
 * nothing is cleared.  d_lambda may be NULL iff d_logprior_out is NULL; any output may be NULL. */
#define SSPS_SCORE_STATIC_MAX_PARENTS   64
#define SSPS_SCORE_STATIC_MAX_SETS      (1 << 24)
#define SSPS_SCORE_STATIC_SCRATCH_BYTES (16u << 20)
int  ssps_score_static(ssps_handle* h, int32_t n_sets, int32_t max_parents,
                       const int32_t* d_child, const int32_t* d_set_offsets /* n_sets + 1 */, const int32_t* d_set_members,
                       const double* d_lambda, double* d_lml_out, double* d_logprior_out, double* d_min_pivot_out);

#ifdef __cplusplus
}
#endif
#endif
