/*
 * ssps_oracle.h -- CPU ORACLE for the SSPS parent-set scoring / Metropolis-Hastings hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (ssps_b200/, include/ssps.h) never links, imports or calls anything in oracle/.
 *
 * PARITY UNPINNED: the reference (gitter-lab/ssps, Julia 1.6 + Gen 0.4.5) cannot run in
 * this image (no julia binary) and its only test (tests/ssps_csv.sh:11) ignores the score
 * column, so no reference-produced number pins this oracle.  It is a restatement of the
 * reference's source semantics, cross-checked against closed forms (k=0, saturated sets),
 * against LAPACK (numpy lstsq) and against the Hill-2012 formula
 * (hill-method/dynamic_network_inference.m:606-648).
 *
 * Reference files restated (paths relative to /root/reference):
 *   SSPS/src/dbn_preprocess.jl:70-98      combine_X / standardize_X      -> orc_standardize
 *   SSPS/src/dbn_distributions.jl:168-194 compute_B_col / construct_B    -> orc_construct_B
 *   SSPS/src/dbn_distributions.jl:201-229 safe_solve (Julia `\`)         -> orc_qr_project
 *   SSPS/src/dbn_distributions.jl:235-248 compute_lml                    -> orc_lml_qr
 *   SSPS/src/dbn_distributions.jl:96-153  ConfParentPrior.logpdf         -> orc_logprior_ref
 *   SSPS/src/dbn_distributions.jl:75-89   ConfParentPrior.random         -> orc_chain_init
 *   SSPS/src/dbn_proposals.jl:11-16       compute_move_probs             -> orc_move_probs
 *   SSPS/src/dbn_proposals.jl:19-49       ith_nonparent / searchsorted_exclude
 *   SSPS/src/dbn_proposals.jl:60-130      smart_proposal / smart_involution -> orc_chain_sweep
 *   SSPS/src/dbn_proposals.jl:244-248     lambda_vec_proposal            -> orc_chain_sweep
 *   SSPS/src/state_updates.jl:35-51       vertex_lambda_update_loop      -> orc_chain_sweep
 *   SSPS/src/state_updates.jl:161-227     update_results_storediff       -> orc_recorder_*
 *   SSPS/src/SSPS.jl:164-169              prob_param_i, lambda_min       -> orc_problem_create
 *   scripts/postprocess_samples.jl:99-191 PSRF / summary                 -> orc_psrf_binary, ...
 *   Gen 0.4.5 (Manifest.toml:189-193, NOT in tree): mh acceptance alpha = w - fwd + bwd,
 *       accept iff log(rand()) < alpha                                   -> orc_chain_sweep
 */
#ifndef SSPS_ORACLE_H
#define SSPS_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* scorer selection for the chain */
enum { ORC_SCORER_CHOL = 0,   /* Gram-Cholesky, same arithmetic order as the CUDA kernel (level-2 parity) */
       ORC_SCORER_QR   = 1 }; /* reference-faithful: build B(S), column-pivoted Householder QR */

/* chain mode flags */
enum { ORC_FLAG_LML_CACHE   = 1,  /* memoise lml on [i;S] like lml_cache (dbn_distributions.jl:161,255-264) */
       ORC_FLAG_STALE_PRIOR = 2,  /* emulate the lambda-blind prior_cache (quirk Q1, dbn_distributions.jl:102) */
       ORC_FLAG_UNIFORM     = 4   /* the deprecated uniform graph proposal (uniform_update_loop, state_updates.jl:98-112) */ };

enum { ORC_ACT_ADD = 1, ORC_ACT_SWAP = 2, ORC_ACT_REM = 3 };

typedef struct orc_problem orc_problem;
typedef struct orc_chain orc_chain;

/* ---------- stateless helpers (each cites its reference lines in the .c) ---------- */
void   orc_standardize(const double* X, int n, int V, double* Z);
void   orc_gram(const double* Z, const double* Y, int n, int V, double* G, double* R, double* yty);
int    orc_n_design_cols(int k, int deg);
void   orc_construct_B(const double* Xminus, int n, const int* S, int k, int deg, double* B);
double orc_qr_project(double* B, int n, int width, const double* y, int* rank_out);
double orc_lml_qr(const double* Xminus, const double* y, int n, const int* S, int k, int deg, int* rank_out);
double orc_lml_chol(const double* G, const double* Ri, double yty, int n, int V,
                    const int* S, int k, double* min_pivot_ratio, int* n_skipped);
double orc_lml_hill(const double* Xminus, const double* y, int n, const int* S, int k);
double orc_logprior_ref(const int* S, int k, int V, const double* conf_col, double lambda);
double orc_logprior_closed(const int* S, int k, int V, const double* conf_col, double lambda);
void   orc_move_probs(int k, int V, double prob_param, double out[3]);
int    orc_ith_nonparent(int i, const int* S, int k);           /* 1-based in, 1-based out */
int    orc_searchsorted_exclude(const int* S, int k, int x);   /* 1-based */
void   orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void   orc_step_draws(uint64_t seed, uint32_t chain, uint32_t vertex, uint64_t sweep, double out[6]);

/* ---------- problem (data + derived parameters, SSPS.jl:155-169) ---------- */
/* Xminus, Xplus: n x V column-major RAW (Xplus is standardised inside, SSPS.jl:156).
 * prior: V x V column-major, entry [parent j + V*child i] (dbn_preprocess.jl:33-40). */
orc_problem* orc_problem_create(const double* Xminus, const double* Xplus, int n, int V,
                                const double* prior, double large_indeg, double lambda_max,
                                double lambda_prop_std, int regression_deg);
void   orc_problem_destroy(orc_problem*);
int    orc_problem_n(const orc_problem*);
int    orc_problem_V(const orc_problem*);
double orc_problem_lambda_min(const orc_problem*);
const double* orc_problem_G(const orc_problem*);
const double* orc_problem_R(const orc_problem*);
const double* orc_problem_Z(const orc_problem*);
const double* orc_problem_Y(const orc_problem*);
const double* orc_problem_yty(const orc_problem*);
const double* orc_problem_prob_param(const orc_problem*);
/* score one (child, S) -- 0-based child and members; scorer = ORC_SCORER_* */
double orc_problem_lml(const orc_problem*, int child, const int* S, int k, int scorer);
double orc_problem_logprior(const orc_problem*, int child, const int* S, int k, double lambda, int closed_form);

/* ---------- chain ---------- */
orc_chain* orc_chain_create(const orc_problem*, int scorer, int flags);
void   orc_chain_destroy(orc_chain*);
/* initial state from the Philox stream (sweep 0) == Gen.generate (mcmc_inference.jl:40) */
void   orc_chain_init(orc_chain*, uint64_t seed, uint32_t chain_id);
/* or an explicit state (resume, SSPS.jl:24-43): parents CSR 0-based sorted */
void   orc_chain_set_state(orc_chain*, const int* offsets, const int* members, const double* lambda);
/* one sweep (state_updates.jl:43-48).  draws: V x 6 doubles (z, u_lam, u_act, u_add, u_rem, u_acc)
 * or NULL => Philox stream keyed by (seed, chain_id, vertex, sweep index).
 * acc_out (nullable): V x 3 ints: lambda accepted, parent move accepted, action. */
void   orc_chain_sweep(orc_chain*, const double* draws, int* acc_out);
void   orc_chain_run(orc_chain*, int64_t n_sweeps);
int64_t orc_chain_sweeps_done(const orc_chain*);
int    orc_chain_indegree(const orc_chain*, int child);
void   orc_chain_parents(const orc_chain*, int child, int* out);
double orc_chain_lambda(const orc_chain*, int child);
double orc_chain_lml(const orc_chain*, int child);
int64_t orc_chain_n_proposals(const orc_chain*);
int64_t orc_chain_n_accepted(const orc_chain*);
int64_t orc_chain_cache_hits(const orc_chain*);
void   orc_chain_k_hist(const orc_chain*, int64_t* hist, int len); /* histogram of proposed |S'| */

/* ---------- recorder: update_results_storediff (state_updates.jl:161-227) ---------- */
/* Attach a recorder; call orc_chain_record() after every `thinning` sweeps.  Events are
 * (sample n, parent j 0-based, value 0/1) per child, and (n, lambda) per child. */
void   orc_chain_recorder_enable(orc_chain*);
void   orc_chain_record(orc_chain*);
int64_t orc_chain_recorded_n(const orc_chain*);
int64_t orc_chain_n_parent_events(const orc_chain*, int child);
void   orc_chain_parent_events(const orc_chain*, int child, int64_t* n_idx, int* parent, int* value);
int64_t orc_chain_n_lambda_events(const orc_chain*, int child);
void   orc_chain_lambda_events(const orc_chain*, int child, int64_t* n_idx, double* value);
/* dense tallies over recorded samples 1..n (increment_counts!, state_updates.jl:240-249) */
void   orc_chain_edge_counts(const orc_chain*, int64_t* counts_child_major /* V*V: [child*V+parent] */);

/* ---------- post-processing (scripts/postprocess_samples.jl) ---------- */
/* Binary split-PSRF for one edge from dense 0/1 sequences: seqs = n_chains pointers of length len
 * (samples 1..len).  Returns psrf (NaN if W==0).  (postprocess_samples.jl:99-138,155-180) */
double orc_psrf_binary_dense(const uint8_t* const* seqs, int n_chains, int64_t len, double burnin);
double orc_psrf_real_dense(const double* const* seqs, int n_chains, int64_t len, double burnin);
/* PSRF from per-half-sequence counts (what the device reduces): counts1[c], counts2[c] over
 * half lengths N1, N2 */
double orc_psrf_from_counts(const int64_t* c1, const int64_t* c2, int n_chains, int64_t N1, int64_t N2);
/* N_eff (postprocess_samples.jl:58-94) on dense sequences */
double orc_neff_binary_dense(const uint8_t* const* seqs, int n_chains, int64_t len, double burnin);

/* ---------- multi-threaded reference-faithful run for bench.py's cpu_baseline ---------- */
/* Runs n_chains independent chains (one per thread, up to n_threads at a time) for n_sweeps
 * each; returns wall seconds; fills edge-count totals (V*V child-major, nullable). */
double orc_run_chains_mt(const orc_problem*, int scorer, int flags, int n_chains, int n_threads,
                         int64_t n_sweeps, uint64_t seed, int64_t* edge_counts_total,
                         int64_t* n_proposals_total);

#ifdef __cplusplus
}
#endif
#endif
