// ssps_mh.cuh -- K3: the persistent fused Metropolis-Hastings kernel (sm_100a).
//
// Reference semantics restated here (paths relative to the gitter-lab/ssps checkout):
//   vertex_lambda_update_loop  SSPS/src/state_updates.jl:35-51     one sweep = V (lambda move, parent-set move) pairs
//   lambda_vec_proposal        SSPS/src/dbn_proposals.jl:244-248   lambda' = lambda + sigma z
//   smart_proposal/involution  SSPS/src/dbn_proposals.jl:60-130    add / swap / remove on the sorted parent list
//   compute_move_probs         SSPS/src/dbn_proposals.jl:11-16     (tabulated by move_table_kernel)
//   Gen.mh                     Gen 0.4.5 (not in tree)             accept iff log(u) < w - fwd + bwd
//   compute_lml                SSPS/src/dbn_distributions.jl:235-248
//   update_results_storediff   SSPS/src/state_updates.jl:161-227   thinned recording / change points
//
// Work decomposition.  A micro-chain = (chain, vertex): the posterior factorises over vertices
// (dbn_models.jl:94-96), so micro-chains never communicate.  One warp-task = (vertex i, 32
// chains), lane = chain; a task runs ALL sweeps of the launch.  Tasks are pulled from an atomic
// counter, vertices ordered by the cycles they cost in the previous launch (longest first), so
// the persistent grid drains evenly.
//
// How a parent-set proposal is scored.  lml(i, S) needs q = r_S' G_SS^-1 r_S.  The device
// definition of q (shared with the oracle's kernel-faithful scorer, orc_lml_chol) is the row
// Cholesky of G_SS with column skipping, taken over the parents in ARRIVAL order (the order in
// which they joined the set; the initial state arrives sorted).  Every micro-chain keeps that
// factor (L, 1/d, w = L^-1 r) of its current set, so
//      add j    : ONE new row against the cached factor, O(k^2) -- the row a from-scratch
//                 factorisation of (S, j) would compute last, bit for bit;
//      remove p : rows after p recomputed into scratch, rows before p reused -- again exactly the
//                 rows of a from-scratch factorisation of S \ p;
//      swap     : remove p, then add j against the rows just recomputed.
// Nothing is re-factorised from scratch and nothing is approximated: q of every proposal equals
// the from-scratch value bitwise (`verify` launches re-derive it with score_set() and count
// differences).  Only the comparison  log(u) < alpha  is SCREENED: both sides are first formed
// with fp32 MUFU logarithms and compared with a margin that bounds the screening error; if they
// are closer than the margin the test is redone in the reference order with fp64 logarithms.
//
// Storage tiers for (ids, factor) of a lane: |S| <= SSPS_KS in shared memory, each lane on its own;
// larger sets in a global arena, processed by the WHOLE WARP one lane at a time (see "Tier storage").
#pragma once
#include "ssps_kernels.cuh"

#ifndef SSPS_KS
#define SSPS_KS 5
#endif
#ifndef SSPS_LOCKSTEP
#define SSPS_LOCKSTEP 1      // 1: the warps of a scheduler walk the sweep in step (named barrier per sweep); 2: the whole CTA; 0: off
#endif
#ifndef SSPS_EARLY_GATHER
#define SSPS_EARLY_GATHER 1   // small_move issues the arriving parent's G / R gathers before the factor work (+3.5 % measured)
#endif
#ifndef SSPS_LS_EVERY
#define SSPS_LS_EVERY 1      // sweeps between the lock-step barriers (power of two)
#endif
#ifndef SSPS_BLOCK
#define SSPS_BLOCK 384       // one CTA per SM: three warps per scheduler
#endif
#ifndef SSPS_MIN_CTAS
#define SSPS_MIN_CTAS 1
#endif
#define SSPS_LN2 0.6931471805599453

enum { MODE_SMALL = 0, MODE_BIG = 1, MODE_PROMOTE = 2 };
// stats slots (unsigned long long each)
enum { ST_ACCP = 0, ST_ACCL = 1, ST_SUMK = 2, ST_F3 = 3, ST_BYTES = 4, ST_KCAP_REJECT = 5, ST_PROP = 6, ST_TIER1 = 7,
       ST_MISMATCH = 10, ST_BIG_CYCLES = 11, ST_LAM_EXACT = 12, ST_BIG_STEPS = 13, ST_MID_STEPS = 14, ST_MID_CYCLES = 15, ST_PROF = 16,
       ST_DEC_MISMATCH = 24, ST_LAM_DEC_MISMATCH = 25,      // VERIFY: screened decision != reference-order fp64 decision
       ST_MARGIN_USE = 26, ST_LAM_MARGIN_USE = 27,          // VERIFY: max |screening error| / margin as float bits (low word)
       ST_SCR_TRY = 40, ST_SCR_REJ = 41,                    // BIG moves with a removal: screened but not rejected, rejected by the screening
       ST_MID_PROF = 32,                                    // profiling build: MID tier cycles by section (8 slots)
       ST_N = 48 };

__device__ __forceinline__ float fast_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_sin(float x) { float y; asm("sin.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_cos(float x) { float y; asm("cos.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// ------------------------------------------------------------------------------------------------
// Per-step random numbers.  (z, u_lam, u_act, u_add, u_rem, u_acc): the proposal's standard normal
// and the five uniforms the reference draws (Gen.normal, rand() of the lambda move, categorical,
// two uniform_discrete, rand() of the parent-set move).
//
// Philox4x32-10 streams keyed (seed; chain, vertex, sweep, block):
//   block 1 of sweep s      -> u_act, u_add, u_rem, u_acc  ((w + 1/2) 2^-32, exact in fp64)
//   block 0 of sweep s | 1  -> one Box-Muller pair shared by sweeps (s & ~1, s | 1): the even sweep
//                              takes r cos(theta) and word 2, the odd one r sin(theta) and word 3
// The normal is formed in fp32 (MUFU lg2 / sqrt / sin / cos) with the sign taken from a separate
// bit, so the proposal kernel is exactly symmetric; it is a proposal, not a score, and
// ssps_export_draws returns the very numbers the kernel uses, so parity tests replay them
// through the oracle bit for bit.
// ------------------------------------------------------------------------------------------------
struct StepDraws { double z, u_lam, u_act, u_add, u_rem, u_acc; };
struct DrawCarry { float z; uint32_t w; int valid; };

struct PhiloxSource {
    uint64_t seed; uint32_t chain_offset;
    __device__ __forceinline__ StepDraws get(int chain_local, int vertex, uint64_t sweep, int64_t /*local*/, DrawCarry& cy) const {
        const uint32_t chain = chain_offset + (uint32_t)chain_local;
        uint32_t b[4];
        philox_block(seed, chain, (uint32_t)vertex, sweep, 1, b);
        float z; uint32_t wl;
        if ((sweep & 1ull) && cy.valid) { z = cy.z; wl = cy.w; cy.valid = 0; }
        else {
            uint32_t a[4];
            philox_block(seed, chain, (uint32_t)vertex, sweep | 1ull, 0, a);
            const float u1 = fmaf((float)a[0], 2.3283064365386963e-10f, 1.1641532182693481e-10f);
            const float r = fast_sqrt(-1.3862943611198906f * fast_lg2(u1));           // sqrt(-2 ln u1)
            const float th = ((float)(a[1] & 0x7fffffffu) + 0.5f) * 1.4629180792671596e-09f;   // pi 2^-31
            float c = fast_cos(th), s = fast_sin(th);
            if (a[1] >> 31) { c = -c; s = -s; }
            if (sweep & 1ull) { z = r * s; wl = a[3]; }
            else { z = r * c; wl = a[2]; cy.z = r * s; cy.w = a[3]; cy.valid = 1; }
        }
        StepDraws d;
        d.z = (double)z;
        d.u_lam = u32_to_unit(wl);
        d.u_act = u32_to_unit(b[0]);
        d.u_add = u32_to_unit(b[1]);
        d.u_rem = u32_to_unit(b[2]);
        d.u_acc = u32_to_unit(b[3]);
        return d;
    }
};

struct ReplaySource {
    const double* draws; int n_chains; int V;   // [local sweep][chain][vertex][6]
    __device__ __forceinline__ StepDraws get(int chain_local, int vertex, uint64_t /*sweep*/, int64_t local, DrawCarry&) const {
        const double* p = draws + ((((size_t)local * n_chains + chain_local) * V) + vertex) * 6;
        StepDraws d;
        d.z = p[0]; d.u_lam = p[1]; d.u_act = p[2]; d.u_add = p[3]; d.u_rem = p[4]; d.u_acc = p[5];
        return d;
    }
};

// the numbers PhiloxSource hands the kernel, for sweeps sweep0+1 .. sweep0+n: out[local][chain][vertex][6]
__global__ void export_draws_kernel(PhiloxSource src, int n_chains, int V, int64_t sweep0, int n_sweeps, double* __restrict__ out) {
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t total = (size_t)n_sweeps * n_chains * V;
    if (e >= total) return;
    const int vertex = (int)(e % V), chain = (int)((e / V) % n_chains), local = (int)(e / ((size_t)V * n_chains));
    DrawCarry cy; cy.valid = 0; cy.z = 0.f; cy.w = 0u;
    const StepDraws d = src.get(chain, vertex, (uint64_t)(sweep0 + local + 1), local, cy);
    double* p = out + e * 6;
    p[0] = d.z; p[1] = d.u_lam; p[2] = d.u_act; p[3] = d.u_add; p[4] = d.u_rem; p[5] = d.u_acc;
}

// ------------------------------------------------------------------------------------------------
// Tier storage.  Slot a (arrival order) of a micro-chain holds: word = id | rel << 16 (`rel` =
// recorded samples of this launch that precede the parent's arrival: lazy tally), rank = position
// of the id in the SORTED parent list (the reference's moves are defined on the sorted list,
// dbn_proposals.jl:19-49), and row a of the factor: L(a, b) for b < a, 1/d_a, w_a.
//
//   SMALL  |S| <= SSPS_KS : shared memory, element-major [e][thread]; every lane works on its own
//          micro-chain, statically unrolled (small_move).
//   BIG    |S| >  SSPS_KS : a global arena, one contiguous block per lane; the WHOLE WARP works on
//          one such lane at a time (coop_eval / coop_commit) while the others wait -- big sets are
//          rare (< 1 % of the proposals at the metric point) but cost O(k^2)..O(k^3), so one lane
//          grinding through them alone would hold its 31 neighbours for the whole step.
// ------------------------------------------------------------------------------------------------
#define SSPS_ROWLEN(K) (SSPS_TRI(K) + (K))          // rows incl. the 1/d slots: TRI(K) doubles, then w: K doubles
#define SSPS_FULL 0xffffffffu
#define SSPS_KU 4                                    // a lane owns columns lane + 32 u, u < SSPS_KU: kcap <= 128

struct SmallAcc {   // shared memory, element-major: [e][thread]
    double* d; uint32_t* w; uint8_t* r;
    __device__ __forceinline__ double& Lo(int a, int b) const { return d[(SSPS_TRI(a) + b) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Li(int a) const { return d[(SSPS_TRI(a) + a) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Wv(int a) const { return d[(SSPS_TRI(SSPS_KS) + a) * SSPS_BLOCK]; }
    __device__ __forceinline__ uint32_t& W(int a) const { return w[a * SSPS_BLOCK]; }
    __device__ __forceinline__ int rank(int a) const { return (int)r[a * SSPS_BLOCK]; }
    __device__ __forceinline__ void set_rank(int a, int v) const { r[a * SSPS_BLOCK] = (uint8_t)v; }
    __device__ __forceinline__ SmallAcc of_lane(int delta) const { SmallAcc o; o.d = d + delta; o.w = w + delta; o.r = r + delta; return o; }
};

// one lane's block of the BIG arena (the master copy of its ids, ranks and factor between steps)
struct BigRef {
    double* f; uint32_t* w; uint16_t* r; int kcap;
    __device__ __forceinline__ double* fw() const { return f + SSPS_TRI(kcap); }
    __device__ __forceinline__ int id(int a) const { SSPS_CHK(a >= 0 && a < kcap); return (int)(w[a] & 0xffffu); }
};

// The warp's working copy of ONE lane's BIG-tier state while the warp works on it: shared memory when
// the proposed set fits SSPS_KST rows, else the warp's global scratch.  A move is applied to the copy
// in place (ids, ranks, factor -> those of the proposed set); the copy is written back only on accept.
#ifndef SSPS_KST
#define SSPS_KST 32
#endif
struct Stage {
    double* t;          // factor rows, TRI-packed, 1/d in the diagonal slots
    double* tw;         // w
    uint32_t* id;       // id | rel << 16, arrival order
    uint16_t* rk;       // sorted rank
    __device__ __forceinline__ int idof(int a) const { SSPS_CHK(a >= 0 && a < 32 * SSPS_KU); return (int)(id[a] & 0xffffu); }
};

// Entry e of a Stage's factor rows / w vector.  SH = the Stage is the warp's shared-memory one: the accesses index the
// ssps_smem symbol (LDS / STS); otherwise plain pointers into the warp's global scratch.
extern __shared__ double ssps_smem[];
template <bool SH>
struct StageAcc {
    double* tp; double* wp; unsigned toff, woff;
    __device__ __forceinline__ explicit StageAcc(const Stage& T) : tp(T.t), wp(T.tw), toff(SH ? (unsigned)(T.t - ssps_smem) : 0u), woff(SH ? (unsigned)(T.tw - ssps_smem) : 0u) {}
    __device__ __forceinline__ double& t(int e) const { SSPS_CHK(e >= 0 && (!SH || e < SSPS_TRI(SSPS_KST))); if (SH) return ssps_smem[toff + e]; return tp[e]; }
    __device__ __forceinline__ double& w(int e) const { SSPS_CHK(e >= 0 && (!SH || e < SSPS_KST)); if (SH) return ssps_smem[woff + e]; return wp[e]; }
};

// dynamic shared memory of one CTA: SMALL-tier slots of every thread + one Stage per warp
#define SSPS_SMEM_DOUBLES (SSPS_ROWLEN(SSPS_KS) * SSPS_BLOCK + (SSPS_BLOCK / 32) * SSPS_ROWLEN(SSPS_KST))
#define SSPS_SMEM_WORDS (SSPS_KS * SSPS_BLOCK + (SSPS_BLOCK / 32) * SSPS_KST)
#define SSPS_SMEM_BYTES ((size_t)SSPS_SMEM_DOUBLES * 8 + (size_t)SSPS_SMEM_WORDS * 4 + (size_t)(SSPS_BLOCK / 32) * SSPS_KST * 2 + (size_t)SSPS_KS * SSPS_BLOCK)

// The CTA's dynamic shared memory.  Device functions that work on a slot of it index THIS symbol (offset = pointer - ssps_smem):
// the compiler then knows the address space and emits LDS / STS; through a `double*` parameter the same accesses are generic
// LD / ST with a descriptor set up before each (measured on the BIG-tier refactorisation: 24 instructions per inner iteration).
extern __shared__ double ssps_smem[];

struct Lane {
    double lam, somc, ip, nrmf;      // nrmf: screening value of the prior normaliser at lam
    float lg_ip;                     // lg2(ip), screening
    int k, mode, n_pev;
};

struct MhParams {
    int64_t sweep0;          // sweeps already done (multiple of thinning)
    int n_local;             // sweeps in this launch
    int thinning;
    int n_tasks, n_slabs;
    int* task_counter;
    const int* order;        // vertices, longest tasks first
    const int* start;        // first task id per rank; start[V] = number of tasks
    const int* split;        // per vertex: warps per slab (1, 2, 4, 8)
    int V;
    unsigned long long* vcost;   // per-vertex cycles of this launch
    // BIG tier: per resident thread [gtid][ROWLEN(kcap)] / [gtid][kcap]; the same per resident WARP as the
    // working copy of sets too large for the shared-memory Stage
    double* big_d; uint32_t* big_w; uint16_t* big_r;
    double* scr_d; uint32_t* scr_w; uint16_t* scr_r;
    double* x_d; uint16_t* x_t;                            // per-thread id lists (x_t) and from-scratch scorer scratch (x_d: verify only)
    int kcap; int stride;
    unsigned long long* stats;
    unsigned long long* khist;   // nullable, 64 bins
    // change-point log (nullable): per micro-chain mc = chain*V + vertex
    uint32_t* pev; int* pev_n; int pev_cap;        // parent events: id | rel_rec<<16 | flag<<31
    double* lev_val; uint16_t* lev_idx; int* lev_n; int lev_cap;
};

struct KCtx {       // what the tiers need besides the lane state.  Pointers of the rare tiers (arena, global Stage, verify
                    // scratch) are formed on demand from the launch parameters: carried in registers they spilled the hot loop
    const MhParams* prm;   // the kernel's __grid_constant__ parameter block
    int wbase;             // global thread index of lane 0 of this warp
    Stage st_smem;         // this warp's shared-memory Stage
    double yty;            // y_i' y_i of this task's vertex
    uint32_t* cnt_row;     // counts + mc * V
    uint32_t* pev;         // pev + mc * pev_cap (nullable)
    int pev_cap, kcap;
    unsigned long long* stats;
    __device__ __forceinline__ BigRef big(int ln) const {
        BigRef b;
        const size_t t = (size_t)(wbase + ln);
        b.f = prm->big_d + t * SSPS_ROWLEN(kcap);
        b.w = prm->big_w + t * kcap; b.r = prm->big_r + t * kcap; b.kcap = kcap;
        return b;
    }
    __device__ __forceinline__ Stage stage(int rows) const {
        if (rows <= SSPS_KST) return st_smem;
        Stage g;
        const size_t w = (size_t)(wbase >> 5);
        g.t = prm->scr_d + w * SSPS_ROWLEN(kcap); g.tw = g.t + SSPS_TRI(kcap);
        g.id = prm->scr_w + w * kcap; g.rk = prm->scr_r + w * kcap;
        return g;
    }
    __device__ __forceinline__ Scratch xs_of(int ln) const {
        Scratch x;
        x.d = prm->x_d + (wbase + ln); x.t = prm->x_t + (wbase + ln); x.stride = prm->stride; x.kcap = kcap;
        return x;
    }
};

__device__ __forceinline__ int ceil_div_i(int a, int b) { return (a + b - 1) / b; }

// screening value of the prior normaliser: sum over groups with 1-c != 0 of mult ln2 lg2(1 + 2^(-(1-c) log2e lambda))
// the same sum with the vertex's first two groups held in registers (0/1 priors have one group): no table loads per step
struct NrmGroups { float2 a, b; int g2, g1; };      // groups g0, g0 + 1 (zero multiplicity if absent); the rest are [g2, g1)
__device__ __forceinline__ NrmGroups normaliser_groups(const SspsDev& P, int child) {
    NrmGroups n;
    const int g0 = __ldg(P.fgrp_off + child);
    n.g1 = __ldg(P.fgrp_off + child + 1); n.g2 = min(g0 + 2, n.g1);
    n.a = make_float2(0.f, 0.f); n.b = n.a;
    if (g0 < n.g1) n.a = __ldg(P.fgrp + g0);
    if (g0 + 1 < n.g1) n.b = __ldg(P.fgrp + g0 + 1);
    return n;
}
__device__ __forceinline__ double normaliser_fast(const SspsDev& P, const NrmGroups& n, double lambda) {
    const float lf = (float)lambda;
    double acc = 0.0;
    {      // group 0, group 1 in table order; an absent group adds (double)(0 * lg2(..)) = +0
        acc += (double)(n.a.y * fast_lg2(1.0f + fast_ex2(n.a.x * lf)));
        acc += (double)(n.b.y * fast_lg2(1.0f + fast_ex2(n.b.x * lf)));
    }
    for (int g = n.g2; g < n.g1; ++g) {
        const float2 am = __ldg(P.fgrp + g);
        acc += (double)(am.y * fast_lg2(1.0f + fast_ex2(am.x * lf)));
    }
    return acc;
}
__device__ __forceinline__ double normaliser_fast(const SspsDev& P, int child, double lambda) {
    const float lf = (float)lambda;
    double acc = 0.0;
    const int g1 = __ldg(P.fgrp_off + child + 1);
    for (int g = __ldg(P.fgrp_off + child); g < g1; ++g) {
        const float2 am = __ldg(P.fgrp + g);
        acc += (double)(am.y * fast_lg2(1.0f + fast_ex2(am.x * lf)));
    }
    return acc;
}

// the SORTED current set of this lane -> xs.S(0..k-1)
__device__ __noinline__ void write_sorted(const int mode, const SmallAcc& S, const BigRef& B, const Scratch& xs, const int k) {
    if (mode == MODE_BIG) { for (int b = 0; b < k; ++b) xs.S((int)B.r[b]) = (uint16_t)B.id(b); }
    else { for (int b = 0; b < k; ++b) xs.S(S.rank(b)) = (uint16_t)(S.W(b) & 0xffffu); }
}

// sum of (1 - c) over the current set, in the reference's (sorted) order
__device__ __noinline__ double sorted_somc(const SspsDev& P, const int child, const int mode, const SmallAcc& S, const BigRef& B,
                                           const Scratch& xs, const int k) {
    write_sorted(mode, S, B, xs, k);
    const double* __restrict__ omc_i = P.omc + (size_t)P.V * child;
    double somc = 0.0;
    for (int a = 0; a < k; ++a) somc += __ldg(omc_i + xs.S(a));
    return somc;
}

// lambda move in the reference order (the oracle's chain_step restates exactly this); xs.S(0..k-1)
// holds the SORTED current set
__device__ __noinline__ int lambda_exact_decide(const SspsDev& P, const int child, const Scratch& xs, const int k, const double lam,
                                                const double lam_new, const double u_lam, double* diff_out) {
    const double* __restrict__ omc_i = P.omc + (size_t)P.V * child;
    double somc = 0.0;
    for (int a = 0; a < k; ++a) somc += __ldg(omc_i + xs.S(a));
    const double nrm = prior_normaliser(P, child, lam), nrm_new = prior_normaliser(P, child, lam_new);
    const double w = -(lam_new - lam) * somc - (nrm_new - nrm);
    const double lu = log(u_lam);
    if (diff_out) *diff_out = lu - w;
    return lu < w ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// Pieces shared by both tiers
// ------------------------------------------------------------------------------------------------
// one Cholesky pivot (orc_lml_chol / score_set): accepted -> (1/d, t/d), q += w^2; numerically
// dependent column -> contributes nothing; NaN data (zero-variance column, quirk Q4) -> NaN
// 1 / sqrt(x), IEEE sqrt then IEEE division (the oracle's bits).  One out-of-line copy: inlined, the two
// software sequences are ~45 instructions at each of the 8 pivots of the hot loop (instruction fetch)
__device__ __noinline__ double inv_sqrt_ieee(const double x) { return 1.0 / sqrt(x); }

__device__ __forceinline__ void chol_pivot(const double d2, const double gjj, const double t, double& inv, double& w, double& q) {
    if (d2 > SSPS_PIVOT_TOL * gjj) { inv = inv_sqrt_ieee(d2); w = t * inv; q = fma(w, w, q); }
    else if (d2 != d2 || gjj != gjj) { inv = __longlong_as_double(0x7ff8000000000000LL); w = inv; q = inv; }
    else { inv = 0.0; w = 0.0; }
}

// smart_proposal (dbn_proposals.jl:60-80): the action and the two ordinal draws (1-based; 0 = not drawn)
__device__ __forceinline__ void select_move(const SspsDev& P, const int child, const int k, const StepDraws& d, int& action, int& a_ord,
                                            int& r_ord) {
    const int V = P.V;
    const size_t trow = ((size_t)child * (V + 1) + k) * 3;
    const double p_add = __ldg(P.mvP + trow), p_swp = __ldg(P.mvP + trow + 1);
    action = d.u_act < p_add ? ACT_ADD : (d.u_act < p_add + p_swp ? ACT_SWAP : ACT_REM);
    a_ord = 0; r_ord = 0;
    if (action != ACT_REM) { const int m = V - k; int a = 1 + (int)floor(d.u_add * (double)m); if (a > m) a = m; a_ord = a; }
    if (action != ACT_ADD) { int r = 1 + (int)floor(d.u_rem * (double)k); if (r > k) r = k; r_ord = r; }
}

// log proposal probabilities of the move and of its reverse (smart_involution, dbn_proposals.jl:83-130); swaps cancel
__device__ __forceinline__ void move_ratio(const SspsDev& P, const int child, const int k, const int action, double& fwd, double& bwd) {
    const int V = P.V;
    const size_t trow = ((size_t)child * (V + 1) + k) * 3;
    fwd = 0.0; bwd = 0.0;
    if (action == ACT_ADD) {
        fwd = __ldg(P.mvLogP + trow) - __ldg(P.logint + (V - k));
        bwd = __ldg(P.mvLogP + ((size_t)child * (V + 1) + (k + 1)) * 3 + 2) - __ldg(P.logint + (k + 1));
    } else if (action == ACT_REM) {
        fwd = __ldg(P.mvLogP + trow + 2) - __ldg(P.logint + k);
        bwd = __ldg(P.mvLogP + ((size_t)child * (V + 1) + (k - 1)) * 3 + 0) - __ldg(P.logint + (V - k + 1));
    }
}

// Gen.mh: accept iff log(u) < alpha.  Both sides are first formed with fp32 MUFU logarithms and compared
// with a margin that bounds the screening error; closer than the margin -> the reference-order fp64 test.
// the reference-order fp64 test (rare: out of line)
__device__ __noinline__ int mh_accept_exact(const double log1pn, const double dn, const int action, const int k, const int kn, const double dprior,
                                            const double fwd, const double bwd, const double ip_cur, const double ip_new, const double u_acc) {
    const double lml_cur = -0.5 * (double)k * log1pn - 0.5 * dn * log(ip_cur);
    const double lml_new = -0.5 * (double)kn * log1pn - 0.5 * dn * log(ip_new);
    const double w = dprior + (lml_new - lml_cur);
    const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
    return log(u_acc) < alpha ? 1 : 0;
}

template <bool VERIFY>
__device__ __forceinline__ bool mh_accept(const SspsDev& P, const int action, const int k, const int kn, const double dprior,
                                          const double fwd, const double bwd, const double ip_cur, const float lg_ip_cur,
                                          const double ip_new, const float lg_new, const double u_acc, unsigned long long* stats) {
    const float lgu = fast_lg2((float)u_acc);
    const double hn = 0.5 * P.dn;
    const double base = ((dprior - fwd) + bwd) - 0.5 * (double)(kn - k) * P.log1pn;
    const double diff = SSPS_LN2 * ((double)lgu + hn * ((double)lg_new - (double)lg_ip_cur)) - base;
    const double marg = 1.2e-6 * ((double)fabsf(lgu) + 1.0 + hn * ((double)(fabsf(lg_new) + fabsf(lg_ip_cur)) + 2.0)) + 1e-9 * fabs(base) + 1e-6;
    bool accept = diff < 0.0;
    if (!(u_acc > 1e-30 && fabs(diff) > marg)) {        // too close (or NaN): the reference-order fp64 test
        atomicAdd(stats + ST_TIER1, 1ull);      // rare (5e-6 per proposal): no counter register in the hot loop
#ifdef SSPS_ACCEPT_OOL
        accept = mh_accept_exact(P.log1pn, P.dn, action, k, kn, dprior, fwd, bwd, ip_cur, ip_new, u_acc) != 0;
#else
        const double lml_cur = -0.5 * (double)k * P.log1pn - 0.5 * P.dn * log(ip_cur);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        const double w = dprior + (lml_new - lml_cur);
        const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
        accept = log(u_acc) < alpha;
#endif
        if (VERIFY) return accept;          // already the reference-order test
    }
    if (VERIFY) {       // the DECISION re-derived in the reference order (Gen.mh: log(rand()) < alpha) for every screened test
        const double lml_cur = -0.5 * (double)k * P.log1pn - 0.5 * P.dn * log(ip_cur);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        const double w = dprior + (lml_new - lml_cur);
        const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
        const double lu = log(u_acc);
        if ((lu < alpha) != accept) atomicAdd(stats + ST_DEC_MISMATCH, 1ull);
        // how much of its margin the screening used: a flip needs |error| > |diff| > margin, i.e. a ratio above 1
        const float use = (float)(fabs(diff - (lu - alpha)) / marg);
        atomicMax((unsigned*)(stats + ST_MARGIN_USE), __float_as_uint(use));
    }
    return accept;
}

// the rare tiers (mid_pair, big_phase): the screening inline (60 instructions), the reference-order test out of line.
// (A fully out-of-line copy of mh_accept cost 2 300 cycles per MID pass: 320 cold instructions for a test that the
// screening decides 99.9995 % of the time.)
template <bool VERIFY>
__device__ __forceinline__ bool mh_accept_rare(const SspsDev& P, const int action, const int k, const int kn, const double dprior,
                                               const double fwd, const double bwd, const double ip_cur, const float lg_ip_cur,
                                               const double ip_new, const float lg_new, const double u_acc, unsigned long long* stats) {
    const float lgu = fast_lg2((float)u_acc);
    const double hn = 0.5 * P.dn;
    const double base = ((dprior - fwd) + bwd) - 0.5 * (double)(kn - k) * P.log1pn;
    const double diff = SSPS_LN2 * ((double)lgu + hn * ((double)lg_new - (double)lg_ip_cur)) - base;
    const double marg = 1.2e-6 * ((double)fabsf(lgu) + 1.0 + hn * ((double)(fabsf(lg_new) + fabsf(lg_ip_cur)) + 2.0)) + 1e-9 * fabs(base) + 1e-6;
    bool accept = diff < 0.0;
    const bool close = !(u_acc > 1e-30 && fabs(diff) > marg);
    if (close || VERIFY) {
        const bool ex = mh_accept_exact(P.log1pn, P.dn, action, k, kn, dprior, fwd, bwd, ip_cur, ip_new, u_acc) != 0;
        if (close) { atomicAdd(stats + ST_TIER1, 1ull); accept = ex; }
        else {
            if (ex != accept) atomicAdd(stats + ST_DEC_MISMATCH, 1ull);
            // how much of its margin the screening used (a flip needs a ratio above 1); the exact difference is re-formed here
            const double lml_cur = -0.5 * (double)k * P.log1pn - 0.5 * P.dn * log(ip_cur);
            const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
            const double w = dprior + (lml_new - lml_cur);
            const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
            atomicMax((unsigned*)(stats + ST_MARGIN_USE), __float_as_uint((float)(fabs(diff - (log(u_acc) - alpha)) / marg)));
        }
    }
    return accept;
}

template <bool VERIFY>
__device__ __noinline__ bool mh_accept_ool(const SspsDev& P, const int action, const int k, const int kn, const double dprior,
                                           const double fwd, const double bwd, const double ip_cur, const float lg_ip_cur,
                                           const double ip_new, const float lg_new, const double u_acc, unsigned long long* stats) {
    return mh_accept<VERIFY>(P, action, k, kn, dprior, fwd, bwd, ip_cur, lg_ip_cur, ip_new, lg_new, u_acc, stats);
}

__device__ __forceinline__ void lane_set_q(const SspsDev& P, int child, Lane& L, double q) {
    L.ip = __ldg(P.yty + child) - P.shrink * q; L.lg_ip = fast_lg2((float)L.ip);
}

// ------------------------------------------------------------------------------------------------
// SMALL tier: factor of the current set from its id list (task start, arrivals from the BIG tier).
// Operation order = orc_lml_chol: l_b = (G_ab - sum_{m<b} l_m L_bm) / d_b, d2 = G_aa - sum l_m^2,
// t = r_a - sum l_m w_m, each sum in ascending m with fma.
// ------------------------------------------------------------------------------------------------
__device__ __noinline__ double small_build(const SspsDev& P, const double* __restrict__ Ri, const SmallAcc& S, const int k) {
    double q = 0.0;
    for (int a = 0; a < k; ++a) {
        const int ja = (int)(S.W(a) & 0xffffu);
        const double* __restrict__ Ga = P.G + (size_t)ja * P.V;
        const double gaa = __ldg(Ga + ja);
        double d2 = gaa, t = __ldg(Ri + ja);
        for (int b = 0; b < a; ++b) {
            double s = __ldg(Ga + (int)(S.W(b) & 0xffffu));
            for (int m = 0; m < b; ++m) s = fma(-S.Lo(a, m), S.Lo(b, m), s);
            const double l = s * S.Li(b);
            S.Lo(a, b) = l;
            d2 = fma(-l, l, d2);
            t = fma(-l, S.Wv(b), t);
        }
        double inv, w;
        chol_pivot(d2, gaa, t, inv, w, q);
        S.Li(a) = inv; S.Wv(a) = w;
    }
    return q;
}

// ------------------------------------------------------------------------------------------------
// The parent-set move of a SMALL lane (|S| <= SSPS_KS), statically unrolled over SSPS_KS slots with
// the proposed set's factor held in REGISTERS, so that the lanes of a warp (different k, different
// removal position p) walk one instruction stream: rows/entries a lane does not need are
// predicated off, rows nobody needs are skipped by a warp-uniform branch.  NL(a, b) = entry (a, b)
// of the factor in the PROPOSED set's coordinates (arrival positions after the removal): rows < p
// and the entries < p of later rows are the cached values, the rest is recomputed -- exactly the
// rows a from-scratch factorisation of the proposed set (in its arrival order) would compute; the
// appended parent's row is jl[].
// An accepted arrival that outgrows the tier leaves the lane in MODE_PROMOTE (ids and factor still
// those of the current set, L.pend_j = the arriving parent); the warp finishes it in big_phase.
// ------------------------------------------------------------------------------------------------
// the step's proposal as drawn (select_move) with its proposal-ratio terms (move_ratio): formed once per step for
// every live lane in the main loop, so that the rare tiers do not carry their own copies of that code
struct Move { int action, a_ord, r_ord; double fwd, bwd; };

#define NL(a, b) nl[SSPS_TRI(a) - (a) + (b)]       // strictly lower triangle: row a starts at a(a-1)/2

template <bool VERIFY>
__device__ __forceinline__ int small_move(const SspsDev& P, const int child, const StepDraws& d, const Move& mv, Lane& L, const SmallAcc& S,
                                          const KCtx& C, const int rel_now, const unsigned peers, int& kn_out) {
    constexpr int KS = SSPS_KS;
    const int V = P.V, k = L.k;
    const int action = mv.action, a_ord = mv.a_ord, r_ord = mv.r_ord;
    const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;

    uint32_t idw[KS]; int rk[KS];
#pragma unroll
    for (int b = 0; b < KS; ++b) {
        idw[b] = 0xffffu; rk[b] = -1;
        if (b < k) { idw[b] = S.W(b); rk[b] = S.rank(b); }
    }
    int j_add = -1, j_rem = -1, p = k, rp = 1 << 20;
    uint32_t w_rem = 0;
    if (has_add) {          // ith_nonparent on the sorted list: the a-th id not in S
        int c = 0;
#pragma unroll
        for (int b = 0; b < KS; ++b) c += (b < k && (int)(idw[b] & 0xffffu) - rk[b] < a_ord) ? 1 : 0;
        j_add = a_ord - 1 + c;
    }
    if (has_rem) {          // the r-th element of the sorted list
        rp = r_ord - 1;
#pragma unroll
        for (int b = 0; b < KS; ++b) if (rk[b] == rp) { p = b; w_rem = idw[b]; j_rem = (int)(w_rem & 0xffffu); }      // (no dynamic index: idw stays in registers)
    }
    const int kr = k - (has_rem ? 1 : 0);
    const int kn = kr + (has_add ? 1 : 0);
    kn_out = kn;
    const double fwd = mv.fwd, bwd = mv.bwd;
    const double* __restrict__ omc_i = P.omc + (size_t)V * child;
    const double* __restrict__ Ri = P.R + (size_t)V * child;
    double omc_add = 0.0, omc_rem = 0.0, dprior = 0.0;
    if (has_add) { omc_add = __ldg(omc_i + j_add); dprior = -L.lam * omc_add; }
    if (has_rem) { omc_rem = __ldg(omc_i + j_rem); dprior = dprior + L.lam * omc_rem; }

    // ids of the reduced set in its own arrival coordinates (a valid dummy id beyond kr keeps gathers in range)
    const int dummy = has_add ? j_add : 0;
    int nid[KS];
#pragma unroll
    for (int a = 0; a < KS; ++a) {
        const int ida = (int)(idw[a] & 0xffffu);
        const int idn = (a + 1 < KS) ? (int)(idw[(a + 1 < KS) ? a + 1 : a] & 0xffffu) : dummy;
        nid[a] = (a < kr) ? ((a >= p) ? idn : ida) : dummy;
    }

#if SSPS_EARLY_GATHER
    // the arriving parent's gathers, issued now: their latency hides behind the factor loads and the recomputed rows
    double gj[KS], gjj_e = 0.0, rj_e = 0.0;
#pragma unroll
    for (int b = 0; b < KS; ++b) gj[b] = 0.0;
    if (has_add) {
        const double* __restrict__ Gj = P.G + (size_t)j_add * V;
        gjj_e = __ldg(Gj + j_add); rj_e = __ldg(Ri + j_add);
#pragma unroll
        for (int b = 0; b < KS; ++b) gj[b] = __ldg(Gj + nid[b]);
    }
#endif
    // the cached factor in the reduced set's coordinates; entries to recompute start as 0
    double nl[SSPS_TRI(KS) - KS + 1], ninv[KS], nw[KS];
    const int kr_max = __reduce_max_sync(peers, kr);
    const int p_min = __reduce_min_sync(peers, p);
    if (kn > C.kcap) { atomicAdd(C.stats + ST_KCAP_REJECT, 1ull); return 0; }      // (after the warp collectives)
#pragma unroll
    for (int a = 0; a < KS; ++a) {
        ninv[a] = 0.0; nw[a] = 0.0;
#pragma unroll
        for (int m = 0; m < a; ++m) NL(a, m) = 0.0;
        if (a < kr_max) {
            if (a < kr) {
                if (a < p) {
                    ninv[a] = S.Li(a); nw[a] = S.Wv(a);
#pragma unroll
                    for (int m = 0; m < a; ++m) NL(a, m) = S.Lo(a, m);
                } else if (a + 1 < KS) {
#pragma unroll
                    for (int m = 0; m < a; ++m) if (m < p) NL(a, m) = S.Lo(a + 1, m);
                }
            }
        }
    }

    // rows [p, kr) of the reduced set: recomputed exactly as a from-scratch factorisation would
#pragma unroll
    for (int a = 0; a < KS - 1; ++a) {
        if (a >= p_min && a < kr_max) {
            if (a >= p && a < kr) {
                const double* __restrict__ Ga = P.G + (size_t)nid[a] * V;
                const double gaa = __ldg(Ga + nid[a]);
                double d2 = gaa, t = __ldg(Ri + nid[a]);
#pragma unroll
                for (int b = 0; b < a; ++b) {
                    double s = __ldg(Ga + nid[b]);
#pragma unroll
                    for (int m = 0; m < b; ++m) s = fma(-NL(a, m), NL(b, m), s);
                    const double l = (b >= p) ? s * ninv[b] : NL(a, b);
                    NL(a, b) = l;
                    d2 = fma(-l, l, d2);
                    t = fma(-l, nw[b], t);
                }
                if (d2 > SSPS_PIVOT_TOL * gaa) {
                    const double invd = inv_sqrt_ieee(d2);
                    ninv[a] = invd; nw[a] = t * invd;
                } else if (d2 != d2 || gaa != gaa) {
                    ninv[a] = __longlong_as_double(0x7ff8000000000000LL); nw[a] = ninv[a];
                } else { ninv[a] = 0.0; nw[a] = 0.0; }
            }
        }
    }
    double q_new = 0.0;
#pragma unroll
    for (int m = 0; m < KS; ++m) if (m < kr) q_new = fma(nw[m], nw[m], q_new);

    // the appended parent's row
    double jl[KS], jinv = 0.0, jw = 0.0;
#pragma unroll
    for (int b = 0; b < KS; ++b) jl[b] = 0.0;
    if (has_add) {
#if SSPS_EARLY_GATHER
        const double gjj = gjj_e;
        double d2 = gjj, t = rj_e;
#else
        const double* __restrict__ Gj = P.G + (size_t)j_add * V;
        const double gjj = __ldg(Gj + j_add);
        double d2 = gjj, t = __ldg(Ri + j_add);
#endif
#pragma unroll
        for (int b = 0; b < KS; ++b) {
            if (b < kr_max) {
#if SSPS_EARLY_GATHER
                double s = gj[b];
#else
                double s = __ldg(Gj + nid[b]);
#endif
#pragma unroll
                for (int m = 0; m < b; ++m) s = fma(-jl[m], NL(b, m), s);
                const double l = (b < kr) ? s * ninv[b] : 0.0;
                jl[b] = l;
                if (b < kr) { d2 = fma(-l, l, d2); t = fma(-l, nw[b], t); }
            }
        }
        chol_pivot(d2, gjj, t, jinv, jw, q_new);
    }
    const double ip_new = C.yty - P.shrink * q_new;
    const float lg_new = fast_lg2((float)ip_new);
    const bool accept = mh_accept<VERIFY>(P, action, k, kn, dprior, fwd, bwd, L.ip, L.lg_ip, ip_new, lg_new, d.u_acc, C.stats);

    if (VERIFY) {       // the proposed set's q from scratch, in its arrival order: must be the same bits
        int t = 0;
        const Scratch xs = C.xs_of(threadIdx.x & 31);
#pragma unroll
        for (int b = 0; b < KS; ++b) if (b < k && b != p) xs.T(t++) = (uint16_t)(idw[b] & 0xffffu);
        if (j_add >= 0) xs.T(t++) = (uint16_t)j_add;
        const double lml_ref = score_set(P, child, xs, t, nullptr);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        if (!(lml_ref == lml_new || (lml_ref != lml_ref && lml_new != lml_new))) atomicAdd(C.stats + ST_MISMATCH, 1ull);
    }
    if (!accept) return 0;

    // ---- accepted: tallies / change points ----
    if (has_rem) {
        const int credit = rel_now - (int)(w_rem >> 16);
        if (credit > 0) atomicAdd(C.cnt_row + j_rem, (uint32_t)credit);     // RED: the lane owns the counter, nothing waits for the old value
        if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_rem | ((uint32_t)(rel_now + 1) << 16);
    }
    if (has_add && C.pev && L.n_pev < C.pev_cap)
        C.pev[L.n_pev++] = (uint32_t)j_add | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
    L.somc = (L.somc - omc_rem) + omc_add;
    L.ip = ip_new; L.lg_ip = lg_new;
    if (kn > KS) {      // an arrival that outgrows shared memory (k == KS, pure add): the lane moves itself to the BIG tier
        const BigRef B = C.big(threadIdx.x & 31);
        for (int e = 0; e < SSPS_TRI(KS); ++e) B.f[e] = S.d[e * SSPS_BLOCK];
        double* __restrict__ fw = B.fw();
        int rj = 0;
        for (int a = 0; a < KS; ++a) {
            const uint32_t w = S.W(a);
            int r = S.rank(a);
            if ((int)(w & 0xffffu) < j_add) ++rj; else ++r;
            B.w[a] = w; B.r[a] = (uint16_t)r; fw[a] = S.Wv(a);
        }
#pragma unroll
        for (int m = 0; m < KS; ++m) B.f[SSPS_TRI(KS) + m] = jl[m];
        B.f[SSPS_TRI(KS) + KS] = jinv; fw[KS] = jw;
        B.w[KS] = (uint32_t)j_add | ((uint32_t)rel_now << 16); B.r[KS] = (uint16_t)rj;
        L.mode = MODE_BIG; L.k = kn;
        return 1;
    }
    // ---- factor and id list back to shared memory ----
    int rj = 0;
#pragma unroll
    for (int a = 0; a < KS; ++a) {
        if (a < kr) {
            // slot a of the reduced set = old slot a (a < p) or a + 1
            // (read back from shared memory rather than held in registers across the factor work; slot a + 1 is still the old one)
            const uint32_t w_old = (a >= p && a + 1 < KS) ? S.W((a + 1 < KS) ? a + 1 : a) : S.W(a);
            int r_old = (a >= p && a + 1 < KS) ? S.rank((a + 1 < KS) ? a + 1 : a) : S.rank(a);
            r_old -= (r_old > rp) ? 1 : 0;
            if (has_add) { if ((int)(w_old & 0xffffu) < j_add) ++rj; else ++r_old; }
            if (a >= p) {
                S.W(a) = w_old;
                S.Li(a) = ninv[a]; S.Wv(a) = nw[a];
#pragma unroll
                for (int m = 0; m < a; ++m) S.Lo(a, m) = NL(a, m);
            }
            S.set_rank(a, r_old);
        }
    }
    if (has_add) {      // row kr (dynamic position)
#pragma unroll
        for (int m = 0; m < KS; ++m) if (m < kr) S.Lo(kr, m) = jl[m];
        S.Li(kr) = jinv; S.Wv(kr) = jw;
        S.W(kr) = (uint32_t)j_add | ((uint32_t)rel_now << 16);
        S.set_rank(kr, rj);
    }
    L.k = kn;
    return 1;
}
#undef NL

// ------------------------------------------------------------------------------------------------
// MID: the parent-set move of BIG-tier lanes whose current set has fewer than SSPS_KM members (6 .. 15 parents: 1.1 % of the
// proposals at the metric point, most of them k = 6 or 7).  TWO such lanes per call, one per HALF-WARP: the 16 lanes of a half
// copy their subject's factor from its arena block into a slot of the warp's shared-memory Stage and recompute the rows
// [p, kn) of the proposed set there IN PLACE (new row a overwrites old row a, whose only reader -- new row a - 1 -- is done),
// row by row, a lane per COLUMN: at step b lane b hands  l_ab = s_b / d_b  to the half by a 16-wide shuffle and every lane
// c > b subtracts  l_ab l_cb  from its s_c -- a forward substitution, each entry receiving its terms in ascending b
// through fma: the operation order of small_move / fac_append / the oracle, the same bits.  Entries left of the removed
// column p are the cached ones.  ids, ranks and the arena are touched only on accept.
// Round 1 / 2a let every MID lane do its own move serially (mid_move): 13 500 cycles per move measured (300 instructions per
// recomputed row from ONE lane while 31 wait), a quarter to a third of all task cycles for 1 % of the proposals; a row costs
// the half-warp  ~50 cycles per column + one 1/sqrt  instead.
// Slot layout: TRI-packed rows (1/d on the diagonal), w behind SSPS_TRI(SSPS_KM), then ids (words) and ranks (halves).
// All arguments but the lane state are warp-uniform; every lane of the warp must call.  s0 / s1 = subject lanes of half
// 0 / 1 (s1 = -1: none).  acc_kn of a subject lane receives  accepted | |S'| << 1.
// ------------------------------------------------------------------------------------------------
#ifndef SSPS_KM
#define SSPS_KM 16
#endif
#define SSPS_MID_SLOT (SSPS_ROWLEN(SSPS_KM) + SSPS_KM)            // doubles: factor, w, then ids (words) and ranks (halves)
template <bool VERIFY>
__device__ __noinline__ Lane mid_pair(const SspsDev& P, const int child, const StepDraws& d, const Move& mv, Lane L, const SmallAcc& S, const KCtx& C,
                                      const int rel_now, const int s0, const int s1, int* acc_kn) {
    const int lane = threadIdx.x & 31, half = lane >> 4, hl = lane & 15;
    const int V = P.V;
    const int src = half ? s1 : s0;                     // this half's subject lane
    const bool on = src >= 0;
    const int srcl = on ? src : s0;
    const unsigned foff = (unsigned)(C.st_smem.t - ssps_smem) + (unsigned)half * SSPS_MID_SLOT;
#ifdef SSPS_BOUNDS
#define FS(e) (*[&](int e_) -> double* { SSPS_CHK(e_ >= 0 && e_ < SSPS_ROWLEN(SSPS_KM)); return &ssps_smem[foff + e_]; }(e))
#else
#define FS(e) ssps_smem[foff + (e)]
#endif
#define FW(e) FS(SSPS_TRI(SSPS_KM) + (e))
    uint32_t* const Fid = (uint32_t*)(ssps_smem + foff + SSPS_ROWLEN(SSPS_KM));      // id | rel << 16, arrival order (the CURRENT set until the commit)
    uint16_t* const Frk = (uint16_t*)(Fid + SSPS_KM);                                 // sorted ranks
    const double* __restrict__ Ri = P.R + (size_t)V * child;
    const int k = __shfl_sync(SSPS_FULL, L.k, srcl);
    const int action = __shfl_sync(SSPS_FULL, mv.action, srcl);
    const int a_ord = __shfl_sync(SSPS_FULL, mv.a_ord, srcl);
    const int r_ord = __shfl_sync(SSPS_FULL, mv.r_ord, srcl);
    const BigRef B = C.big(srcl);
#ifdef SSPS_PROFILE_BIG
    long long tp_ = clock64();
#endif
    if (on) {       // the subject's factor, ids and ranks into the slot
        for (int e = hl; e < SSPS_TRI(k); e += 16) FS(e) = B.f[e];
        if (hl < k) { FW(hl) = B.fw()[hl]; Fid[hl] = B.w[hl]; Frk[hl] = B.r[hl]; }
    }
#ifdef SSPS_PROFILE_BIG
    { const long long tn_ = clock64(); if (lane == 0) atomicAdd(C.stats + ST_MID_PROF + 0, (unsigned long long)(tn_ - tp_)); tp_ = tn_; }
#endif
    __syncwarp();
    const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;
    const unsigned hshift = 16u * (unsigned)half;
    const bool mem = on && hl < k;
    const uint32_t myw = mem ? Fid[hl] : 0u;
    const int myid = (int)(myw & 0xffffu), myrk = mem ? (int)Frk[hl] : -1;
    int j_add = -1, j_rem = -1, p = k, rp = 1 << 20;
    uint32_t w_rem = 0;
    {   // ith_nonparent on the sorted list: the a-th id not in S; the r-th element of the sorted list
        const unsigned ba = (__ballot_sync(SSPS_FULL, mem && myid - myrk < a_ord) >> hshift) & 0xffffu;
        const unsigned br = (__ballot_sync(SSPS_FULL, mem && myrk == r_ord - 1) >> hshift) & 0xffffu;
        if (has_add) j_add = a_ord - 1 + __popc(ba);
        if (has_rem && br) { rp = r_ord - 1; p = __ffs(br) - 1; w_rem = Fid[p]; j_rem = (int)(w_rem & 0xffffu); }
    }
    const int kr = k - (has_rem ? 1 : 0);
    const int kn = kr + (has_add ? 1 : 0);
    const bool work = on && kn <= C.kcap;
    // q over the unchanged rows
    double q_new = 0.0;
    if (work) {
#pragma unroll 1
        for (int m = 0; m < p && m < kr; ++m) { const double wm = FW(m); q_new = fma(wm, wm, q_new); }
    }
#ifdef SSPS_PROFILE_BIG
    { const long long tn_ = clock64(); if (lane == 0) atomicAdd(C.stats + ST_MID_PROF + 1, (unsigned long long)(tn_ - tp_)); tp_ = tn_; }
#endif
    // All gathers of the pass first, back to back (an L2 round trip each when they miss: one latency instead of one per row):
    // lane r of the half fetches G_jj and r_j of row p + r into registers; lane c fetches the seeds G(j_a, id_c) of its column
    // for every row below it -- into the row's own slot entries (dead: they are about to be recomputed), the arriving row's
    // into the spare third slot of the Stage (its slot still holds cached entries the row above it reads).
#define SCR(e) ssps_smem[(unsigned)(C.st_smem.t - ssps_smem) + 2u * SSPS_MID_SLOT + 16u * (unsigned)half + (e)]
    double ga_own = 1.0, t_own = 0.0;
    if (work) {
        const int a_own = p + hl;
        if (a_own < kn) {
            const int ja = (a_own == kr) ? j_add : (int)(Fid[a_own + 1] & 0xffffu);
            ga_own = __ldg(P.G + (size_t)ja * V + ja); t_own = __ldg(Ri + ja);
        }
        if (hl < kr) {
            const int idc = (int)(Fid[hl + (hl >= p ? 1 : 0)] & 0xffffu);
#pragma unroll 1
            for (int a = max(p, hl + 1); a < kn; ++a) {
                const bool arr = a == kr;
                if (!arr && hl < p) continue;                   // left of the removed column: the cached entry
                const int ja = arr ? j_add : (int)(Fid[a + 1] & 0xffffu);
                const double g = __ldg(P.G + (size_t)ja * V + idc);
                if (arr) SCR(hl) = g; else FS(SSPS_TRI(a) + hl) = g;
            }
        }
    }
    __syncwarp();
    // rows [p, kn) of the proposed set; the row loop is warp-uniform (shuffles inside), a half works on the rows of its own range
    const unsigned myrow = foff + (unsigned)SSPS_TRI(hl);           // lane c = hl owns column c: l_cb is entry b of (new) row c
    double inv_own = (work && hl < p) ? ssps_smem[myrow + hl] : 0.0;        // 1/d of row hl; rows >= p get theirs as they are recomputed
    const int a_lo = work ? p : (1 << 20), a_hi = work ? kn : 0;
    const int a_min = min(a_lo, __shfl_xor_sync(SSPS_FULL, a_lo, 16)), a_max = max(a_hi, __shfl_xor_sync(SSPS_FULL, a_hi, 16));
#pragma unroll 1
    for (int a = a_min; a < a_max; ++a) {
        const bool act = a >= a_lo && a < a_hi;
        const bool arr = a == kr;                               // the arriving parent's row (nothing cached)
        const int c0 = arr ? 0 : p;
        const int xa = SSPS_TRI(a), olda = SSPS_TRI(a + 1);
        const double gaa = __shfl_sync(SSPS_FULL, ga_own, (a - p) & 15, 16);
        double t = __shfl_sync(SSPS_FULL, t_own, (a - p) & 15, 16);
        double sc = 0.0;
        if (act && hl >= c0 && hl < a) sc = arr ? SCR(hl) : FS(xa + hl);
        double d2 = gaa;
        // Step b: lane b scales its own s_b by ITS row's 1/d (held in a register: no load on the chain) and hands it round;
        // the dependent chain per step is  multiply -> shuffle -> fma.
#pragma unroll 1
        for (int b = 0; b < a; ++b) {
            const double lsh = __shfl_sync(SSPS_FULL, sc * inv_own, b, 16);
            if (act) {
                const double l = (b < c0) ? FS(olda + b) : lsh;
                d2 = fma(-l, l, d2); t = fma(-l, FW(b), t);
                if (hl > b && hl >= c0 && hl < a) sc = fma(-l, ssps_smem[myrow + b], sc);
                if (hl == b) FS(xa + b) = l;
            }
        }
        if (act) {
            double inv = 0.0, w = 0.0;
            if (d2 > SSPS_PIVOT_TOL * gaa) { inv = inv_sqrt_ieee(d2); w = t * inv; }
            else if (d2 != d2 || gaa != gaa) { inv = __longlong_as_double(0x7ff8000000000000LL); w = inv; }
            q_new = fma(w, w, q_new);
            if (hl == 0) { FS(xa + a) = inv; FW(a) = w; }
            if (hl == a) inv_own = inv;
        }
        __syncwarp();
    }
#ifdef SSPS_PROFILE_BIG
    { const long long tn_ = clock64(); if (lane == 0) atomicAdd(C.stats + ST_MID_PROF + 3, (unsigned long long)(tn_ - tp_)); tp_ = tn_; }
#endif
    // ---- the subject lanes: acceptance (they may sit in either half: the move's numbers come from lane 0 of the half that did the work) ----
    const int wh = (lane == s1) ? 16 : 0;
    const double q_s = __shfl_sync(SSPS_FULL, q_new, wh);
    const int jadd_s = __shfl_sync(SSPS_FULL, j_add, wh), jrem_s = __shfl_sync(SSPS_FULL, j_rem, wh);
    const int kn_s = __shfl_sync(SSPS_FULL, kn, wh);
    const uint32_t wrem_s = __shfl_sync(SSPS_FULL, w_rem, wh);
    const int work_s = __shfl_sync(SSPS_FULL, work ? 1 : 0, wh);
    int acc = 0;
    if (lane == s0 || lane == s1) {
        const int ks = L.k, act_s = mv.action;
        const bool add_s = act_s != ACT_REM, rem_s = act_s != ACT_ADD;
        if (!work_s) { atomicAdd(C.stats + ST_KCAP_REJECT, 1ull); *acc_kn = kn_s << 1; }
        else {
            const double* __restrict__ omc_i = P.omc + (size_t)V * child;
            double omc_add = 0.0, omc_rem = 0.0, dprior = 0.0;
            if (add_s) { omc_add = __ldg(omc_i + jadd_s); dprior = -L.lam * omc_add; }
            if (rem_s) { omc_rem = __ldg(omc_i + jrem_s); dprior = dprior + L.lam * omc_rem; }
            const double ip_new = C.yty - P.shrink * q_s;
            const float lg_new = fast_lg2((float)ip_new);
            acc = mh_accept_rare<VERIFY>(P, act_s, ks, kn_s, dprior, mv.fwd, mv.bwd, L.ip, L.lg_ip, ip_new, lg_new, d.u_acc, C.stats) ? 1 : 0;
            if (VERIFY) {       // the proposed set's q from scratch, in its arrival order: must be the same bits
                const uint32_t* ids_s = (const uint32_t*)(ssps_smem + (unsigned)(C.st_smem.t - ssps_smem) + (unsigned)(wh >> 4) * SSPS_MID_SLOT + SSPS_ROWLEN(SSPS_KM));
                int tt = 0;
                const Scratch xs = C.xs_of(lane);
#pragma unroll 1
                for (int b = 0; b < ks; ++b) if (!rem_s || (int)(ids_s[b] & 0xffffu) != jrem_s) xs.T(tt++) = (uint16_t)(ids_s[b] & 0xffffu);
                if (add_s) xs.T(tt++) = (uint16_t)jadd_s;
                const double lml_ref = score_set(P, child, xs, tt, nullptr);
                const double lml_new = -0.5 * (double)kn_s * P.log1pn - 0.5 * P.dn * log(ip_new);
                if (!(lml_ref == lml_new || (lml_ref != lml_ref && lml_new != lml_new))) atomicAdd(C.stats + ST_MISMATCH, 1ull);
            }
            if (acc) {
                if (rem_s) {
                    const int credit = rel_now - (int)(wrem_s >> 16);
                    if (credit > 0) atomicAdd(C.cnt_row + jrem_s, (uint32_t)credit);     // RED: the lane owns the counter, nothing waits for the old value
                    if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)jrem_s | ((uint32_t)(rel_now + 1) << 16);
                }
                if (add_s && C.pev && L.n_pev < C.pev_cap)
                    C.pev[L.n_pev++] = (uint32_t)jadd_s | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
                L.somc = (L.somc - omc_rem) + omc_add;
                L.ip = ip_new; L.lg_ip = lg_new;
                L.k = kn_s;
                if (kn_s <= SSPS_KS) L.mode = MODE_SMALL;
            }
            *acc_kn = acc | (kn_s << 1);
        }
    }
#ifdef SSPS_PROFILE_BIG
    { const long long tn_ = clock64(); if (lane == 0) atomicAdd(C.stats + ST_MID_PROF + 4, (unsigned long long)(tn_ - tp_)); tp_ = tn_; }
#endif
    // ---- commit (the half that did the work): ids with the gap at p closed and the sorted ranks shifted, rows [p, kn) ----
    const int acc_h = __shfl_sync(SSPS_FULL, acc, srcl);
    if (on && acc_h) {
        uint32_t wv = 0; int rv = 0;
        if (hl < kr) {
            const int oa = hl + (hl >= p ? 1 : 0);
            wv = Fid[oa]; rv = (int)Frk[oa];
            rv -= (rv > rp) ? 1 : 0;
        }
        const unsigned lt = (__ballot_sync(0xffffu << hshift, hl < kr && (int)(wv & 0xffffu) < j_add) >> hshift) & 0xffffu;
        if (has_add) { if (hl < kr && !((int)(wv & 0xffffu) < j_add)) ++rv; }
        if (hl == kr && has_add) { wv = (uint32_t)j_add | ((uint32_t)rel_now << 16); rv = __popc(lt); }
        if (kn <= SSPS_KS) {        // a removal that brings the set back into the SMALL tier
            const SmallAcc Ss = S.of_lane(src - lane);
            for (int e = hl; e < SSPS_TRI(kn); e += 16) Ss.d[e * SSPS_BLOCK] = FS(e);
            if (hl < kn) { Ss.Wv(hl) = FW(hl); Ss.W(hl) = wv; Ss.set_rank(hl, rv); }
        } else {
            if (hl < kn) { B.w[hl] = wv; B.r[hl] = (uint16_t)rv; }
            double* __restrict__ fw = B.fw();
            for (int e = SSPS_TRI(p) + hl; e < SSPS_TRI(kn); e += 16) B.f[e] = FS(e);
            if (hl >= p && hl < kn) fw[hl] = FW(hl);
        }
    }
#ifdef SSPS_PROFILE_BIG
    { const long long tn_ = clock64(); if (lane == 0) atomicAdd(C.stats + ST_MID_PROF + 5, (unsigned long long)(tn_ - tp_)); tp_ = tn_; }
#endif
    __syncwarp();
#ifdef SSPS_PROFILE_BIG
    if (lane == 0) { atomicAdd(C.stats + ST_MID_PROF + 6, (unsigned long long)max(a_max - a_min, 0)); atomicAdd(C.stats + ST_MID_PROF + 7, 1ull); }
#endif
    if (lane == 0) atomicAdd(C.stats + ST_MID_STEPS, (unsigned long long)(1 + (s1 >= 0 ? 1 : 0)));
    return L;
#undef FS
#undef FW
#undef SCR
}

// ------------------------------------------------------------------------------------------------
// BIG tier, warp-cooperative.  All arguments are warp-uniform; every lane of the warp must call.
//
// The warp copies the lane's state into its Stage, applies the move to the copy IN PLACE
// (ids_remove / fac_remove, then fac_append / ids_append) and reads off q of the proposed set; the
// copy goes back to the arena (or to the lane's SMALL slots) only if the move is accepted.
//
// The rows after a removal are refactorised RIGHT-LOOKING: once column b is final, every lane
// subtracts its product from the entries (a, c), c > b, it owns.  Each entry therefore receives
// the terms of  G_ac - sum_m l_am l_cm  in ascending m through fma -- the operation order of the
// row-by-row factorisation in orc_lml_chol -- so the result is the same bit pattern.
// ------------------------------------------------------------------------------------------------
// The factor in the Stage stays in the coordinates it was loaded in.  A removal of arrival slot p only
// marks row p and column p DEAD (pd = p; pd = SSPS_NODEAD when nothing was removed): position a of the
// proposed set lives in stage row / column a + (a >= pd).  Nothing is moved until the copy leaves the
// Stage (stage_store / stage_to_small compact it on the way out).  The id list IS compacted at once.
#define SSPS_NODEAD (1 << 20)
__device__ __forceinline__ int st_map(const int a, const int pd) { return a + (a >= pd ? 1 : 0); }

__device__ __forceinline__ void stage_load(const Stage T, const BigRef B, const int k, const int lane) {
    for (int e = lane; e < SSPS_TRI(k); e += 32) T.t[e] = B.f[e];
    const double* __restrict__ fw = B.fw();
    for (int a = lane; a < k; a += 32) { T.tw[a] = fw[a]; T.id[a] = B.w[a]; T.rk[a] = B.r[a]; }
    __syncwarp();
}
// kn rows of the proposed set -> the lane's arena block
__device__ __forceinline__ void stage_store(const Stage T, const BigRef B, const int kn, const int pd, const int lane) {
    for (int a = 0; a < kn; ++a) {
        const double* __restrict__ row = T.t + SSPS_TRI(st_map(a, pd));
        double* __restrict__ dst = B.f + SSPS_TRI(a);
        for (int m = lane; m <= a; m += 32) dst[m] = row[st_map(m, pd)];
    }
    double* __restrict__ fw = B.fw();
    for (int a = lane; a < kn; a += 32) { fw[a] = T.tw[st_map(a, pd)]; B.w[a] = T.id[a]; B.r[a] = T.rk[a]; }
    __syncwarp();
}
// SMALL-tier slots of one lane <-> Stage
__device__ __forceinline__ void stage_from_small(const Stage T, const SmallAcc Ss, const int k, const int lane) {
    for (int e = lane; e < SSPS_TRI(k); e += 32) T.t[e] = Ss.d[e * SSPS_BLOCK];
    if (lane < k) { T.tw[lane] = Ss.Wv(lane); T.id[lane] = Ss.W(lane); T.rk[lane] = (uint16_t)Ss.rank(lane); }
    __syncwarp();
}
__device__ __forceinline__ void stage_to_small(const Stage T, const SmallAcc Ss, const int kn, const int pd, const int lane) {
    for (int a = 0; a < kn; ++a) {
        const double* __restrict__ row = T.t + SSPS_TRI(st_map(a, pd));
        if (lane <= a) Ss.d[(SSPS_TRI(a) + lane) * SSPS_BLOCK] = row[st_map(lane, pd)];
    }
    if (lane < kn) { Ss.Wv(lane) = T.tw[st_map(lane, pd)]; Ss.W(lane) = T.id[lane]; Ss.set_rank(lane, (int)T.rk[lane]); }
    __syncwarp();
}

// id list: drop arrival slot p, close the gap in the sorted ranks
__device__ __forceinline__ void ids_remove(const Stage T, const int k, const int p, const int lane) {
    const int rp = (int)T.rk[p];
    __syncwarp();
    for (int base = 0; base < k; base += 32) {
        const int b = base + lane;
        uint32_t wv = 0; int rv = 0;
        if (b < k - 1) {
            const int from = b >= p ? b + 1 : b;
            wv = T.id[from]; rv = (int)T.rk[from];
            rv -= (rv > rp) ? 1 : 0;
        }
        __syncwarp();
        if (b < k - 1) { T.id[b] = wv; T.rk[b] = (uint16_t)rv; }
        __syncwarp();
    }
}
// id list: parent j arrives in slot kr
__device__ __forceinline__ void ids_append(const Stage T, const int kr, const int j, const int rel_now, const int lane) {
    int rj = 0;
    for (int base = 0; base < kr; base += 32) {
        const int b = base + lane;
        const bool lt = b < kr && T.idof(b) < j;
        rj += __popc(__ballot_sync(SSPS_FULL, lt));
        if (b < kr && !lt) T.rk[b] = (uint16_t)(T.rk[b] + 1);
    }
    if (lane == 0) { T.id[kr] = (uint32_t)j | ((uint32_t)rel_now << 16); T.rk[kr] = (uint16_t)rj; }
    __syncwarp();
}

// pair number idx -> (u, v), 0 <= u <= v, idx = v (v + 1) / 2 + u: the pairs of a triangle with n rows are idx < n (n + 1) / 2
__device__ __forceinline__ void tri_decode(const int idx, int& u, int& v) {
    int vv = (int)((sqrtf((float)(8 * idx + 1)) - 1.0f) * 0.5f);
    while (vv * (vv + 1) / 2 > idx) --vv;
    while ((vv + 1) * (vv + 2) / 2 <= idx) ++vv;
    v = vv; u = idx - vv * (vv + 1) / 2;
}

// Factor after the removal of arrival slot p of k: rows p .. k-2 of the reduced set (stage rows p+1 .. k-1;
// T.id already holds the reduced list).  q_in = sum of w_m^2 over m < p; returns q over all k - 1 rows.
//
// Phase A: every entry (a, c), p <= c <= a, of the trailing triangle is seeded with  G_ac - sum_{m<p} l_am l_cm  (the
// unchanged columns, ascending m), the diagonal pair of row a also with  t_a = r_a - sum_{m<p} l_am w_m.  The pairs are
// spread over the lanes; the G / R gathers of all of a lane's pairs are issued first, back to back (an L2 round trip
// each when they miss: one latency instead of one per pair).
// Phase B: right-looking elimination of the trailing triangle, a lane per ROW.  Step b: every lane forms the pivot
// (1 / d_b, w_b) from d_b^2 and t_b, which the owner of row b handed to the warp by shuffle at the end of the previous
// step; lane a scales its entry l_ab (loaded BEFORE the 1/sqrt chain, so the load hides behind it) and the owner of
// row b + 1 forms the next pivot's inputs at once -- the dependent chain from one pivot to the next is
// 1/sqrt -> multiply -> fma -> shuffle and nothing else.  Then the column of scaled entries goes round by shuffle: for
// c = b + 1 ..: l_cb comes from lane c, and lane a subtracts l_ab l_cb from its entry (a, c).  No barrier, no shared-
// memory traffic other than a lane's own row.  Every entry receives its terms in ascending b through fma -- the bits
// of the row-by-row factorisation.  (Round 1: three barriers per step, 700 cycles; round 2, first version: one barrier,
// every lane re-reading and re-scaling the other rows' column entries from shared memory: 1 680 cycles per step
// measured in the time-to-convergence regime, 68 % of a BIG-tier pass.)
// Rows beyond the first 32 of the trailing block (sets of more than ~40 parents) take the general path below.
template <bool SH>
__device__ __noinline__ double fac_remove_impl(const SspsDev& P, const double* __restrict__ Ri, const Stage T, const int k, const int p,
                                               double q, const int lane
#ifdef SSPS_PROFILE_BIG
                                               , unsigned long long* prof
#endif
                                               ) {
    const int V = P.V, kr = k - 1, nt = kr - p;
    const StageAcc<SH> A(T);
    if (nt <= 0) return q;
#ifdef SSPS_PROFILE_BIG
    const long long tp0_ = clock64();
#endif
#define SROW(a) SSPS_TRI((a) + 1)            /* a >= p: stage row a + 1 (offset of its first entry) */
#define SCOL(c) ((c) + 1)                    /* c >= p: stage column c + 1 */
    // ---- phase A ----
    const int npairs = nt * (nt + 1) / 2;
    for (int idx = lane; idx < npairs; idx += 32) {         // gathers first
        int u, v;
        tri_decode(idx, u, v);
        const int a = kr - 1 - u, c = kr - 1 - v;
        const int ja = T.idof(a);
        A.t(SROW(a) + SCOL(c)) = __ldg(P.G + (size_t)ja * V + T.idof(c));
        if (c == a) A.w(a + 1) = __ldg(Ri + ja);
    }
    for (int idx = lane; idx < npairs; idx += 32) {         // the lane re-reads its own seeds: no barrier needed
        int u, v;
        tri_decode(idx, u, v);
        const int a = kr - 1 - u, c = kr - 1 - v;
        const int ra = SROW(a), rc = SROW(c);
        double x = A.t(ra + SCOL(c));
        if (c == a) {
            double xt = A.w(a + 1);
            for (int m = 0; m < p; ++m) { const double l = A.t(ra + m); x = fma(-l, l, x); xt = fma(-l, A.w(m), xt); }
            A.w(a + 1) = xt;
        } else {
            for (int m = 0; m < p; ++m) x = fma(-A.t(ra + m), A.t(rc + m), x);
        }
        A.t(ra + SCOL(c)) = x;
    }
    __syncwarp();
#ifdef SSPS_PROFILE_BIG
    const long long tp1_ = clock64();
#endif
    // ---- phase B ----
    if (nt <= 32) {
        const int a = p + lane;
        const bool mine = a < kr;
        const int ra = SROW(mine ? a : p);
        double gd = 0.0, tw_a = 0.0;
        if (mine) { const int ja = T.idof(a); gd = __ldg(P.G + (size_t)ja * V + ja); tw_a = A.w(a + 1); }
        double d2 = A.t(SROW(p) + SCOL(p)), tb = A.w(p + 1);
        const int a_eff = mine ? a : -1;                    // rows of the other lanes: no entry passes  c <= a_eff
        for (int b = p; b < kr; ++b) {
            const bool below = a_eff > b;
            int rb = ra + SCOL(b);                                              // entry (a, b); (a, b + 1), .. follow
            const double xab = below ? A.t(rb) : 0.0;                           // before the 1/sqrt chain
            const double xnn = (a_eff == b + 1) ? A.t(rb + 1) : 0.0;            // the next diagonal, too
            const double gbb = __shfl_sync(SSPS_FULL, gd, b - p);
            double inv, wb;
            chol_pivot(d2, gbb, tb, inv, wb, q);
            const double lab = xab * inv;                                       // 0 for the lanes at or above the pivot row
            const double nlab = -lab;
            tw_a = fma(nlab, wb, tw_a);
            // the next pivot's inputs, from the owner of row b + 1 (its diagonal entry is updated again, identically, in the loop below)
            const double d2n = __shfl_sync(SSPS_FULL, fma(nlab, lab, xnn), (b + 1 - p) & 31);
            const double tn = __shfl_sync(SSPS_FULL, tw_a, (b + 1 - p) & 31);
            const int a_lim = below ? a : -1;
            int src = b + 1 - p;
#pragma unroll 4
            for (int c = b + 1; c < kr; ++c, ++src) {
                const double lcb = __shfl_sync(SSPS_FULL, lab, src);
                ++rb;
                if (c <= a_lim) A.t(rb) = fma(nlab, lcb, A.t(rb));
            }
            rb = ra + SCOL(b);
            if (below) A.t(rb) = lab;
            if (a_eff == b) { A.t(rb) = inv; A.w(b + 1) = wb; }
            d2 = d2n; tb = tn;
        }
        __syncwarp();
    } else {
        // general path: a lane owns rows p + lane + 32 u; one warp barrier per step (the scaled column goes to shared memory after it)
        double gd = 0.0;                                    // G_aa of this lane's first row: the pivots' tolerance reference
        if (p + lane < kr) { const int ja = T.idof(p + lane); gd = __ldg(P.G + (size_t)ja * V + ja); }
        double d2 = A.t(SROW(p) + SCOL(p)), tb = A.w(p + 1);
        for (int b = p; b < kr; ++b) {
            double gbb = __shfl_sync(SSPS_FULL, gd, (b - p) & 31);
            if (b - p >= 32) { const int jb = T.idof(b); gbb = __ldg(P.G + (size_t)jb * V + jb); }
            double inv, wb;
            chol_pivot(d2, gbb, tb, inv, wb, q);
            const int own = (b + 1 - p) & 31;
            double d2n = 0.0, tn = 0.0;
            if (b + 1 < kr && lane == own) {
                const int rn = SROW(b + 1);
                const double l = A.t(rn + SCOL(b)) * inv;
                d2n = fma(-l, l, A.t(rn + SCOL(b + 1)));
                tn = fma(-l, wb, A.w(b + 2));
            }
            d2n = __shfl_sync(SSPS_FULL, d2n, own);
            tn = __shfl_sync(SSPS_FULL, tn, own);
            double lab_keep[SSPS_KU];
#pragma unroll
            for (int u = 0; u < SSPS_KU; ++u) {
                lab_keep[u] = 0.0;
                const int a = p + lane + 32 * u;
                if (a > b && a < kr) {
                    const int ra = SROW(a);
                    const double lab = A.t(ra + SCOL(b)) * inv;
                    for (int c = b + 1; c < a; ++c) A.t(ra + SCOL(c)) = fma(-lab, A.t(SROW(c) + SCOL(b)) * inv, A.t(ra + SCOL(c)));
                    A.t(ra + SCOL(a)) = fma(-lab, lab, A.t(ra + SCOL(a)));
                    A.w(a + 1) = fma(-lab, wb, A.w(a + 1));
                    lab_keep[u] = lab;
                }
            }
            __syncwarp();                               // nobody reads the unscaled column b any more
#pragma unroll
            for (int u = 0; u < SSPS_KU; ++u) {
                const int a = p + lane + 32 * u;
                if (a > b && a < kr) A.t(SROW(a) + SCOL(b)) = lab_keep[u];
            }
            if (lane == 0) { A.t(SROW(b) + SCOL(b)) = inv; A.w(b + 1) = wb; }
            d2 = d2n; tb = tn;
        }
        __syncwarp();
    }
#ifdef SSPS_PROFILE_BIG
    if (lane == 0) { atomicAdd(prof + 28, (unsigned long long)(tp1_ - tp0_)); atomicAdd(prof + 29, (unsigned long long)(clock64() - tp1_)); atomicAdd(prof + 30, (unsigned long long)nt); atomicAdd(prof + 31, 1ull); }
#endif
#undef SROW
#undef SCOL
    return q;
}

__device__ __forceinline__ double fac_remove(const SspsDev& P, const double* __restrict__ Ri, const Stage T, const int k, const int p,
                                             double q, const int lane
#ifdef SSPS_PROFILE_BIG
                                             , unsigned long long* prof
#endif
                                             ) {
#ifdef SSPS_PROFILE_BIG
    return __isShared(T.t) ? fac_remove_impl<true>(P, Ri, T, k, p, q, lane, prof) : fac_remove_impl<false>(P, Ri, T, k, p, q, lane, prof);
#else
    return __isShared(T.t) ? fac_remove_impl<true>(P, Ri, T, k, p, q, lane) : fac_remove_impl<false>(P, Ri, T, k, p, q, lane);
#endif
}

// Row kr of the proposed set's factor: parent j appended to its kr rows (forward substitution, lane owns
// columns lane + 32 u).  pd = dead stage row / column (SSPS_NODEAD: none).  Returns q + w_kr^2.
template <int KU>
__device__ __noinline__ double fac_append_ku(const SspsDev& P, const double* __restrict__ Ri, const Stage T, const int kr, const int pd,
                                             const int j, double q, const int lane) {
    const int V = P.V;
    double* t = T.t; double* tw = T.tw;
    const double* __restrict__ Gj = P.G + (size_t)j * V;
    const double gjj = __ldg(Gj + j);
    double d2 = gjj, tj = __ldg(Ri + j);
    double sc[KU];
    const double* myrow[KU];
#pragma unroll
    for (int u = 0; u < KU; ++u) {
        const int c = lane + 32 * u;
        sc[u] = (c < kr) ? __ldg(Gj + T.idof(c)) : 0.0;
        myrow[u] = t + SSPS_TRI(st_map(c < kr ? c : 0, pd));
    }
    double* rj = t + SSPS_TRI(st_map(kr, pd));
    for (int b = 0; b < kr; ++b) {
        const int sb = st_map(b, pd);
        const double invb = t[SSPS_TRI(sb) + sb];
        const double wb = tw[sb];
        double mine = sc[0];
        if (KU > 1) {
#pragma unroll
            for (int u = 1; u < KU; ++u) if ((b >> 5) == u) mine = sc[u];
        }
        const double l = __shfl_sync(SSPS_FULL, mine * invb, b & 31);
        d2 = fma(-l, l, d2);
        tj = fma(-l, wb, tj);
        if (lane == (b & 31)) rj[sb] = l;
#pragma unroll
        for (int u = 0; u < KU; ++u) {
            const int c = lane + 32 * u;
            if (c > b && c < kr) sc[u] = fma(-l, myrow[u][sb], sc[u]);
        }
    }
    double inv, wj;
    chol_pivot(d2, gjj, tj, inv, wj, q);
    if (lane == 0) { rj[st_map(kr, pd)] = inv; tw[st_map(kr, pd)] = wj; }
    __syncwarp();
    return q;
}

// up to 32 rows: a lane owns one column, a quarter of the code per step (the common case by far)
__device__ __forceinline__ double fac_append(const SspsDev& P, const double* __restrict__ Ri, const Stage T, const int kr, const int pd,
                                             const int j, double q, const int lane) {
    return kr <= 32 ? fac_append_ku<1>(P, Ri, T, kr, pd, j, q, lane) : fac_append_ku<SSPS_KU>(P, Ri, T, kr, pd, j, q, lane);
}

// ------------------------------------------------------------------------------------------------
// Screening of a BIG-tier move WITH a removal.  Recomputing the rows after the removed column is the expensive part of a
// pass (13 000 of 17 600 cycles) and 60-75 % of those proposals are rejected (measured).  From the cached factor L of T (the
// current set, plus the arriving parent's row for a swap) the drop of q when column p leaves costs ONE forward substitution:
//      L z = e_p,   q(T \ p) = q(T) - (z . w)^2 / (z . z)            ((z . w) = beta_p, (z . z) = (G_TT^-1)_pp)
// That value is not the bit pattern of the from-scratch factorisation, so it never decides an acceptance and never enters
// the state: it only lets the screened test (mh_accept's fp32 comparison, margin widened by a bound on the error of q)
// REJECT early.  Anything else -- accepted, too close, an ill-conditioned or rank-deficient factor -- takes the exact path
// as before.  `verify = 1` runs both and counts a screened rejection that the exact test accepts as a decision mismatch.
// removal_gain: lane r = row r of the factor at ssps_smem[off + TRI(a) + b] (1/d on the diagonal), w at ssps_smem[woff + a];
// ok = every pivot ratio d^2 / G_aa of T is above 1e-6 and finite (G_aa = n - 1 for the standardised design): the relative
// error of the result is then below 1e-8.  (Tried in the MID tier too and dropped: sets of 6-7 parents are excursions, half
// of their removals are accepted, and the screening cost more than it saved: 5.7e9 -> 5.3e9 proposals/s.)
// ------------------------------------------------------------------------------------------------
__device__ __noinline__ double removal_gain(const unsigned off, const unsigned woff, const int kt, const int p, const int r,
                                            const double gref, bool& ok) {
    const bool row = r < kt;
    const double inv_r = row ? ssps_smem[off + SSPS_TRI(r) + r] : 1.0;
    const double w_r = row ? ssps_smem[woff + r] : 0.0;
    ok = !__any_sync(SSPS_FULL, row && !(inv_r > 0.0 && inv_r * inv_r * gref < 1e6));
    double acc = 0.0, z = 0.0;
#pragma unroll 1
    for (int b = p; b < kt; ++b) {
        const double zb = __shfl_sync(SSPS_FULL, (r == p) ? inv_r : -acc * inv_r, b);
        if (r == b) z = zb;
        if (row && r > b) acc = fma(ssps_smem[off + SSPS_TRI(r) + b], zb, acc);
    }
    double s1 = z * w_r, s2 = z * z;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(SSPS_FULL, s1, o); s2 += __shfl_xor_sync(SSPS_FULL, s2, o); }
    return s2 > 0.0 ? s1 * s1 / s2 : 0.0;
}

// the screened test on an approximate q of the proposed set (|error| <= dq_abs): true = rejected for certain
__device__ __noinline__ bool screen_reject(const SspsDev& P, const int action, const int k, const int kn, const double dprior,
                                           const double fwd, const double bwd, const float lg_ip_cur, const double ip_apx,
                                           const double dq_abs, const double u_acc) {
    if (!(ip_apx > 0.0) || !(u_acc > 1e-30)) return false;
    const float lgu = fast_lg2((float)u_acc), lg_new = fast_lg2((float)ip_apx);
    const double hn = 0.5 * P.dn;
    const double base = ((dprior - fwd) + bwd) - 0.5 * (double)(kn - k) * P.log1pn;
    const double diff = SSPS_LN2 * ((double)lgu + hn * ((double)lg_new - (double)lg_ip_cur)) - base;
    const double marg = 1.2e-6 * ((double)fabsf(lgu) + 1.0 + hn * ((double)(fabsf(lg_new) + fabsf(lg_ip_cur)) + 2.0)) + 1e-9 * fabs(base) + 1e-6
                        + 2.0 * hn * P.shrink * dq_abs / ip_apx;
    return diff > 2.0 * marg;
}

// ------------------------------------------------------------------------------------------------
// The parent-set move of the BIG lanes of a warp (mask `bigm`), one lane at a time with the whole
// warp working on it, then the pending arrivals into the tier (mask `pend`).  Same draws, same
// decisions, same bits as the SMALL path and the oracle: smart_proposal + smart_involution + Gen.mh.
// acc_kn (per lane): accepted | |S'| << 1 of this lane's move.
// ------------------------------------------------------------------------------------------------
template <bool VERIFY>
__device__ __noinline__ Lane big_phase(const SspsDev& P, const int child, const StepDraws& d, const Move& mv, Lane L, const SmallAcc& S, const KCtx& C,
                                       const int rel_now, unsigned bigm, unsigned pend, int* acc_kn) {
    const int lane = threadIdx.x & 31;
    const int V = P.V;
    const double* __restrict__ Ri = P.R + (size_t)V * child;
    unsigned n_big = 0;
    const long long t_in = clock64();
#ifdef SSPS_PROFILE_BIG
    long long prof_[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long tp_ = clock64();
#endif
    while (bigm) {
        const int src = __ffs(bigm) - 1;
        bigm &= bigm - 1;
        const BigRef B = C.big(src);
        const int k = __shfl_sync(SSPS_FULL, L.k, src);
        const int action = __shfl_sync(SSPS_FULL, mv.action, src);
        const int a_ord = __shfl_sync(SSPS_FULL, mv.a_ord, src);
        const int r_ord = __shfl_sync(SSPS_FULL, mv.r_ord, src);
        const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;
        const int kn = k + (has_add ? 1 : 0) - (has_rem ? 1 : 0);
        ++n_big;
        if (kn > C.kcap) {
            if (lane == src) { atomicAdd(C.stats + ST_KCAP_REJECT, 1ull); *acc_kn = kn << 1; }
            continue;
        }

#ifdef SSPS_PROFILE_BIG
        { const long long tn_ = clock64(); prof_[0] += tn_ - tp_; tp_ = tn_; }
#endif
        const Stage T = C.stage(k + 1);
        stage_load(T, B, k, lane);
#ifdef SSPS_PROFILE_BIG
        { const long long tn_ = clock64(); prof_[1] += tn_ - tp_; tp_ = tn_; }
#endif

        int j_add = -1, j_rem = -1, p = k;
        uint32_t w_rem = 0;
        if (has_add) {          // ith_nonparent on the sorted list: the a-th id not in S
            int c = 0;
            for (int base = 0; base < k; base += 32) {
                const int b = base + lane;
                c += __popc(__ballot_sync(SSPS_FULL, b < k && T.idof(b) - (int)T.rk[b] < a_ord));
            }
            j_add = a_ord - 1 + c;
        }
        if (has_rem) {          // the r-th element of the sorted list
            for (int base = 0; base < k; base += 32) {
                const int b = base + lane;
                const unsigned hit = __ballot_sync(SSPS_FULL, b < k && (int)T.rk[b] == r_ord - 1);
                if (hit) p = base + __ffs(hit) - 1;
            }
            w_rem = T.id[p];
            j_rem = (int)(w_rem & 0xffffu);
        }

#ifdef SSPS_PROFILE_BIG
        { const long long tn_ = clock64(); prof_[2] += tn_ - tp_; tp_ = tn_; }
#endif
        // ---- moves with a removal: try to reject from the cached factor before anything is recomputed (removal_gain) ----
        bool scr_rej = false;
        if (has_rem && k + (has_add ? 1 : 0) <= 32 && __isShared(T.t)) {
            double qT = 0.0;
            for (int m = 0; m < k; ++m) { const double wm = T.tw[m]; qT = fma(wm, wm, qT); }
            int kt = k;
            if (has_add) { qT = fac_append(P, Ri, T, k, SSPS_NODEAD, j_add, qT, lane); kt = k + 1; }        // row k of the Stage: rewritten by the exact path
            bool ok = false;
            const double dq = removal_gain((unsigned)(T.t - ssps_smem), (unsigned)(T.tw - ssps_smem), kt, p, lane, P.dn - 1.0, ok);
            int rej = 0;
            if (ok && lane == src) {
                const double* __restrict__ omc_i = P.omc + (size_t)V * child;
                double dprior = L.lam * __ldg(omc_i + j_rem);
                if (has_add) dprior = -L.lam * __ldg(omc_i + j_add) + dprior;
                const double q_red = qT - dq;
                rej = screen_reject(P, action, k, kn, dprior, mv.fwd, mv.bwd, L.lg_ip, C.yty - P.shrink * q_red, 1e-8 * fabs(q_red) + 1e-12, d.u_acc) ? 1 : 0;
                atomicAdd(C.stats + (rej ? ST_SCR_REJ : ST_SCR_TRY), 1ull);
            }
            scr_rej = __shfl_sync(SSPS_FULL, rej, src) != 0;
            if (scr_rej && !VERIFY) {
                if (lane == src) *acc_kn = kn << 1;
                continue;
            }
        }
        // ---- apply the move to the working copy ----
        double q_new = 0.0;
        {
            const int pm = p < k ? p : k;
            for (int m = 0; m < pm; ++m) { const double wm = T.tw[m]; q_new = fma(wm, wm, q_new); }
        }
        int kr = k;
        const int pd = has_rem ? p : SSPS_NODEAD;
        if (has_rem) {
            ids_remove(T, k, p, lane);
#ifdef SSPS_PROFILE_BIG
            q_new = fac_remove(P, Ri, T, k, p, q_new, lane, C.stats);
#else
            q_new = fac_remove(P, Ri, T, k, p, q_new, lane);
#endif
            kr = k - 1;
        }

#ifdef SSPS_PROFILE_BIG
        { const long long tn_ = clock64(); prof_[3] += tn_ - tp_; tp_ = tn_; }
#endif
        if (has_add) {
            q_new = fac_append(P, Ri, T, kr, pd, j_add, q_new, lane);
            ids_append(T, kr, j_add, rel_now, lane);
        }

#ifdef SSPS_PROFILE_BIG
        { const long long tn_ = clock64(); prof_[4] += tn_ - tp_; tp_ = tn_; }
#endif
        int acc = 0;
        if (lane == src) {
            const double fwd = mv.fwd, bwd = mv.bwd;
            const double* __restrict__ omc_i = P.omc + (size_t)V * child;
            double omc_add = 0.0, omc_rem = 0.0, dprior = 0.0;
            if (has_add) { omc_add = __ldg(omc_i + j_add); dprior = -L.lam * omc_add; }
            if (has_rem) { omc_rem = __ldg(omc_i + j_rem); dprior = dprior + L.lam * omc_rem; }
            const double ip_new = C.yty - P.shrink * q_new;
            const float lg_new = fast_lg2((float)ip_new);
            acc = mh_accept_rare<VERIFY>(P, action, k, kn, dprior, fwd, bwd, L.ip, L.lg_ip, ip_new, lg_new, d.u_acc, C.stats) ? 1 : 0;
            if (VERIFY && scr_rej && acc) atomicAdd(C.stats + ST_DEC_MISMATCH, 1ull);      // the screening may only reject what the exact test rejects
            if (VERIFY) {       // the proposed set's q from scratch, in its arrival order: must be the same bits
                const Scratch xs = C.xs_of(lane);
                for (int b = 0; b < kn; ++b) xs.T(b) = (uint16_t)T.idof(b);
                const double lml_ref = score_set(P, child, xs, kn, nullptr);
                const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
                if (!(lml_ref == lml_new || (lml_ref != lml_ref && lml_new != lml_new))) atomicAdd(C.stats + ST_MISMATCH, 1ull);
            }
            if (acc) {
                if (has_rem) {
                    const int credit = rel_now - (int)(w_rem >> 16);
                    if (credit > 0) atomicAdd(C.cnt_row + j_rem, (uint32_t)credit);     // RED: the lane owns the counter, nothing waits for the old value
                    if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_rem | ((uint32_t)(rel_now + 1) << 16);
                }
                if (has_add && C.pev && L.n_pev < C.pev_cap)
                    C.pev[L.n_pev++] = (uint32_t)j_add | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
                L.somc = (L.somc - omc_rem) + omc_add;
                L.ip = ip_new; L.lg_ip = lg_new;
                L.k = kn;
                if (kn <= SSPS_KS) L.mode = MODE_SMALL;
            }
            *acc_kn = acc | (kn << 1);
        }

#ifdef SSPS_PROFILE_BIG
        { const long long tn_ = clock64(); prof_[5] += tn_ - tp_; tp_ = tn_; }
#endif
        acc = __shfl_sync(SSPS_FULL, acc, src);
        if (acc) {
            if (kn <= SSPS_KS) stage_to_small(T, S.of_lane(src - lane), kn, pd, lane);
            else stage_store(T, B, kn, pd, lane);
        }
    }
#ifdef SSPS_PROFILE_BIG
    { const long long tn_ = clock64(); prof_[6] += tn_ - tp_; tp_ = tn_; }
#endif
    (void)pend;             // arrivals into the tier are written by the arriving lane itself (small_move)
#ifdef SSPS_PROFILE_BIG
    if (lane == 0) { prof_[7] += clock64() - tp_; for (int e_ = 0; e_ < 8; ++e_) atomicAdd(C.stats + ST_PROF + e_, (unsigned long long)prof_[e_]); }
#endif
    if (lane == 0) {
        if (n_big) atomicAdd(C.stats + ST_BIG_STEPS, (unsigned long long)n_big);
        atomicAdd(C.stats + ST_BIG_CYCLES, (unsigned long long)(clock64() - t_in));
    }
    return L;
}

// task start: ids of the lanes are in place (SMALL: shared memory, BIG: arena); build the factors
__device__ __noinline__ Lane task_settle(const SspsDev& P, const int child, Lane L, const bool live, const SmallAcc& S, const KCtx& C) {
    const int lane = threadIdx.x & 31;
    const double* __restrict__ Ri = P.R + (size_t)P.V * child;
    if (live) {
        L.somc = sorted_somc(P, child, L.mode, S, C.big(lane), C.xs_of(lane), L.k);
        if (L.mode == MODE_SMALL) lane_set_q(P, child, L, small_build(P, Ri, S, L.k));
    }
    unsigned bigm = __ballot_sync(SSPS_FULL, live && L.mode == MODE_BIG);
    while (bigm) {          // rows appended in arrival order
        const int src = __ffs(bigm) - 1;
        bigm &= bigm - 1;
        const int k = __shfl_sync(SSPS_FULL, L.k, src);
        const BigRef B = C.big(src);
        const Stage T = C.stage(k);
        for (int a = lane; a < k; a += 32) { T.id[a] = B.w[a]; T.rk[a] = B.r[a]; }
        __syncwarp();
        double q = 0.0;
        for (int a = 0; a < k; ++a) q = fac_append(P, Ri, T, a, SSPS_NODEAD, T.idof(a), q, lane);
        stage_store(T, B, k, SSPS_NODEAD, lane);
        if (lane == src) lane_set_q(P, child, L, q);
    }
    return L;
}

// ------------------------------------------------------------------------------------------------
// The kernel
// ------------------------------------------------------------------------------------------------
template <class Source, bool VERIFY>
__global__ void __launch_bounds__(SSPS_BLOCK, SSPS_MIN_CTAS) mh_sweeps_kernel(const __grid_constant__ SspsDev P, const __grid_constant__ ChainState st,
                                                                              const __grid_constant__ MhParams prm, const __grid_constant__ Source src) {
    double* const smem_d = ssps_smem;
    const int lane = threadIdx.x & 31;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    SmallAcc S;
    double* stage_smem_t; uint32_t* stage_smem_id; uint16_t* stage_smem_rk;
    {   // [SMALL d | Stage d per warp | SMALL w | Stage id per warp | Stage rk per warp | SMALL r]
        const int wic = threadIdx.x >> 5;
        double* const stage_d = smem_d + (size_t)SSPS_ROWLEN(SSPS_KS) * SSPS_BLOCK;
        uint32_t* const words = (uint32_t*)(smem_d + SSPS_SMEM_DOUBLES);
        uint16_t* const halves = (uint16_t*)(words + SSPS_SMEM_WORDS);
        uint8_t* const bytes = (uint8_t*)(halves + (SSPS_BLOCK / 32) * SSPS_KST);
        S.d = smem_d + threadIdx.x;
        S.w = words + threadIdx.x;
        S.r = bytes + threadIdx.x;
        stage_smem_t = stage_d + (size_t)wic * SSPS_ROWLEN(SSPS_KST);
        stage_smem_id = words + SSPS_KS * SSPS_BLOCK + wic * SSPS_KST;
        stage_smem_rk = halves + wic * SSPS_KST;
    }
    KCtx C;
    C.kcap = prm.kcap;
    C.prm = &prm; C.wbase = gtid - lane;
    C.st_smem.t = stage_smem_t; C.st_smem.tw = stage_smem_t + SSPS_TRI(SSPS_KST); C.st_smem.id = stage_smem_id; C.st_smem.rk = stage_smem_rk;
    C.stats = prm.stats; C.pev_cap = prm.pev_cap;
    const int V = P.V;
    const int thin = prm.thinning;

    unsigned long long s_f3 = 0;
    unsigned s_accp = 0, s_accl = 0, s_k = 0, s_b = 0, s_prop = 0;

    // A launch with fewer tasks than resident warps (few chains: the time-to-convergence runs) is spread over the SMs:
    // every CTA keeps only the warps its share of the tasks needs -- a warp alone on its scheduler runs a step about
    // twice as fast as one of three.
    const int n_total = __ldg(prm.start + prm.V);
    const int warps_needed = max(1, (n_total + (int)gridDim.x - 1) / (int)gridDim.x);
#if SSPS_LOCKSTEP
    // The warps that share a scheduler (warp id mod 4) pull their tasks together and meet at a named barrier before
    // every sweep, so that they walk the step's instruction stream side by side and share its fetches.
    __shared__ int sh_task[4];
#if SSPS_LOCKSTEP == 2      // the whole CTA in step
    const int ls_sched = 0, ls_member = threadIdx.x >> 5;
    const int LS_N = min(SSPS_BLOCK / 32, warps_needed);
#else
    const int ls_sched = (threadIdx.x >> 5) & 3, ls_member = threadIdx.x >> 7;
    const int LS_N = min((SSPS_BLOCK / 32 - ls_sched + 3) / 4, (warps_needed + 3) / 4);     // warps on this scheduler (block sizes that are not a multiple of 128 too)
#endif
    if (ls_member >= LS_N) return;
#define LS_BAR() asm volatile("bar.sync %0, %1;" ::"r"(1 + ls_sched), "r"(32 * LS_N) : "memory")
#else
    if ((int)(threadIdx.x >> 5) >= warps_needed) return;
#define LS_BAR()
#endif
    for (;;) {
        int task = 0;
#if SSPS_LOCKSTEP
        if (ls_member == 0 && lane == 0) sh_task[ls_sched] = atomicAdd(prm.task_counter, LS_N);
        LS_BAR();
        const int task_base = sh_task[ls_sched];
        LS_BAR();
        if (task_base >= n_total) break;
        task = task_base + ls_member;
        const bool task_ok = task < n_total;
        if (!task_ok) task = task_base;
#else
        if (lane == 0) task = atomicAdd(prm.task_counter, 1);
        task = __shfl_sync(SSPS_FULL, task, 0);
        if (task >= n_total) break;
        const bool task_ok = true;
#endif
        const long long t_start = clock64();
        int lo = 0, hi = prm.V - 1;                     // the rank whose task range holds `task`
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (__ldg(prm.start + mid) <= task) lo = mid; else hi = mid - 1; }
        const int i = __ldg(prm.order + lo);
        const int sp = __ldg(prm.split + i);
        const int width = 32 / sp;
        const int rel = task - __ldg(prm.start + lo);
        const int chain = (rel / sp) * 32 + (rel % sp) * width + lane;
        bool live = task_ok && lane < width && chain < st.n_chains;
        const size_t sidx = (size_t)i * st.nch_pad + (live ? chain : 0);
        const size_t mc = (size_t)(live ? chain : 0) * V + i;
        C.cnt_row = st.counts + mc * V;
        C.yty = __ldg(P.yty + i);
        C.pev = prm.pev ? prm.pev + mc * prm.pev_cap : nullptr;
        Lane L;
        L.k = 0; L.lam = 0.0; L.n_pev = 0;
        L.ip = 0.0; L.lg_ip = 0.f; L.somc = 0.0; L.nrmf = 0.0; L.mode = MODE_SMALL;
        if (live) {
            L.k = st.k[sidx]; L.lam = st.lambda[sidx];
            if (L.k > prm.kcap) live = false;       // cannot happen: ssps_init_chains refuses initial sets above max_indegree
        }
        if (live) {     // parent list: id | rank << 16, arrival order
            if (L.k <= SSPS_KS) {
                for (int a = 0; a < L.k; ++a) {
                    const uint32_t sl = slot_ref(st, V, i, a, chain);
                    S.W(a) = sl & 0xffffu; S.set_rank(a, (int)(sl >> 16));
                }
            } else {
                L.mode = MODE_BIG;
                const BigRef B = C.big(lane);
                for (int a = 0; a < L.k; ++a) {
                    const uint32_t sl = slot_ref(st, V, i, a, chain);
                    B.w[a] = sl & 0xffffu; B.r[a] = (uint16_t)(sl >> 16);
                }
            }
        }
        __syncwarp();
        L = task_settle(P, i, L, live, S, C);
        float lmarg = 0.f;
        const NrmGroups ng = normaliser_groups(P, i);
        if (live) { L.nrmf = normaliser_fast(P, ng, L.lam); lmarg = __ldg(P.lam_margin + i); }
        double lam_sum = 0.0, lam_sq = 0.0, lam_logged = L.lam;
        int n_lev = 0;
        DrawCarry cy; cy.valid = 0; cy.z = 0.f; cy.w = 0u;

        int rec_done = 0, phase = 0;                        // recorded samples so far, sweeps since the last one (no division per step)
        for (int l = 1; l <= prm.n_local; ++l) {
#if SSPS_LOCKSTEP
            if (((l - 1) & (SSPS_LS_EVERY - 1)) == 0) LS_BAR();
#endif
            StepDraws d;
            d.z = 0.0; d.u_lam = 0.5; d.u_act = 0.5; d.u_add = 0.5; d.u_rem = 0.5; d.u_acc = 0.5;
            if (live) {
                d = src.get(chain, i, (uint64_t)(prm.sweep0 + l), (int64_t)(l - 1), cy);

                // ---------------- lambda move (lambda_vec_proposal + Gen.mh) ----------------
                const double lam_new = L.lam + P.sigma * d.z;
                if (lam_new >= P.lambda_min && lam_new <= P.lambda_max) {
                    const double nrm_new = normaliser_fast(P, ng, lam_new);
                    const double w = -(lam_new - L.lam) * L.somc - (nrm_new - L.nrmf);
                    const float lgu = fast_lg2((float)d.u_lam);
                    const double diff = SSPS_LN2 * (double)lgu - w;
                    bool acc = diff < 0.0;
                    const double lmg = (double)lmarg + 1.2e-6 * ((double)fabsf(lgu) + 1.0) + 1e-9 * fabs(w);
                    if (!(d.u_lam > 1e-30 && fabs(diff) > lmg)) {
                        const Scratch xs = C.xs_of(lane);
                        write_sorted(L.mode, S, C.big(lane), xs, L.k);
                        acc = lambda_exact_decide(P, i, xs, L.k, L.lam, lam_new, d.u_lam, nullptr) != 0;
                        atomicAdd(prm.stats + ST_LAM_EXACT, 1ull);
                    } else if (VERIFY) {    // the screened decision against the reference-order fp64 test
                        const Scratch xs = C.xs_of(lane);
                        write_sorted(L.mode, S, C.big(lane), xs, L.k);
                        double dx = 0.0;
                        const bool ex = lambda_exact_decide(P, i, xs, L.k, L.lam, lam_new, d.u_lam, &dx) != 0;
                        if (ex != acc) atomicAdd(prm.stats + ST_LAM_DEC_MISMATCH, 1ull);
                        atomicMax((unsigned*)(prm.stats + ST_LAM_MARGIN_USE), __float_as_uint((float)(fabs(diff - dx) / lmg)));
                    }
                    if (acc) { L.lam = lam_new; L.nrmf = nrm_new; ++s_accl; }
                }
            }

            // ---------------- parent-set move ----------------
            const int rel_now = rec_done;                  // recorded samples of this launch that precede the change (= ceil(l / thin) - 1)
            int acc_kn = 0;
            Move mv;
            mv.action = ACT_ADD; mv.a_ord = 0; mv.r_ord = 0; mv.fwd = 0.0; mv.bwd = 0.0;
            if (live) {
                select_move(P, i, L.k, d, mv.action, mv.a_ord, mv.r_ord);
                move_ratio(P, i, L.k, mv.action, mv.fwd, mv.bwd);
            }
            const bool is_big = live && L.mode == MODE_BIG;
            const bool is_mid = is_big && L.k < SSPS_KM;        // current and proposed set within SSPS_KM rows: the lane on its own
            const unsigned bigm = __ballot_sync(SSPS_FULL, is_big && !is_mid);
            const unsigned smallm = __ballot_sync(SSPS_FULL, live && L.mode == MODE_SMALL);
            if (live && L.mode == MODE_SMALL) {
                int kn = 0;
                const int acc = small_move<VERIFY>(P, i, d, mv, L, S, C, rel_now, smallm, kn);
                acc_kn = acc | (kn << 1);
            }
            unsigned midm = __ballot_sync(SSPS_FULL, is_mid);
#ifdef SSPS_PROFILE_BIG
            const long long t_mid_ = midm ? clock64() : 0;
            const unsigned midm0_ = midm;
#endif
            while (midm) {      // two MID lanes per pass, one per half-warp
                const int s0 = __ffs(midm) - 1;
                midm &= midm - 1;
                const int s1 = midm ? __ffs(midm) - 1 : -1;
                midm &= midm - 1;                   // (0 & anything = 0)
                L = mid_pair<VERIFY>(P, i, d, mv, L, S, C, rel_now, s0, s1, &acc_kn);
            }
#ifdef SSPS_PROFILE_BIG
            if (midm0_ && lane == 0) { atomicAdd(prm.stats + ST_MID_CYCLES, (unsigned long long)(clock64() - t_mid_)); atomicAdd(prm.stats + ST_MID_PROF + 2, (unsigned long long)((__popc(midm0_) + 1) / 2)); }
#endif
            if (bigm) L = big_phase<VERIFY>(P, i, d, mv, L, S, C, rel_now, bigm, 0u, &acc_kn);
            const bool record = ++phase == thin;
            if (record) { phase = 0; ++rec_done; }
            if (live) {
                const int kn = acc_kn >> 1;
                s_accp += (unsigned)(acc_kn & 1);
                s_k += (unsigned)kn;
                s_f3 += (unsigned long long)((unsigned)kn * ((unsigned)kn * ((unsigned)kn + 6u) + 6u));   // k^3 + 6 k^2 + 6 k < 2^32 for k <= 1600
                s_b += (unsigned)(kn * (kn + 1) / 2 + kn);
                if (prm.khist) atomicAdd(prm.khist + (kn < 63 ? kn : 63), 1ull);

                // ---------------- thinned recording ----------------
                if (record) {
                    lam_sum += L.lam; lam_sq = fma(L.lam, L.lam, lam_sq);
                    if (prm.lev_val && L.lam != lam_logged) {
                        if (n_lev < prm.lev_cap) {
                            prm.lev_val[mc * prm.lev_cap + n_lev] = L.lam;
                            prm.lev_idx[mc * prm.lev_cap + n_lev] = (uint16_t)rec_done;
                            ++n_lev;
                        }
                        lam_logged = L.lam;
                    }
                }
            }
        }

        // flush lazy tallies, store the parent list (arrival order, id | rank << 16) and the state
        if (live) {
            s_prop += (unsigned)prm.n_local;
            const int rec_local = prm.n_local / thin;
            const BigRef B = C.big(lane);
            for (int b = 0; b < L.k; ++b) {
                const uint32_t slw = L.mode == MODE_SMALL ? S.W(b) : B.w[b];
                const int rk = L.mode == MODE_SMALL ? S.rank(b) : (int)B.r[b];
                const int credit = rec_local - (int)(slw >> 16);
                if (credit > 0) atomicAdd(C.cnt_row + (slw & 0xffffu), (uint32_t)credit);
                slot_ref(st, V, i, b, chain) = (slw & 0xffffu) | ((uint32_t)rk << 16);
            }
            st.k[sidx] = L.k;
            st.lambda[sidx] = L.lam;
            st.lam_sum[mc] += lam_sum;
            st.lam_sq[mc] += lam_sq;
            if (prm.pev_n) prm.pev_n[mc] = L.n_pev;
            if (prm.lev_n) prm.lev_n[mc] = n_lev;
        }
        __syncwarp();
        if (lane == 0 && task_ok) atomicAdd(prm.vcost + i, (unsigned long long)(clock64() - t_start));
    }

    // stats: warp reduce, one atomic per warp
    unsigned long long v[6] = {s_accp, s_accl, s_k, s_f3, s_b, s_prop};
    const int slot[6] = {ST_ACCP, ST_ACCL, ST_SUMK, ST_F3, ST_BYTES, ST_PROP};
#pragma unroll 1
    for (int e = 0; e < 6; ++e) {
        unsigned long long x = v[e];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(SSPS_FULL, x, o);
        if (lane == 0 && x) atomicAdd(prm.stats + slot[e], x);
    }
}

// ------------------------------------------------------------------------------------------------
// Task schedule of one launch.  A task = (vertex, slab of 32 chains, part): the 32 chains of a slab
// are run by `split` warps, 32 / split lanes each.  A vertex whose micro-chains sit in the BIG tier
// costs a warp one cooperative pass per such lane and step; spreading its slab over more warps
// shortens its tasks, so that the persistent grid drains evenly (longest tasks first) and, when
// there are fewer tasks than resident warps, so that a sweep takes as long as ONE cooperative pass.
//   cost[v]        warp cycles of all tasks of v in the previous launch (split_prev[v] parts per slab)
//   order[rank]    vertex, by estimated cycles per task, descending
//   start[rank]    first task id of that vertex; start[V] = number of tasks
// Model: a task of v at split s costs  base + big_v / s  (base = a task without BIG lanes); the launch
// takes about  all task cycles / resident warps + SSPS_TAIL_WEIGHT x longest task  (list scheduling, longest
// first: the bound has weight 1, a full grid hides most of the tail: see SSPS_TAIL_WEIGHT).  Greedy: keep halving the lanes of the
// longest task while that lowers this estimate.  One CTA; V is at most a few thousand.
// ------------------------------------------------------------------------------------------------
#define SSPS_MAX_SPLIT 32
#ifndef SSPS_TAIL_WEIGHT
#define SSPS_TAIL_WEIGHT 0.045     // measured plateau 0.035 .. 0.06 with the MID tier (6.1e9 proposals/s; 0.02 and 0.1 give 5.6e9)
#endif
__global__ void __launch_bounds__(256) schedule_kernel(const unsigned long long* __restrict__ cost, int V, int n_slabs, int n_warps,
                                                       int have_cycles, int* __restrict__ split /* in: previous, out: this launch */,
                                                       int* __restrict__ order, int* __restrict__ start,
                                                       double* __restrict__ est /* 2 V scratch: task estimate, big part */) {
    __shared__ double sh_sum, sh_val[8];
    __shared__ int sh_idx[8], sh_stop;
    __shared__ unsigned long long sh_bits;
    double* __restrict__ big = est + V;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) { sh_sum = 0.0; sh_bits = 0x7ff0000000000000ull; sh_stop = 0; }
    __syncthreads();
    // mean cycles per task of each vertex in the previous launch
    for (int v = tid; v < V; v += blockDim.x) {
        const double t = (double)cost[v] / ((double)n_slabs * (double)split[v]);
        est[v] = t;
        if (t > 0.0) atomicMin(&sh_bits, (unsigned long long)__double_as_longlong(t));    // positive doubles order like their bit patterns
    }
    __syncthreads();
    const double base = __longlong_as_double((long long)sh_bits);
    const bool model = have_cycles && base < 1e300;
    double lsum = 0.0;
    for (int v = tid; v < V; v += blockDim.x) {
        const double bg = model ? fmax(est[v] - base, 0.0) * (double)split[v] : 0.0;
        big[v] = bg;
        if (model) est[v] = base + bg;          // at split 1
        split[v] = 1;
        lsum += est[v];
    }
    atomicAdd(&sh_sum, lsum);
    __syncthreads();
    // a launch with far fewer tasks than resident warps (few chains: time-to-convergence runs) lasts as long as its longest
    // task: the tail gets full weight there, so that every heavy slab is spread until that stops shortening it
    const double tail_w = (2 * (long long)V * n_slabs <= (long long)n_warps) ? 1.0 : SSPS_TAIL_WEIGHT;
    if (model) {
        double load = sh_sum * (double)n_slabs / (double)n_warps;      // all task cycles / resident warps
        for (int it = 0; it < 5 * V; ++it) {
            // longest task
            double bv = -1.0; int bi = -1;
            for (int v = tid; v < V; v += blockDim.x) { const double e = est[v]; if (e > bv) { bv = e; bi = v; } }
            for (int o = 16; o > 0; o >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi >= 0 && (bi < 0 || oi < bi))) { bv = ov; bi = oi; }
            }
            if (lane == 0) { sh_val[wid] = bv; sh_idx[wid] = bi; }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
                    if (sh_val[w] > bv || (sh_val[w] == bv && sh_idx[w] >= 0 && (bi < 0 || sh_idx[w] < bi))) { bv = sh_val[w]; bi = sh_idx[w]; }
                const int sp = split[bi];
                const double cur = tail_w * bv + load;
                const double e_new = base + big[bi] / (double)(2 * sp);
                const double load_new = load + (double)n_slabs * (double)sp * base / (double)n_warps;   // sp more tasks, each + base
                // the runner-up is not known here; e_new bounds the new longest task from below, which is enough to reject
                if (sp >= SSPS_MAX_SPLIT || big[bi] <= 0.0 || tail_w * e_new + load_new >= cur) sh_stop = 1;
                else { split[bi] = 2 * sp; est[bi] = e_new; sh_sum = load_new; }
            }
            __syncthreads();
            if (sh_stop) break;
            load = sh_sum;
            __syncthreads();
        }
    }
    __syncthreads();
    for (int v = tid; v < V; v += blockDim.x) {
        const double ev = est[v];
        int rank = 0;
        for (int j = 0; j < V; ++j) {
            const double ej = est[j];
            rank += (ej > ev || (ej == ev && j < v)) ? 1 : 0;
        }
        order[rank] = v;
    }
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int r = 0; r < V; ++r) { start[r] = acc; acc += n_slabs * split[order[r]]; }
        start[V] = acc;
    }
}

// first launch: no measured cost yet; estimate from the state, sum over chains of (k+1)^2
__global__ void cost_from_state_kernel(ChainState st, int V, unsigned long long* __restrict__ cost) {
    const int i = blockIdx.x;
    unsigned long long acc = 0;
    for (int c = threadIdx.x; c < st.n_chains; c += blockDim.x) {
        const unsigned long long k = (unsigned long long)st.k[(size_t)i * st.nch_pad + c];
        acc += (k + 1) * (k + 1);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(cost + i, acc);
}
