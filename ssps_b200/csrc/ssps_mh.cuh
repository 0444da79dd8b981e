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
// Storage tiers for (ids, factor, scratch) of a lane: |S| <= SSPS_KS in shared memory, larger
// sets in a coalesced global arena (L1/L2-resident, touched only by the few lanes that need it).
#pragma once
#include "ssps_kernels.cuh"

#ifndef SSPS_KS
#define SSPS_KS 6
#endif
#ifndef SSPS_BLOCK
#define SSPS_BLOCK 128
#endif
#ifndef SSPS_MIN_CTAS
#define SSPS_MIN_CTAS 3
#endif
#define SSPS_LN2 0.6931471805599453

enum { MODE_SMALL = 0, MODE_BIG = 1 };
// stats slots (unsigned long long each)
enum { ST_ACCP = 0, ST_ACCL = 1, ST_SUMK = 2, ST_F3 = 3, ST_BYTES = 4, ST_OVERFLOW = 5, ST_PROP = 6, ST_TIER1 = 7,
       ST_MISMATCH = 10, ST_LAM_EXACT = 12, ST_BIG_STEPS = 13, ST_N = 16 };

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
// Tier storage.  Slot a (arrival order) holds: word = id | rel << 16 (`rel` = recorded samples of
// this launch that precede the parent's arrival: lazy tally), rank = position of the id in the
// SORTED parent list (the reference's moves are defined on the sorted list,
// dbn_proposals.jl:19-49), row a of the factor: Lo(a, b) for b < a, Li(a) = 1/d_a, Wv(a) = w_a.
// The scratch factor So/Si/Sw is indexed by the PROPOSED set's arrival positions.
// ------------------------------------------------------------------------------------------------
#define SSPS_ROWLEN(K) (SSPS_TRI(K) + (K))          // off-diagonals + diagonal, then w: TRI(K) + K doubles

struct SmallAcc {   // shared memory, element-major: [e][thread]
    double* d; uint32_t* w; uint8_t* r;
    static constexpr bool is_small = true;
    __device__ __forceinline__ int cap() const { return SSPS_KS; }
    __device__ __forceinline__ double& D(int e) const { return d[e * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Lo(int a, int b) const { return d[(SSPS_TRI(a) + b) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Li(int a) const { return d[(SSPS_TRI(a) + a) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Wv(int a) const { return d[(SSPS_TRI(SSPS_KS) + a) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& So(int a, int b) const { return d[(SSPS_ROWLEN(SSPS_KS) + SSPS_TRI(a) + b) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Si(int a) const { return d[(SSPS_ROWLEN(SSPS_KS) + SSPS_TRI(a) + a) * SSPS_BLOCK]; }
    __device__ __forceinline__ double& Sw(int a) const { return d[(SSPS_ROWLEN(SSPS_KS) + SSPS_TRI(SSPS_KS) + a) * SSPS_BLOCK]; }
    __device__ __forceinline__ uint32_t& W(int a) const { return w[a * SSPS_BLOCK]; }
    __device__ __forceinline__ int rank(int a) const { return (int)r[a * SSPS_BLOCK]; }
    __device__ __forceinline__ void set_rank(int a, int v) const { r[a * SSPS_BLOCK] = (uint8_t)v; }
};
struct BigAcc {     // global arena, element-major over all resident threads: [e][gtid]
    double* d; uint32_t* w; uint16_t* r; int stride, kcap;
    static constexpr bool is_small = false;
    __device__ __forceinline__ int cap() const { return kcap; }
    __device__ __forceinline__ double& D(int e) const { return d[(unsigned)e * (unsigned)stride]; }
    __device__ __forceinline__ double& Lo(int a, int b) const { return d[(unsigned)(SSPS_TRI(a) + b) * (unsigned)stride]; }
    __device__ __forceinline__ double& Li(int a) const { return d[(unsigned)(SSPS_TRI(a) + a) * (unsigned)stride]; }
    __device__ __forceinline__ double& Wv(int a) const { return d[(unsigned)(SSPS_TRI(kcap) + a) * (unsigned)stride]; }
    __device__ __forceinline__ double& So(int a, int b) const { return d[(unsigned)(SSPS_ROWLEN(kcap) + SSPS_TRI(a) + b) * (unsigned)stride]; }
    __device__ __forceinline__ double& Si(int a) const { return d[(unsigned)(SSPS_ROWLEN(kcap) + SSPS_TRI(a) + a) * (unsigned)stride]; }
    __device__ __forceinline__ double& Sw(int a) const { return d[(unsigned)(SSPS_ROWLEN(kcap) + SSPS_TRI(kcap) + a) * (unsigned)stride]; }
    __device__ __forceinline__ uint32_t& W(int a) const { return w[(unsigned)a * (unsigned)stride]; }
    __device__ __forceinline__ int rank(int a) const { return (int)r[(unsigned)a * (unsigned)stride]; }
    __device__ __forceinline__ void set_rank(int a, int v) const { r[(unsigned)a * (unsigned)stride] = (uint16_t)v; }
};

struct Lane {
    double lam, somc, q, ip, nrmf;   // nrmf: screening value of the prior normaliser at lam
    float lg_ip;                     // lg2(ip), screening
    int k, mode, n_pev;
};

struct MhParams {
    int64_t sweep0;          // sweeps already done (multiple of thinning)
    int n_local;             // sweeps in this launch
    int thinning;
    int n_tasks, n_slabs;
    int* task_counter;
    const int* order;        // vertices, most expensive first
    unsigned long long* vcost;   // per-vertex cycles of this launch
    // global arenas, per resident thread
    double* big_d; uint32_t* big_w; uint16_t* big_r;      // factor + scratch of lanes with |S| > SSPS_KS
    double* x_d; uint16_t* x_t;                            // from-scratch scorer scratch: L, w | T, S
    int kcap; int stride;
    unsigned long long* stats;
    unsigned long long* khist;   // nullable, 64 bins
    // change-point log (nullable): per micro-chain mc = chain*V + vertex
    uint32_t* pev; int* pev_n; int pev_cap;        // parent events: id | rel_rec<<16 | flag<<31
    double* lev_val; uint16_t* lev_idx; int* lev_n; int lev_cap;
};

struct KCtx {       // per-lane pointers
    BigAcc big; Scratch xs;
    uint32_t* cnt_row;     // counts + mc * V
    uint32_t* pev;         // pev + mc * pev_cap (nullable)
    int pev_cap, kcap;
    unsigned long long* stats;
};

__device__ __forceinline__ int ceil_div_i(int a, int b) { return (a + b - 1) / b; }

// screening value of the prior normaliser: sum over groups with 1-c != 0 of mult ln2 lg2(1 + 2^(-(1-c) log2e lambda))
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

// lambda move in the reference order (the oracle's chain_step restates exactly this); xs.S(0..k-1)
// holds the SORTED current set
__device__ __noinline__ int lambda_exact_decide(const SspsDev& P, const int child, const Scratch& xs, const int k, const double lam,
                                                const double lam_new, const double u_lam) {
    const double* __restrict__ omc_i = P.omc + (size_t)P.V * child;
    double somc = 0.0;
    for (int a = 0; a < k; ++a) somc += __ldg(omc_i + xs.S(a));
    const double nrm = prior_normaliser(P, child, lam), nrm_new = prior_normaliser(P, child, lam_new);
    const double w = -(lam_new - lam) * somc - (nrm_new - nrm);
    return log(u_lam) < w ? 1 : 0;
}

__device__ __noinline__ void write_sorted(const int mode, const SmallAcc& S, const BigAcc& B, const Scratch& xs, const int k) {
    if (mode == MODE_SMALL) { for (int b = 0; b < k; ++b) xs.S(S.rank(b)) = (uint16_t)(S.W(b) & 0xffffu); }
    else { for (int b = 0; b < k; ++b) xs.S(B.rank(b)) = (uint16_t)(B.W(b) & 0xffffu); }
}

// ------------------------------------------------------------------------------------------------
// One row of the arrival-order Cholesky, written to the scratch factor at (new) position a.
//   p      arrival position being removed (p >= k: none); rows < p of the cached factor are reused
//   b0     entries b < b0 of this row are taken from cached row `old_row` (unchanged by the removal)
// Operation order = orc_lml_chol / score_set: l_b = (G - sum_{m<b} l_m L_bm) / d_b,
// d2 = G_jj - sum l_m^2, t = r_j - sum l_m w_m, each sum in ascending m with fma.
// ------------------------------------------------------------------------------------------------
// Where a freshly computed row lives in the scratch factor: element offsets of its off-diagonals, 1/d and w.
struct RowAt { int lo, inv, w; };
template <class A>
__device__ __forceinline__ RowAt scratch_row(const A& F, int a) {       // row a of the scratch factor
    const int K = F.cap();
    RowAt r; r.lo = SSPS_ROWLEN(K) + SSPS_TRI(a); r.inv = r.lo + a; r.w = SSPS_ROWLEN(K) + SSPS_TRI(K) + a;
    return r;
}
template <class A>
__device__ __forceinline__ RowAt scratch_flat(const A& F) {             // a lone appended row (pure add): start of the scratch
    const int K = F.cap();
    RowAt r; r.lo = SSPS_ROWLEN(K); r.inv = r.lo + K; r.w = r.inv + 1;
    return r;
}

template <class A>
__device__ __forceinline__ double factor_row(const SspsDev& P, const double* __restrict__ Ri, const A& F, const int p, const int a,
                                             const int ja, const int b0, const int old_row, const RowAt at, double q) {
    const double* __restrict__ Gj = P.G + (size_t)ja * P.V;
    const double gjj = __ldg(Gj + ja);
    double d2 = gjj, t = __ldg(Ri + ja);
    for (int b = 0; b < a; ++b) {
        double l;
        if (b < b0) l = F.Lo(old_row, b);
        else {
            const int jb = (int)(F.W(b >= p ? b + 1 : b) & 0xffffu);
            double s = __ldg(Gj + jb);
            if (b < p) {                         // row b of the cached factor
                for (int m = 0; m < b; ++m) s = fma(-((m < b0) ? F.Lo(old_row, m) : F.D(at.lo + m)), F.Lo(b, m), s);
                l = s * F.Li(b);
            } else {                             // row b was recomputed: entries < p from cached row b+1, the rest from scratch
                for (int m = 0; m < b; ++m)
                    s = fma(-((m < b0) ? F.Lo(old_row, m) : F.D(at.lo + m)), (m < p) ? F.Lo(b + 1, m) : F.So(b, m), s);
                l = s * F.Si(b);
            }
            F.D(at.lo + b) = l;
        }
        d2 = fma(-l, l, d2);
        t = fma(-l, (b < p) ? F.Wv(b) : F.Sw(b), t);
    }
    if (d2 > SSPS_PIVOT_TOL * gjj) {
        const double invd = 1.0 / sqrt(d2);
        const double wa = t * invd;
        F.D(at.inv) = invd; F.D(at.w) = wa;
        q = fma(wa, wa, q);
    } else if (d2 != d2 || gjj != gjj) {           // NaN data: zero-variance column (quirk Q4)
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        F.D(at.inv) = nan; F.D(at.w) = nan; q = nan;
    } else {                                       // numerically dependent column: contributes nothing
        F.D(at.inv) = 0.0; F.D(at.w) = 0.0;
    }
    return q;
}

// q of the proposed set (remove arrival position p if p < k, then append j_add if >= 0); the new
// rows p .. kn-1 are left in the scratch factor
template <class A>
__device__ __forceinline__ double eval_move(const SspsDev& P, const double* __restrict__ Ri, const A& F, const int k, const int p,
                                            const int j_add) {
    double q = 0.0;
    const int pm = p < k ? p : k;
    for (int m = 0; m < pm; ++m) { const double wm = F.Wv(m); q = fma(wm, wm, q); }
    const int kr = p < k ? k - 1 : k;
    for (int a = p; a < kr; ++a) q = factor_row(P, Ri, F, p, a, (int)(F.W(a + 1) & 0xffffu), p, a + 1, scratch_row(F, a), q);
    if (j_add >= 0) q = factor_row(P, Ri, F, p, kr, j_add, 0, 0, p < k ? scratch_row(F, kr) : scratch_flat(F), q);
    return q;
}

// accepted: fold the scratch rows into the cached factor, edit the id list (tallies, change points, ranks)
template <class A>
__device__ __forceinline__ void commit_move(const A& F, const KCtx& C, Lane& L, const int p, const int j_add, const int j_rem,
                                            const int rel_now) {
    int k = L.k;
    if (p < k) {
        const uint32_t slw = F.W(p);
        const int credit = rel_now - (int)(slw >> 16);
        if (credit > 0) C.cnt_row[j_rem] += (uint32_t)credit;
        if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_rem | ((uint32_t)(rel_now + 1) << 16);
        const int rp = F.rank(p);
        for (int a = 0; a < k; ++a) { const int ra = F.rank(a); if (ra > rp) F.set_rank(a, ra - 1); }
        for (int a = p; a < k - 1; ++a) {
            for (int m = 0; m < p; ++m) F.Lo(a, m) = F.Lo(a + 1, m);
            for (int m = p; m < a; ++m) F.Lo(a, m) = F.So(a, m);
            F.Li(a) = F.Si(a); F.Wv(a) = F.Sw(a);
            F.W(a) = F.W(a + 1); F.set_rank(a, F.rank(a + 1));
        }
        --k;
    }
    if (j_add >= 0) {
        const RowAt at = p < L.k ? scratch_row(F, k) : scratch_flat(F);     // k is already the post-removal size
        for (int m = 0; m < k; ++m) F.Lo(k, m) = F.D(at.lo + m);
        F.Li(k) = F.D(at.inv); F.Wv(k) = F.D(at.w);
        int rj = 0;
        for (int a = 0; a < k; ++a) {
            if ((int)(F.W(a) & 0xffffu) < j_add) ++rj; else F.set_rank(a, F.rank(a) + 1);
        }
        F.W(k) = (uint32_t)j_add | ((uint32_t)rel_now << 16);
        F.set_rank(k, rj);
        if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_add | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
        ++k;
    }
    L.k = k;
}

__device__ __forceinline__ void lane_set_q(const SspsDev& P, int child, Lane& L, double q) {
    L.q = q; L.ip = __ldg(P.yty + child) - P.shrink * q; L.lg_ip = fast_lg2((float)L.ip);
}

// factor of the current set from its id list (task start, tier changes): the same rows, in order
template <class A>
__device__ __forceinline__ double build_factor(const SspsDev& P, const double* __restrict__ Ri, const A& F, const int k) {
    double q = 0.0;
    for (int a = 0; a < k; ++a) {
        // p = a: "rows < a cached, nothing removed yet" -- factor_row reads rows < a from the cached factor and writes row a
        // to scratch; fold it in at once
        const RowAt at = scratch_flat(F);
        q = factor_row(P, Ri, F, k, a, (int)(F.W(a) & 0xffffu), 0, 0, at, q);
        for (int m = 0; m < a; ++m) F.Lo(a, m) = F.D(at.lo + m);
        F.Li(a) = F.D(at.inv); F.Wv(a) = F.D(at.w);
    }
    return q;
}

// ids (word, rank) of the lane are in the arena; build its factor in the tier its size selects
__device__ __noinline__ Lane lane_settle(const SspsDev& P, const int child, Lane L, const SmallAcc& S, const KCtx& C) {
    const int k = L.k;
    const double* __restrict__ omc_i = P.omc + (size_t)P.V * child;
    const double* __restrict__ Ri = P.R + (size_t)P.V * child;
    for (int b = 0; b < k; ++b) C.xs.S(C.big.rank(b)) = (uint16_t)(C.big.W(b) & 0xffffu);
    double somc = 0.0;
    for (int a = 0; a < k; ++a) somc += __ldg(omc_i + C.xs.S(a));      // the reference-order (sorted) sum
    L.somc = somc;
    if (k <= SSPS_KS) {
        for (int b = 0; b < k; ++b) { S.W(b) = C.big.W(b); S.set_rank(b, C.big.rank(b)); }
        L.mode = MODE_SMALL;
        lane_set_q(P, child, L, build_factor<SmallAcc>(P, Ri, S, k));
    } else {
        L.mode = MODE_BIG;
        lane_set_q(P, child, L, build_factor<BigAcc>(P, Ri, C.big, k));
    }
    return L;
}

// a SMALL lane accepted a move that outgrows shared memory: hand ids over to the arena, rebuild there
__device__ __noinline__ Lane small_to_big(const SspsDev& P, const int child, Lane L, const SmallAcc& S, const KCtx& C, const int p,
                                          const int j_add, const int j_rem, const int rel_now) {
    const int k = L.k;
    for (int b = 0; b < k; ++b) { C.big.W(b) = S.W(b); C.big.set_rank(b, S.rank(b)); }
    // edit the id list only (the factor is rebuilt): reuse commit_move's bookkeeping on a dummy factor region
    if (p < k) {
        const uint32_t slw = C.big.W(p);
        const int credit = rel_now - (int)(slw >> 16);
        if (credit > 0) C.cnt_row[j_rem] += (uint32_t)credit;
        if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_rem | ((uint32_t)(rel_now + 1) << 16);
        const int rp = C.big.rank(p);
        for (int a = 0; a < k; ++a) { const int ra = C.big.rank(a); if (ra > rp) C.big.set_rank(a, ra - 1); }
        for (int a = p; a < k - 1; ++a) { C.big.W(a) = C.big.W(a + 1); C.big.set_rank(a, C.big.rank(a + 1)); }
        L.k = k - 1;
    }
    if (j_add >= 0) {
        const int kk = L.k;
        int rj = 0;
        for (int a = 0; a < kk; ++a) {
            if ((int)(C.big.W(a) & 0xffffu) < j_add) ++rj; else C.big.set_rank(a, C.big.rank(a) + 1);
        }
        C.big.W(kk) = (uint32_t)j_add | ((uint32_t)rel_now << 16);
        C.big.set_rank(kk, rj);
        if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_add | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
        L.k = kk + 1;
    }
    return lane_settle(P, child, L, S, C);
}

// ------------------------------------------------------------------------------------------------
// The parent-set move (smart_proposal + smart_involution + Gen.mh), generic over the tier.
// Returns 1 if accepted.  kn_out = |S'|.
// ------------------------------------------------------------------------------------------------
template <class A, bool VERIFY>
__device__ __forceinline__ int parent_move(const SspsDev& P, const int child, const StepDraws& d, Lane& L, const A& F, const SmallAcc& S,
                                           const KCtx& C, const int rel_now, unsigned& n_t1, int& kn_out) {
    const int V = P.V, k = L.k;
    const size_t trow = ((size_t)child * (V + 1) + k) * 3;
    const double p_add = __ldg(P.mvP + trow), p_swp = __ldg(P.mvP + trow + 1);
    const int action = d.u_act < p_add ? ACT_ADD : (d.u_act < p_add + p_swp ? ACT_SWAP : ACT_REM);
    const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;
    int j_add = -1, j_rem = -1, p = k;
    if (has_add) {          // ith_nonparent on the sorted list: the a-th id not in S
        const int m = V - k;
        int a = 1 + (int)floor(d.u_add * (double)m); if (a > m) a = m;
        int c = 0;
        for (int b = 0; b < k; ++b) c += ((int)(F.W(b) & 0xffffu) - F.rank(b) < a) ? 1 : 0;
        j_add = a - 1 + c;
    }
    if (has_rem) {          // the r-th element of the sorted list
        int r = 1 + (int)floor(d.u_rem * (double)k); if (r > k) r = k;
        for (int b = 0; b < k; ++b) if (F.rank(b) == r - 1) { p = b; j_rem = (int)(F.W(b) & 0xffffu); }
    }
    const int kn = k + (has_add ? 1 : 0) - (has_rem ? 1 : 0);
    kn_out = kn;
    double fwd = 0.0, bwd = 0.0;
    if (action == ACT_ADD) {
        fwd = __ldg(P.mvLogP + trow) - __ldg(P.logint + (V - k));
        bwd = __ldg(P.mvLogP + ((size_t)child * (V + 1) + (k + 1)) * 3 + 2) - __ldg(P.logint + (k + 1));
    } else if (action == ACT_REM) {
        fwd = __ldg(P.mvLogP + trow + 2) - __ldg(P.logint + k);
        bwd = __ldg(P.mvLogP + ((size_t)child * (V + 1) + (k - 1)) * 3 + 0) - __ldg(P.logint + (V - k + 1));
    }
    if (kn > C.kcap) { atomicExch(C.stats + ST_OVERFLOW, 1ull); return 0; }
    const double* __restrict__ omc_i = P.omc + (size_t)V * child;
    double omc_add = 0.0, omc_rem = 0.0, dprior = 0.0;
    if (has_add) { omc_add = __ldg(omc_i + j_add); dprior = -L.lam * omc_add; }
    if (has_rem) { omc_rem = __ldg(omc_i + j_rem); dprior = dprior + L.lam * omc_rem; }

    const double q_new = eval_move(P, P.R + (size_t)V * child, F, k, p, j_add);
    const double ip_new = __ldg(P.yty + child) - P.shrink * q_new;

    // ---- screened acceptance test: accept iff log(u) < alpha ----
    const float lgu = fast_lg2((float)d.u_acc);
    const float lg_new = fast_lg2((float)ip_new);
    const double hn = 0.5 * P.dn;
    const double base = ((dprior - fwd) + bwd) - 0.5 * (double)(kn - k) * P.log1pn;
    const double diff = SSPS_LN2 * ((double)lgu + hn * ((double)lg_new - (double)L.lg_ip)) - base;
    const double marg = 1.2e-6 * ((double)fabsf(lgu) + 1.0 + hn * ((double)(fabsf(lg_new) + fabsf(L.lg_ip)) + 2.0)) + 1e-9 * fabs(base) + 1e-6;
    bool accept = diff < 0.0;
    if (!(d.u_acc > 1e-30 && fabs(diff) > marg)) {        // too close (or NaN): the reference-order fp64 test
        ++n_t1;
        const double lml_cur = -0.5 * (double)k * P.log1pn - 0.5 * P.dn * log(L.ip);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        const double w = dprior + (lml_new - lml_cur);
        const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
        accept = log(d.u_acc) < alpha;
    }
    if (VERIFY) {       // the proposed set's q from scratch, in its arrival order: must be the same bits
        int t = 0;
        for (int b = 0; b < k; ++b) if (b != p) C.xs.T(t++) = (uint16_t)(F.W(b) & 0xffffu);
        if (j_add >= 0) C.xs.T(t++) = (uint16_t)j_add;
        const double lml_ref = score_set(P, child, C.xs, t, nullptr);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        if (!(lml_ref == lml_new || (lml_ref != lml_ref && lml_new != lml_new))) atomicAdd(C.stats + ST_MISMATCH, 1ull);
    }
    if (!accept) return 0;

    if (A::is_small && kn > SSPS_KS) {     // SMALL lane outgrows shared memory
        L = small_to_big(P, child, L, S, C, p, j_add, j_rem, rel_now);
        return 1;
    }
    commit_move(F, C, L, p, j_add, j_rem, rel_now);
    L.somc = (L.somc - omc_rem) + omc_add;
    L.q = q_new; L.ip = ip_new; L.lg_ip = lg_new;
    return 1;
}

// ------------------------------------------------------------------------------------------------
// The parent-set move of a SMALL lane (|S| <= SSPS_KS): same arithmetic as parent_move<SmallAcc>,
// but statically unrolled over SSPS_KS slots with the proposed set's factor held in REGISTERS, so
// that the 32 lanes of a warp (different k, different removal position p) walk one instruction
// stream: rows/entries a lane does not need are predicated off, rows nobody needs are skipped by a
// warp-uniform branch.  NL(a, b) = entry (a, b) of the factor in the PROPOSED set's coordinates
// (arrival positions after the removal): rows < p and the entries < p of later rows are the
// cached values, the rest is recomputed; the appended parent's row is jl[].
// ------------------------------------------------------------------------------------------------
#define NL(a, b) nl[SSPS_TRI(a) - (a) + (b)]       // strictly lower triangle: row a starts at a(a-1)/2

template <bool VERIFY>
__device__ __forceinline__ int small_move(const SspsDev& P, const int child, const StepDraws& d, Lane& L, const SmallAcc& S,
                                          const KCtx& C, const int rel_now, unsigned& n_t1, int& kn_out) {
    constexpr int KS = SSPS_KS;
    const int V = P.V, k = L.k;
    const size_t trow = ((size_t)child * (V + 1) + k) * 3;
    const double p_add = __ldg(P.mvP + trow), p_swp = __ldg(P.mvP + trow + 1);
    const int action = d.u_act < p_add ? ACT_ADD : (d.u_act < p_add + p_swp ? ACT_SWAP : ACT_REM);
    const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;

    uint32_t idw[KS]; int rk[KS];
#pragma unroll
    for (int b = 0; b < KS; ++b) {
        idw[b] = 0xffffu; rk[b] = -1;
        if (b < k) { idw[b] = S.W(b); rk[b] = S.rank(b); }
    }
    int j_add = -1, j_rem = -1, p = k, rp = 1 << 20;
    if (has_add) {          // ith_nonparent on the sorted list: the a-th id not in S
        const int m = V - k;
        int a = 1 + (int)floor(d.u_add * (double)m); if (a > m) a = m;
        int c = 0;
#pragma unroll
        for (int b = 0; b < KS; ++b) c += (b < k && (int)(idw[b] & 0xffffu) - rk[b] < a) ? 1 : 0;
        j_add = a - 1 + c;
    }
    if (has_rem) {          // the r-th element of the sorted list
        int r = 1 + (int)floor(d.u_rem * (double)k); if (r > k) r = k;
        rp = r - 1;
#pragma unroll
        for (int b = 0; b < KS; ++b) if (rk[b] == rp) { p = b; j_rem = (int)(idw[b] & 0xffffu); }
    }
    const int kr = k - (has_rem ? 1 : 0);
    const int kn = kr + (has_add ? 1 : 0);
    kn_out = kn;
    double fwd = 0.0, bwd = 0.0;
    if (action == ACT_ADD) {
        fwd = __ldg(P.mvLogP + trow) - __ldg(P.logint + (V - k));
        bwd = __ldg(P.mvLogP + ((size_t)child * (V + 1) + (k + 1)) * 3 + 2) - __ldg(P.logint + (k + 1));
    } else if (action == ACT_REM) {
        fwd = __ldg(P.mvLogP + trow + 2) - __ldg(P.logint + k);
        bwd = __ldg(P.mvLogP + ((size_t)child * (V + 1) + (k - 1)) * 3 + 0) - __ldg(P.logint + (V - k + 1));
    }
    if (kn > C.kcap) { atomicExch(C.stats + ST_OVERFLOW, 1ull); return 0; }
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

    // the cached factor in the reduced set's coordinates; entries to recompute start as 0
    double nl[SSPS_TRI(KS) - KS + 1], ninv[KS], nw[KS];
    const int kr_max = __reduce_max_sync(__activemask(), kr);
    const int p_min = __reduce_min_sync(__activemask(), p);
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
                    const double invd = 1.0 / sqrt(d2);
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
        const double* __restrict__ Gj = P.G + (size_t)j_add * V;
        const double gjj = __ldg(Gj + j_add);
        double d2 = gjj, t = __ldg(Ri + j_add);
#pragma unroll
        for (int b = 0; b < KS; ++b) {
            if (b < kr_max) {
                double s = __ldg(Gj + nid[b]);
#pragma unroll
                for (int m = 0; m < b; ++m) s = fma(-jl[m], NL(b, m), s);
                const double l = (b < kr) ? s * ninv[b] : 0.0;
                jl[b] = l;
                if (b < kr) { d2 = fma(-l, l, d2); t = fma(-l, nw[b], t); }
            }
        }
        if (d2 > SSPS_PIVOT_TOL * gjj) {
            jinv = 1.0 / sqrt(d2); jw = t * jinv;
            q_new = fma(jw, jw, q_new);
        } else if (d2 != d2 || gjj != gjj) {
            jinv = __longlong_as_double(0x7ff8000000000000LL); jw = jinv; q_new = jinv;
        }
    }
    const double ip_new = __ldg(P.yty + child) - P.shrink * q_new;

    // ---- screened acceptance test: accept iff log(u) < alpha ----
    const float lgu = fast_lg2((float)d.u_acc);
    const float lg_new = fast_lg2((float)ip_new);
    const double hn = 0.5 * P.dn;
    const double base = ((dprior - fwd) + bwd) - 0.5 * (double)(kn - k) * P.log1pn;
    const double diff = SSPS_LN2 * ((double)lgu + hn * ((double)lg_new - (double)L.lg_ip)) - base;
    const double marg = 1.2e-6 * ((double)fabsf(lgu) + 1.0 + hn * ((double)(fabsf(lg_new) + fabsf(L.lg_ip)) + 2.0)) + 1e-9 * fabs(base) + 1e-6;
    bool accept = diff < 0.0;
    if (!(d.u_acc > 1e-30 && fabs(diff) > marg)) {        // too close (or NaN): the reference-order fp64 test
        ++n_t1;
        const double lml_cur = -0.5 * (double)k * P.log1pn - 0.5 * P.dn * log(L.ip);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        const double w = dprior + (lml_new - lml_cur);
        const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
        accept = log(d.u_acc) < alpha;
    }
    if (VERIFY) {       // the proposed set's q from scratch, in its arrival order: must be the same bits
        int t = 0;
#pragma unroll
        for (int b = 0; b < KS; ++b) if (b < k && b != p) C.xs.T(t++) = (uint16_t)(idw[b] & 0xffffu);
        if (j_add >= 0) C.xs.T(t++) = (uint16_t)j_add;
        const double lml_ref = score_set(P, child, C.xs, t, nullptr);
        const double lml_new = -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip_new);
        if (!(lml_ref == lml_new || (lml_ref != lml_ref && lml_new != lml_new))) atomicAdd(C.stats + ST_MISMATCH, 1ull);
    }
    if (!accept) return 0;

    if (kn > KS) {      // outgrows shared memory
        L = small_to_big(P, child, L, S, C, p, j_add, j_rem, rel_now);
        return 1;
    }
    // ---- accepted: tallies / change points, then factor and id list back to shared memory ----
    if (has_rem) {
        const int credit = rel_now - (int)(S.W(p) >> 16);
        if (credit > 0) C.cnt_row[j_rem] += (uint32_t)credit;
        if (C.pev && L.n_pev < C.pev_cap) C.pev[L.n_pev++] = (uint32_t)j_rem | ((uint32_t)(rel_now + 1) << 16);
    }
    if (has_add && C.pev && L.n_pev < C.pev_cap)
        C.pev[L.n_pev++] = (uint32_t)j_add | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
    int rj = 0;
#pragma unroll
    for (int a = 0; a < KS; ++a) {
        if (a < kr) {
            // slot a of the reduced set = old slot a (a < p) or a + 1
            const uint32_t w_old = (a >= p && a + 1 < KS) ? idw[(a + 1 < KS) ? a + 1 : a] : idw[a];
            int r_old = (a >= p && a + 1 < KS) ? rk[(a + 1 < KS) ? a + 1 : a] : rk[a];
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
    L.somc = (L.somc - omc_rem) + omc_add;
    L.q = q_new; L.ip = ip_new; L.lg_ip = lg_new;
    return 1;
}
#undef NL

// lanes whose set lives in the arena
template <bool VERIFY>
__device__ __noinline__ Lane big_move(const SspsDev& P, const int child, const StepDraws& d, Lane L, const SmallAcc& S, const KCtx& C,
                                      const int rel_now, int* acc_kn) {
    unsigned n_t1 = 0; int kn = 0;
    const int acc = parent_move<BigAcc, VERIFY>(P, child, d, L, C.big, S, C, rel_now, n_t1, kn);
    if (n_t1) atomicAdd(C.stats + ST_TIER1, 1ull);
    atomicAdd(C.stats + ST_BIG_STEPS, 1ull);
    if (acc && L.k <= SSPS_KS) L = lane_settle(P, child, L, S, C);     // back into shared memory
    *acc_kn = acc | (kn << 1);
    return L;
}

// ------------------------------------------------------------------------------------------------
// The kernel
// ------------------------------------------------------------------------------------------------
template <class Source, bool VERIFY>
__global__ void __launch_bounds__(SSPS_BLOCK, SSPS_MIN_CTAS) mh_sweeps_kernel(const __grid_constant__ SspsDev P, const __grid_constant__ ChainState st,
                                                                              const __grid_constant__ MhParams prm, const __grid_constant__ Source src) {
    extern __shared__ double smem_d[];
    const int lane = threadIdx.x & 31;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    SmallAcc S;
    S.d = smem_d + threadIdx.x;
    S.w = (uint32_t*)(smem_d + (size_t)2 * SSPS_ROWLEN(SSPS_KS) * SSPS_BLOCK) + threadIdx.x;
    S.r = (uint8_t*)((uint32_t*)(smem_d + (size_t)2 * SSPS_ROWLEN(SSPS_KS) * SSPS_BLOCK) + (size_t)SSPS_KS * SSPS_BLOCK) + threadIdx.x;
    KCtx C;
    C.big.d = prm.big_d + gtid; C.big.w = prm.big_w + gtid; C.big.r = prm.big_r + gtid; C.big.stride = prm.stride; C.big.kcap = prm.kcap;
    C.xs.d = prm.x_d + gtid; C.xs.t = prm.x_t + gtid; C.xs.stride = prm.stride; C.xs.kcap = prm.kcap;
    C.kcap = prm.kcap; C.stats = prm.stats; C.pev_cap = prm.pev_cap;
    const int V = P.V;
    const int thin = prm.thinning;

    unsigned long long s_f3 = 0;
    unsigned s_accp = 0, s_accl = 0, s_k = 0, s_b = 0, s_prop = 0, s_t1 = 0, s_lx = 0;

    for (;;) {
        int task = 0;
        if (lane == 0) task = atomicAdd(prm.task_counter, 1);
        task = __shfl_sync(0xffffffffu, task, 0);
        if (task >= prm.n_tasks) break;
        const long long t_start = clock64();
        const int i = __ldg(prm.order + task / prm.n_slabs);
        const int chain = (task % prm.n_slabs) * 32 + lane;
        if (chain < st.n_chains) {
            const size_t sidx = (size_t)i * st.nch_pad + chain;
            const size_t mc = (size_t)chain * V + i;
            C.cnt_row = st.counts + mc * V;
            C.pev = prm.pev ? prm.pev + mc * prm.pev_cap : nullptr;
            Lane L;
            L.k = st.k[sidx]; L.lam = st.lambda[sidx]; L.n_pev = 0;
            L.q = 0.0; L.ip = 0.0; L.lg_ip = 0.f; L.somc = 0.0; L.mode = MODE_BIG;
            bool dead = false;
            if (L.k > prm.kcap) { atomicExch(prm.stats + ST_OVERFLOW, 1ull); dead = true; }
            else {
                for (int a = 0; a < L.k; ++a) {
                    const uint32_t sl = slot_ref(st, V, i, a, chain);       // id | rank << 16, arrival order
                    C.big.W(a) = sl & 0xffffu; C.big.set_rank(a, (int)(sl >> 16));
                }
                L = lane_settle(P, i, L, S, C);
            }
            L.nrmf = normaliser_fast(P, i, L.lam);
            const float lmarg = __ldg(P.lam_margin + i);
            double lam_sum = 0.0, lam_sq = 0.0, lam_logged = L.lam;
            int n_lev = 0;
            DrawCarry cy; cy.valid = 0; cy.z = 0.f; cy.w = 0u;

            for (int l = 1; l <= prm.n_local && !dead; ++l) {
                const StepDraws d = src.get(chain, i, (uint64_t)(prm.sweep0 + l), (int64_t)(l - 1), cy);

                // ---------------- lambda move (lambda_vec_proposal + Gen.mh) ----------------
                const double lam_new = L.lam + P.sigma * d.z;
                if (lam_new >= P.lambda_min && lam_new <= P.lambda_max) {
                    const double nrm_new = normaliser_fast(P, i, lam_new);
                    const double w = -(lam_new - L.lam) * L.somc - (nrm_new - L.nrmf);
                    const float lgu = fast_lg2((float)d.u_lam);
                    const double diff = SSPS_LN2 * (double)lgu - w;
                    bool acc = diff < 0.0;
                    if (!(d.u_lam > 1e-30 && fabs(diff) > (double)lmarg + 1.2e-6 * ((double)fabsf(lgu) + 1.0) + 1e-9 * fabs(w))) {
                        ++s_lx;
                        write_sorted(L.mode, S, C.big, C.xs, L.k);
                        acc = lambda_exact_decide(P, i, C.xs, L.k, L.lam, lam_new, d.u_lam) != 0;
                    }
                    if (acc) { L.lam = lam_new; L.nrmf = nrm_new; ++s_accl; }
                }

                // ---------------- parent-set move ----------------
                const int rel_now = ceil_div_i(l, thin) - 1;   // recorded samples of this launch that precede the change
                int kn = 0;
                if (L.mode == MODE_SMALL) {
                    s_accp += (unsigned)small_move<VERIFY>(P, i, d, L, S, C, rel_now, s_t1, kn);
                } else {
                    int acc_kn = 0;
                    L = big_move<VERIFY>(P, i, d, L, S, C, rel_now, &acc_kn);
                    s_accp += (unsigned)(acc_kn & 1);
                    kn = acc_kn >> 1;
                }
                ++s_prop;
                s_k += (unsigned)kn;
                s_f3 += (unsigned long long)kn * kn * kn + 6ull * kn * kn + 6ull * kn;
                s_b += (unsigned)(kn * (kn + 1) / 2 + kn);
                if (prm.khist) atomicAdd(prm.khist + (kn < 63 ? kn : 63), 1ull);

                // ---------------- thinned recording ----------------
                if (l % thin == 0) {
                    lam_sum += L.lam; lam_sq = fma(L.lam, L.lam, lam_sq);
                    if (prm.lev_val && L.lam != lam_logged) {
                        if (n_lev < prm.lev_cap) {
                            prm.lev_val[mc * prm.lev_cap + n_lev] = L.lam;
                            prm.lev_idx[mc * prm.lev_cap + n_lev] = (uint16_t)(l / thin);
                            ++n_lev;
                        }
                        lam_logged = L.lam;
                    }
                }
            }

            // flush lazy tallies, store the parent list (arrival order, id | rank << 16) and the state
            if (!dead) {
                const int rec_local = prm.n_local / thin;
                for (int b = 0; b < L.k; ++b) {
                    const uint32_t slw = L.mode == MODE_SMALL ? S.W(b) : C.big.W(b);
                    const int rk = L.mode == MODE_SMALL ? S.rank(b) : C.big.rank(b);
                    const int credit = rec_local - (int)(slw >> 16);
                    if (credit > 0) C.cnt_row[slw & 0xffffu] += (uint32_t)credit;
                    slot_ref(st, V, i, b, chain) = (slw & 0xffffu) | ((uint32_t)rk << 16);
                }
                st.k[sidx] = L.k;
                st.lambda[sidx] = L.lam;
                st.lam_sum[mc] += lam_sum;
                st.lam_sq[mc] += lam_sq;
            }
            if (prm.pev_n) prm.pev_n[mc] = L.n_pev;
            if (prm.lev_n) prm.lev_n[mc] = n_lev;
        }
        __syncwarp();
        if (lane == 0) atomicAdd(prm.vcost + i, (unsigned long long)(clock64() - t_start));
    }

    // stats: warp reduce, one atomic per warp
    unsigned long long v[8] = {s_accp, s_accl, s_k, s_f3, s_b, s_prop, s_t1, s_lx};
    const int slot[8] = {ST_ACCP, ST_ACCL, ST_SUMK, ST_F3, ST_BYTES, ST_PROP, ST_TIER1, ST_LAM_EXACT};
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        unsigned long long x = v[e];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0 && x) atomicAdd(prm.stats + slot[e], x);
    }
}

// vertices by cost, most expensive first (ties by index): order[rank] = vertex
__global__ void order_kernel(const unsigned long long* __restrict__ cost, int V, int* __restrict__ order) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    const unsigned long long ci = cost[i];
    int rank = 0;
    for (int j = 0; j < V; ++j) {
        const unsigned long long cj = cost[j];
        rank += (cj > ci || (cj == ci && j < i)) ? 1 : 0;
    }
    order[rank] = i;
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
