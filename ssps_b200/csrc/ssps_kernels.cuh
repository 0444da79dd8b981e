// ssps_kernels.cuh -- sm_100a device code of the SSPS hot path.
//
//   K1  standardize_kernel / lag_gather_kernel / gram_kernel : lagged design -> Z, Y -> G, R, yty
//   K1' tables_kernel                                        : prob_param, move tables, log table
//   K2  score_kernel                                         : batched (child, S) -> lml, logprior
//   K3  init_chains_kernel, mh_sweeps_kernel<Source>         : persistent fused MH sweep kernel
//   K4  psrf_sums_kernel, psrf_finish_kernel, edge_sum_kernel: tallies -> posterior / split-PSRF
//
// Everything is IEEE fp64.  The file is compiled with -fmad=false and every fused multiply-add
// is an explicit fma(), so the operation order is exactly the one the CPU oracle's
// kernel-faithful path (oracle/ssps_oracle.c: orc_lml_chol, chain_step) restates.
//
// Reference semantics (paths relative to the gitter-lab/ssps checkout):
//   compute_lml                SSPS/src/dbn_distributions.jl:235-248
//   ConfParentPrior.logpdf     SSPS/src/dbn_distributions.jl:96-153   (closed form)
//   compute_move_probs         SSPS/src/dbn_proposals.jl:11-16
//   smart_proposal/involution  SSPS/src/dbn_proposals.jl:60-130
//   lambda_vec_proposal        SSPS/src/dbn_proposals.jl:244-248
//   vertex_lambda_update_loop  SSPS/src/state_updates.jl:35-51
//   update_results_storediff   SSPS/src/state_updates.jl:161-227
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SSPS_PIVOT_TOL 9.094947017729282e-13  // 2^-40
#define SSPS_TRI(k) (((k) * ((k) + 1)) / 2)

enum { ACT_ADD = 1, ACT_SWAP = 2, ACT_REM = 3 };

struct SspsDev {
    int n, V;
    double dn, shrink, log1pn;           // n, n/(n+1), log(1+n)
    double lambda_min, lambda_max, sigma;
    const double* G;                     // V x V (symmetric)
    const double* R;                     // V x V col-major: R[j + V*i] = z_j . y_i
    const double* yty;                   // V
    const double* omc;                   // V x V col-major: 1 - c_ji at [j + V*i]
    const double* conf;                  // V x V col-major
    const int* grp_off;                  // V+1  : distinct (1-c) groups per child
    const double* grp_omc;
    const double* grp_mult;
    const double* mvP;                   // [i][k][3]   k = 0..V : P(add), P(swap), P(remove)
    const double* mvLogP;                // [i][k][3]
    const double* logint;                // log(m), m = 0..V+1
    // fp32 screening tables for the lambda move (K3): per child the groups with 1-c != 0 as
    // (-(1-c) log2 e, mult ln 2), and the screening margin of the child
    const int* fgrp_off;                 // V+1
    const float2* fgrp;
    const float* lam_margin;             // V
};

// ------------------------------------------------------------------------------------------------
// Philox4x32-10 and the per-step draw convention (shared with the oracle)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4]) {
    // rolled on purpose: the MH kernel is bound by instruction fetch, not issue (profiles/): 140 fewer static
    // instructions per call site measured +2.5 % proposals/s
#pragma unroll 1
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ void philox_block(uint64_t seed, uint32_t chain, uint32_t vertex, uint64_t sweep,
                                             uint32_t block, uint32_t out[4]) {
    philox4x32_10(block | (vertex << 8), (uint32_t)sweep, (uint32_t)(sweep >> 32), chain,
                  (uint32_t)seed, (uint32_t)(seed >> 32), out);
}
__device__ __forceinline__ double u32_to_unit(uint32_t w) { return ((double)w + 0.5) * 2.3283064365386963e-10; }
__device__ __forceinline__ double u53_to_unit(uint32_t hi, uint32_t lo) {
    uint64_t x = (((uint64_t)hi << 32) | lo) >> 11;
    return (double)x * 1.1102230246251565e-16;
}

// ------------------------------------------------------------------------------------------------
// Per-thread scratch: lower-triangular factor (diagonal slot = 1/d), w, and the proposed set,
// interleaved across threads so that a warp touching the same logical index is conflict-free /
// coalesced.  Two homes: shared memory (|S'| <= SSPS_KS) and a global arena (larger sets).
// ------------------------------------------------------------------------------------------------
// -DSSPS_BOUNDS: index checks on the scratch / Stage / slot accessors, counted in a device symbol (ssps_bounds_failures).
// compute-sanitizer is closed on the build pool; this is the memcheck stand-in that the GPU test suite is run under
// (tools/bounds_run.sh, profiles/r2_bounds_check.txt).
#ifdef SSPS_BOUNDS
__device__ unsigned long long g_ssps_bounds_fail = 0ull;
#define SSPS_CHK(cond) do { if (!(cond)) atomicAdd(&g_ssps_bounds_fail, 1ull); } while (0)
#else
#define SSPS_CHK(cond) do { } while (0)
#endif

struct Scratch {
    double* d;        // &arena[tid], logical index e lives at d[e * stride]
    uint16_t* t;      // &tarena[tid]
    int stride;
    int kcap;         // capacity in rows
    __device__ __forceinline__ double& L(int e) const { SSPS_CHK(e >= 0 && e < SSPS_TRI(kcap)); return d[(size_t)e * stride]; }
    __device__ __forceinline__ double& w(int a) const { SSPS_CHK(a >= 0 && a < kcap); return d[(size_t)(SSPS_TRI(kcap) + a) * stride]; }
    __device__ __forceinline__ uint16_t& T(int a) const { SSPS_CHK(a >= 0 && a < kcap); return t[(size_t)a * stride]; }
    __device__ __forceinline__ uint16_t& S(int a) const { SSPS_CHK(a >= 0 && a < kcap); return t[(size_t)(kcap + a) * stride]; }   // second id list (K3 only)
};

// lml(child, S') for the sorted set held in sc.T(0..kn-1).  Mirrors orc_lml_chol operation for
// operation.  min_ratio (nullable) receives the smallest accepted pivot ratio.
__device__ __forceinline__ double score_set(const SspsDev& P, int child, const Scratch& sc, int kn,
                                            double* min_ratio) {
    const double* __restrict__ Ri = P.R + (size_t)P.V * child;
    double q = 0.0, minr = __longlong_as_double(0x7ff0000000000000LL);
    for (int a = 0; a < kn; ++a) {
        const int ja = sc.T(a);
        const double* __restrict__ Gj = P.G + (size_t)ja * P.V;
        const int ra = SSPS_TRI(a);
        for (int b = 0; b < a; ++b) {
            const int rb = SSPS_TRI(b);
            double s = __ldg(Gj + sc.T(b));
            for (int m = 0; m < b; ++m) s = fma(-sc.L(ra + m), sc.L(rb + m), s);
            sc.L(ra + b) = s * sc.L(rb + b);
        }
        const double gjj = __ldg(Gj + ja);
        double d2 = gjj;
        double t = __ldg(Ri + ja);
        for (int m = 0; m < a; ++m) {
            const double l = sc.L(ra + m);
            d2 = fma(-l, l, d2);
            t = fma(-l, sc.w(m), t);
        }
        if (d2 > SSPS_PIVOT_TOL * gjj) {
            const double ratio = d2 / gjj;
            if (ratio < minr) minr = ratio;
            const double invd = 1.0 / sqrt(d2);
            sc.L(ra + a) = invd;
            const double wa = t * invd;
            sc.w(a) = wa;
            q = fma(wa, wa, q);
        } else if (d2 != d2 || gjj != gjj) {          // NaN data: zero-variance column (quirk Q4)
            const double nan = __longlong_as_double(0x7ff8000000000000LL);
            sc.L(ra + a) = nan; sc.w(a) = nan; q = nan;
        } else {                                      // numerically dependent column: contributes nothing
            sc.L(ra + a) = 0.0; sc.w(a) = 0.0;
            for (int m = 0; m < a; ++m) sc.L(ra + m) = 0.0;
        }
    }
    if (min_ratio) *min_ratio = minr;
    const double ip = P.yty[child] - P.shrink * q;
    return -0.5 * (double)kn * P.log1pn - 0.5 * P.dn * log(ip);
}

// normaliser of the edge prior: sum_j log(1 + exp(-lambda (1 - c_ji))), grouped by distinct value
__device__ __forceinline__ double prior_normaliser(const SspsDev& P, int child, double lambda) {
    double acc = 0.0;
    const int g1 = P.grp_off[child + 1];
    for (int g = P.grp_off[child]; g < g1; ++g)
        acc = fma(__ldg(P.grp_mult + g), log1p(exp(-lambda * __ldg(P.grp_omc + g))), acc);
    return acc;
}

// ------------------------------------------------------------------------------------------------
// K1: standardise columns (mean, corrected std) -- one thread per column keeps the oracle's
// sequential summation order; n is at most a few hundred rows and this runs once per data set.
// src is n x ncols column-major with leading dimension ld; rows [row0, row0+n).
// ------------------------------------------------------------------------------------------------
__global__ void standardize_kernel(const double* __restrict__ src, int n, int ncols, double* __restrict__ dst) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ncols) return;
    const double* x = src + (size_t)j * n;
    double* z = dst + (size_t)j * n;
    double s = 0.0;
    for (int t = 0; t < n; ++t) s += x[t];
    const double mean = s / (double)n;
    double ss = 0.0;
    for (int t = 0; t < n; ++t) { const double d = x[t] - mean; ss = fma(d, d, ss); }
    const double sd = sqrt(ss / (double)(n - 1));
    for (int t = 0; t < n; ++t) z[t] = (x[t] - mean) / sd;
}

// combine_X on the device: row_map[t] = source row of lagged row t (Xminus); +1 gives Xplus.
__global__ void lag_gather_kernel(const double* __restrict__ X, int total_rows, int V, const int* __restrict__ row_map,
                                  int n, double* __restrict__ Xminus, double* __restrict__ Xplus) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (size_t)n * V) return;
    int t = (int)(e % n), j = (int)(e / n);
    int r = row_map[t];
    Xminus[e] = X[(size_t)j * total_rows + r];
    Xplus[e] = X[(size_t)j * total_rows + r + 1];
}

// G = A'B for column-major A (n x Va), B (n x Vb); out is Va x Vb column-major.  32x32 output tile
// per CTA, the time axis staged through shared memory in slabs of 32 rows; every output element
// is accumulated by ONE thread in ascending-t order with fma (bit-identical to the oracle).
#define GRAM_TILE 32
__global__ void __launch_bounds__(GRAM_TILE * 8) gram_kernel(const double* __restrict__ A, const double* __restrict__ B,
                                                             int n, int Va, int Vb, double* __restrict__ out) {
    __shared__ double sa[GRAM_TILE][GRAM_TILE + 1];   // [t][col]
    __shared__ double sb[GRAM_TILE][GRAM_TILE + 1];
    const int a0 = blockIdx.x * GRAM_TILE, b0 = blockIdx.y * GRAM_TILE;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8 threads, 4 outputs each
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int t0 = 0; t0 < n; t0 += GRAM_TILE) {
        // coalesced along t (columns are contiguous)
        for (int c = ty; c < GRAM_TILE; c += 8) {
            const int t = t0 + tx;
            sa[tx][c] = (t < n && a0 + c < Va) ? A[(size_t)(a0 + c) * n + t] : 0.0;
            sb[tx][c] = (t < n && b0 + c < Vb) ? B[(size_t)(b0 + c) * n + t] : 0.0;
        }
        __syncthreads();
        const int tmax = min(GRAM_TILE, n - t0);
        for (int t = 0; t < tmax; ++t) {
            const double av = sa[t][tx];
#pragma unroll
            for (int r = 0; r < 4; ++r) acc[r] = fma(av, sb[t][ty + 8 * r], acc[r]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int a = a0 + tx, b = b0 + ty + 8 * r;
        if (a < Va && b < Vb) out[(size_t)a + (size_t)Va * b] = acc[r];
    }
}

__global__ void yty_kernel(const double* __restrict__ Y, int n, int V, double* __restrict__ yty) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    const double* y = Y + (size_t)i * n;
    double acc = 0.0;
    for (int t = 0; t < n; ++t) acc = fma(y[t], y[t], acc);
    yty[i] = acc;
}

// prior -> conf / omc (values <= 0 mean "absent", dbn_preprocess.jl:36)
__global__ void prior_kernel(const double* __restrict__ prior, size_t vv, double* __restrict__ conf, double* __restrict__ omc) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= vv) return;
    const double c = prior[e] > 0.0 ? prior[e] : 0.0;
    conf[e] = c; omc[e] = 1.0 - c;
}

// prob_param_i = 1 / log2(V / max(sum_j c_ji, 2))   (SSPS.jl:164-165)
__global__ void prob_param_kernel(const double* __restrict__ conf, int V, double* __restrict__ prob_param) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    double s = 0.0;
    for (int j = 0; j < V; ++j) s += conf[(size_t)j + (size_t)V * i];
    const double cnt = fmax(s, 2.0);
    prob_param[i] = 1.0 / log2((double)V / cnt);
}

// compute_move_probs (dbn_proposals.jl:11-16) tabulated for every (child, in-degree)
__global__ void move_table_kernel(const double* __restrict__ prob_param, int V, double* __restrict__ mvP,
                                  double* __restrict__ mvLogP, double* __restrict__ logint) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < (size_t)V + 2) logint[e] = log((double)e);
    if (e >= (size_t)V * (V + 1)) return;
    const int i = (int)(e / (V + 1)), k = (int)(e % (V + 1));
    const double frac = pow((double)k / (double)V, prob_param[i]);
    const double comp = 1.0 - frac;
    const double swp = 2.0 * frac * comp;
    const double tot = frac + swp + comp;
    const double p0 = comp / tot, p1 = swp / tot, p2 = frac / tot;
    mvP[e * 3 + 0] = p0; mvP[e * 3 + 1] = p1; mvP[e * 3 + 2] = p2;
    mvLogP[e * 3 + 0] = log(p0); mvLogP[e * 3 + 1] = log(p1); mvLogP[e * 3 + 2] = log(p2);
}

// ------------------------------------------------------------------------------------------------
// K2: batched scorer (the level-1 parity hook).  One thread per (child, S); scratch from the arena.
// ------------------------------------------------------------------------------------------------
__global__ void score_kernel(SspsDev P, int n_sets, const int* __restrict__ child, const int* __restrict__ offsets,
                             const int* __restrict__ members, const double* __restrict__ lambda,
                             double* __restrict__ arena_d, uint16_t* __restrict__ arena_t, int kcap,
                             double* __restrict__ lml_out, double* __restrict__ lp_out, double* __restrict__ minp_out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_sets) return;
    const int stride = gridDim.x * blockDim.x;
    Scratch sc{arena_d + s, arena_t + s, stride, kcap};
    const int i = child[s], k = offsets[s + 1] - offsets[s];
    const int* mem = members + offsets[s];
    for (int a = 0; a < k; ++a) sc.T(a) = (uint16_t)mem[a];
    double minr;
    const double lml = score_set(P, i, sc, k, &minr);
    if (lml_out) lml_out[s] = lml;
    if (minp_out) minp_out[s] = minr;
    if (lp_out) {
        const double lam = lambda[s];
        double somc = 0.0;
        for (int a = 0; a < k; ++a) somc += P.omc[(size_t)mem[a] + (size_t)P.V * i];
        lp_out[s] = -lam * somc - prior_normaliser(P, i, lam);
    }
}

// ------------------------------------------------------------------------------------------------
// Chain state (device-resident between launches), laid out chain-fastest so that a warp = 32
// chains of the same vertex reads and writes it coalesced:
//   k      [vertex][chain]            int32
//   lambda [vertex][chain]            fp64
//   slots  [vertex][a][chain]         uint32 = parent id | (rank << 16), a < V: ARRIVAL order (the order in
//                                     which the parents joined the set; the initial state arrives sorted),
//                                     rank = position of the id in the sorted parent list
// ------------------------------------------------------------------------------------------------
struct ChainState {
    int n_chains, nch_pad;
    int* k; double* lambda; uint32_t* slots;
    uint32_t* counts;       // [chain][child][parent] cumulative tallies over recorded samples
    double* lam_sum;        // [chain][vertex]
    double* lam_sq;
};

__device__ __forceinline__ uint32_t& slot_ref(const ChainState& st, int V, int vertex, int a, int chain) {
    SSPS_CHK(vertex >= 0 && vertex < V && a >= 0 && a < V && chain >= 0 && chain < st.nch_pad);
    return st.slots[((size_t)vertex * V + a) * st.nch_pad + chain];
}

// Gen.generate: lambda_i ~ U(lmin, lmax); S_i ~ ConfParentPrior.random (dbn_distributions.jl:75-89)
__global__ void init_chains_kernel(SspsDev P, ChainState st, uint64_t seed, uint32_t chain_offset) {
    const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int chain = (int)(m % st.nch_pad), vertex = (int)(m / st.nch_pad);
    if (vertex >= P.V || chain >= st.n_chains) return;
    uint32_t w[4];
    const uint32_t cid = chain_offset + (uint32_t)chain;
    philox_block(seed, cid, (uint32_t)vertex, 0, 0, w);
    const double u = u53_to_unit(w[0], w[1]);
    const double lam = u * (P.lambda_max - P.lambda_min) + P.lambda_min;
    const double enl = exp(-lam);
    int k = 0;
    for (int j = 0; j < P.V; ++j) {
        if ((j & 3) == 0) philox_block(seed, cid, (uint32_t)vertex, 0, 1u + (uint32_t)(j >> 2), w);
        const double conf = P.conf[(size_t)j + (size_t)P.V * vertex];
        const double p = enl / (enl + exp(-conf * lam));
        if (u32_to_unit(w[j & 3]) < p) { slot_ref(st, P.V, vertex, k, chain) = (uint32_t)j | ((uint32_t)k << 16); ++k; }
    }
    st.k[(size_t)vertex * st.nch_pad + chain] = k;
    st.lambda[(size_t)vertex * st.nch_pad + chain] = lam;
}

// largest in-degree over all micro-chains (ssps_init_chains checks it against max_indegree before any sweep runs)
__global__ void max_indegree_kernel(ChainState st, int V, int* __restrict__ out) {
    const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int chain = (int)(m % st.nch_pad), vertex = (int)(m / st.nch_pad);
    int k = 0;
    if (vertex < V && chain < st.n_chains) k = st.k[(size_t)vertex * st.nch_pad + chain];
    k = __reduce_max_sync(0xffffffffu, k);
    if ((threadIdx.x & 31) == 0 && k > 0) atomicMax(out, k);
}

// ------------------------------------------------------------------------------------------------
// state read-back helpers
// ------------------------------------------------------------------------------------------------
__global__ void export_chain_kernel(SspsDev P, ChainState st, int chain, int* __restrict__ indeg,
                                    int* __restrict__ members_padded /* V x V */, double* __restrict__ lambda) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.V) return;
    const size_t sidx = (size_t)i * st.nch_pad + chain;
    const int k = st.k[sidx];
    indeg[i] = k; lambda[i] = st.lambda[sidx];
    for (int a = 0; a < k; ++a) {       // slots are in arrival order; the word carries the sorted rank
        const uint32_t sl = slot_ref(st, P.V, i, a, chain);
        members_padded[(size_t)i * P.V + (sl >> 16)] = (int)(sl & 0xffffu);
    }
}

__global__ void import_state_kernel(SspsDev P, ChainState st, const int* __restrict__ offsets,
                                    const int* __restrict__ members, const double* __restrict__ lambda) {
    const size_t m = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int chain = (int)(m % st.nch_pad), vertex = (int)(m / st.nch_pad);
    if (vertex >= P.V || chain >= st.n_chains) return;
    const size_t mc = (size_t)chain * P.V + vertex;
    const int k = offsets[mc + 1] - offsets[mc];
    for (int a = 0; a < k; ++a) slot_ref(st, P.V, vertex, a, chain) = (uint32_t)members[offsets[mc] + a] | ((uint32_t)a << 16);
    st.k[(size_t)vertex * st.nch_pad + chain] = k;
    st.lambda[(size_t)vertex * st.nch_pad + chain] = lambda[mc];
}

// lml of the current state of one chain (for ssps_get_state)
__global__ void state_lml_kernel(SspsDev P, ChainState st, int chain, double* __restrict__ arena_d,
                                 uint16_t* __restrict__ arena_t, int kcap, int stride, double* __restrict__ lml) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.V) return;
    Scratch sc{arena_d + i, arena_t + i, stride, kcap};
    const int k = st.k[(size_t)i * st.nch_pad + chain];
    if (k > kcap) { lml[i] = __longlong_as_double(0x7ff8000000000000LL); return; }
    for (int a = 0; a < k; ++a) sc.T(a) = (uint16_t)(slot_ref(st, P.V, i, a, chain) & 0xffffu);
    lml[i] = score_set(P, i, sc, k, nullptr);
}

// ------------------------------------------------------------------------------------------------
// K4: tallies -> posterior / split-PSRF sufficient statistics
// counts snapshots are [chain][V*V] uint32 cumulative; c_b / c_s may be NULL (= all zero).
// ------------------------------------------------------------------------------------------------
// Chains are split into gridDim.y chunks (a thread per edge alone leaves V^2 threads walking all the chains: 11 ms at
// V = 100 x 4096 chains).  The tallies are integers below 2^53, so the double atomics are exact and order-independent.
// `out` must be zeroed.
__global__ void edge_sum_kernel(const uint32_t* __restrict__ c_hi, const uint32_t* __restrict__ c_lo, int n_chains,
                                size_t vv, int chunk_len, double* __restrict__ out) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= vv) return;
    const int c0 = blockIdx.y * chunk_len, c1 = min(n_chains, c0 + chunk_len);
    unsigned long long s = 0;
    for (int c = c0; c < c1; ++c) {
        uint32_t hi = c_hi[(size_t)c * vv + e];
        uint32_t lo = c_lo ? c_lo[(size_t)c * vv + e] : 0u;
        s += (unsigned long long)(hi - lo);
    }
    if (s) atomicAdd(out + e, (double)s);
}

// block-wide sum of up to three values per thread, fixed tree (deterministic); result valid in thread 0.  blockDim.x = 256
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c) {
    __shared__ double sh[3][8];
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); c += __shfl_down_sync(0xffffffffu, c, o);
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { sh[0][w] = a; sh[1][w] = b; sh[2][w] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = sh[0][0]; b = sh[1][0]; c = sh[2][0];
        for (int e = 1; e < (int)(blockDim.x >> 5); ++e) { a += sh[0][e]; b += sh[1][e]; c += sh[2][e]; }
    }
}

// one block per vertex, threads stride over the chains (a thread per vertex walked 4096 chains alone: 2 ms)
__global__ void __launch_bounds__(256) lam_sum_kernel(const double* __restrict__ s_hi, const double* __restrict__ s_lo, int n_chains, int V,
                                                      double* __restrict__ out) {
    const int i = blockIdx.x;
    double s = 0.0, z1 = 0.0, z2 = 0.0;
    for (int c = threadIdx.x; c < n_chains; c += blockDim.x) s += s_hi[(size_t)c * V + i] - (s_lo ? s_lo[(size_t)c * V + i] : 0.0);
    block_sum3(s, z1, z2);
    if (threadIdx.x == 0) out[i] = s;
}

// per edge: sum over half-chains of m, m^2, m(1-m)N/(N-1)   (postprocess_samples.jl:99-138)
__global__ void psrf_sums_kernel(const uint32_t* __restrict__ c_len, const uint32_t* __restrict__ c_s,
                                 const uint32_t* __restrict__ c_b, int n_chains, size_t vv, double N1, double N2,
                                 int chunk_len, double* __restrict__ partial /* [chunk][3][vv]; one chunk: the sums themselves */) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= vv) return;
    const int c0 = blockIdx.y * chunk_len, c1 = min(n_chains, c0 + chunk_len);
    double sm = 0.0, sm2 = 0.0, sv = 0.0;
    for (int c = c0; c < c1; ++c) {
        const uint32_t xl = c_len[(size_t)c * vv + e];
        const uint32_t xs = c_s[(size_t)c * vv + e];
        const uint32_t xb = c_b ? c_b[(size_t)c * vv + e] : 0u;
        const double m1 = (double)(xs - xb) / N1, m2 = (double)(xl - xs) / N2;
        sm += m1; sm += m2;
        sm2 = fma(m1, m1, sm2); sm2 = fma(m2, m2, sm2);
        sv += m1 * (1.0 - m1) * N1 / (N1 - 1.0);
        sv += m2 * (1.0 - m2) * N2 / (N2 - 1.0);
    }
    double* __restrict__ dst = partial + (size_t)blockIdx.y * 3 * vv;
    dst[e] = sm; dst[vv + e] = sm2; dst[2 * vv + e] = sv;
}
// chunk partials -> sums, chunks added in ascending order (deterministic)
__global__ void psrf_sums_reduce_kernel(const double* __restrict__ partial, int chunks, size_t n3 /* 3 vv */, double* __restrict__ sums) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n3) return;
    double s = 0.0;
    for (int ch = 0; ch < chunks; ++ch) s += partial[(size_t)ch * n3 + e];
    sums[e] = s;
}

// lambda_i: per half-chain mean and corrected variance from cumulative sums / sums of squares
__global__ void __launch_bounds__(256) psrf_lambda_sums_kernel(const double* __restrict__ s_len, const double* __restrict__ s_s,
                                        const double* __restrict__ s_b, const double* __restrict__ q_len,
                                        const double* __restrict__ q_s, const double* __restrict__ q_b, int n_chains,
                                        int V, double N1, double N2, double* __restrict__ sums /* 3*V */) {
    const int i = blockIdx.x;
    double sm = 0.0, sm2 = 0.0, sv = 0.0;
    for (int c = threadIdx.x; c < n_chains; c += blockDim.x) {
        const size_t e = (size_t)c * V + i;
        const double sb = s_b ? s_b[e] : 0.0, qb = q_b ? q_b[e] : 0.0;
        const double a1 = s_s[e] - sb, a2 = s_len[e] - s_s[e];
        const double b1 = q_s[e] - qb, b2 = q_len[e] - q_s[e];
        const double m1 = a1 / N1, m2 = a2 / N2;
        const double v1 = (b1 - N1 * m1 * m1) / (N1 - 1.0), v2 = (b2 - N2 * m2 * m2) / (N2 - 1.0);
        sm += m1 + m2; sm2 += m1 * m1 + m2 * m2; sv += fmax(v1, 0.0) + fmax(v2, 0.0);
    }
    block_sum3(sm, sm2, sv);
    if (threadIdx.x == 0) { sums[i] = sm; sums[V + i] = sm2; sums[2 * V + i] = sv; }
}

// psrf = sqrt(((n-1)/n W + B/n) / W),  B = n * sum (m - mbar)^2 / H (1/H: quirk Q6), W = sum var / H
__global__ void psrf_finish_kernel(const double* __restrict__ sums, size_t count, size_t stride, double H, double n,
                                   double* __restrict__ psrf, unsigned long long* __restrict__ max_bits) {
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= count) return;
    const double sm = sums[e], sm2 = sums[stride + e], sv = sums[2 * stride + e];
    const double mbar = sm / H;
    double bs = sm2 - H * mbar * mbar;       // sum (m - mbar)^2
    if (bs < 0.0) bs = 0.0;
    const double B = n * bs / H;
    const double W = sv / H;
    const double vp = (n - 1.0) * W / n + B / n;
    const double r = sqrt(vp / W);           // W == 0 -> NaN or inf -> excluded from the max
    psrf[e] = r;
    if (max_bits && r == r && W > 0.0) atomicMax(max_bits, (unsigned long long)__double_as_longlong(r));
}

// ------------------------------------------------------------------------------------------------
// FP64 peak microbenchmark (roofline denominator; MEASURED_PEAKS.json has no fp64 figure)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, int iters) {
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ------------------------------------------------------------------------------------------------
// Gather-bandwidth microbenchmarks (roofline denominators of SURVEY 8d's `achieved_gather`): random 8-byte loads
// from a table of the size of G and R (160 kB at V = 100), eight independent loads in flight per thread.
//   MODE 0  ld.global.nc  (the kernel's __ldg path: L1-resident table)
//   MODE 1  ld.global.cg  (cached in L2 only: the L2-resident figure)
//   MODE 2  shared memory (the table's first 48 kB staged in shared memory)
// The figure returned is USEFUL bytes (8 per load) per second, the same unit as the kernel's gathered bytes.
// ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) gather_peak_kernel(const double* __restrict__ table, unsigned n_elems, int iters, double* __restrict__ out) {
    __shared__ double sh[MODE == 2 ? 6144 : 1];
    unsigned ne = n_elems;
    if (MODE == 2) {
        ne = n_elems < 6144u ? n_elems : 6144u;
        for (unsigned e = threadIdx.x; e < ne; e += blockDim.x) sh[e] = table[e];
        __syncthreads();
    }
    unsigned s[8];
    double acc[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { s[u] = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 40503u * (u + 1); acc[u] = 0.0; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            s[u] = s[u] * 1664525u + 1013904223u;
            const unsigned idx = __umulhi(s[u], ne);
            if (MODE == 0) acc[u] += __ldg(table + idx);
            else if (MODE == 1) acc[u] += __ldcg(table + idx);
            else acc[u] += sh[idx];
        }
    }
    double r = 0.0;
#pragma unroll
    for (int u = 0; u < 8; ++u) r += acc[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
