// ssps_uniform.cuh -- the deprecated uniform graph proposal (--proposal uniform), sm_100a.
//
// Reference semantics (paths relative to the gitter-lab/ssps checkout):
//   uniform_update_loop   SSPS/src/state_updates.jl:98-112   for i = 1..V: lambda move of vertex i, then ONE graph move
//   unpaired_edges        SSPS/src/dbn_proposals.jl:142-155  edges (p -> child), p != child, whose reverse is not an edge,
//                                                            sorted as (p, child) tuples
//   uniform_proposal      SSPS/src/dbn_proposals.jl:161-179  idx ~ uniform_discrete(1, V^2 + |ue|); idx <= V^2: the pair
//                                                            (u, v) = (div(idx-1, V), (idx-1) % V): REMOVE u from the parents of v
//                                                            if it is one, else ADD it; above: REVERSE the unpaired edge ue[idx - V^2]
//   uniform_involution    SSPS/src/dbn_proposals.jl:184-230  Gen.update weight = change of (log prior + lml) of the one or two
//                                                            vertices whose parent set changed
//   Gen.mh: accept iff log(u) < weight + log(bound) - log(bound'),  bound' = V^2 + |ue'| of the proposed graph.
// The reference deprecates this proposal (SSPS.jl:88) and --multithread overrides it (SSPS.jl:98); it is kept because it is
// part of the CLI.  A graph move may touch ANY vertex, so the per-vertex factorisation the main kernel rests on is gone: a
// chain is strictly sequential here, one THREAD per chain, every changed set scored from scratch in its arrival order with
// score_set() -- the oracle's Cholesky scorer, bit for bit.  Correct, not tuned.
// Per chain and launch: adjacency bitmaps (parents of v / children of u, rebuilt from the slots at launch start), the cached
// lml per vertex, the number of unpaired edges.  Draw convention: u_act picks idx = 1 + floor(u_act bound), u_acc accepts;
// u_add / u_rem are unused.
#pragma once
#include "ssps_mh.cuh"

struct UniformScratch {
    uint32_t* par;      // [(v W + w) stride + chain]  bit u of word w of row v: u -> v is an edge
    uint32_t* chi;      // the transpose: children of u
    double* lml;        // [v stride + chain]
    double* lam_logged; // [v stride + chain]
    int W, stride;
};

__device__ __forceinline__ uint32_t& ubit_word(uint32_t* base, const UniformScratch& U, int row, int w, int chain) {
    return base[((size_t)row * U.W + w) * U.stride + chain];
}

// lml(v, arrival list of v with parent u added at the end / removed in place); the proposed list stays in xs.T
__device__ inline double uniform_score(const SspsDev& P, const ChainState& st, const Scratch& xs, const int v, const int chain, const int k,
                                       const int u, const bool add, int& kn_out) {
    int kn = 0;
    for (int a = 0; a < k; ++a) {
        const int id = (int)(slot_ref(st, P.V, v, a, chain) & 0xffffu);
        if (add || id != u) xs.T(kn++) = (uint16_t)id;
    }
    if (add) xs.T(kn++) = (uint16_t)u;
    kn_out = kn;
    return score_set(P, v, xs, kn, nullptr);
}

// commit: slots (id | sorted rank << 16, arrival order), k, bitmaps
__device__ inline void uniform_apply(const SspsDev& P, const ChainState& st, const UniformScratch& U, const int v, const int chain, const int u,
                                     const bool add) {
    const int V = P.V;
    const size_t sidx = (size_t)v * st.nch_pad + chain;
    const int k = st.k[sidx];
    if (add) {
        int rank = 0;
        for (int a = 0; a < k; ++a) {
            uint32_t& sl = slot_ref(st, V, v, a, chain);
            const int id = (int)(sl & 0xffffu);
            if (id < u) ++rank; else sl += 1u << 16;
        }
        slot_ref(st, V, v, k, chain) = (uint32_t)u | ((uint32_t)rank << 16);
        st.k[sidx] = k + 1;
        ubit_word(U.par, U, v, u >> 5, chain) |= 1u << (u & 31);
        ubit_word(U.chi, U, u, v >> 5, chain) |= 1u << (v & 31);
    } else {
        int pos = k, rank_u = 1 << 20;
        for (int a = 0; a < k; ++a) { const uint32_t sl = slot_ref(st, V, v, a, chain); if ((int)(sl & 0xffffu) == u) { pos = a; rank_u = (int)(sl >> 16); } }
        for (int a = 0; a + 1 < k; ++a) {
            uint32_t sl = slot_ref(st, V, v, a + (a >= pos ? 1 : 0), chain);
            if ((int)(sl >> 16) > rank_u) sl -= 1u << 16;
            slot_ref(st, V, v, a, chain) = sl;
        }
        st.k[sidx] = k - 1;
        ubit_word(U.par, U, v, u >> 5, chain) &= ~(1u << (u & 31));
        ubit_word(U.chi, U, u, v >> 5, chain) &= ~(1u << (v & 31));
    }
}

template <class Source>
__global__ void __launch_bounds__(128) mh_uniform_kernel(const __grid_constant__ SspsDev P, const __grid_constant__ ChainState st,
                                                         const __grid_constant__ MhParams prm, const __grid_constant__ Source src,
                                                         const __grid_constant__ UniformScratch U, double* __restrict__ x_d,
                                                         uint16_t* __restrict__ x_t) {
    const int chain = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int V = P.V, W = U.W, thin = prm.thinning;
    unsigned long long s_accp = 0, s_accl = 0, s_k = 0, s_f3 = 0, s_prop = 0;
    if (chain < st.n_chains) {
        Scratch xs;
        xs.d = x_d + chain; xs.t = x_t + chain; xs.stride = prm.stride; xs.kcap = prm.kcap;
        // ---- launch start: bitmaps, lml and log state from the slots ----
        for (int v = 0; v < V; ++v)
            for (int w = 0; w < W; ++w) { ubit_word(U.par, U, v, w, chain) = 0u; ubit_word(U.chi, U, v, w, chain) = 0u; }
        for (int v = 0; v < V; ++v) {
            const int k = st.k[(size_t)v * st.nch_pad + chain];
            for (int a = 0; a < k; ++a) {
                const int u = (int)(slot_ref(st, V, v, a, chain) & 0xffffu);
                ubit_word(U.par, U, v, u >> 5, chain) |= 1u << (u & 31);
                ubit_word(U.chi, U, u, v >> 5, chain) |= 1u << (v & 31);
                xs.T(a) = (uint16_t)u;
            }
            U.lml[(size_t)v * U.stride + chain] = score_set(P, v, xs, k, nullptr);
            U.lam_logged[(size_t)v * U.stride + chain] = st.lambda[(size_t)v * st.nch_pad + chain];
            if (prm.pev_n) prm.pev_n[(size_t)chain * V + v] = 0;
            if (prm.lev_n) prm.lev_n[(size_t)chain * V + v] = 0;
        }
        long long n_ue = 0;
        for (int p = 0; p < V; ++p)
            for (int w = 0; w < W; ++w) {
                uint32_t m = ubit_word(U.chi, U, p, w, chain) & ~ubit_word(U.par, U, p, w, chain);
                if ((p >> 5) == w) m &= ~(1u << (p & 31));
                n_ue += __popc(m);
            }
        const long long vv = (long long)V * V;
        for (int l = 1; l <= prm.n_local; ++l) {
            const int rel_now = ceil_div_i(l, thin) - 1;
            for (int i = 0; i < V; ++i) {
                DrawCarry cy; cy.valid = 0; cy.z = 0.f; cy.w = 0u;
                const StepDraws d = src.get(chain, i, (uint64_t)(prm.sweep0 + l), (int64_t)(l - 1), cy);
                // ---- lambda move of vertex i (lambda_vec_proposal + Gen.mh), reference order; the sum runs over the sorted parents ----
                {
                    const size_t li = (size_t)i * st.nch_pad + chain;
                    const double lam = st.lambda[li];
                    const double lam_new = lam + P.sigma * d.z;
                    if (lam_new >= P.lambda_min && lam_new <= P.lambda_max) {
                        const double* __restrict__ omc_i = P.omc + (size_t)V * i;
                        double somc = 0.0;
                        for (int w = 0; w < W; ++w) {
                            uint32_t m = ubit_word(U.par, U, i, w, chain);
                            while (m) { const int b = __ffs(m) - 1; m &= m - 1; somc += __ldg(omc_i + (w << 5) + b); }
                        }
                        const double nrm = prior_normaliser(P, i, lam), nrm_new = prior_normaliser(P, i, lam_new);
                        const double wgt = -(lam_new - lam) * somc - (nrm_new - nrm);
                        if (log(d.u_lam) < wgt) { st.lambda[li] = lam_new; ++s_accl; }
                    }
                }
                // ---- the graph move ----
                const long long bound = vv + n_ue;
                long long idx = 1 + (long long)floor(d.u_act * (double)bound);
                if (idx > bound) idx = bound;
                ++s_prop;
                if (idx <= vv) {
                    const int u = (int)((idx - 1) / V), v = (int)((idx - 1) % V);
                    const bool is_parent = (ubit_word(U.par, U, v, u >> 5, chain) >> (u & 31)) & 1u;
                    const bool rev = (ubit_word(U.par, U, u, v >> 5, chain) >> (v & 31)) & 1u;      // is v -> u an edge?
                    const size_t sv = (size_t)v * st.nch_pad + chain;
                    const int k = st.k[sv];
                    if (!is_parent && k + 1 > prm.kcap) { atomicAdd(prm.stats + ST_KCAP_REJECT, 1ull); s_k += (unsigned)(k + 1); continue; }
                    int kn = 0;
                    const double lml_new = uniform_score(P, st, xs, v, chain, k, u, !is_parent, kn);
                    s_k += (unsigned)kn; s_f3 += (unsigned long long)((unsigned)kn * ((unsigned)kn * ((unsigned)kn + 6u) + 6u));
                    const double lam_v = st.lambda[sv], omc = __ldg(P.omc + (size_t)u + (size_t)V * v);
                    const double dprior = is_parent ? lam_v * omc : -lam_v * omc;
                    const double wgt = dprior + (lml_new - U.lml[(size_t)v * U.stride + chain]);
                    long long n_ue_new = n_ue;
                    if (u != v) n_ue_new += is_parent ? (rev ? 1 : -1) : (rev ? -1 : 1);
                    const double alpha = wgt + (log((double)bound) - log((double)(vv + n_ue_new)));
                    if (log(d.u_acc) < alpha) {
                        ++s_accp;
                        uniform_apply(P, st, U, v, chain, u, !is_parent);
                        U.lml[(size_t)v * U.stride + chain] = lml_new;
                        n_ue = n_ue_new;
                        if (prm.pev) {
                            const size_t mc = (size_t)chain * V + v;
                            int np = prm.pev_n[mc];
                            if (np < prm.pev_cap) prm.pev[mc * prm.pev_cap + np++] = (uint32_t)u | ((uint32_t)(rel_now + 1) << 16) | (is_parent ? 0u : 0x80000000u);
                            prm.pev_n[mc] = np;
                        }
                    }
                } else {
                    // the r-th unpaired edge in (p, child) order: u -> v becomes v -> u
                    long long r = idx - vv - 1;
                    int u = -1, v = -1;
                    for (int p = 0; p < V && u < 0; ++p)
                        for (int w = 0; w < W; ++w) {
                            uint32_t m = ubit_word(U.chi, U, p, w, chain) & ~ubit_word(U.par, U, p, w, chain);
                            if ((p >> 5) == w) m &= ~(1u << (p & 31));
                            const int c = __popc(m);
                            if (r < c) { for (long long e = 0; e < r; ++e) m &= m - 1; u = p; v = (w << 5) + __ffs(m) - 1; break; }
                            r -= c;
                        }
                    if (u < 0) continue;        // cannot happen: n_ue counts exactly these bits
                    const size_t sv = (size_t)v * st.nch_pad + chain, su = (size_t)u * st.nch_pad + chain;
                    const int kv = st.k[sv], ku = st.k[su];
                    if (ku + 1 > prm.kcap) { atomicAdd(prm.stats + ST_KCAP_REJECT, 1ull); s_k += (unsigned)(ku + 1); continue; }
                    int kn1 = 0, kn2 = 0;
                    const double lml1 = uniform_score(P, st, xs, v, chain, kv, u, false, kn1);
                    const double w1 = st.lambda[sv] * __ldg(P.omc + (size_t)u + (size_t)V * v) + (lml1 - U.lml[(size_t)v * U.stride + chain]);
                    const double lml2 = uniform_score(P, st, xs, u, chain, ku, v, true, kn2);
                    const double w2 = -st.lambda[su] * __ldg(P.omc + (size_t)v + (size_t)V * u) + (lml2 - U.lml[(size_t)u * U.stride + chain]);
                    s_k += (unsigned)kn2; s_f3 += (unsigned long long)((unsigned)kn2 * ((unsigned)kn2 * ((unsigned)kn2 + 6u) + 6u));
                    const double alpha = w1 + w2;       // |ue| unchanged: the proposal ratio is 1
                    if (log(d.u_acc) < alpha) {
                        ++s_accp;
                        uniform_apply(P, st, U, v, chain, u, false);
                        uniform_apply(P, st, U, u, chain, v, true);
                        U.lml[(size_t)v * U.stride + chain] = lml1;
                        U.lml[(size_t)u * U.stride + chain] = lml2;
                        if (prm.pev) {
                            size_t mc = (size_t)chain * V + v;
                            int np = prm.pev_n[mc];
                            if (np < prm.pev_cap) prm.pev[mc * prm.pev_cap + np++] = (uint32_t)u | ((uint32_t)(rel_now + 1) << 16);
                            prm.pev_n[mc] = np;
                            mc = (size_t)chain * V + u;
                            np = prm.pev_n[mc];
                            if (np < prm.pev_cap) prm.pev[mc * prm.pev_cap + np++] = (uint32_t)v | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
                            prm.pev_n[mc] = np;
                        }
                    }
                }
            }
            // ---- thinned recording ----
            if (l % thin == 0) {
                for (int v = 0; v < V; ++v) {
                    const size_t mc = (size_t)chain * V + v;
                    uint32_t* cnt_row = st.counts + mc * V;
                    for (int w = 0; w < W; ++w) {
                        uint32_t m = ubit_word(U.par, U, v, w, chain);
                        while (m) { const int b = __ffs(m) - 1; m &= m - 1; cnt_row[(w << 5) + b] += 1u; }
                    }
                    const double lam = st.lambda[(size_t)v * st.nch_pad + chain];
                    st.lam_sum[mc] += lam; st.lam_sq[mc] = fma(lam, lam, st.lam_sq[mc]);
                    if (prm.lev_val && lam != U.lam_logged[(size_t)v * U.stride + chain]) {
                        int nl = prm.lev_n[mc];
                        if (nl < prm.lev_cap) {
                            prm.lev_val[mc * prm.lev_cap + nl] = lam;
                            prm.lev_idx[mc * prm.lev_cap + nl] = (uint16_t)(l / thin);
                            prm.lev_n[mc] = nl + 1;
                        }
                        U.lam_logged[(size_t)v * U.stride + chain] = lam;
                    }
                }
            }
        }
    }
    unsigned long long vals[5] = {s_accp, s_accl, s_k, s_f3, s_prop};
    const int slot[5] = {ST_ACCP, ST_ACCL, ST_SUMK, ST_F3, ST_PROP};
#pragma unroll
    for (int e = 0; e < 5; ++e) {
        unsigned long long x = vals[e];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0 && x) atomicAdd(prm.stats + slot[e], x);
    }
}
