// ssps_poly.cuh -- regression_deg != 1: interaction columns (sm_100a).
//
// Reference semantics (paths relative to the gitter-lab/ssps checkout):
//   compute_B_col / construct_B   SSPS/src/dbn_distributions.jl:168-194
//       for deg = 1..regression_deg, for each Combinatorics.combinations(sorted S, deg) (lexicographic):
//       elementwise product of the RAW Xminus columns, then (x - mean) / std (corrected, ddof 1)
//   compute_lml                   SSPS/src/dbn_distributions.jl:235-248
//       lml = -1/2 width log(1 + n) - 1/2 n log( y'y - n/(n+1) |P_B y|^2 )
//   regression_deg <= 0 (the CLI default -1, "full") builds an EMPTY design in the reference (quirk Q2):
//       width = 0, lml = -1/2 n log(y'y) for every parent set -- reproduced as is.
//
// The product columns are re-standardised per set, so their Gram matrix cannot be tabulated once
// like G = Z'Z of the degree-1 path: every proposal is scored FROM SCRATCH.  One thread scores one
// set: two passes over the n lagged rows (column means; then the Gram matrix of the centred columns
// and their products with y), a scaling to the standardised design, and the same pivot-skipping
// Cholesky as the degree-1 path (column order = the reference's).  Scratch lives in a global arena,
// interleaved across threads so that a warp touching the same logical index is coalesced.
//
// Cost per scored set: n (m + m (m + 1) / 2 + m) multiply-adds for m = sum_{d <= deg} C(k, d) columns
// (SURVEY 8d); the arithmetic is fp64 on the DFMA pipe.
#pragma once
#include "ssps_kernels.cuh"

#define SSPS_POLY_KMAX 12          // parents per set the interaction path handles
#define SSPS_POLY_MMAX 48          // design columns per set

struct PolyDev {
    const double* Xminus;          // n x V column-major, RAW
    const double* Y;               // n x V column-major, standardised
    int deg;                       // regression_deg (<= 0: empty design)
};

// per-thread scratch, element e at d[e * stride]
struct PolyScratch {
    double* d; int stride;
    __device__ __forceinline__ double& at(int e) const { return d[(size_t)e * stride]; }
    // layout: mean[M] | cval[M] | rhs[M] | L[TRI(M)]  with M = SSPS_POLY_MMAX
    __device__ __forceinline__ double& mean(int c) const { return at(c); }
    __device__ __forceinline__ double& cval(int c) const { return at(SSPS_POLY_MMAX + c); }
    __device__ __forceinline__ double& rhs(int c) const { return at(2 * SSPS_POLY_MMAX + c); }
    __device__ __forceinline__ double& L(int a, int b) const { return at(3 * SSPS_POLY_MMAX + SSPS_TRI(a) + b); }
};
#define SSPS_POLY_SCRATCH_DOUBLES (3 * SSPS_POLY_MMAX + SSPS_TRI(SSPS_POLY_MMAX))

// number of design columns: sum_{d = 1..min(deg, k)} C(k, d); -1 if it exceeds SSPS_POLY_MMAX
__host__ __device__ inline int poly_width(int k, int deg) {
    int m = 0;
    long long c = 1;
    const int dmax = deg < k ? deg : k;
    for (int d = 1; d <= dmax; ++d) {
        c = c * (k - d + 1) / d;
        if (c > SSPS_POLY_MMAX) return -1;
        m += (int)c;
        if (m > SSPS_POLY_MMAX) return -1;
    }
    return m;
}

// bit masks (over positions of the sorted set) of the design columns, in the reference's order
__device__ inline int poly_masks(int k, int deg, unsigned short* masks) {
    int m = 0;
    const int dmax = deg < k ? deg : k;
    int idx[SSPS_POLY_KMAX];
    for (int d = 1; d <= dmax; ++d) {
        for (int a = 0; a < d; ++a) idx[a] = a;
        for (;;) {
            unsigned mk = 0;
            for (int a = 0; a < d; ++a) mk |= 1u << idx[a];
            masks[m++] = (unsigned short)mk;
            int i = d - 1;
            while (i >= 0 && idx[i] == k - d + i) --i;
            if (i < 0) break;
            ++idx[i];
            for (int a = i + 1; a < d; ++a) idx[a] = idx[a - 1] + 1;
        }
    }
    return m;
}

// lml(child, S) for the SORTED set S[0..k).  Returns NaN when the set is outside the path's limits
// (the callers check the limits first and fail loudly).
__device__ inline double poly_score_set(const SspsDev& P, const PolyDev& Q, const int child, const int* S, const int k,
                                        const PolyScratch& sc) {
    const int n = P.n;
    const double yty = P.yty[child];
    const int m = (Q.deg <= 0) ? 0 : poly_width(k, Q.deg);
    if (m < 0 || k > SSPS_POLY_KMAX) return __longlong_as_double(0x7ff8000000000000LL);
    if (m == 0) return -0.5 * P.dn * log(yty);               // empty design (k = 0, or quirk Q2)
    unsigned short masks[SSPS_POLY_MMAX];
    poly_masks(k, Q.deg, masks);
    const double* __restrict__ y = Q.Y + (size_t)child * n;
    double xs[SSPS_POLY_KMAX];
    // pass 1: column means
    for (int c = 0; c < m; ++c) sc.mean(c) = 0.0;
    for (int t = 0; t < n; ++t) {
        for (int a = 0; a < k; ++a) xs[a] = __ldg(Q.Xminus + (size_t)S[a] * n + t);
        for (int c = 0; c < m; ++c) {
            double pr = 1.0;
            for (int a = 0; a < k; ++a) if (masks[c] >> a & 1) pr *= xs[a];
            sc.mean(c) += pr;
        }
    }
    for (int c = 0; c < m; ++c) sc.mean(c) = sc.mean(c) / (double)n;
    // pass 2: Gram matrix of the centred columns, and their products with y
    for (int c = 0; c < m; ++c) { sc.rhs(c) = 0.0; for (int b = 0; b <= c; ++b) sc.L(c, b) = 0.0; }
    for (int t = 0; t < n; ++t) {
        for (int a = 0; a < k; ++a) xs[a] = __ldg(Q.Xminus + (size_t)S[a] * n + t);
        const double yt = __ldg(y + t);
        for (int c = 0; c < m; ++c) {
            double pr = 1.0;
            for (int a = 0; a < k; ++a) if (masks[c] >> a & 1) pr *= xs[a];
            const double cv = pr - sc.mean(c);
            sc.cval(c) = cv;
            sc.rhs(c) = fma(cv, yt, sc.rhs(c));
            for (int b = 0; b <= c; ++b) sc.L(c, b) = fma(cv, sc.cval(b), sc.L(c, b));
        }
    }
    // standardise: B_c = centred column / sd_c, sd_c = sqrt(C_cc / (n - 1)); mean(c) is reused for 1 / sd_c
    for (int c = 0; c < m; ++c) sc.mean(c) = 1.0 / sqrt(sc.L(c, c) / (double)(n - 1));
    for (int c = 0; c < m; ++c) {
        const double ic = sc.mean(c);
        sc.rhs(c) = sc.rhs(c) * ic;
        for (int b = 0; b <= c; ++b) sc.L(c, b) = sc.L(c, b) * ic * sc.mean(b);
    }
    // pivot-skipping Cholesky of B'B in column order; cval(b) is reused for 1 / d_b, rhs -> w
    double q = 0.0;
    for (int a = 0; a < m; ++a) {
        const double gaa = sc.L(a, a);
        double d2 = gaa, tt = sc.rhs(a);
        for (int b = 0; b < a; ++b) {
            double s = sc.L(a, b);
            for (int mm = 0; mm < b; ++mm) s = fma(-sc.L(a, mm), sc.L(b, mm), s);
            const double l = s * sc.cval(b);
            sc.L(a, b) = l;
            d2 = fma(-l, l, d2);
            tt = fma(-l, sc.rhs(b), tt);
        }
        double inv, w;
        if (d2 > SSPS_PIVOT_TOL * gaa) { inv = 1.0 / sqrt(d2); w = tt * inv; q = fma(w, w, q); }
        else if (d2 != d2 || gaa != gaa) { inv = __longlong_as_double(0x7ff8000000000000LL); w = inv; q = inv; }
        else { inv = 0.0; w = 0.0; }
        sc.cval(a) = inv; sc.rhs(a) = w;
    }
    const double ip = yty - P.shrink * q;
    return -0.5 * (double)m * P.log1pn - 0.5 * P.dn * log(ip);
}

// K2 for the interaction path: one thread per (child, S)
__global__ void poly_score_kernel(SspsDev P, PolyDev Q, int n_sets, const int* __restrict__ child, const int* __restrict__ offsets,
                                  const int* __restrict__ members, const double* __restrict__ lambda, double* __restrict__ arena,
                                  double* __restrict__ lml_out, double* __restrict__ lp_out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_sets) return;
    PolyScratch sc{arena + s, (int)(gridDim.x * blockDim.x)};
    const int i = child[s], k = offsets[s + 1] - offsets[s];
    const int* mem = members + offsets[s];
    int S[SSPS_POLY_KMAX];
    for (int a = 0; a < k && a < SSPS_POLY_KMAX; ++a) S[a] = mem[a];
    if (lml_out) lml_out[s] = poly_score_set(P, Q, i, S, k, sc);
    if (lp_out) {
        const double lam = lambda[s];
        double somc = 0.0;
        for (int a = 0; a < k; ++a) somc += P.omc[(size_t)mem[a] + (size_t)P.V * i];
        lp_out[s] = -lam * somc - prior_normaliser(P, i, lam);
    }
}

// ------------------------------------------------------------------------------------------------
// K3 for the interaction path: the same sweep (lambda move, then add / swap / remove on the sorted
// parent list, Gen.mh) as mh_sweeps_kernel, one thread per micro-chain, every proposal scored from
// scratch by poly_score_set, the acceptance test in the reference order with fp64 logarithms.
// Chain state uses the degree-1 layout (slots = id | rank << 16); this kernel keeps the list sorted,
// so rank = position.  Tallies are incremented at every recorded sample (no lazy credit here).
// Included after ssps_mh.cuh (StepDraws, sources, select_move, move_ratio, MhParams).
// ------------------------------------------------------------------------------------------------
template <class Source>
__global__ void __launch_bounds__(128) mh_poly_kernel(const __grid_constant__ SspsDev P, const __grid_constant__ PolyDev Q,
                                                      const __grid_constant__ ChainState st, const __grid_constant__ MhParams prm,
                                                      const __grid_constant__ Source src, double* __restrict__ arena) {
    const int lane = threadIdx.x & 31;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const PolyScratch sc{arena + gtid, (int)(gridDim.x * blockDim.x)};
    const int V = P.V, thin = prm.thinning;
    unsigned long long s_f3 = 0;
    unsigned s_accp = 0, s_accl = 0, s_k = 0, s_prop = 0;
    for (;;) {
        int task = 0;
        if (lane == 0) task = atomicAdd(prm.task_counter, 1);
        task = __shfl_sync(0xffffffffu, task, 0);
        if (task >= prm.n_tasks) break;
        const int i = task / prm.n_slabs;
        const int chain = (task % prm.n_slabs) * 32 + lane;
        if (chain >= st.n_chains) continue;
        const size_t sidx = (size_t)i * st.nch_pad + chain;
        const size_t mc = (size_t)chain * V + i;
        uint32_t* cnt_row = st.counts + mc * V;
        uint32_t* pev = prm.pev ? prm.pev + mc * prm.pev_cap : nullptr;
        int k = st.k[sidx];
        double lam = st.lambda[sidx];
        if (k > SSPS_POLY_KMAX || (Q.deg > 0 && poly_width(k, Q.deg) < 0)) continue;    // cannot happen: refused by ssps_init_chains
        int S[SSPS_POLY_KMAX], T[SSPS_POLY_KMAX];
        for (int a = 0; a < k; ++a) S[(int)(slot_ref(st, V, i, a, chain) >> 16)] = (int)(slot_ref(st, V, i, a, chain) & 0xffffu);
        const double* __restrict__ omc_i = P.omc + (size_t)V * i;
        double lml = poly_score_set(P, Q, i, S, k, sc);
        double lam_sum = 0.0, lam_sq = 0.0, lam_logged = lam;
        int n_pev = 0, n_lev = 0;
        DrawCarry cy; cy.valid = 0; cy.z = 0.f; cy.w = 0u;
        for (int l = 1; l <= prm.n_local; ++l) {
            const StepDraws d = src.get(chain, i, (uint64_t)(prm.sweep0 + l), (int64_t)(l - 1), cy);
            // ---- lambda move (lambda_vec_proposal + Gen.mh), reference order ----
            const double lam_new = lam + P.sigma * d.z;
            if (lam_new >= P.lambda_min && lam_new <= P.lambda_max) {
                double somc = 0.0;
                for (int a = 0; a < k; ++a) somc += __ldg(omc_i + S[a]);
                const double nrm = prior_normaliser(P, i, lam), nrm_new = prior_normaliser(P, i, lam_new);
                const double w = -(lam_new - lam) * somc - (nrm_new - nrm);
                if (log(d.u_lam) < w) { lam = lam_new; ++s_accl; }
            }
            // ---- parent-set move ----
            int action, a_ord, r_ord;
            select_move(P, i, k, d, action, a_ord, r_ord);
            const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;
            int j_add = -1, j_rem = -1, p = k;
            if (has_add) {          // ith_nonparent: the a-th id not in S
                int c = 0;
                for (int b = 0; b < k; ++b) c += (S[b] - b < a_ord) ? 1 : 0;
                j_add = a_ord - 1 + c;
            }
            if (has_rem) { p = r_ord - 1; j_rem = S[p]; }
            int kn = 0;
            bool placed = !has_add;
            for (int b = 0; b < k; ++b) {           // the proposed set, sorted
                if (b == p) continue;
                if (!placed && j_add < S[b]) { T[kn++] = j_add; placed = true; }
                T[kn++] = S[b];
            }
            if (!placed) T[kn++] = j_add;
            ++s_prop; s_k += (unsigned)kn;
            const int mw = Q.deg > 0 ? poly_width(kn, Q.deg) : 0;
            if (mw > 0) {       // 3 x (n m (m + 5) + m^3/3 + 2 m^2 + 2 m) flops (deg > 1); the host divides by 3
                const unsigned long long m_ = (unsigned long long)mw;
                s_f3 += 3ull * (unsigned long long)P.n * m_ * (m_ + 5ull) + m_ * m_ * m_ + 6ull * m_ * m_ + 6ull * m_;
            }
            // a proposal beyond the path's bounds is rejected and counted (ssps.h: max_indegree), the sweep goes on
            const bool in_range = !(kn > SSPS_POLY_KMAX || kn > prm.kcap || mw < 0);
            if (!in_range) atomicAdd(prm.stats + ST_KCAP_REJECT, 1ull);
            double fwd, bwd;
            move_ratio(P, i, k, action, fwd, bwd);
            double dprior = 0.0;
            if (has_add) dprior = -lam * __ldg(omc_i + j_add);
            if (has_rem) dprior = dprior + lam * __ldg(omc_i + j_rem);
            const double lml_new = in_range ? poly_score_set(P, Q, i, T, kn, sc) : lml;
            const double w = dprior + (lml_new - lml);
            const double alpha = (action == ACT_SWAP) ? w : ((w - fwd) + bwd);
            const int rel_now = ceil_div_i(l, thin) - 1;
            if (in_range && log(d.u_acc) < alpha) {
                ++s_accp;
                if (pev) {
                    if (has_rem && n_pev < prm.pev_cap) pev[n_pev++] = (uint32_t)j_rem | ((uint32_t)(rel_now + 1) << 16);
                    if (has_add && n_pev < prm.pev_cap) pev[n_pev++] = (uint32_t)j_add | ((uint32_t)(rel_now + 1) << 16) | 0x80000000u;
                }
                for (int b = 0; b < kn; ++b) S[b] = T[b];
                k = kn; lml = lml_new;
            }
            // ---- thinned recording ----
            if (l % thin == 0) {
                for (int b = 0; b < k; ++b) cnt_row[S[b]] += 1u;
                lam_sum += lam; lam_sq = fma(lam, lam, lam_sq);
                if (prm.lev_val && lam != lam_logged) {
                    if (n_lev < prm.lev_cap) {
                        prm.lev_val[mc * prm.lev_cap + n_lev] = lam;
                        prm.lev_idx[mc * prm.lev_cap + n_lev] = (uint16_t)(l / thin);
                        ++n_lev;
                    }
                    lam_logged = lam;
                }
            }
        }
        for (int b = 0; b < k; ++b) slot_ref(st, V, i, b, chain) = (uint32_t)S[b] | ((uint32_t)b << 16);
        st.k[sidx] = k;
        st.lambda[sidx] = lam;
        st.lam_sum[mc] += lam_sum;
        st.lam_sq[mc] += lam_sq;
        if (prm.pev_n) prm.pev_n[mc] = n_pev;
        if (prm.lev_n) prm.lev_n[mc] = n_lev;
    }
    unsigned long long v[5] = {s_accp, s_accl, s_k, s_f3, s_prop};
    const int slot[5] = {ST_ACCP, ST_ACCL, ST_SUMK, ST_F3, ST_PROP};
#pragma unroll
    for (int e = 0; e < 5; ++e) {
        unsigned long long x = v[e];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0 && x) atomicAdd(prm.stats + slot[e], x);
    }
}
