// ssps_poly.cuh -- regression_deg >= 2: interaction columns (sm_100a).
//
// Reference semantics (paths relative to the gitter-lab/ssps checkout):
//   compute_B_col / construct_B   SSPS/src/dbn_distributions.jl:168-194
//       for deg = 1..regression_deg, for each Combinatorics.combinations(sorted S, deg) (lexicographic):
//       elementwise product of the RAW Xminus columns, then (x - mean) / std (corrected, ddof 1)
//   safe_solve / compute_lml      SSPS/src/dbn_distributions.jl:201-248
//       lml = -1/2 width log(1 + n) - 1/2 n log( y'y - n/(n+1) |P_B y|^2 ),  P_B = projection on span(B)
//       (Julia's `\` on a non-square B is pivoted QR with a rank cut: ANY width is legal, also width > n)
//   regression_deg <= 0 (the CLI default -1, "full") builds an EMPTY design in the reference (quirk Q2):
//       that case runs through the degree-1 cached-factor kernel with q = 0 (ssps_capi.cu), not through this file.
//   "full" of BASELINE.json's config C5 = regression_deg >= |S| (all non-empty product subsets, Hill's 'full').
//
// The product columns are re-standardised per set, so their Gram matrix cannot be tabulated once like G = Z'Z of
// the degree-1 path: every proposal is scored FROM SCRATCH.  Three classes of (k, deg, n), by design width
// m = sum_{d <= min(deg, k)} C(k, d):
//
//   SATURATED  m >= n - 1.  The columns are centred, so they span at most the (n - 1)-dimensional centred subspace,
//              which holds y: the projection is y itself, |P_B y|^2 = y'y, and
//                  lml = -1/2 m log(1 + n) - 1/2 n log( y'y / (n + 1) )
//              in closed form -- no design is built ("|S| >= n - 1 the projection saturates").  This is what
//              the reference's rank-cut QR returns for a design of full row rank; it removes every limit on k for
//              such sets (2^k - 1 >= n - 1 at "full" from k = log2 n on; the reference itself cannot allocate B there).
//   THREAD     m <= 7: one thread scores one set (poly_score_set): two passes over the n lagged rows (column
//              means; Gram matrix of the centred columns and their products with y), scaling to the standardised
//              design, and the same pivot-skipping Cholesky as the degree-1 path (column order = the reference's).
//   WARP       7 < m < n - 1 (k <= 64): the whole warp scores one set (poly_score_warp): a lane owns COLUMNS, every
//              Gram entry is accumulated by one lane in ascending row order and the Cholesky is right-looking with every
//              entry receiving its updates in ascending pivot order -- the operation order of the THREAD tier, so the
//              two tiers return the same bits and the tier choice (a function of k, deg, n alone) is not visible in
//              the results.  The n m^2 Gram of the product columns is the one place on this path where a single
//              proposal is large enough for 32 lanes (1e5..1e7 fma).
//
// Cost per scored set (deg > 1): n (3 m + m^2 + 2 m) + m^3/3 + 2 m^2 + 2 m flops; fp64 on the DFMA pipe.
#pragma once
#include "ssps_kernels.cuh"

#define SSPS_POLY_KMAX 7           // THREAD tier: parents per set
#ifndef SSPS_POLY_MMAX
#define SSPS_POLY_MMAX 7           // THREAD tier: design columns per set (measured on C5: 48 -> 1.17e6, 15 -> 1.49e6, 7 -> 1.56e6 proposals/s:
                                   // the per-thread Gram accumulators live in global memory, 3 memory operations per fma)
#endif
#define SSPS_POLY_KW 64            // WARP tier: parents per set (64-bit column masks)
#define SSPS_POLY_KLIST 128        // per-thread parent lists of the MH kernel (= the library's hard max_indegree ceiling)

enum { PC_NONE = 0, PC_EMPTY = 1, PC_SAT = 2, PC_THREAD = 3, PC_WARP = 4, PC_REFUSE = 5 };

struct PolyDev {
    const double* Xminus;          // n x V column-major, RAW
    const double* Y;               // n x V column-major, standardised
    int deg;                       // regression_deg (>= 2 on this path)
    int force_warp;                // test hook (SSPS_POLY_FORCE_WARP): sets of the THREAD class go through the WARP tier -- same bits
};

// design width sum_{d = 1..min(deg, k)} C(k, d), exact while <= limit; limit + 1 as soon as it exceeds the limit
__host__ __device__ inline long long poly_width_capped(int k, int deg, long long limit) {
    long long m = 0, c = 1;
    const int dmax = deg < k ? deg : k;
    for (int d = 1; d <= dmax; ++d) {
        c = c * (long long)(k - d + 1) / (long long)d;      // exact: c is C(k, d - 1) <= limit, times at most k
        if (c > limit) return limit + 1;
        m += c;
        if (m > limit) return limit + 1;
    }
    return m;
}
// the same width as a double (exact up to 2^53; the reference's Int64 sum overflows long before it could build such a B)
__host__ __device__ inline double poly_width_d(int k, int deg) {
    double m = 0.0, c = 1.0;
    const int dmax = deg < k ? deg : k;
    for (int d = 1; d <= dmax; ++d) { c = c * (double)(k - d + 1) / (double)d; m += c; }
    return m;
}
// class of a set of k parents; m_out = its width when it is scored numerically (THREAD / WARP)
__host__ __device__ inline int poly_class(int k, int deg, int n, int& m_out) {
    m_out = 0;
    if (deg <= 0 || k == 0) return PC_EMPTY;
    const long long lim = (long long)n - 2;
    const long long m = poly_width_capped(k, deg, lim < 0 ? 0 : lim);
    if (m > lim) return PC_SAT;
    m_out = (int)m;
    if (m <= SSPS_POLY_MMAX && k <= SSPS_POLY_KMAX) return PC_THREAD;
    return k <= SSPS_POLY_KW ? PC_WARP : PC_REFUSE;
}
__host__ __device__ inline int poly_class(int k, const PolyDev& Q, int n, int& m_out) {
    const int c = poly_class(k, Q.deg, n, m_out);
    return (c == PC_THREAD && Q.force_warp) ? PC_WARP : c;
}
// widest numerically scored design any set of at most kcap parents can have (sizes the WARP tier's scratch)
inline int poly_max_scored_width(int kcap, int deg, int n) {
    int best = 0;
    for (int k = 1; k <= kcap && k <= SSPS_POLY_KW; ++k) {
        int m = 0;
        const int c = poly_class(k, deg, n, m);
        if (c == PC_THREAD || c == PC_WARP) best = m > best ? m : best;
    }
    return best;
}

// closed forms
__device__ __forceinline__ double poly_lml_empty(const SspsDev& P, const int child) { return -0.5 * P.dn * log(__ldg(P.yty + child)); }
__device__ __forceinline__ double poly_lml_saturated(const SspsDev& P, const PolyDev& Q, const int child, const int k) {
    const double yty = __ldg(P.yty + child);
    const double ip = yty - P.shrink * yty;             // |P_B y|^2 = y'y
    return -0.5 * poly_width_d(k, Q.deg) * P.log1pn - 0.5 * P.dn * log(ip);
}

// ------------------------------------------------------------------------------------------------
// THREAD tier
// ------------------------------------------------------------------------------------------------
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

// bit masks (over positions of the sorted set) of the design columns, in the reference's order:
// for d = 1..min(deg, k), Combinatorics.combinations(S, d) in lexicographic order
template <class Mask, class Emit>
__device__ inline int poly_enumerate(int k, int deg, Emit emit) {
    int m = 0;
    const int dmax = deg < k ? deg : k;
    int idx[SSPS_POLY_KW];
    for (int d = 1; d <= dmax; ++d) {
        for (int a = 0; a < d; ++a) idx[a] = a;
        for (;;) {
            Mask mk = 0;
            for (int a = 0; a < d; ++a) mk |= (Mask)1 << idx[a];
            emit(m++, mk);
            int i = d - 1;
            while (i >= 0 && idx[i] == k - d + i) --i;
            if (i < 0) break;
            ++idx[i];
            for (int a = i + 1; a < d; ++a) idx[a] = idx[a - 1] + 1;
        }
    }
    return m;
}

// lml(child, S) for the SORTED set S[0..k) of class PC_THREAD with m columns
__device__ inline double poly_score_set(const SspsDev& P, const PolyDev& Q, const int child, const uint16_t* S, const int k, const int m,
                                        const PolyScratch& sc) {
    const int n = P.n;
    const double yty = P.yty[child];
    unsigned short masks[SSPS_POLY_MMAX];
    {
        int idx[SSPS_POLY_KMAX];
        int c = 0;
        const int dmax = Q.deg < k ? Q.deg : k;
        for (int d = 1; d <= dmax; ++d) {
            for (int a = 0; a < d; ++a) idx[a] = a;
            for (;;) {
                unsigned mk = 0;
                for (int a = 0; a < d; ++a) mk |= 1u << idx[a];
                masks[c++] = (unsigned short)mk;
                int i = d - 1;
                while (i >= 0 && idx[i] == k - d + i) --i;
                if (i < 0) break;
                ++idx[i];
                for (int a = i + 1; a < d; ++a) idx[a] = idx[a - 1] + 1;
            }
        }
    }
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

// ------------------------------------------------------------------------------------------------
// WARP tier.  Scratch of one warp (global memory; the working set of one set is n mp + m mp doubles, L1/L2-resident
// for the widths that occur in a stationary chain): cv[t][c] = the centred product columns (row-major, mp = m rounded
// up to 32 columns, padding columns zero), g[a][b] = the Gram matrix / the factor in place (b <= a), and a few vectors.
// ------------------------------------------------------------------------------------------------
struct PolyWarpScratch {
    double* base;           // this warp's block
    int n, mcap_p;          // rows, column capacity (multiple of 32)
    __device__ __forceinline__ double* cv() const { return base; }                                         // n * mcap_p
    __device__ __forceinline__ double* g() const { return base + (size_t)n * mcap_p; }                    // mcap_p * mcap_p
    __device__ __forceinline__ double* vec(int e) const { return g() + (size_t)mcap_p * mcap_p + (size_t)e * mcap_p; }   // e < 4: isd, rhs, gd, colp
    __device__ __forceinline__ unsigned long long* masks() const { return (unsigned long long*)vec(4); }   // mcap_p
    __device__ __forceinline__ int* ids() const { return (int*)vec(5); }                                    // SSPS_POLY_KW
};
__host__ __device__ inline size_t poly_warp_scratch_doubles(int n, int mcap_p) { return (size_t)n * mcap_p + (size_t)mcap_p * mcap_p + 5 * (size_t)mcap_p + SSPS_POLY_KW; }

// 1 / sqrt(x), IEEE sqrt then IEEE division: the THREAD tier's bits
__device__ __noinline__ double poly_inv_sqrt(const double x) { return 1.0 / sqrt(x); }

// lml(child, S) for the sorted set W.ids()[0..k) of class PC_WARP with m columns.  All arguments warp-uniform, every lane
// calls.  Same operation order, same bits as poly_score_set.
__device__ __noinline__ double poly_score_warp(const SspsDev& P, const PolyDev& Q, const int child, const int k, const int m,
                                               const PolyWarpScratch& W, const int lane) {
    const int n = P.n;
    const int mp = (m + 31) & ~31, nj = mp >> 5;
    double* cv = W.cv();
    double* g = W.g();
    double* isd = W.vec(0);
    double* rhs = W.vec(1);
    double* gd = W.vec(2);
    double* colp = W.vec(3);
    unsigned long long* masks = W.masks();     // (no __restrict__ on the scratch: lanes hand data to each other through it)
    const int* ids = W.ids();
    const double* __restrict__ y = Q.Y + (size_t)child * n;

    if (lane == 0) poly_enumerate<unsigned long long>(k, Q.deg, [&](int c, unsigned long long mk) { masks[c] = mk; });
    __syncwarp();
    // pass 1: products and column means (a lane owns columns lane + 32 j; ascending t)
    for (int j = 0; j < nj; ++j) {
        const int c = lane + 32 * j;
        const unsigned long long mk = c < m ? masks[c] : 0ull;
        double mean = 0.0;
        for (int t = 0; t < n; ++t) {
            double pr = 0.0;
            if (c < m) {
                pr = 1.0;
                unsigned long long mm = mk;
                while (mm) { const int a = __ffsll((long long)mm) - 1; mm &= mm - 1; pr *= __ldg(Q.Xminus + (size_t)ids[a] * n + t); }
                mean += pr;
            }
            cv[(size_t)t * mp + c] = pr;
        }
        mean = mean / (double)n;
        // pass 2: centre; products with y
        double r = 0.0;
        if (c < m) {
            for (int t = 0; t < n; ++t) {
                const double v = cv[(size_t)t * mp + c] - mean;
                cv[(size_t)t * mp + c] = v;
                r = fma(v, __ldg(y + t), r);
            }
        }
        rhs[c] = r;
    }
    __syncwarp();
    // Gram matrix of the centred columns.  Tiles of 16 rows x 32 columns: every entry is one lane's ascending-t fma chain
    // (the THREAD tier's bits); per t a lane loads ONE value of its own column and the tile's 16 row values come as eight
    // warp-uniform 16-byte loads, so the centred columns are streamed from L2 once per 16 rows of the matrix (the first
    // version, 4 x 128 tiles, read them once per 4 rows: 33 MB per 255-column set, the bound of the whole scorer).
    for (int c0 = 0; c0 < m; c0 += 16) {
        const int jend = (min(c0 + 15, m - 1) >> 5) + 1;            // column chunks that hold some b <= c
        for (int jb = 0; jb < jend; ++jb) {
            double acc[16];
#pragma unroll
            for (int r = 0; r < 16; ++r) acc[r] = 0.0;
            const double* row = cv;
            const int b = lane + 32 * jb;
            // eight rows of t in flight: the scratch of ~1 000 resident warps (1.3 MB each at C5) lives in HBM, and two rows
            // in flight per warp left the loop waiting on DRAM latency (measured: 40e6 cycles per 255-column set)
            int t = 0;
            for (; t + 8 <= n; t += 8, row += 8 * (size_t)mp) {
                double col[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) col[u] = row[(size_t)u * mp + b];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const double2* bc = reinterpret_cast<const double2*>(row + (size_t)u * mp + c0);     // c0 + 15 < mp: padding columns are zero
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const double2 v = bc[r];
                        acc[2 * r] = fma(v.x, col[u], acc[2 * r]);
                        acc[2 * r + 1] = fma(v.y, col[u], acc[2 * r + 1]);
                    }
                }
            }
            for (; t < n; ++t, row += mp) {
                const double col = row[b];
                const double2* bc = reinterpret_cast<const double2*>(row + c0);
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const double2 v = bc[r];
                    acc[2 * r] = fma(v.x, col, acc[2 * r]);
                    acc[2 * r + 1] = fma(v.y, col, acc[2 * r + 1]);
                }
            }
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int c = c0 + r;
                if (c < m && b <= c) g[(size_t)c * mp + b] = acc[r];
            }
        }
    }
    __syncwarp();
    // standardise: B_c = centred column / sd_c, sd_c = sqrt(C_cc / (n - 1))
    for (int c = lane; c < m; c += 32) isd[c] = 1.0 / sqrt(g[(size_t)c * mp + c] / (double)(n - 1));
    __syncwarp();
    for (int c = lane; c < m; c += 32) rhs[c] = rhs[c] * isd[c];
    for (int c = 0; c < m; ++c) {
        const double ic = isd[c];
        for (int b = lane; b <= c; b += 32) g[(size_t)c * mp + b] = g[(size_t)c * mp + b] * ic * isd[b];
    }
    __syncwarp();
    for (int c = lane; c < m; c += 32) gd[c] = g[(size_t)c * mp + c];
    __syncwarp();
    // Pivot-skipping Cholesky in block columns of 32.  Entry (a, b), b <= a, receives  - l_ap l_bp  for p = 0, 1, .. b - 1 in
    // ascending order through fma (the diagonal and the rhs entry of a row are entries of that row: d2_a -= l_ap^2,
    // t_a -= l_ap w_p) -- the THREAD tier's order.  For the block column [b0, b0 + 32):
    //   (1) the pivots p < b0, LEFT-looking: a lane owns column b0 + lane, sixteen rows per pass in registers, the scaled
    //       columns l_.p read from the transposed copy kept in the otherwise unused upper triangle of g (row p: contiguous);
    //   (2) the block's own 32 pivots, right-looking inside the block column ((m - b0) x 32 entries: L1-resident).
    // The first version was right-looking over the whole trailing matrix: m^3/3 read-modify-writes through L2 (44 MB for a
    // 255-column set); this one reads the finished columns once per 16 rows (3 MB).
    double q = 0.0;
    double* wv = colp;                                  // w of the finished pivots (0 for a skipped one)
    for (int b0 = 0; b0 < m; b0 += 32) {
        const int b = b0 + lane;
        const bool bcol = b < m;
        const int bend = min(b0 + 32, m);
        if (b0 > 0) {
            for (int a0 = b0; a0 < m; a0 += 16) {
                double acc[16];
#pragma unroll
                for (int r = 0; r < 16; ++r) { const int a = a0 + r; acc[r] = (bcol && a < m && b <= a) ? g[(size_t)a * mp + b] : 0.0; }
#pragma unroll 2
                for (int p = 0; p < b0; ++p) {
                    const double* Lp = g + (size_t)p * mp;
                    const double nlb = bcol ? -Lp[b] : 0.0;
                    const double2* la = reinterpret_cast<const double2*>(Lp + a0);          // a0 is a multiple of 16
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const double2 v = la[r];
                        acc[2 * r] = fma(v.x, nlb, acc[2 * r]);
                        acc[2 * r + 1] = fma(v.y, nlb, acc[2 * r + 1]);
                    }
                }
#pragma unroll
                for (int r = 0; r < 16; ++r) { const int a = a0 + r; if (bcol && a < m && b <= a) g[(size_t)a * mp + b] = acc[r]; }
            }
            if (bcol) {
                double t = rhs[b];
                for (int p = 0; p < b0; ++p) t = fma(-g[(size_t)p * mp + b], wv[p], t);
                rhs[b] = t;
            }
            __syncwarp();
        }
        for (int p = b0; p < bend; ++p) {
            const double d2 = g[(size_t)p * mp + p], tt = rhs[p], gaa = gd[p];
            double inv = 0.0, w = 0.0;
            bool used = true;
            if (d2 > SSPS_PIVOT_TOL * gaa) { inv = poly_inv_sqrt(d2); w = tt * inv; q = fma(w, w, q); }
            else if (d2 != d2 || gaa != gaa) { inv = __longlong_as_double(0x7ff8000000000000LL); w = inv; q = inv; }
            else used = false;                          // numerically dependent column: l_ap = 0 for every a
            if (lane == 0) wv[p] = w;
            for (int a = p + 1 + lane; a < m; a += 32) g[(size_t)p * mp + a] = used ? g[(size_t)a * mp + p] * inv : 0.0;
            __syncwarp();
            if (used && bcol && b > p) {
                const double* Lp = g + (size_t)p * mp;
                const double nlb = -Lp[b];
                rhs[b] = fma(nlb, w, rhs[b]);
                int a = b;
                for (; a + 3 < m; a += 4) {
                    double* e = g + (size_t)a * mp + b;
                    const double x0 = fma(Lp[a], nlb, e[0]), x1 = fma(Lp[a + 1], nlb, e[mp]);
                    const double x2 = fma(Lp[a + 2], nlb, e[2 * (size_t)mp]), x3 = fma(Lp[a + 3], nlb, e[3 * (size_t)mp]);
                    e[0] = x0; e[mp] = x1; e[2 * (size_t)mp] = x2; e[3 * (size_t)mp] = x3;
                }
                for (; a < m; ++a) g[(size_t)a * mp + b] = fma(Lp[a], nlb, g[(size_t)a * mp + b]);
            }
            __syncwarp();
        }
    }
    const double ip = __ldg(P.yty + child) - P.shrink * q;
    return -0.5 * (double)m * P.log1pn - 0.5 * P.dn * log(ip);
}

// lml of this lane's sorted set S[0..k) whatever its class; warp-collective (every lane calls; dead lanes pass live = false).
// A PC_REFUSE set (more than SSPS_POLY_KW parents whose design is still narrower than n - 1 rows) returns NaN: the callers refuse
// such sets before they get here.
__device__ inline double poly_score_any(const SspsDev& P, const PolyDev& Q, const int child, const uint16_t* S, const int k, const bool live,
                                        const PolyScratch& sc, const PolyWarpScratch& W, const int lane, unsigned long long* prof = nullptr) {
    int m = 0;
    const int cls = live ? poly_class(k, Q, P.n, m) : PC_NONE;
    double v = __longlong_as_double(0x7ff8000000000000LL);
#ifdef SSPS_PROFILE_BIG
    const long long tq0_ = clock64();
#endif
    if (cls == PC_EMPTY) v = poly_lml_empty(P, child);
    else if (cls == PC_SAT) v = poly_lml_saturated(P, Q, child, k);
    else if (cls == PC_THREAD) v = poly_score_set(P, Q, child, S, k, m, sc);
    unsigned wm = __ballot_sync(0xffffffffu, cls == PC_WARP);
#ifdef SSPS_PROFILE_BIG
    const long long tq1_ = clock64();
    const unsigned wm0_ = wm;
#endif
    while (wm) {
        const int src = __ffs(wm) - 1;
        wm &= wm - 1;
        const int ks = __shfl_sync(0xffffffffu, k, src), ms = __shfl_sync(0xffffffffu, m, src), cs = __shfl_sync(0xffffffffu, child, src);
        if (lane == src) { int* ids = W.ids(); for (int a = 0; a < k; ++a) ids[a] = (int)S[a]; }
        __syncwarp();
        const double r = poly_score_warp(P, Q, cs, ks, ms, W, lane);
        if (lane == src) v = r;
        __syncwarp();
    }
#ifdef SSPS_PROFILE_BIG
    if (prof && lane == 0) {        // warp cycles in the closed-form / THREAD part (its slowest lane) and in the WARP tier; sets scored there
        atomicAdd(prof + 42, (unsigned long long)(tq1_ - tq0_)); atomicAdd(prof + 43, (unsigned long long)(clock64() - tq1_));
        atomicAdd(prof + 44, (unsigned long long)__popc(wm0_)); atomicAdd(prof + 45, 1ull);
    }
#endif
    return v;
}

// K2 for the interaction path: one thread per (child, S), grid-stride over the sets (the warp walks 32 sets at a time)
__global__ void __launch_bounds__(128) poly_score_kernel(SspsDev P, PolyDev Q, int n_sets, const int* __restrict__ child,
                                                         const int* __restrict__ offsets, const int* __restrict__ members,
                                                         const double* __restrict__ lambda, double* __restrict__ arena,
                                                         double* __restrict__ warp_arena, int mcap_p, double* __restrict__ lml_out,
                                                         double* __restrict__ lp_out) {
    const int lane = threadIdx.x & 31;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    const PolyScratch sc{arena + gtid, nthreads};
    const PolyWarpScratch W{warp_arena + (size_t)(gtid >> 5) * poly_warp_scratch_doubles(P.n, mcap_p), P.n, mcap_p};
    for (int s0 = gtid - lane; s0 < n_sets; s0 += nthreads) {
        const int s = s0 + lane;
        const bool live = s < n_sets;
        const int i = live ? child[s] : 0, k = live ? offsets[s + 1] - offsets[s] : 0;
        const int* mem = members + (live ? offsets[s] : 0);
        uint16_t S[SSPS_POLY_KLIST];
        int m = 0;
        const bool scored = live && k <= SSPS_POLY_KLIST && poly_class(k, Q, P.n, m) != PC_REFUSE;
        if (scored) for (int a = 0; a < k; ++a) S[a] = (uint16_t)mem[a];
        double v = poly_score_any(P, Q, i, S, scored ? k : 0, scored, sc, W, lane);
        if (live && !scored) {      // saturated sets of any size are closed form; anything else this large is refused by the host
            v = poly_class(k, Q, P.n, m) == PC_SAT ? poly_lml_saturated(P, Q, i, k) : __longlong_as_double(0x7ff8000000000000LL);
        }
        if (live && lml_out) lml_out[s] = v;
        if (live && lp_out) {
            const double lam = lambda[s];
            double somc = 0.0;
            for (int a = 0; a < k; ++a) somc += P.omc[(size_t)mem[a] + (size_t)P.V * i];
            lp_out[s] = -lam * somc - prior_normaliser(P, i, lam);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K3 for the interaction path: the same sweep (lambda move, then add / swap / remove on the sorted
// parent list, Gen.mh) as mh_sweeps_kernel, one thread per micro-chain, every proposal scored from
// scratch (closed form / THREAD tier by the lane, WARP tier by the whole warp one lane at a time), the
// acceptance test in the reference order with fp64 logarithms.
// Chain state uses the degree-1 layout (slots = id | rank << 16); this kernel keeps the list sorted,
// so rank = position.  Tallies are incremented at every recorded sample (no lazy credit here).
// A proposal beyond max_indegree, or one of class PC_REFUSE, is rejected and counted (ssps.h: max_indegree).
// Included after ssps_mh.cuh (StepDraws, sources, select_move, move_ratio, MhParams).
// ------------------------------------------------------------------------------------------------
template <class Source>
__global__ void __launch_bounds__(128) mh_poly_kernel(const __grid_constant__ SspsDev P, const __grid_constant__ PolyDev Q,
                                                      const __grid_constant__ ChainState st, const __grid_constant__ MhParams prm,
                                                      const __grid_constant__ Source src, double* __restrict__ arena,
                                                      double* __restrict__ warp_arena, const int mcap_p) {
    const int lane = threadIdx.x & 31;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const PolyScratch sc{arena + gtid, (int)(gridDim.x * blockDim.x)};
    const PolyWarpScratch W{warp_arena + (size_t)(gtid >> 5) * poly_warp_scratch_doubles(P.n, mcap_p), P.n, mcap_p};
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
        const bool live = chain < st.n_chains;
        const size_t sidx = (size_t)i * st.nch_pad + (live ? chain : 0);
        const size_t mc = (size_t)(live ? chain : 0) * V + i;
        uint32_t* cnt_row = st.counts + mc * V;
        uint32_t* pev = prm.pev ? prm.pev + mc * prm.pev_cap : nullptr;
        int k = live ? st.k[sidx] : 0;
        double lam = live ? st.lambda[sidx] : 0.0;
        uint16_t S[SSPS_POLY_KLIST], T[SSPS_POLY_KLIST];
        if (live) for (int a = 0; a < k; ++a) { const uint32_t sl = slot_ref(st, V, i, a, chain); S[(int)(sl >> 16)] = (uint16_t)(sl & 0xffffu); }
        const double* __restrict__ omc_i = P.omc + (size_t)V * i;
        double lml = poly_score_any(P, Q, i, S, k, live, sc, W, lane);
        double lam_sum = 0.0, lam_sq = 0.0, lam_logged = lam;
        int n_pev = 0, n_lev = 0;
        DrawCarry cy; cy.valid = 0; cy.z = 0.f; cy.w = 0u;
        for (int l = 1; l <= prm.n_local; ++l) {
            StepDraws d;
            d.z = 0.0; d.u_lam = 0.5; d.u_act = 0.5; d.u_add = 0.5; d.u_rem = 0.5; d.u_acc = 0.5;
            int action = ACT_ADD, kn = 0, j_add = -1, j_rem = -1;
            bool in_range = false;
            if (live) {
                d = src.get(chain, i, (uint64_t)(prm.sweep0 + l), (int64_t)(l - 1), cy);
                // ---- lambda move (lambda_vec_proposal + Gen.mh), reference order ----
                const double lam_new = lam + P.sigma * d.z;
                if (lam_new >= P.lambda_min && lam_new <= P.lambda_max) {
                    double somc = 0.0;
                    for (int a = 0; a < k; ++a) somc += __ldg(omc_i + S[a]);
                    const double nrm = prior_normaliser(P, i, lam), nrm_new = prior_normaliser(P, i, lam_new);
                    const double w = -(lam_new - lam) * somc - (nrm_new - nrm);
                    if (log(d.u_lam) < w) { lam = lam_new; ++s_accl; }
                }
                // ---- parent-set move: the proposed set, sorted ----
                int a_ord, r_ord;
                select_move(P, i, k, d, action, a_ord, r_ord);
                const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;
                int p = k;
                if (has_add) {          // ith_nonparent: the a-th id not in S
                    int c = 0;
                    for (int b = 0; b < k; ++b) c += ((int)S[b] - b < a_ord) ? 1 : 0;
                    j_add = a_ord - 1 + c;
                }
                if (has_rem) { p = r_ord - 1; j_rem = (int)S[p]; }
                kn = k + (has_add ? 1 : 0) - (has_rem ? 1 : 0);
                in_range = kn <= prm.kcap && kn <= SSPS_POLY_KLIST;
                if (in_range) {
                    int t = 0;
                    bool placed = !has_add;
                    for (int b = 0; b < k; ++b) {
                        if (b == p) continue;
                        if (!placed && j_add < (int)S[b]) { T[t++] = (uint16_t)j_add; placed = true; }
                        T[t++] = S[b];
                    }
                    if (!placed) T[t++] = (uint16_t)j_add;
                }
                ++s_prop; s_k += (unsigned)kn;
                int mw = 0;
                const int cls = poly_class(kn, Q, P.n, mw);
                if (cls == PC_REFUSE) in_range = false;
                if (in_range && mw > 0) {   // 3 x (n m (m + 5) + m^3/3 + 2 m^2 + 2 m) flops (deg > 1); the host divides by 3
                    const unsigned long long m_ = (unsigned long long)mw;
                    s_f3 += 3ull * (unsigned long long)P.n * m_ * (m_ + 5ull) + m_ * m_ * m_ + 6ull * m_ * m_ + 6ull * m_;
                }
                if (!in_range) atomicAdd(prm.stats + ST_KCAP_REJECT, 1ull);
            }
            const double lml_new = poly_score_any(P, Q, i, T, kn, live && in_range, sc, W, lane, prm.stats);
            if (live) {
                const bool has_add = action != ACT_REM, has_rem = action != ACT_ADD;
                double fwd, bwd;
                move_ratio(P, i, k, action, fwd, bwd);
                double dprior = 0.0;
                if (has_add) dprior = -lam * __ldg(omc_i + j_add);
                if (has_rem) dprior = dprior + lam * __ldg(omc_i + j_rem);
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
        }
        if (live) {
            for (int b = 0; b < k; ++b) slot_ref(st, V, i, b, chain) = (uint32_t)S[b] | ((uint32_t)b << 16);
            st.k[sidx] = k;
            st.lambda[sidx] = lam;
            st.lam_sum[mc] += lam_sum;
            st.lam_sq[mc] += lam_sq;
            if (prm.pev_n) prm.pev_n[mc] = n_pev;
            if (prm.lev_n) prm.lev_n[mc] = n_lev;
        }
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
