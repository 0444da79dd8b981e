/*
 * ssps_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).  See ssps_oracle.h.
 *
 * PARITY UNPINNED (no Julia in this image; the reference's one test ignores scores).
 *
 * Two scorers live here on purpose:
 *   - orc_lml_qr   : reference-faithful.  Builds B(S) from the RAW Xminus exactly like
 *                    construct_B / compute_B_col (dbn_distributions.jl:168-194) and solves the
 *                    least-squares problem with a column-pivoted Householder QR, which is what
 *                    Julia 1.6's `B \ Xp` does for a non-square B (dbn_distributions.jl:203).
 *   - orc_lml_chol : kernel-faithful.  Same value computed from the lagged Gram matrix
 *                    G = Z'Z, R = Z'Y by a pivot-skipping Cholesky in the SAME operation order
 *                    (explicit fma, sorted S) as ssps_b200/csrc -- used for the bit-exact
 *                    replay test (level-2 parity).  Build with -ffp-contract=off.
 *
 * Random stream convention (the reference never seeds Julia's MersenneTwister, so any stream
 * is a convention): per (chain, vertex, sweep) six numbers (z, u_lam, u_act, u_add, u_rem,
 * u_acc); categorical by inverse CDF with `draw < cumsum` (Distributions.jl
 * DiscreteNonParametric sampler), uniform_discrete(1,m) = 1 + floor(u*m), normal = mu + sd*z.
 * Default source: Philox4x32-10, key = 64-bit seed, counter = (block | vertex<<8, sweep_lo,
 * sweep_hi, chain).
 */
#define _GNU_SOURCE
#include "ssps_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define ORC_PIVOT_TOL 9.094947017729282e-13 /* 2^-40: skip column when d^2 <= tol * G_jj */

/* ------------------------------------------------------------------------------------------
 * Pre-processing: standardize_X (dbn_preprocess.jl:90-98), corrected std (ddof = 1).
 * Column-major n x V.  Operation order is the contract the CUDA Gram builder follows.
 * ---------------------------------------------------------------------------------------- */
void orc_standardize(const double* X, int n, int V, double* Z) {
    for (int j = 0; j < V; ++j) {
        const double* x = X + (size_t)j * n;
        double* z = Z + (size_t)j * n;
        double s = 0.0;
        for (int t = 0; t < n; ++t) s += x[t];
        double mean = s / (double)n;
        double ss = 0.0;
        for (int t = 0; t < n; ++t) { double d = x[t] - mean; ss = fma(d, d, ss); }
        double sd = sqrt(ss / (double)(n - 1));
        for (int t = 0; t < n; ++t) z[t] = (x[t] - mean) / sd;
    }
}

/* G = Z'Z (V x V), R = Z'Y (V x V, column i = Z' y_i), yty_i = y_i'y_i. */
void orc_gram(const double* Z, const double* Y, int n, int V, double* G, double* R, double* yty) {
    for (int b = 0; b < V; ++b)
        for (int a = 0; a < V; ++a) {
            const double* za = Z + (size_t)a * n; const double* zb = Z + (size_t)b * n;
            double acc = 0.0;
            for (int t = 0; t < n; ++t) acc = fma(za[t], zb[t], acc);
            G[(size_t)a + (size_t)V * b] = acc;
        }
    for (int i = 0; i < V; ++i) {
        const double* yi = Y + (size_t)i * n;
        for (int j = 0; j < V; ++j) {
            const double* zj = Z + (size_t)j * n;
            double acc = 0.0;
            for (int t = 0; t < n; ++t) acc = fma(zj[t], yi[t], acc);
            R[(size_t)j + (size_t)V * i] = acc;
        }
        double acc = 0.0;
        for (int t = 0; t < n; ++t) acc = fma(yi[t], yi[t], acc);
        yty[i] = acc;
    }
}

/* ------------------------------------------------------------------------------------------
 * construct_B / compute_B_col (dbn_distributions.jl:168-194)
 * ---------------------------------------------------------------------------------------- */
static double binom(int n, int k) {
    if (k < 0 || k > n) return 0.0;
    double r = 1.0;
    for (int i = 1; i <= k; ++i) r = r * (double)(n - k + i) / (double)i;
    return floor(r + 0.5);
}

int orc_n_design_cols(int k, int deg) {
    /* n_cols = sum(binomial(k, d) for d = 1:min(deg, k)); deg <= 0 => empty sum (quirk Q2) */
    int m = deg < k ? deg : k;
    double s = 0.0;
    for (int d = 1; d <= m; ++d) s += binom(k, d);
    return (int)s;
}

static void b_col(const double* Xminus, int n, const int* inds, int d, double* col) {
    /* prod(Xminus[:,inds], dims=2) then (col - mean) / std (ddof = 1) */
    for (int t = 0; t < n; ++t) {
        double p = Xminus[(size_t)inds[0] * n + t];
        for (int a = 1; a < d; ++a) p *= Xminus[(size_t)inds[a] * n + t];
        col[t] = p;
    }
    double s = 0.0;
    for (int t = 0; t < n; ++t) s += col[t];
    double mean = s / (double)n;
    double ss = 0.0;
    for (int t = 0; t < n; ++t) { double e = col[t] - mean; ss = fma(e, e, ss); }
    double sd = sqrt(ss / (double)(n - 1));
    for (int t = 0; t < n; ++t) col[t] = (col[t] - mean) / sd;
}

void orc_construct_B(const double* Xminus, int n, const int* S, int k, int deg, double* B) {
    int col = 0;
    int comb[64];
    int inds[64];
    for (int d = 1; d <= deg && d <= k && d <= 64; ++d) {
        /* Combinatorics.combinations: lexicographic over positions of the (sorted) S */
        for (int a = 0; a < d; ++a) comb[a] = a;
        for (;;) {
            for (int a = 0; a < d; ++a) inds[a] = S[comb[a]];
            b_col(Xminus, n, inds, d, B + (size_t)col * n);
            ++col;
            int a = d - 1;
            while (a >= 0 && comb[a] == k - d + a) --a;
            if (a < 0) break;
            ++comb[a];
            for (int b = a + 1; b < d; ++b) comb[b] = comb[b - 1] + 1;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * safe_solve (dbn_distributions.jl:201-229): Julia 1.6 `B \ y`, n != width => qr(B, Val(true)):
 * LAPACK geqp3 + numerical-rank cut at rcond = min(n,width)*eps + minimum-norm solution; the
 * fitted values B*beta are then the orthogonal projection of y onto the span of the leading
 * `rank` pivoted columns.  Returns ||P_B y||^2 (so that ip = y'y - n/(n+1) * result).
 * n == width goes through LU in Julia and is rounding-dominated there; the oracle
 * uses the projection for every shape.  B is overwritten.
 * ---------------------------------------------------------------------------------------- */
double orc_qr_project(double* B, int n, int width, const double* y, int* rank_out) {
    if (rank_out) *rank_out = 0;
    if (width == 0) return 0.0;
    for (size_t e = 0; e < (size_t)n * width; ++e)
        if (isnan(B[e])) return NAN;
    double* qty = (double*)malloc(sizeof(double) * (size_t)n);
    memcpy(qty, y, sizeof(double) * (size_t)n);
    int r = n < width ? n : width;
    double rcond = (double)r * 2.220446049250313e-16;
    double r00 = 0.0;
    int rank = 0;
    for (int j = 0; j < r; ++j) {
        /* pivot: largest remaining partial column norm (recomputed, not downdated) */
        int piv = j; double best = -1.0;
        for (int c = j; c < width; ++c) {
            double s = 0.0;
            for (int t = j; t < n; ++t) { double v = B[(size_t)c * n + t]; s = fma(v, v, s); }
            if (s > best) { best = s; piv = c; }
        }
        if (piv != j)
            for (int t = 0; t < n; ++t) {
                double tmp = B[(size_t)j * n + t]; B[(size_t)j * n + t] = B[(size_t)piv * n + t];
                B[(size_t)piv * n + t] = tmp;
            }
        double* x = B + (size_t)j * n;
        double norm = sqrt(best);
        if (j == 0) r00 = norm;
        if (norm <= rcond * r00 || norm == 0.0) break;
        double alpha = x[j] > 0.0 ? -norm : norm;
        double v0 = x[j] - alpha;
        /* v = (v0, x[j+1..]) ; H = I - 2 v v' / (v'v) */
        double vtv = v0 * v0;
        for (int t = j + 1; t < n; ++t) vtv = fma(x[t], x[t], vtv);
        if (vtv > 0.0) {
            for (int c = j + 1; c < width; ++c) {
                double* cc = B + (size_t)c * n;
                double dot = v0 * cc[j];
                for (int t = j + 1; t < n; ++t) dot = fma(x[t], cc[t], dot);
                double f = 2.0 * dot / vtv;
                cc[j] -= f * v0;
                for (int t = j + 1; t < n; ++t) cc[t] -= f * x[t];
            }
            double dot = v0 * qty[j];
            for (int t = j + 1; t < n; ++t) dot = fma(x[t], qty[t], dot);
            double f = 2.0 * dot / vtv;
            qty[j] -= f * v0;
            for (int t = j + 1; t < n; ++t) qty[t] -= f * x[t];
        }
        x[j] = alpha;
        rank = j + 1;
    }
    double q = 0.0;
    for (int t = 0; t < rank; ++t) q = fma(qty[t], qty[t], q);
    free(qty);
    if (rank_out) *rank_out = rank;
    return q;
}

/* compute_lml (dbn_distributions.jl:235-248):
 *   ip = dot(Xp, Xp - n/(n+1) * B * (B \ Xp));  lml = -0.5*Bwidth*log(1+n) - 0.5*n*log(ip) */
double orc_lml_qr(const double* Xminus, const double* y, int n, const int* S, int k, int deg,
                  int* rank_out) {
    int width = orc_n_design_cols(k, deg);
    double q = 0.0;
    if (width > 0) {
        double* B = (double*)malloc(sizeof(double) * (size_t)n * width);
        orc_construct_B(Xminus, n, S, k, deg, B);
        q = orc_qr_project(B, n, width, y, rank_out);
        free(B);
    } else if (rank_out) *rank_out = 0;
    double yty = 0.0;
    for (int t = 0; t < n; ++t) yty = fma(y[t], y[t], yty);
    double dn = (double)n;
    double ip = yty - (dn / (dn + 1.0)) * q;
    return -0.5 * (double)width * log(1.0 + dn) - 0.5 * dn * log(ip);
}

/* ------------------------------------------------------------------------------------------
 * Kernel-faithful scorer (deg 1 only): B = Z[:,S]  =>  q = r_S' G_SS^-1 r_S, computed by a
 * row-by-row Cholesky over S IN THE ORDER GIVEN (the chain passes the arrival order, ssps_score
 * callers the sorted list) with column skipping: a column whose Schur pivot
 * d^2 <= 2^-40 * G_jj is numerically dependent on the earlier ones and contributes nothing
 * (its row of L and its w are zeroed), so q stays the squared norm of the projection -- the
 * rank-deficient semantics of the reference's pivoted QR ("hard parts").
 * The penalty uses the column COUNT k (dbn_distributions.jl:241,246).
 * Operation order here is the contract for the CUDA scorer (csrc/ssps_score.cuh).
 * ---------------------------------------------------------------------------------------- */
double orc_lml_chol(const double* G, const double* Ri, double yty, int n, int V,
                    const int* S, int k, double* min_pivot_ratio, int* n_skipped) {
    double* L = (double*)malloc(sizeof(double) * ((size_t)k * (k + 1) / 2 + 1));
    double* w = (double*)malloc(sizeof(double) * (size_t)(k + 1));
    double* invd = (double*)malloc(sizeof(double) * (size_t)(k + 1));
    double q = 0.0, minr = INFINITY;
    int skipped = 0;
    for (int a = 0; a < k; ++a) {
        int ja = S[a];
        double* La = L + (size_t)a * (a + 1) / 2;
        for (int b = 0; b < a; ++b) {
            const double* Lb = L + (size_t)b * (b + 1) / 2;
            double s = G[(size_t)ja + (size_t)V * S[b]];
            for (int m = 0; m < b; ++m) s = fma(-La[m], Lb[m], s);
            La[b] = s * invd[b];
        }
        double gjj = G[(size_t)ja + (size_t)V * ja];
        double d2 = gjj;
        for (int m = 0; m < a; ++m) d2 = fma(-La[m], La[m], d2);
        double t = Ri[ja];
        for (int m = 0; m < a; ++m) t = fma(-La[m], w[m], t);
        if (d2 > ORC_PIVOT_TOL * gjj) {
            double ratio = d2 / gjj;
            if (ratio < minr) minr = ratio;
            invd[a] = 1.0 / sqrt(d2);
            w[a] = t * invd[a];
            q = fma(w[a], w[a], q);
        } else if (d2 != d2 || gjj != gjj) { /* NaN data (zero-variance column, quirk Q4) */
            invd[a] = NAN; w[a] = NAN; q = NAN;
        } else {
            ++skipped;
            invd[a] = 0.0; w[a] = 0.0;
            for (int m = 0; m < a; ++m) La[m] = 0.0;
        }
    }
    free(L); free(w); free(invd);
    if (min_pivot_ratio) *min_pivot_ratio = minr;
    if (n_skipped) *n_skipped = skipped;
    double dn = (double)n;
    double ip = yty - (dn / (dn + 1.0)) * q;
    return -0.5 * (double)k * log(1.0 + dn) - 0.5 * dn * log(ip);
}

/* Third, independent statement: Hill et al. 2012 (hill-method/dynamic_network_inference.m:621)
 *   LL = (-d/2) log(1+n) + (-n/2) log(Y'Y - n/(n+1) Y'X (X'X)^-1 X'Y)
 * solved by Gauss-Jordan with partial pivoting on the normal equations of the standardised
 * design built from RAW Xminus.  Well-conditioned sets only. */
double orc_lml_hill(const double* Xminus, const double* y, int n, const int* S, int k) {
    double dn = (double)n;
    double yty = 0.0;
    for (int t = 0; t < n; ++t) yty += y[t] * y[t];
    if (k == 0) return -0.5 * dn * log(yty);
    double* B = (double*)malloc(sizeof(double) * (size_t)n * k);
    orc_construct_B(Xminus, n, S, k, 1, B);
    int m = k + 1;
    double* A = (double*)calloc((size_t)k * m, sizeof(double)); /* [X'X | X'y], row-major */
    for (int a = 0; a < k; ++a) {
        for (int b = 0; b < k; ++b) {
            double s = 0.0;
            for (int t = 0; t < n; ++t) s += B[(size_t)a * n + t] * B[(size_t)b * n + t];
            A[(size_t)a * m + b] = s;
        }
        double s = 0.0;
        for (int t = 0; t < n; ++t) s += B[(size_t)a * n + t] * y[t];
        A[(size_t)a * m + k] = s;
    }
    double* xty = (double*)malloc(sizeof(double) * (size_t)k);
    for (int a = 0; a < k; ++a) xty[a] = A[(size_t)a * m + k];
    for (int c = 0; c < k; ++c) {
        int p = c; double best = fabs(A[(size_t)c * m + c]);
        for (int r2 = c + 1; r2 < k; ++r2)
            if (fabs(A[(size_t)r2 * m + c]) > best) { best = fabs(A[(size_t)r2 * m + c]); p = r2; }
        if (p != c)
            for (int e = 0; e < m; ++e) {
                double tmp = A[(size_t)c * m + e]; A[(size_t)c * m + e] = A[(size_t)p * m + e];
                A[(size_t)p * m + e] = tmp;
            }
        double piv = A[(size_t)c * m + c];
        for (int e = 0; e < m; ++e) A[(size_t)c * m + e] /= piv;
        for (int r2 = 0; r2 < k; ++r2) {
            if (r2 == c) continue;
            double f = A[(size_t)r2 * m + c];
            for (int e = 0; e < m; ++e) A[(size_t)r2 * m + e] -= f * A[(size_t)c * m + e];
        }
    }
    double q = 0.0;
    for (int a = 0; a < k; ++a) q += xty[a] * A[(size_t)a * m + k];
    free(A); free(B); free(xty);
    return (-0.5 * (double)k) * log(1.0 + dn) + (-0.5 * dn) * log(yty - (dn / (dn + 1.0)) * q);
}

/* ------------------------------------------------------------------------------------------
 * ConfParentPrior.logpdf (dbn_distributions.jl:96-153), both branches, cache ignored.
 * conf_col[j] = confidence of parent j for this child (0 when absent from the Dict).
 * ---------------------------------------------------------------------------------------- */
static int in_sorted(const int* S, int k, int x) {
    int lo = 0, hi = k;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (S[mid] < x) lo = mid + 1; else hi = mid; }
    return lo < k && S[lo] == x;
}

double orc_logprior_ref(const int* S, int k, int V, const double* conf_col, double lambda) {
    double enl = exp(-lambda);
    int nnz = 0;
    for (int j = 0; j < V; ++j) if (conf_col[j] > 0.0) ++nnz;
    double lp = 0.0;
    if (nnz > 50) { /* :110-128 */
        lp += -log(1.0 + enl) * (double)(V - nnz);
        double s = 0.0;
        for (int j = 0; j < V; ++j)
            if (conf_col[j] > 0.0) { double x = -lambda * conf_col[j]; s += x - log(enl + exp(x)); }
        lp += s;
        for (int a = 0; a < k; ++a) {
            double c = conf_col[S[a]];
            if (c > 0.0) lp += -lambda * (1.0 - c); else lp += -lambda;
        }
    } else { /* :131-147 */
        for (int a = 0; a < k; ++a) {
            double c = conf_col[S[a]];
            lp += -lambda - log(enl + exp(-c * lambda));
        }
        int n_union = k;
        for (int j = 0; j < V; ++j) {
            if (conf_col[j] > 0.0 && !in_sorted(S, k, j)) {
                double npl = -conf_col[j] * lambda;
                lp += npl - log(enl + exp(npl));
                ++n_union;
            }
        }
        double zero_n_lp = -log(1.0 + enl);
        lp += zero_n_lp * (double)(V - n_union);
    }
    return lp;
}

/* closed form: -lambda * sum_{j in S}(1-c_j) - sum_j log1p(exp(-lambda (1-c_j))) */
double orc_logprior_closed(const int* S, int k, int V, const double* conf_col, double lambda) {
    double s = 0.0;
    for (int a = 0; a < k; ++a) s += 1.0 - conf_col[S[a]];
    double nrm = 0.0;
    for (int j = 0; j < V; ++j) nrm += log1p(exp(-lambda * (1.0 - conf_col[j])));
    return -lambda * s - nrm;
}

/* compute_move_probs (dbn_proposals.jl:11-16) */
void orc_move_probs(int k, int V, double prob_param, double out[3]) {
    double frac = pow((double)k / (double)V, prob_param);
    double comp = 1.0 - frac;
    double swp = 2.0 * frac * comp;
    double tot = frac + swp + comp;
    out[0] = comp / tot; out[1] = swp / tot; out[2] = frac / tot;
}

/* ith_nonparent (dbn_proposals.jl:19-34), 1-based i, S holds 1-based sorted ids here */
int orc_ith_nonparent(int i, const int* S, int k) {
    if (k == 0) return i;
    int p = 1;
    while (p <= k) {
        if (S[p - 1] - p >= i) break;
        ++p;
    }
    return i + p - 1;
}

/* searchsorted_exclude (dbn_proposals.jl:42-49), 1-based values */
int orc_searchsorted_exclude(const int* S, int k, int x) {
    int start = 0; while (start < k && S[start] < x) ++start;   /* first idx (0-based) with S >= x */
    int stop = start; while (stop < k && S[stop] == x) ++stop;  /* one past last idx with S <= x */
    /* Julia: r.start = start+1, r.stop = stop (1-based) */
    if (stop < start + 1) return x - stop;
    return x - stop + 1;
}

/* ------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al. 2011) -- the stream convention shared with the CUDA kernels.
 * ---------------------------------------------------------------------------------------- */
void orc_philox4x32_10(const uint32_t ctr_in[4], const uint32_t key_in[2], uint32_t out[4]) {
    uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
    uint32_t k0 = key_in[0], k1 = key_in[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static void philox_block(uint64_t seed, uint32_t chain, uint32_t vertex, uint64_t sweep,
                         uint32_t block, uint32_t out[4]) {
    uint32_t ctr[4] = { block | (vertex << 8), (uint32_t)sweep, (uint32_t)(sweep >> 32), chain };
    uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    orc_philox4x32_10(ctr, key, out);
}

static inline double u32_to_unit(uint32_t w) { return ((double)w + 0.5) * 2.3283064365386963e-10; }
static inline double u53_to_unit(uint32_t hi, uint32_t lo) {
    uint64_t x = (((uint64_t)hi << 32) | lo) >> 11;
    return (double)x * 1.1102230246251565e-16;
}

/* the six per-step numbers: z, u_lam, u_act, u_add, u_rem, u_acc */
void orc_step_draws(uint64_t seed, uint32_t chain, uint32_t vertex, uint64_t sweep, double out[6]) {
    uint32_t a[4], b[4], c[4];
    philox_block(seed, chain, vertex, sweep, 0, a);
    philox_block(seed, chain, vertex, sweep, 1, b);
    philox_block(seed, chain, vertex, sweep, 2, c);
    double u1 = u32_to_unit(a[0]), u2 = u32_to_unit(a[1]);
    out[0] = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2); /* Box-Muller */
    out[1] = u53_to_unit(a[2], a[3]);
    out[2] = u32_to_unit(b[0]);
    out[3] = u32_to_unit(b[1]);
    out[4] = u32_to_unit(b[2]);
    out[5] = u53_to_unit(c[0], c[1]);
}

/* ------------------------------------------------------------------------------------------
 * Problem
 * ---------------------------------------------------------------------------------------- */
struct orc_problem {
    int n, V, deg;
    double lambda_min, lambda_max, lambda_prop_std;
    double *Xminus, *Z, *Y, *G, *R, *yty;
    double *conf;       /* V x V col-major [parent + V*child] */
    double *omc;        /* 1 - conf */
    double *prob_param; /* V */
    /* distinct (1-c) groups per child, order of first appearance */
    int *grp_off;       /* V+1 */
    double *grp_omc; double *grp_mult;
    /* move tables per child, k = 0..V : P[3], logP[3] */
    double *mvP, *mvLogP;
    double *logint;     /* log(m), m = 0..V+1 */
};

orc_problem* orc_problem_create(const double* Xminus, const double* Xplus, int n, int V,
                                const double* prior, double large_indeg, double lambda_max,
                                double lambda_prop_std, int regression_deg) {
    orc_problem* P = (orc_problem*)calloc(1, sizeof(orc_problem));
    P->n = n; P->V = V; P->deg = regression_deg;
    P->lambda_max = lambda_max; P->lambda_prop_std = lambda_prop_std;
    /* SSPS.jl:169  lambda_min = log(max(V/large_indeg - 1.0, exp(0.5))) */
    P->lambda_min = log(fmax((double)V / large_indeg - 1.0, exp(0.5)));
    size_t nv = (size_t)n * V, vv = (size_t)V * V;
    P->Xminus = (double*)malloc(sizeof(double) * nv); memcpy(P->Xminus, Xminus, sizeof(double) * nv);
    P->Z = (double*)malloc(sizeof(double) * nv);
    P->Y = (double*)malloc(sizeof(double) * nv);
    P->G = (double*)malloc(sizeof(double) * vv);
    P->R = (double*)malloc(sizeof(double) * vv);
    P->yty = (double*)malloc(sizeof(double) * (size_t)V);
    orc_standardize(Xminus, n, V, P->Z);   /* quirk Q4: done per column inside compute_B_col */
    orc_standardize(Xplus, n, V, P->Y);    /* SSPS.jl:156 */
    orc_gram(P->Z, P->Y, n, V, P->G, P->R, P->yty);
    P->conf = (double*)malloc(sizeof(double) * vv);
    P->omc = (double*)malloc(sizeof(double) * vv);
    for (size_t e = 0; e < vv; ++e) {
        double c = prior[e] > 0.0 ? prior[e] : 0.0; /* dbn_preprocess.jl:36: only > 0 enters the Dict */
        P->conf[e] = c; P->omc[e] = 1.0 - c;
    }
    /* SSPS.jl:164-165  prob_param_i = 1 / log2(V / max(sum_j c_ji, 2)) */
    P->prob_param = (double*)malloc(sizeof(double) * (size_t)V);
    for (int i = 0; i < V; ++i) {
        double s = 0.0;
        for (int j = 0; j < V; ++j) s += P->conf[(size_t)j + (size_t)V * i];
        double cnt = fmax(s, 2.0);
        P->prob_param[i] = 1.0 / log2((double)V / cnt);
    }
    /* groups of distinct (1-c) */
    P->grp_off = (int*)malloc(sizeof(int) * (size_t)(V + 1));
    P->grp_omc = (double*)malloc(sizeof(double) * vv);
    P->grp_mult = (double*)malloc(sizeof(double) * vv);
    int off = 0;
    for (int i = 0; i < V; ++i) {
        P->grp_off[i] = off;
        int ng = 0;
        for (int j = 0; j < V; ++j) {
            double o = P->omc[(size_t)j + (size_t)V * i];
            int g = 0;
            for (; g < ng; ++g) if (P->grp_omc[off + g] == o) break;
            if (g == ng) { P->grp_omc[off + ng] = o; P->grp_mult[off + ng] = 0.0; ++ng; }
            P->grp_mult[off + g] += 1.0;
        }
        off += ng;
    }
    P->grp_off[V] = off;
    /* move tables */
    P->mvP = (double*)malloc(sizeof(double) * (size_t)V * (V + 1) * 3);
    P->mvLogP = (double*)malloc(sizeof(double) * (size_t)V * (V + 1) * 3);
    for (int i = 0; i < V; ++i)
        for (int k = 0; k <= V; ++k) {
            double pr[3];
            orc_move_probs(k, V, P->prob_param[i], pr);
            for (int a = 0; a < 3; ++a) {
                P->mvP[((size_t)i * (V + 1) + k) * 3 + a] = pr[a];
                P->mvLogP[((size_t)i * (V + 1) + k) * 3 + a] = log(pr[a]);
            }
        }
    P->logint = (double*)malloc(sizeof(double) * (size_t)(V + 2));
    for (int m = 0; m <= V + 1; ++m) P->logint[m] = log((double)m);
    return P;
}

void orc_problem_destroy(orc_problem* P) {
    if (!P) return;
    free(P->Xminus); free(P->Z); free(P->Y); free(P->G); free(P->R); free(P->yty);
    free(P->conf); free(P->omc); free(P->prob_param); free(P->grp_off); free(P->grp_omc);
    free(P->grp_mult); free(P->mvP); free(P->mvLogP); free(P->logint); free(P);
}
int orc_problem_n(const orc_problem* P) { return P->n; }
int orc_problem_V(const orc_problem* P) { return P->V; }
double orc_problem_lambda_min(const orc_problem* P) { return P->lambda_min; }
const double* orc_problem_G(const orc_problem* P) { return P->G; }
const double* orc_problem_R(const orc_problem* P) { return P->R; }
const double* orc_problem_Z(const orc_problem* P) { return P->Z; }
const double* orc_problem_Y(const orc_problem* P) { return P->Y; }
const double* orc_problem_yty(const orc_problem* P) { return P->yty; }
const double* orc_problem_prob_param(const orc_problem* P) { return P->prob_param; }

double orc_problem_lml(const orc_problem* P, int child, const int* S, int k, int scorer) {
    if (scorer == ORC_SCORER_QR)
        return orc_lml_qr(P->Xminus, P->Y + (size_t)child * P->n, P->n, S, k, P->deg, NULL);
    return orc_lml_chol(P->G, P->R + (size_t)P->V * child, P->yty[child], P->n, P->V, S, k, NULL, NULL);
}

/* kernel-faithful normaliser: sum over distinct-(1-c) groups, order of first appearance */
static double normaliser(const orc_problem* P, int child, double lambda) {
    double acc = 0.0;
    for (int g = P->grp_off[child]; g < P->grp_off[child + 1]; ++g)
        acc = fma(P->grp_mult[g], log1p(exp(-lambda * P->grp_omc[g])), acc);
    return acc;
}
static double sum_omc(const orc_problem* P, int child, const int* S, int k) {
    double s = 0.0;
    for (int a = 0; a < k; ++a) s += P->omc[(size_t)S[a] + (size_t)P->V * child];
    return s;
}

double orc_problem_logprior(const orc_problem* P, int child, const int* S, int k, double lambda,
                            int closed_form) {
    const double* cc = P->conf + (size_t)P->V * child;
    if (closed_form == 2) /* the grouped form the kernel uses */
        return -lambda * sum_omc(P, child, S, k) - normaliser(P, child, lambda);
    return closed_form ? orc_logprior_closed(S, k, P->V, cc, lambda)
                       : orc_logprior_ref(S, k, P->V, cc, lambda);
}

/* ------------------------------------------------------------------------------------------
 * fingerprint hash map  [i;S] -> double   (stand-in for LRU{Vector{Int64},Float64})
 * ---------------------------------------------------------------------------------------- */
typedef struct { uint64_t* keys; double* vals; size_t cap, used; } fpmap;
static void fpmap_init(fpmap* m, size_t cap) {
    m->cap = cap; m->used = 0;
    m->keys = (uint64_t*)calloc(cap, sizeof(uint64_t));
    m->vals = (double*)malloc(cap * sizeof(double));
}
static void fpmap_free(fpmap* m) { free(m->keys); free(m->vals); m->keys = NULL; m->vals = NULL; }
static uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31;
    return x;
}
static uint64_t key_of(int child, const int* S, int k) {
    uint64_t h = mix64(0x9E3779B97F4A7C15ULL + (uint64_t)child);
    for (int a = 0; a < k; ++a) h = mix64(h ^ ((uint64_t)S[a] + 0x632BE59BD9B4E019ULL + (h << 6)));
    return h | 1ULL; /* 0 marks empty */
}
static int fpmap_get(const fpmap* m, uint64_t key, double* v) {
    size_t i = (size_t)(key >> 1) & (m->cap - 1);
    while (m->keys[i]) { if (m->keys[i] == key) { *v = m->vals[i]; return 1; } i = (i + 1) & (m->cap - 1); }
    return 0;
}
static void fpmap_put(fpmap* m, uint64_t key, double v) {
    if (m->used * 10 > m->cap * 7) { memset(m->keys, 0, m->cap * sizeof(uint64_t)); m->used = 0; }
    size_t i = (size_t)(key >> 1) & (m->cap - 1);
    while (m->keys[i] && m->keys[i] != key) i = (i + 1) & (m->cap - 1);
    if (!m->keys[i]) ++m->used;
    m->keys[i] = key; m->vals[i] = v;
}

/* ------------------------------------------------------------------------------------------
 * Chain
 * ---------------------------------------------------------------------------------------- */
typedef struct { int64_t* n; int* parent; int* value; size_t len, cap; } pev_list;
typedef struct { int64_t* n; double* value; size_t len, cap; } lev_list;

struct orc_chain {
    const orc_problem* P;
    int scorer, flags;
    uint64_t seed; uint32_t chain_id;
    int64_t sweeps;
    int* k; int* S;          /* S: V rows of capacity V, sorted ascending */
    int* A;                  /* the same sets in ARRIVAL order (initial state arrives sorted; an accepted add
                                appends, a remove deletes in place, a swap deletes then appends): the order the
                                kernel-faithful Cholesky scorer factorises in (csrc/ssps_mh.cuh) */
    double* lambda; double* lml; double* nrm; double* somc;
    double* lp;              /* ref-faithful: cached full logprior of the current state */
    int64_t n_prop, n_acc, cache_hits;
    int64_t* khist; int khist_len;
    fpmap lml_cache, prior_cache;
    int* tmp;                /* V scratch */
    /* recorder */
    int rec_on; int64_t rec_n;
    pev_list* pev; lev_list* lev; int* prevS; int* prevk; int64_t* counts;
};

/* S: sorted list (the reference's key / design-column order); A: the same set in arrival order.
 * The Cholesky scorer factorises in arrival order (what the device caches incrementally); the
 * reference-faithful QR scorer builds B(S) from the sorted list like construct_B does. */
static double chain_lml(orc_chain* C, int child, const int* S, const int* A, int k) {
    if (C->scorer == ORC_SCORER_CHOL) return orc_problem_lml(C->P, child, A, k, ORC_SCORER_CHOL);
    if (C->flags & ORC_FLAG_LML_CACHE) {
        uint64_t key = key_of(child, S, k);
        double v;
        if (fpmap_get(&C->lml_cache, key, &v)) { ++C->cache_hits; return v; }
        v = orc_problem_lml(C->P, child, S, k, C->scorer);
        fpmap_put(&C->lml_cache, key, v);
        return v;
    }
    return orc_problem_lml(C->P, child, S, k, C->scorer);
}

/* full log prior as the reference evaluates it (with optional stale cache, quirk Q1) */
static double chain_logprior_full(orc_chain* C, int child, const int* S, int k, double lambda) {
    const orc_problem* P = C->P;
    if (C->flags & ORC_FLAG_STALE_PRIOR) {
        uint64_t key = key_of(child, S, k);
        double v;
        if (fpmap_get(&C->prior_cache, key, &v)) return v;
        v = orc_logprior_ref(S, k, P->V, P->conf + (size_t)P->V * child, lambda);
        fpmap_put(&C->prior_cache, key, v);
        return v;
    }
    return orc_logprior_ref(S, k, P->V, P->conf + (size_t)P->V * child, lambda);
}

orc_chain* orc_chain_create(const orc_problem* P, int scorer, int flags) {
    orc_chain* C = (orc_chain*)calloc(1, sizeof(orc_chain));
    int V = P->V;
    C->P = P; C->scorer = scorer; C->flags = flags;
    C->k = (int*)calloc((size_t)V, sizeof(int));
    C->S = (int*)calloc((size_t)V * V, sizeof(int));
    C->A = (int*)calloc((size_t)V * V, sizeof(int));
    C->lambda = (double*)calloc((size_t)V, sizeof(double));
    C->lml = (double*)calloc((size_t)V, sizeof(double));
    C->nrm = (double*)calloc((size_t)V, sizeof(double));
    C->somc = (double*)calloc((size_t)V, sizeof(double));
    C->lp = (double*)calloc((size_t)V, sizeof(double));
    C->tmp = (int*)calloc(2 * (size_t)V + 2, sizeof(int));
    C->khist_len = V + 1;
    C->khist = (int64_t*)calloc((size_t)V + 1, sizeof(int64_t));
    if (flags & ORC_FLAG_LML_CACHE) fpmap_init(&C->lml_cache, (size_t)1 << 18);
    if (flags & ORC_FLAG_STALE_PRIOR) fpmap_init(&C->prior_cache, (size_t)1 << 18);
    return C;
}

void orc_chain_destroy(orc_chain* C) {
    if (!C) return;
    int V = C->P->V;
    free(C->k); free(C->S); free(C->A); free(C->lambda); free(C->lml); free(C->nrm); free(C->somc); free(C->lp);
    free(C->tmp); free(C->khist);
    if (C->lml_cache.keys) fpmap_free(&C->lml_cache);
    if (C->prior_cache.keys) fpmap_free(&C->prior_cache);
    if (C->rec_on) {
        for (int i = 0; i < V; ++i) {
            free(C->pev[i].n); free(C->pev[i].parent); free(C->pev[i].value);
            free(C->lev[i].n); free(C->lev[i].value);
        }
        free(C->pev); free(C->lev); free(C->prevS); free(C->prevk); free(C->counts);
    }
    free(C);
}

static void chain_refresh_vertex(orc_chain* C, int i) {
    const orc_problem* P = C->P;
    int* S = C->S + (size_t)i * P->V;
    C->lml[i] = chain_lml(C, i, S, C->A + (size_t)i * P->V, C->k[i]);
    C->nrm[i] = normaliser(P, i, C->lambda[i]);
    C->somc[i] = sum_omc(P, i, S, C->k[i]);
    C->lp[i] = chain_logprior_full(C, i, S, C->k[i], C->lambda[i]);
}

/* Gen.generate (mcmc_inference.jl:40): lambda_i ~ U(lmin,lmax) (dbn_models.jl:66-70),
 * S_i ~ ConfParentPrior.random (dbn_distributions.jl:75-89). Stream: sweep 0. */
void orc_chain_init(orc_chain* C, uint64_t seed, uint32_t chain_id) {
    const orc_problem* P = C->P;
    int V = P->V;
    C->seed = seed; C->chain_id = chain_id; C->sweeps = 0;
    for (int i = 0; i < V; ++i) {
        uint32_t w[4];
        philox_block(seed, chain_id, (uint32_t)i, 0, 0, w);
        double u = u53_to_unit(w[0], w[1]);
        double lam = u * (P->lambda_max - P->lambda_min) + P->lambda_min;
        C->lambda[i] = lam;
        double enl = exp(-lam);
        int* S = C->S + (size_t)i * V;
        int k = 0;
        for (int j = 0; j < V; ++j) {
            if ((j & 3) == 0) philox_block(seed, chain_id, (uint32_t)i, 0, 1u + (uint32_t)(j >> 2), w);
            double conf = P->conf[(size_t)j + (size_t)V * i];
            double p = enl / (enl + exp(-conf * lam));
            if (u32_to_unit(w[j & 3]) < p) S[k++] = j;
        }
        C->k[i] = k;
        memcpy(C->A + (size_t)i * V, S, sizeof(int) * (size_t)k);
        chain_refresh_vertex(C, i);
    }
}

void orc_chain_set_state(orc_chain* C, const int* offsets, const int* members, const double* lambda) {
    int V = C->P->V;
    for (int i = 0; i < V; ++i) {
        int k = offsets[i + 1] - offsets[i];
        C->k[i] = k;
        memcpy(C->S + (size_t)i * V, members + offsets[i], sizeof(int) * (size_t)k);
        memcpy(C->A + (size_t)i * V, members + offsets[i], sizeof(int) * (size_t)k);
        C->lambda[i] = lambda[i];
        chain_refresh_vertex(C, i);
    }
}

/* One (lambda move, parent-set move) pair for vertex i -- state_updates.jl:44-47 */
static void chain_step(orc_chain* C, int i, const double d[6], int* acc_out) {
    const orc_problem* P = C->P;
    const int V = P->V;
    int* S = C->S + (size_t)i * V;
    int k = C->k[i];
    const double z = d[0], u_lam = d[1], u_act = d[2], u_add = d[3], u_rem = d[4], u_acc = d[5];
    const int ref_mode = (C->scorer == ORC_SCORER_QR);
    int acc_l = 0, acc_p = 0;

    /* ---- lambda move: lambda_vec_proposal (dbn_proposals.jl:244-248) + Gen.mh ---- */
    double lam = C->lambda[i];
    double lam_new = lam + P->lambda_prop_std * z;
    if (lam_new >= P->lambda_min && lam_new <= P->lambda_max) { /* Gen.uniform logpdf finite */
        if (C->flags & ORC_FLAG_STALE_PRIOR) {
            /* quirk Q1: the cached prior ignores lambda => weight 0 => always accepted */
            double w = 0.0;
            if (log(u_lam) < w) { lam = lam_new; acc_l = 1; }
        } else if (ref_mode) {
            double lp_new = chain_logprior_full(C, i, S, k, lam_new);
            double w = lp_new - C->lp[i];
            if (log(u_lam) < w) { lam = lam_new; C->lp[i] = lp_new; acc_l = 1; }
        } else {
            double nrm_new = normaliser(P, i, lam_new);
            double w = -(lam_new - lam) * C->somc[i] - (nrm_new - C->nrm[i]);
            if (log(u_lam) < w) { lam = lam_new; C->nrm[i] = nrm_new; acc_l = 1; }
        }
    }
    C->lambda[i] = lam;

    /* ---- parent-set move: smart_proposal / smart_involution (dbn_proposals.jl:60-130) ---- */
    const double* mvP = P->mvP + ((size_t)i * (V + 1) + k) * 3;
    const double* mvL = P->mvLogP + ((size_t)i * (V + 1) + k) * 3;
    double cp1 = mvP[0], cp2 = mvP[0] + mvP[1];
    int action = u_act < cp1 ? ORC_ACT_ADD : (u_act < cp2 ? ORC_ACT_SWAP : ORC_ACT_REM);
    int* T = C->tmp;
    int* TA = C->tmp + V + 1;      /* proposed set in arrival order */
    int* Acur = C->A + (size_t)i * V;
    int kn = k, j_add = -1, j_rem = -1;
    double fwd = 0.0, bwd = 0.0;
    if (action == ORC_ACT_ADD || action == ORC_ACT_SWAP) {
        int m = V - k;
        int a = 1 + (int)floor(u_add * (double)m); if (a > m) a = m;
        /* ith_nonparent on 0-based storage */
        int p = 0;
        while (p < k) { if (S[p] - p >= a) break; ++p; }
        j_add = a + p - 1; /* 0-based id of the a-th (1-based) non-parent */
    }
    if (action == ORC_ACT_SWAP || action == ORC_ACT_REM) {
        int r = 1 + (int)floor(u_rem * (double)k); if (r > k) r = k;
        j_rem = S[r - 1];
    }
    /* build the proposed sorted set */
    {
        int t = 0, placed = (j_add < 0);
        for (int a = 0; a < k; ++a) {
            if (S[a] == j_rem) continue;
            if (!placed && j_add < S[a]) { T[t++] = j_add; placed = 1; }
            T[t++] = S[a];
        }
        if (!placed) T[t++] = j_add;
        kn = t;
        t = 0;
        for (int a = 0; a < k; ++a) if (Acur[a] != j_rem) TA[t++] = Acur[a];
        if (j_add >= 0) TA[t++] = j_add;
    }
    if (action == ORC_ACT_ADD) {
        fwd = mvL[0] - P->logint[V - k];
        bwd = P->mvLogP[((size_t)i * (V + 1) + (k + 1)) * 3 + 2] - P->logint[k + 1];
    } else if (action == ORC_ACT_REM) {
        fwd = mvL[2] - P->logint[k];
        bwd = P->mvLogP[((size_t)i * (V + 1) + (k - 1)) * 3 + 0] - P->logint[V - k + 1];
    }
    ++C->n_prop;
    if (kn < C->khist_len) ++C->khist[kn];
    double lml_new = chain_lml(C, i, T, TA, kn);
    double alpha, lp_new = 0.0;
    if (ref_mode || (C->flags & ORC_FLAG_STALE_PRIOR)) {
        lp_new = chain_logprior_full(C, i, T, kn, lam);
        double lp_old = (C->flags & ORC_FLAG_STALE_PRIOR) ? chain_logprior_full(C, i, S, k, lam) : C->lp[i];
        double w = (lp_new - lp_old) + (lml_new - C->lml[i]);
        alpha = w - fwd + bwd;
    } else {
        double dprior = 0.0;
        if (j_add >= 0) dprior = -lam * P->omc[(size_t)j_add + (size_t)V * i];
        if (j_rem >= 0) dprior = dprior + lam * P->omc[(size_t)j_rem + (size_t)V * i];
        double w = dprior + (lml_new - C->lml[i]);
        alpha = (action == ORC_ACT_SWAP) ? w : ((w - fwd) + bwd);
    }
    if (log(u_acc) < alpha) {
        acc_p = 1; ++C->n_acc;
        memcpy(S, T, sizeof(int) * (size_t)kn);
        memcpy(Acur, TA, sizeof(int) * (size_t)kn);
        C->k[i] = kn;
        C->lml[i] = lml_new;
        C->somc[i] = sum_omc(P, i, S, kn);
        C->lp[i] = lp_new;
    }
    if (acc_out) { acc_out[0] = acc_l; acc_out[1] = acc_p; acc_out[2] = action; }
}

/* ------------------------------------------------------------------------------------------
 * The deprecated uniform graph proposal (--proposal uniform, SSPS.jl:87-88,173-175):
 *   uniform_update_loop   state_updates.jl:98-112   for i = 1..V: lambda move of vertex i, then ONE uniform graph move
 *   unpaired_edges        dbn_proposals.jl:142-155  edges (p -> child), p != child, whose reverse is not an edge,
 *                                                   sorted as (p, child) tuples
 *   uniform_proposal      dbn_proposals.jl:161-179  idx ~ uniform_discrete(1, V^2 + |ue|); idx <= V^2: the pair
 *                                                   (u, v) = (div(idx-1, V), (idx-1) % V), REMOVE if u is a parent of v
 *                                                   else ADD; above: REVERSE the unpaired edge ue[idx - V^2]
 *   uniform_involution    dbn_proposals.jl:184-230  backward choice = the same cell (add <-> remove) or the position of the
 *                                                   reversed edge in the new list; Gen.update weight = change of
 *                                                   (log prior + lml) of the one or two vertices whose parent set changed
 * Gen.mh accepts iff log(u) < weight + log q(backward) - log q(forward) = weight + log(bound) - log(bound'), bound' =
 * V^2 + |ue'| of the proposed graph (add / remove change |ue| by +-1, a self-edge and a reversal leave it).
 * Draw convention (shared with the device): d[2] picks idx = 1 + floor(d[2] bound), d[5] is the acceptance uniform;
 * d[3], d[4] are not used.  Parent lists: S sorted, A in arrival order (add appends, remove deletes in place).
 * ---------------------------------------------------------------------------------------- */
static int set_insert_sorted(int* S, int k, int x) { int a = k; while (a > 0 && S[a - 1] > x) { S[a] = S[a - 1]; --a; } S[a] = x; return k + 1; }
static int set_erase(int* S, int k, int x) { int t = 0; for (int a = 0; a < k; ++a) if (S[a] != x) S[t++] = S[a]; return t; }

static int64_t count_unpaired(const orc_chain* C) {
    const int V = C->P->V;
    int64_t n = 0;
    for (int c = 0; c < V; ++c)
        for (int a = 0; a < C->k[c]; ++a) {
            const int p = C->S[(size_t)c * V + a];
            if (p != c && !in_sorted(C->S + (size_t)p * V, C->k[p], c)) ++n;
        }
    return n;
}
/* the r-th (0-based) unpaired edge in (p, child) order */
static void nth_unpaired(const orc_chain* C, int64_t r, int* p_out, int* c_out) {
    const int V = C->P->V;
    for (int p = 0; p < V; ++p)
        for (int c = 0; c < V; ++c)
            if (p != c && in_sorted(C->S + (size_t)c * V, C->k[c], p) && !in_sorted(C->S + (size_t)p * V, C->k[p], c)) {
                if (r == 0) { *p_out = p; *c_out = c; return; }
                --r;
            }
    *p_out = -1; *c_out = -1;
}

/* lml and prior change of vertex v when parent u is added (add = 1) or removed; the proposed lists go to T / TA */
static double uniform_delta(orc_chain* C, int v, int u, int add, int* T, int* TA, int* kn_out, double* lml_new_out) {
    const orc_problem* P = C->P;
    const int V = P->V, k = C->k[v];
    const int* S = C->S + (size_t)v * V; const int* A = C->A + (size_t)v * V;
    memcpy(T, S, sizeof(int) * (size_t)k); memcpy(TA, A, sizeof(int) * (size_t)k);
    int kn;
    if (add) { kn = set_insert_sorted(T, k, u); TA[k] = u; }
    else { kn = set_erase(T, k, u); set_erase(TA, k, u); }
    const double lml_new = chain_lml(C, v, T, TA, kn);
    const double lam = C->lambda[v], omc = P->omc[(size_t)u + (size_t)V * v];
    const double dprior = add ? -lam * omc : lam * omc;
    *kn_out = kn; *lml_new_out = lml_new;
    return dprior + (lml_new - C->lml[v]);
}
static void uniform_commit(orc_chain* C, int v, const int* T, const int* TA, int kn, double lml_new) {
    const int V = C->P->V;
    memcpy(C->S + (size_t)v * V, T, sizeof(int) * (size_t)kn);
    memcpy(C->A + (size_t)v * V, TA, sizeof(int) * (size_t)kn);
    C->k[v] = kn; C->lml[v] = lml_new;
    C->somc[v] = sum_omc(C->P, v, C->S + (size_t)v * V, kn);
    C->lp[v] = chain_logprior_full(C, v, C->S + (size_t)v * V, kn, C->lambda[v]);
}

static void chain_step_uniform(orc_chain* C, int i, const double d[6], int* acc_out) {
    const orc_problem* P = C->P;
    const int V = P->V;
    const double z = d[0], u_lam = d[1], u_idx = d[2], u_acc = d[5];
    int acc_l = 0, acc_p = 0, action = 0;
    /* lambda move of vertex i (lambda_vec_proposal + Gen.mh), exact prior */
    {
        double lam = C->lambda[i];
        const double lam_new = lam + P->lambda_prop_std * z;
        if (lam_new >= P->lambda_min && lam_new <= P->lambda_max) {
            const double nrm_new = normaliser(P, i, lam_new);
            const double w = -(lam_new - lam) * C->somc[i] - (nrm_new - C->nrm[i]);
            if (log(u_lam) < w) { lam = lam_new; C->nrm[i] = nrm_new; acc_l = 1; }
        }
        C->lambda[i] = lam;
        C->lp[i] = chain_logprior_full(C, i, C->S + (size_t)i * V, C->k[i], lam);
    }
    /* the graph move */
    const int64_t n_ue = count_unpaired(C);
    const int64_t vv = (int64_t)V * V, bound = vv + n_ue;
    int64_t idx = 1 + (int64_t)floor(u_idx * (double)bound); if (idx > bound) idx = bound;
    int* T1 = (int*)malloc(sizeof(int) * 4 * ((size_t)V + 1));
    int* TA1 = T1 + V + 1; int* T2 = TA1 + V + 1; int* TA2 = T2 + V + 1;
    ++C->n_prop;
    if (idx <= vv) {
        const int u = (int)((idx - 1) / V), v = (int)((idx - 1) % V);
        const int is_parent = in_sorted(C->S + (size_t)v * V, C->k[v], u);
        const int rev = in_sorted(C->S + (size_t)u * V, C->k[u], v);       /* is v -> u an edge? */
        int kn; double lml_new;
        action = is_parent ? ORC_ACT_REM : ORC_ACT_ADD;
        const double w = uniform_delta(C, v, u, !is_parent, T1, TA1, &kn, &lml_new);
        int64_t n_ue_new = n_ue;
        if (u != v) n_ue_new += is_parent ? (rev ? 1 : -1) : (rev ? -1 : 1);
        const double alpha = w + (log((double)bound) - log((double)(vv + n_ue_new)));
        if (kn < C->khist_len) ++C->khist[kn];
        if (log(u_acc) < alpha) { acc_p = 1; ++C->n_acc; uniform_commit(C, v, T1, TA1, kn, lml_new); }
    } else {
        int u, v;
        nth_unpaired(C, idx - vv - 1, &u, &v);         /* edge u -> v becomes v -> u */
        int kn1, kn2; double lml1, lml2;
        action = ORC_ACT_SWAP;
        const double w1 = uniform_delta(C, v, u, 0, T1, TA1, &kn1, &lml1);
        const double w2 = uniform_delta(C, u, v, 1, T2, TA2, &kn2, &lml2);
        const double alpha = w1 + w2;                  /* |ue| unchanged: the proposal ratio is 1 */
        if (kn2 < C->khist_len) ++C->khist[kn2];
        if (log(u_acc) < alpha) {
            acc_p = 1; ++C->n_acc;
            uniform_commit(C, v, T1, TA1, kn1, lml1);
            uniform_commit(C, u, T2, TA2, kn2, lml2);
        }
    }
    free(T1);
    if (acc_out) { acc_out[0] = acc_l; acc_out[1] = acc_p; acc_out[2] = action; }
}

void orc_chain_sweep(orc_chain* C, const double* draws, int* acc_out) {
    int V = C->P->V;
    ++C->sweeps;
    for (int i = 0; i < V; ++i) {
        double d[6];
        if (draws) memcpy(d, draws + (size_t)i * 6, sizeof(d));
        else orc_step_draws(C->seed, C->chain_id, (uint32_t)i, (uint64_t)C->sweeps, d);
        if (C->flags & ORC_FLAG_UNIFORM) chain_step_uniform(C, i, d, acc_out ? acc_out + (size_t)i * 3 : NULL);
        else chain_step(C, i, d, acc_out ? acc_out + (size_t)i * 3 : NULL);
    }
}

void orc_chain_run(orc_chain* C, int64_t n_sweeps) {
    for (int64_t s = 0; s < n_sweeps; ++s) {
        orc_chain_sweep(C, NULL, NULL);
        if (C->rec_on) orc_chain_record(C);
    }
}

int64_t orc_chain_sweeps_done(const orc_chain* C) { return C->sweeps; }
int orc_chain_indegree(const orc_chain* C, int child) { return C->k[child]; }
void orc_chain_parents(const orc_chain* C, int child, int* out) {
    memcpy(out, C->S + (size_t)child * C->P->V, sizeof(int) * (size_t)C->k[child]);
}
double orc_chain_lambda(const orc_chain* C, int child) { return C->lambda[child]; }
double orc_chain_lml(const orc_chain* C, int child) { return C->lml[child]; }
int64_t orc_chain_n_proposals(const orc_chain* C) { return C->n_prop; }
int64_t orc_chain_n_accepted(const orc_chain* C) { return C->n_acc; }
int64_t orc_chain_cache_hits(const orc_chain* C) { return C->cache_hits; }
void orc_chain_k_hist(const orc_chain* C, int64_t* hist, int len) {
    for (int a = 0; a < len; ++a) hist[a] = a < C->khist_len ? C->khist[a] : 0;
}

/* ------------------------------------------------------------------------------------------
 * Recorder: update_results_storediff (state_updates.jl:161-227)
 * ---------------------------------------------------------------------------------------- */
static void pev_push(pev_list* L, int64_t n, int parent, int value) {
    if (L->len == L->cap) {
        L->cap = L->cap ? L->cap * 2 : 16;
        L->n = (int64_t*)realloc(L->n, L->cap * sizeof(int64_t));
        L->parent = (int*)realloc(L->parent, L->cap * sizeof(int));
        L->value = (int*)realloc(L->value, L->cap * sizeof(int));
    }
    L->n[L->len] = n; L->parent[L->len] = parent; L->value[L->len] = value; ++L->len;
}
static void lev_push(lev_list* L, int64_t n, double v) {
    if (L->len == L->cap) {
        L->cap = L->cap ? L->cap * 2 : 16;
        L->n = (int64_t*)realloc(L->n, L->cap * sizeof(int64_t));
        L->value = (double*)realloc(L->value, L->cap * sizeof(double));
    }
    L->n[L->len] = n; L->value[L->len] = v; ++L->len;
}

void orc_chain_recorder_enable(orc_chain* C) {
    int V = C->P->V;
    C->rec_on = 1; C->rec_n = 0;
    C->pev = (pev_list*)calloc((size_t)V, sizeof(pev_list));
    C->lev = (lev_list*)calloc((size_t)V, sizeof(lev_list));
    C->prevS = (int*)calloc((size_t)V * V, sizeof(int));
    C->prevk = (int*)calloc((size_t)V, sizeof(int));
    C->counts = (int64_t*)calloc((size_t)V * V, sizeof(int64_t));
    for (int i = 0; i < V; ++i) { /* :166-186, index 0 = initial state */
        lev_push(&C->lev[i], 0, C->lambda[i]);
        const int* S = C->S + (size_t)i * V;
        for (int a = 0; a < C->k[i]; ++a) pev_push(&C->pev[i], 0, S[a], 1);
        memcpy(C->prevS + (size_t)i * V, S, sizeof(int) * (size_t)C->k[i]);
        C->prevk[i] = C->k[i];
    }
}

void orc_chain_record(orc_chain* C) {
    int V = C->P->V;
    int64_t n = C->rec_n + 1; /* :190 */
    for (int i = 0; i < V; ++i) {
        lev_list* L = &C->lev[i];
        if (L->value[L->len - 1] != C->lambda[i]) lev_push(L, n, C->lambda[i]); /* :193-197 */
        const int* cur = C->S + (size_t)i * V; int kc = C->k[i];
        int* prev = C->prevS + (size_t)i * V; int kp = C->prevk[i];
        int a = 0, b = 0;
        /* setdiff(cur, prev) -> (n,1);  setdiff(prev, cur) -> (n,0)   (:205-217) */
        while (a < kc || b < kp) {
            if (b >= kp || (a < kc && cur[a] < prev[b])) { pev_push(&C->pev[i], n, cur[a], 1); ++a; }
            else if (a >= kc || prev[b] < cur[a]) { pev_push(&C->pev[i], n, prev[b], 0); ++b; }
            else { ++a; ++b; }
        }
        memcpy(prev, cur, sizeof(int) * (size_t)kc);
        C->prevk[i] = kc;
        for (int e = 0; e < kc; ++e) ++C->counts[(size_t)i * V + cur[e]]; /* increment_counts! :240-249 */
    }
    C->rec_n = n;
}

int64_t orc_chain_recorded_n(const orc_chain* C) { return C->rec_n; }
int64_t orc_chain_n_parent_events(const orc_chain* C, int child) { return (int64_t)C->pev[child].len; }
void orc_chain_parent_events(const orc_chain* C, int child, int64_t* n_idx, int* parent, int* value) {
    const pev_list* L = &C->pev[child];
    memcpy(n_idx, L->n, L->len * sizeof(int64_t));
    memcpy(parent, L->parent, L->len * sizeof(int));
    memcpy(value, L->value, L->len * sizeof(int));
}
int64_t orc_chain_n_lambda_events(const orc_chain* C, int child) { return (int64_t)C->lev[child].len; }
void orc_chain_lambda_events(const orc_chain* C, int child, int64_t* n_idx, double* value) {
    const lev_list* L = &C->lev[child];
    memcpy(n_idx, L->n, L->len * sizeof(int64_t));
    memcpy(value, L->value, L->len * sizeof(double));
}
void orc_chain_edge_counts(const orc_chain* C, int64_t* out) {
    memcpy(out, C->counts, sizeof(int64_t) * (size_t)C->P->V * C->P->V);
}

/* ------------------------------------------------------------------------------------------
 * Post-processing: split-chain PSRF / N_eff (scripts/postprocess_samples.jl:58-180)
 * ---------------------------------------------------------------------------------------- */
static void split_points(int64_t len, double burnin, int64_t* b, int64_t* s) {
    *b = (int64_t)floor(burnin * (double)len);        /* :167 */
    *s = *b + (len - *b) / 2;                          /* :168 div */
}

static double psrf_from_moments(const double* means, const double* vars, int H, int64_t n, double* vp_out) {
    double msum = 0.0; for (int h = 0; h < H; ++h) msum += means[h];
    double mbar = msum / (double)H;
    double bs = 0.0; for (int h = 0; h < H; ++h) bs += (means[h] - mbar) * (means[h] - mbar);
    double B = (double)n * bs / (double)H;                         /* b_stat :110 (1/M, quirk Q6) */
    double ws = 0.0; for (int h = 0; h < H; ++h) ws += vars[h];
    double W = ws / (double)H;                                      /* w_stat :108 */
    double vp = ((double)n - 1.0) * W / (double)n + B / (double)n;  /* varplus :112 */
    if (vp_out) *vp_out = vp;
    return sqrt(vp / W);                                            /* :114 */
}

double orc_psrf_binary_dense(const uint8_t* const* seqs, int n_chains, int64_t len, double burnin) {
    int64_t b, s; split_points(len, burnin, &b, &s);
    int H = 2 * n_chains;
    double* means = (double*)malloc(sizeof(double) * (size_t)H);
    double* vars = (double*)malloc(sizeof(double) * (size_t)H);
    int64_t N1 = s - b, N2 = len - s;
    for (int c = 0; c < n_chains; ++c) {
        int64_t c1 = 0, c2 = 0;
        for (int64_t t = b; t < s; ++t) c1 += seqs[c][t] != 0;   /* samples b+1..s (1-based) */
        for (int64_t t = s; t < len; ++t) c2 += seqs[c][t] != 0;
        double m1 = (double)c1 / (double)N1, m2 = (double)c2 / (double)N2;
        means[2 * c] = m1; means[2 * c + 1] = m2;
        vars[2 * c] = m1 * (1.0 - m1) * (double)N1 / (double)(N1 - 1);   /* seq_var binary :99-106 */
        vars[2 * c + 1] = m2 * (1.0 - m2) * (double)N2 / (double)(N2 - 1);
    }
    double r = psrf_from_moments(means, vars, H, N1, NULL);
    free(means); free(vars);
    return r;
}

double orc_psrf_real_dense(const double* const* seqs, int n_chains, int64_t len, double burnin) {
    int64_t b, s; split_points(len, burnin, &b, &s);
    int H = 2 * n_chains;
    double* means = (double*)malloc(sizeof(double) * (size_t)H);
    double* vars = (double*)malloc(sizeof(double) * (size_t)H);
    int64_t lo[2] = { b, s }, hi[2] = { s, len };
    for (int c = 0; c < n_chains; ++c)
        for (int h = 0; h < 2; ++h) {
            int64_t N = hi[h] - lo[h];
            double sum = 0.0; for (int64_t t = lo[h]; t < hi[h]; ++t) sum += seqs[c][t];
            double m = sum / (double)N;
            double ss = 0.0; for (int64_t t = lo[h]; t < hi[h]; ++t) ss += (seqs[c][t] - m) * (seqs[c][t] - m);
            means[2 * c + h] = m; vars[2 * c + h] = ss / (double)(N - 1);
        }
    double r = psrf_from_moments(means, vars, H, s - b, NULL);
    free(means); free(vars);
    return r;
}

double orc_psrf_from_counts(const int64_t* c1, const int64_t* c2, int n_chains, int64_t N1, int64_t N2) {
    int H = 2 * n_chains;
    double* means = (double*)malloc(sizeof(double) * (size_t)H);
    double* vars = (double*)malloc(sizeof(double) * (size_t)H);
    for (int c = 0; c < n_chains; ++c) {
        double m1 = (double)c1[c] / (double)N1, m2 = (double)c2[c] / (double)N2;
        means[2 * c] = m1; means[2 * c + 1] = m2;
        vars[2 * c] = m1 * (1.0 - m1) * (double)N1 / (double)(N1 - 1);
        vars[2 * c + 1] = m2 * (1.0 - m2) * (double)N2 / (double)(N2 - 1);
    }
    double r = psrf_from_moments(means, vars, H, N1, NULL);
    free(means); free(vars);
    return r;
}

/* N_eff: correlation_sum + n_eff_stat (postprocess_samples.jl:58-94), dense binary version */
double orc_neff_binary_dense(const uint8_t* const* seqs, int n_chains, int64_t len, double burnin) {
    int64_t b, s; split_points(len, burnin, &b, &s);
    int H = 2 * n_chains;
    int64_t N1 = s - b, N2 = len - s;
    double* means = (double*)malloc(sizeof(double) * (size_t)H);
    double* vars = (double*)malloc(sizeof(double) * (size_t)H);
    const uint8_t** hs = (const uint8_t**)malloc(sizeof(uint8_t*) * (size_t)H);
    int64_t* hl = (int64_t*)malloc(sizeof(int64_t) * (size_t)H);
    for (int c = 0; c < n_chains; ++c) {
        hs[2 * c] = seqs[c] + b; hl[2 * c] = N1;
        hs[2 * c + 1] = seqs[c] + s; hl[2 * c + 1] = N2;
    }
    for (int h = 0; h < H; ++h) {
        int64_t cnt = 0; for (int64_t t = 0; t < hl[h]; ++t) cnt += hs[h][t] != 0;
        double m = (double)cnt / (double)hl[h];
        means[h] = m; vars[h] = m * (1.0 - m) * (double)hl[h] / (double)(hl[h] - 1);
    }
    double vp;
    psrf_from_moments(means, vars, H, N1, &vp);
    double sum = 0.0, prev2 = 0.0, prev1 = 0.0;
    int64_t n = hl[0];
    for (int64_t t = 1; t <= n - 2; ++t) {
        double comb = 0.0;
        for (int h = 0; h < H; ++h) {
            int64_t L = hl[h], diff = 0;
            for (int64_t e = 0; e + t < L; ++e) diff += hs[h][e] != hs[h][e + t];
            comb += (double)diff / (double)(L - t);
        }
        comb /= (double)H;
        double corr = 1.0 - comb / (2.0 * vp);
        if (t > 2) {
            if (corr + prev1 >= 0.0) sum += prev2; else break;
        }
        prev2 = prev1; prev1 = corr;
    }
    free(means); free(vars); free(hs); free(hl);
    return (double)H * (double)n / (1.0 + 2.0 * sum);
}

/* ------------------------------------------------------------------------------------------
 * Multi-threaded driver for the CPU baseline (one chain per thread at a time)
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    const orc_problem* P; int scorer, flags, n_chains, tid, n_threads;
    int64_t n_sweeps; uint64_t seed; int64_t* counts; int64_t n_prop;
} mt_arg;

static void* mt_worker(void* p) {
    mt_arg* A = (mt_arg*)p;
    int V = A->P->V;
    for (int c = A->tid; c < A->n_chains; c += A->n_threads) {
        orc_chain* C = orc_chain_create(A->P, A->scorer, A->flags);
        orc_chain_init(C, A->seed, (uint32_t)c);
        for (int64_t s = 0; s < A->n_sweeps; ++s) {
            orc_chain_sweep(C, NULL, NULL);
            if (A->counts)
                for (int i = 0; i < V; ++i) {
                    const int* S = C->S + (size_t)i * V;
                    for (int e = 0; e < C->k[i]; ++e) ++A->counts[(size_t)i * V + S[e]];
                }
        }
        A->n_prop += C->n_prop;
        orc_chain_destroy(C);
    }
    return NULL;
}

double orc_run_chains_mt(const orc_problem* P, int scorer, int flags, int n_chains, int n_threads,
                         int64_t n_sweeps, uint64_t seed, int64_t* edge_counts_total,
                         int64_t* n_proposals_total) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n_chains) n_threads = n_chains;
    int V = P->V;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
    mt_arg* args = (mt_arg*)calloc((size_t)n_threads, sizeof(mt_arg));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < n_threads; ++t) {
        args[t].P = P; args[t].scorer = scorer; args[t].flags = flags; args[t].n_chains = n_chains;
        args[t].tid = t; args[t].n_threads = n_threads; args[t].n_sweeps = n_sweeps; args[t].seed = seed;
        args[t].counts = edge_counts_total ? (int64_t*)calloc((size_t)V * V, sizeof(int64_t)) : NULL;
        pthread_create(&th[t], NULL, mt_worker, &args[t]);
    }
    int64_t np = 0;
    for (int t = 0; t < n_threads; ++t) {
        pthread_join(th[t], NULL);
        np += args[t].n_prop;
        if (edge_counts_total) {
            for (size_t e = 0; e < (size_t)V * V; ++e) {
                if (t == 0) edge_counts_total[e] = 0;
                edge_counts_total[e] += args[t].counts[e];
            }
            free(args[t].counts);
        }
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (n_proposals_total) *n_proposals_total = np;
    free(th); free(args);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
