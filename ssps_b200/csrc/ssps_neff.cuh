// ssps_neff.cuh -- K5: split-chain PSRF and N_eff of every edge indicator and every lambda_i from the change-point log,
// without dense sequences (sm_100a).
//
// Reference semantics (scripts/postprocess_samples.jl, paths relative to the gitter-lab/ssps checkout):
//   evaluate_convergence  :155-180   burn-in index b = floor(burnin * len), split index s = b + div(len - b, 2); the half
//                                    sequences are x[b+1 .. s] and x[s+1 .. len] of every chain
//   compute_psrf_neff     :124-138   means / variances per half sequence (binary: m (1 - m) N / (N - 1), :99-106),
//                                    B = n sum (m_q - mbar)^2 / M (divides by M: quirk Q6), W = mean variance,
//                                    V+ = (n - 1) W / n + B / n, PSRF = sqrt(V+ / W); n = length of the FIRST half sequence
//   correlation_sum       :58-91     for lag t = 1 .. n - 2: variogram_t = mean over half sequences of
//                                    mean_i (x_i != x_{i+t})   (binary)   or   mean_i (x_i - x_{i+t})^2   (real),
//                                    each over its own length; corr_t = 1 - variogram_t / (2 V+); from t = 3 on: if
//                                    corr_t + corr_{t-1} >= 0 add corr_{t-2}, else stop  (the reference's accumulation rule)
//   n_eff_stat            :93        M n / (1 + 2 corr_sum)
//   evaluate_summary      :183-191   pooled mean over the samples after each chain's own burn-in
//   ChangepointVec        scripts/changepoint_vec.jl: value at position t = value of the last change point with index <= t
//                                    (index 0 = initial state), default 0
//
// The reference densifies every sequence whose change-point density exceeds 0.001 (:66-67) and then walks n lags over n
// positions.  Here a half sequence stays a list of runs: for one lag, x_i and x_{i+t} are both piecewise constant, so the
// lag's disagreement count (or sum of squared differences) is a two-pointer merge over the runs -- O(runs) per lag, and exact
// integers for the edge indicators.  One warp per quantity (edge or lambda_i); lanes stride over the M = 2 x chains half
// sequences; the lag loop is sequential (the stopping rule is data dependent), one warp reduction per lag.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

struct NeffLog {            // change points of Q quantities x C chains: sequence (q, c) owns events [off[q * C + c], off[q * C + c + 1])
    const long long* off;
    const int* pos;         // sample index of the change (0 = initial state), ascending inside a sequence
    const double* val;
    int n_chains;
};

// value at position t (>= 1) and index of the event in force: last event with pos <= t; -1 => default 0
__device__ __forceinline__ long long neff_find(const int* __restrict__ pos, long long lo, long long hi, int t) {
    long long a = lo, b = hi;           // first event with pos > t
    while (a < b) { const long long m = (a + b) >> 1; if (pos[m] <= t) a = m + 1; else b = m; }
    return a - 1;                       // lo - 1 if none
}

// sum over u = 1 .. N - t of f(x_{w0 + u}, x_{w0 + u + t}), f = (a != b) or (a - b)^2, for the window (w0, w0 + N]
template <bool BINARY>
__device__ __forceinline__ double neff_lag_sum(const int* __restrict__ pos, const double* __restrict__ val, long long lo, long long hi,
                                               int w0, int N, int t) {
    const int last = N - t;             // u runs 1 .. last
    if (last < 1) return 0.0;
    long long ia = neff_find(pos, lo, hi, w0 + 1), ib = neff_find(pos, lo, hi, w0 + 1 + t);
    int u = 1;
    double acc = 0.0;
    long long acc_i = 0;
    while (u <= last) {
        const double va = ia >= lo ? val[ia] : 0.0, vb = ib >= lo ? val[ib] : 0.0;
        // next u at which either side changes
        int ua = last + 1, ub = last + 1;
        if (ia + 1 < hi) { const long long x = (long long)pos[ia + 1] - w0; if (x < ua) ua = (int)x; }
        if (ib + 1 < hi) { const long long x = (long long)pos[ib + 1] - w0 - t; if (x < ub) ub = (int)x; }
        const int un = ua < ub ? ua : ub;
        const int len = un - u;
        if (BINARY) { if ((va != 0.0) != (vb != 0.0)) acc_i += len; }
        else { const double dlt = va - vb; acc = fma((double)len, dlt * dlt, acc); }
        u = un;
        if (ua == un && ia + 1 < hi) ++ia;
        if (ub == un && ib + 1 < hi) ++ib;
    }
    return BINARY ? (double)acc_i : acc;
}

// sum of g(x) over the window positions w0 + 1 .. w0 + N: g = (x != 0) [BINARY], x [MODE 1], (x - mean)^2 [MODE 2]
template <int MODE>
__device__ __forceinline__ double neff_window_sum(const int* __restrict__ pos, const double* __restrict__ val, long long lo, long long hi,
                                                  int w0, int N, double mean) {
    long long ia = neff_find(pos, lo, hi, w0 + 1);
    int u = 1;
    double acc = 0.0;
    while (u <= N) {
        const double va = ia >= lo ? val[ia] : 0.0;
        int un = N + 1;
        if (ia + 1 < hi) { const long long x = (long long)pos[ia + 1] - w0; if (x < un) un = (int)x; }
        const double len = (double)(un - u);
        if (MODE == 0) acc += (va != 0.0) ? len : 0.0;
        else if (MODE == 1) acc = fma(len, va, acc);
        else { const double dlt = va - mean; acc = fma(len, dlt * dlt, acc); }
        u = un;
        if (ia + 1 < hi) ++ia;
    }
    return acc;
}

__device__ __forceinline__ double neff_warp_sum(double x) {
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}

// One warp per quantity.  b / s / len: burn-in index, split index, stop point.  out_*: [Q]
template <bool BINARY>
__global__ void __launch_bounds__(128) neff_kernel(NeffLog lg, int Q, int b, int s, int len, double* __restrict__ out_mean,
                                                   double* __restrict__ out_psrf, double* __restrict__ out_neff) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= Q) return;
    const int C = lg.n_chains, M = 2 * C;
    const int N1 = s - b, N2 = len - s;
    // ---- compute_psrf_neff: means, variances ----
    double sum_mean = 0.0, sum_var = 0.0, sum_x = 0.0;
    for (int h = lane; h < M; h += 32) {
        const int c = h >> 1, w0 = (h & 1) ? s : b, N = (h & 1) ? N2 : N1;
        const long long lo = lg.off[(size_t)warp * C + c], hi = lg.off[(size_t)warp * C + c + 1];
        double mean, var;
        if (BINARY) {
            const double ones = neff_window_sum<0>(lg.pos, lg.val, lo, hi, w0, N, 0.0);
            mean = ones / (double)N;
            var = mean * (1.0 - mean) * (double)N / (double)(N - 1);
            sum_x += ones;
        } else {
            const double tot = neff_window_sum<1>(lg.pos, lg.val, lo, hi, w0, N, 0.0);
            mean = tot / (double)N;
            var = neff_window_sum<2>(lg.pos, lg.val, lo, hi, w0, N, mean) / (double)(N - 1);
            sum_x += tot;
        }
        sum_mean += mean; sum_var += var;
    }
    sum_mean = neff_warp_sum(sum_mean); sum_var = neff_warp_sum(sum_var); sum_x = neff_warp_sum(sum_x);
    const double mbar = sum_mean / (double)M;
    double ss = 0.0;
    for (int h = lane; h < M; h += 32) {        // second pass for sum (m_q - mbar)^2 (the reference's two-pass form)
        const int c = h >> 1, w0 = (h & 1) ? s : b, N = (h & 1) ? N2 : N1;
        const long long lo = lg.off[(size_t)warp * C + c], hi = lg.off[(size_t)warp * C + c + 1];
        const double mean = (BINARY ? neff_window_sum<0>(lg.pos, lg.val, lo, hi, w0, N, 0.0) : neff_window_sum<1>(lg.pos, lg.val, lo, hi, w0, N, 0.0)) / (double)N;
        const double d = mean - mbar;
        ss = fma(d, d, ss);
    }
    ss = neff_warp_sum(ss);
    const double n = (double)N1;
    const double B = n * ss / (double)M, W = sum_var / (double)M;
    const double vp = (n - 1.0) * W / n + B / n;
    const double psrf = sqrt(vp / W);
    // ---- correlation_sum ----
    double csum = 0.0, prev2 = 0.0, prev1 = 0.0;
    for (int t = 1; t <= N1 - 2; ++t) {
        double part = 0.0;
        for (int h = lane; h < M; h += 32) {
            const int c = h >> 1, w0 = (h & 1) ? s : b, N = (h & 1) ? N2 : N1;
            const long long lo = lg.off[(size_t)warp * C + c], hi = lg.off[(size_t)warp * C + c + 1];
            const double v = hi > lo ? neff_lag_sum<BINARY>(lg.pos, lg.val, lo, hi, w0, N, t) : 0.0;
            part += v / (double)(N - t);
        }
        const double comb = neff_warp_sum(part) / (double)M;
        const double corr = 1.0 - comb / (2.0 * vp);
        if (t > 2) {
            if (corr + prev1 >= 0.0) csum += prev2;
            else break;
        }
        prev2 = prev1; prev1 = corr;
    }
    if (lane == 0) {
        if (out_mean) out_mean[warp] = sum_x / ((double)C * (double)(len - b));     // evaluate_summary
        if (out_psrf) out_psrf[warp] = psrf;
        if (out_neff) out_neff[warp] = (double)M * n / (1.0 + 2.0 * csum);
    }
}
