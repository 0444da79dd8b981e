// ssps_capi.cu -- host side of libssps_b200.so: the C ABI declared in include/ssps.h.
//
// One handle = one CUDA device.  Every entry point validates its arguments, launches the sm_100a
// kernels of ssps_kernels.cuh on the handle's stream and returns 0 or a negative ssps_status
// (message through ssps_last_error()).  There is NO CPU implementation of any scoring or
// sampling step in this file: without a CUDA device every call fails with SSPS_ERR_CUDA.
//
// Host-side arithmetic is limited to bookkeeping: lag row maps (combine_X,
// SSPS/src/dbn_preprocess.jl:70-87), grouping equal prior confidences (so that the device
// normaliser visits each distinct value once), snapshot scheduling and turning the device's
// per-launch toggle log into the reference's change-point lists
// (update_results_storediff, SSPS/src/state_updates.jl:161-227).
#include "../../include/ssps.h"
#include "ssps_kernels.cuh"
#include "ssps_mh.cuh"
#include "ssps_poly.cuh"
#include "ssps_uniform.cuh"
#include "ssps_neff.cuh"

#include <algorithm>
#include <chrono>
#include <map>
#include <mutex>
#include <utility>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <dlfcn.h>
#include <new>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(SSPS_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)
#define CKS(expr)                    \
    do {                             \
        int s_ = (expr);             \
        if (s_ != SSPS_OK) return s_; \
    } while (0)

// Device blocks are recycled through a process-wide cache keyed by (device, size): cudaMalloc of the 100 MB-class buffers
// (tally snapshots, factor arena) costs 2..100 ms at random, inside run_sweeps and between the handles of back-to-back runs
// (measured: bench steps of 83 ms instead of 39).  A released block goes back to the cache; the same shapes come again.
// Ordering: a block is only ever released by the handle that used it, after that handle's stream work on it was enqueued;
// re-use inside the same handle is stream-ordered, and ssps_destroy synchronises the stream before its blocks are released.
// The cache is BOUNDED: ssps_destroy first returns whatever an earlier handle left behind to the driver, so the cache
// never holds more than the last destroyed handle's footprint (plus the small transient buffers released since), and
// never more than SSPS_CACHE_BYTES (default: a quarter of the device's memory).  A host that shares the GPU (the Julia
// process of INTEGRATION.md) can empty it at any time with ssps_release_cached_memory.
thread_local cudaStream_t g_stream = nullptr;     // stream of the handle the calling thread is working on (bind)
thread_local long long g_device_allocs = 0;       // cudaMalloc calls made through BlockCache::take on this thread

class BlockCache {
public:
    static BlockCache& get() { static BlockCache c; return c; }
    cudaError_t take(size_t bytes, void** out) {
        int dev = 0; cudaGetDevice(&dev);
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = free_.find({dev, bytes});
            if (it != free_.end()) {
                auto& v = it->second;
                for (size_t e = v.size(); e-- > 0;) {
                    // usable once the work that was pending on the block at its release has completed
                    if (v[e].ev && cudaEventQuery(v[e].ev) != cudaSuccess) { cudaGetLastError(); continue; }
                    if (v[e].ev) cudaEventDestroy(v[e].ev);
                    *out = v[e].p; v.erase(v.begin() + (long)e); cached_ -= bytes;
                    return cudaSuccess;
                }
            }
        }
        ++g_device_allocs;
        cudaError_t e = cudaMalloc(out, bytes);
        if (e != cudaSuccess) {         // give the cache back to the driver and try once more
            cudaGetLastError();
            trim(dev);
            e = cudaMalloc(out, bytes);
        }
        return e;
    }
    void give(void* p, size_t bytes) {
        int dev = 0; cudaGetDevice(&dev);
        {
            std::lock_guard<std::mutex> g(mu_);
            if (cached_ + bytes <= cap(dev)) {
                Block b; b.p = p; b.ev = nullptr;
                if (cudaEventCreateWithFlags(&b.ev, cudaEventDisableTiming) == cudaSuccess) {
                    if (cudaEventRecord(b.ev, g_stream) != cudaSuccess) { cudaGetLastError(); cudaEventDestroy(b.ev); b.ev = nullptr; cudaStreamSynchronize(g_stream); }
                } else { cudaGetLastError(); b.ev = nullptr; cudaStreamSynchronize(g_stream); }
                free_[{dev, bytes}].push_back(b); cached_ += bytes;
                return;
            }
        }
        cudaFree(p);
    }
    size_t cached_bytes(int dev) {
        std::lock_guard<std::mutex> g(mu_);
        size_t t = 0;
        for (auto& kv : free_) if (kv.first.first == dev) t += kv.first.second * kv.second.size();
        return t;
    }
    void trim(int dev) {
        std::lock_guard<std::mutex> g(mu_);
        for (auto& kv : free_)
            if (kv.first.first == dev) {
                for (auto& b : kv.second) { if (b.ev) cudaEventDestroy(b.ev); cudaFree(b.p); cached_ -= kv.first.second; }
                kv.second.clear();
            }
    }
private:
    struct Block { void* p; cudaEvent_t ev; };
    std::mutex mu_;
    std::map<std::pair<int, size_t>, std::vector<Block>> free_;
    size_t cached_ = 0;
    size_t cap_ = 0;                    // bytes kept at most; 0 = not yet sized
    size_t cap(int) {                   // (called with mu_ held)
        if (cap_ == 0) {
            if (const char* e = getenv("SSPS_CACHE_BYTES")) cap_ = (size_t)strtoull(e, nullptr, 10);
            if (cap_ == 0) {
                size_t free_b = 0, total_b = 0;
                if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); total_b = (size_t)64 << 30; }
                cap_ = total_b / 4;
            }
        }
        return cap_;
    }
};

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void release() { if (p) BlockCache::get().give(p, n * sizeof(T)); p = nullptr; n = 0; }
    cudaError_t alloc(size_t count) {
        if (count == 0) count = 1;
        if (p && n == count) return cudaSuccess;        // same shape again: keep the block (contents are undefined either way)
        release();
        void* q = nullptr;
        cudaError_t e = BlockCache::get().take(count * sizeof(T), &q);
        if (e == cudaSuccess) { p = (T*)q; n = count; } else p = nullptr;
        return e;
    }
    size_t bytes() const { return n * sizeof(T); }
};

// A copy of the per-chain tallies at recorded sample n: a slot of the handle's snapshot slab (one allocation made by
// ssps_init_chains; ssps_run never allocates)
struct Snapshot {
    int64_t n = 0;                 // recorded samples per chain at the time of the copy
    int slot = -1;
    uint32_t* counts = nullptr;    // [chain][child][parent]
    double* lam_sum = nullptr;     // [chain][vertex]
    double* lam_sq = nullptr;
};

struct ParentEvent { int64_t n; int32_t child, parent, value; };
struct LambdaEvent { int64_t n; int32_t child; double value; };

inline int div_up(size_t a, size_t b) { return (int)((a + b - 1) / b); }

}  // namespace

struct ssps_handle {
    ssps_config cfg{};
    int device = 0;
    cudaStream_t stream = nullptr;
    int sm_count = 0;

    // ---- data (K1) ----
    int n = 0, V = 0;
    bool has_data = false;
    DevBuf<double> Xminus, Xplus, Z, Y, G, R, yty, conf, omc, prob_param, mvP, mvLogP, logint, grp_omc, grp_mult;
    DevBuf<int> grp_off, fgrp_off;
    DevBuf<float2> fgrp;
    DevBuf<float> lam_margin;
    double lambda_min = 0.0;
    SspsDev dev{};
    PolyDev poly{};                 // regression_deg != 1: interaction-column path (csrc/ssps_poly.cuh)
    bool is_poly = false;
    bool is_uniform = false;        // cfg.proposal == 1: the deprecated uniform graph proposal (csrc/ssps_uniform.cuh)
    DevBuf<uint32_t> u_par, u_chi;
    DevBuf<double> u_lml, u_lamlog;
    DevBuf<double> poly_arena, poly_warena;
    int poly_mcap_p = 0;            // WARP tier column capacity (0: tier not needed)
    int poly_grid = 0;

    // ---- chains (K3) ----
    int n_chains = 0, nch_pad = 0;
    uint64_t seed = 0;
    uint32_t chain_offset = 0;
    int kcap = 0;                   // scratch capacity in rows (max |S'|)
    int thinning = 0;               // fixed by the first ssps_run / ssps_replay
    int64_t sweeps_done = 0, samples_recorded = 0;
    DevBuf<int> k;
    DevBuf<double> lambda, lam_sum, lam_sq;
    DevBuf<uint32_t> slots, counts;
    ChainState st{};
    int mh_grid = 0;
    size_t mh_smem = 0;
    DevBuf<double> big_d, scr_d, x_d;   // BIG-tier factors (per resident thread), scratch factor (per resident warp), verify scratch
    DevBuf<uint32_t> big_w, scr_w;
    DevBuf<uint16_t> big_r, scr_r, x_t;
    DevBuf<int> task_counter, order, task_start, task_split;
    DevBuf<double> task_est;
    DevBuf<unsigned long long> stats, khist, vcost;
    bool have_cost = false;         // vcost holds the previous launch's per-vertex cycles

    // ---- snapshots for split-PSRF / burn-in (K4) ----
    int64_t snap_grid = 0;          // in recorded samples; 0 = disabled
    DevBuf<unsigned char> snap_slab;    // snap_slots x snap_per bytes
    size_t snap_per = 0;
    int snap_slots = 0;
    std::vector<int> snap_free;     // unused slots
    std::vector<Snapshot> snaps;    // ascending n, all multiples of snap_grid
    long long allocs_in_run = 0;    // device allocations made inside run_sweeps since ssps_init_chains
    DevBuf<double> sums_scratch, psrf_scratch, psrf_partial;
    DevBuf<unsigned long long> max_bits;

    // ---- change-point log (small runs) ----
    int pev_cap = 0, lev_cap = 0;
    DevBuf<uint32_t> pev;
    DevBuf<int> pev_n, lev_n;
    DevBuf<double> lev_val;
    DevBuf<uint16_t> lev_idx;
    std::vector<std::vector<ParentEvent>> cp_parent;   // per chain
    std::vector<std::vector<LambdaEvent>> cp_lambda;   // per chain
    std::vector<uint8_t> cp_present;                   // [chain][child][parent] at the last recorded sample
    std::vector<double> cp_lambda_last;                // [chain][child]

    // ---- multi-GPU (NCCL communicator owned by the library; see "Multi-GPU" below) ----
    void* comm = nullptr;           // ncclComm_t
    int comm_rank = 0, comm_size = 1;
    DevBuf<double> comm_buf;        // psrf sums / posterior sums staged for the all-reduce

    // ---- stats ----
    double kernel_ms = 0.0;
    int64_t kernel_launches = 0, mh_launches = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pool;

    ~ssps_handle() {
        for (auto& e : ev_pool) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    }
};

namespace {

int bind(ssps_handle* h) {
    if (!h) return fail(SSPS_ERR_INVALID, "null handle");
    CK(cudaSetDevice(h->device));
    g_stream = h->stream;
    return SSPS_OK;
}

template <class T>
int upload(ssps_handle* h, DevBuf<T>& dst, const T* src, size_t count) {
    CK(dst.alloc(count));
    if (count) CK(cudaMemcpyAsync(dst.p, src, count * sizeof(T), cudaMemcpyHostToDevice, h->stream));
    return SSPS_OK;
}

template <class T>
int download(ssps_handle* h, T* dst, const T* src, size_t count) {
    if (count) CK(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyDeviceToHost, h->stream));
    return SSPS_OK;
}

int launch_ok(ssps_handle* h, const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(SSPS_ERR_CUDA, "%s launch: %s", what, cudaGetErrorString(e));
    ++h->kernel_launches;
    return SSPS_OK;
}

// Z, Y, G, R, yty, prior tables, move tables from raw Xminus / Xplus already on the device.
int finish_data(ssps_handle* h, const double* prior_host) {
    const int n = h->n, V = h->V;
    const size_t nv = (size_t)n * V, vv = (size_t)V * V;
    cudaStream_t s = h->stream;
    CK(h->Z.alloc(nv)); CK(h->Y.alloc(nv)); CK(h->G.alloc(vv)); CK(h->R.alloc(vv)); CK(h->yty.alloc(V));
    CK(h->conf.alloc(vv)); CK(h->omc.alloc(vv)); CK(h->prob_param.alloc(V));
    CK(h->mvP.alloc((size_t)V * (V + 1) * 3)); CK(h->mvLogP.alloc((size_t)V * (V + 1) * 3)); CK(h->logint.alloc(V + 2));

    standardize_kernel<<<div_up(V, 128), 128, 0, s>>>(h->Xminus.p, n, V, h->Z.p);   // quirk Q4
    CKS(launch_ok(h, "standardize(Xminus)"));
    standardize_kernel<<<div_up(V, 128), 128, 0, s>>>(h->Xplus.p, n, V, h->Y.p);    // SSPS.jl:156
    CKS(launch_ok(h, "standardize(Xplus)"));
    dim3 grid(div_up(V, GRAM_TILE), div_up(V, GRAM_TILE));
    gram_kernel<<<grid, GRAM_TILE * 8, 0, s>>>(h->Z.p, h->Z.p, n, V, V, h->G.p);
    CKS(launch_ok(h, "gram(G)"));
    gram_kernel<<<grid, GRAM_TILE * 8, 0, s>>>(h->Z.p, h->Y.p, n, V, V, h->R.p);
    CKS(launch_ok(h, "gram(R)"));
    yty_kernel<<<div_up(V, 128), 128, 0, s>>>(h->Y.p, n, V, h->yty.p);
    CKS(launch_ok(h, "yty"));

    DevBuf<double> prior_dev;
    CKS(upload(h, prior_dev, prior_host, vv));
    prior_kernel<<<div_up(vv, 256), 256, 0, s>>>(prior_dev.p, vv, h->conf.p, h->omc.p);
    CKS(launch_ok(h, "prior"));
    prob_param_kernel<<<div_up(V, 128), 128, 0, s>>>(h->conf.p, V, h->prob_param.p);
    CKS(launch_ok(h, "prob_param"));
    move_table_kernel<<<div_up((size_t)V * (V + 1) + 2, 256), 256, 0, s>>>(h->prob_param.p, V, h->mvP.p, h->mvLogP.p, h->logint.p);
    CKS(launch_ok(h, "move_table"));

    // distinct (1 - c) values per child, order of first appearance (bookkeeping only: the
    // normaliser sum_j log1p(exp(-lambda (1 - c_ji))) then costs one exp/log1p per distinct value)
    std::vector<int> goff(V + 1, 0);
    std::vector<double> gomc, gmult;
    gomc.reserve(4 * (size_t)V); gmult.reserve(4 * (size_t)V);
    for (int i = 0; i < V; ++i) {
        goff[i] = (int)gomc.size();
        std::unordered_map<uint64_t, int> seen;
        for (int j = 0; j < V; ++j) {
            const double pr = prior_host[(size_t)j + (size_t)V * i];
            const double o = 1.0 - (pr > 0.0 ? pr : 0.0);
            uint64_t bits;
            memcpy(&bits, &o, sizeof bits);
            auto it = seen.find(bits);
            if (it == seen.end()) {
                seen.emplace(bits, (int)gomc.size());
                gomc.push_back(o); gmult.push_back(1.0);
            } else {
                gmult[it->second] += 1.0;
            }
        }
    }
    goff[V] = (int)gomc.size();
    // fp32 screening tables of the lambda move: groups with 1-c != 0 (the others cancel in a difference)
    std::vector<int> foff(V + 1, 0);
    std::vector<float2> fg;
    std::vector<float> fmarg(V, 0.f);
    for (int i = 0; i < V; ++i) {
        foff[i] = (int)fg.size();
        double msum = 0.0;
        for (int g = goff[i]; g < goff[i + 1]; ++g) {
            if (gomc[g] == 0.0) continue;
            fg.push_back(make_float2((float)(-gomc[g] * 1.4426950408889634), (float)(gmult[g] * 0.6931471805599453)));
            msum += gmult[g];
        }
        fmarg[i] = (float)(1e-5 + 4e-6 * msum);
    }
    foff[V] = (int)fg.size();
    CKS(upload(h, h->fgrp_off, foff.data(), foff.size()));
    CKS(upload(h, h->fgrp, fg.data(), fg.size()));
    CKS(upload(h, h->lam_margin, fmarg.data(), fmarg.size()));
    CKS(upload(h, h->grp_off, goff.data(), goff.size()));
    CKS(upload(h, h->grp_omc, gomc.data(), gomc.size()));
    CKS(upload(h, h->grp_mult, gmult.data(), gmult.size()));
    CK(cudaStreamSynchronize(s));   // host vectors go out of scope

    // SSPS.jl:169  lambda_min = log(max(V / large_indeg - 1, exp(0.5)))
    h->lambda_min = log(fmax((double)V / h->cfg.large_indeg - 1.0, exp(0.5)));
    if (!(h->cfg.lambda_max > h->lambda_min))      // Gen.uniform(lambda_min, lambda_max) would have an empty support (dbn_models.jl:66-70)
        return fail(SSPS_ERR_INVALID, "lambda_max=%g does not exceed lambda_min=log(max(V/large_indeg - 1, e^0.5))=%g (V=%d, large_indeg=%g)",
                    h->cfg.lambda_max, h->lambda_min, V, h->cfg.large_indeg);
    SspsDev& d = h->dev;
    d.n = n; d.V = V; d.dn = (double)n; d.shrink = (double)n / ((double)n + 1.0); d.log1pn = log(1.0 + (double)n);
    d.lambda_min = h->lambda_min; d.lambda_max = h->cfg.lambda_max; d.sigma = h->cfg.lambda_prop_std;
    d.G = h->G.p; d.R = h->R.p; d.yty = h->yty.p; d.omc = h->omc.p; d.conf = h->conf.p;
    d.grp_off = h->grp_off.p; d.grp_omc = h->grp_omc.p; d.grp_mult = h->grp_mult.p;
    d.mvP = h->mvP.p; d.mvLogP = h->mvLogP.p; d.logint = h->logint.p;
    d.fgrp_off = h->fgrp_off.p; d.fgrp = h->fgrp.p; d.lam_margin = h->lam_margin.p;
    if (h->cfg.regression_deg <= 0) {
        // quirk Q2: construct_B loops over 1:min(deg, k) = nothing, the design is n x 0 for EVERY parent set
        // (dbn_distributions.jl:179-194), so lml = -1/2 n log(y'y): no projection (R := 0 makes every w and q vanish) and
        // no width penalty (log(1 + n) := 0).  The ordinary cached-factor kernel then runs the prior-only chain with no
        // limit on |S| (round 1 sent this through the 12-parent interaction path, which the reference's own default
        // flags overflowed at once).
        CK(cudaMemsetAsync(h->R.p, 0, h->R.bytes(), s));
        d.log1pn = 0.0;
    }
    h->poly.Xminus = h->Xminus.p; h->poly.Y = h->Y.p; h->poly.deg = h->cfg.regression_deg;
    h->poly.force_warp = getenv("SSPS_POLY_FORCE_WARP") != nullptr;
    h->has_data = true;
    // any chains belong to the previous data set
    h->n_chains = 0;
    h->snaps.clear(); h->snap_free.clear();
    return SSPS_OK;
}

int check_dims(int n, int V) {
    if (n < 2) return fail(SSPS_ERR_INVALID, "need at least 2 lagged rows, got n=%d", n);
    if (V < 1 || V > 65535) return fail(SSPS_ERR_INVALID, "V=%d outside 1..65535 (parent ids are 16-bit on the device)", V);
    return SSPS_OK;
}

}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

const char* ssps_last_error(void) { return g_err.c_str(); }
int ssps_abi_version(void) { return SSPS_ABI_VERSION; }

int ssps_create(const ssps_config* cfg, ssps_handle** out) {
    if (!cfg || !out) return fail(SSPS_ERR_INVALID, "ssps_create: null argument");
    if (cfg->struct_size != (int32_t)sizeof(ssps_config))
        return fail(SSPS_ERR_INVALID, "ssps_config.struct_size=%d, library expects %d", cfg->struct_size, (int)sizeof(ssps_config));
    if (!(cfg->large_indeg > 0.0) || !(cfg->lambda_prop_std > 0.0))
        return fail(SSPS_ERR_INVALID, "large_indeg and lambda_prop_std must be positive");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 1) return fail(SSPS_ERR_CUDA, "no CUDA device (libssps_b200 has no CPU path)");
    int dev = cfg->device;
    if (dev < 0) CK(cudaGetDevice(&dev));
    if (dev >= ndev) return fail(SSPS_ERR_INVALID, "device %d out of range (%d visible)", dev, ndev);
    ssps_handle* h = new (std::nothrow) ssps_handle();
    if (!h) return fail(SSPS_ERR_NOMEM, "out of host memory");
    h->cfg = *cfg;
    h->is_poly = cfg->regression_deg >= 2;
    h->is_uniform = cfg->proposal == 1;
    if (cfg->proposal != 0 && cfg->proposal != 1) { delete h; return fail(SSPS_ERR_INVALID, "proposal=%d: 0 (sparse) or 1 (uniform)", cfg->proposal); }
    if (h->is_uniform && h->is_poly) { delete h; return fail(SSPS_ERR_UNSUPPORTED, "proposal uniform is implemented for regression_deg <= 1 only"); }
    h->device = dev;
    cudaError_t e = cudaSetDevice(dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) { delete h; return fail(SSPS_ERR_CUDA, "device %d: %s", dev, cudaGetErrorString(e)); }
    double lmin = log(fmax(1.0, exp(0.5)));
    if (!(cfg->lambda_max > lmin)) { delete h; return fail(SSPS_ERR_INVALID, "lambda_max=%g must exceed 0.5", cfg->lambda_max); }
    *out = h;
    return SSPS_OK;
}

int ssps_comm_destroy(ssps_handle* h);

void ssps_destroy(ssps_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->comm) ssps_comm_destroy(h);
    BlockCache::get().trim(h->device);      // what earlier handles left behind goes back to the driver: the cache holds ONE footprint
    g_stream = h->stream;
    delete h;
    g_stream = nullptr;
}

int ssps_set_stream(ssps_handle* h, void* cuda_stream) {
    CKS(bind(h));
    CK(cudaStreamSynchronize(h->stream));
    h->stream = (cudaStream_t)cuda_stream;
    g_stream = h->stream;
    return SSPS_OK;
}

int ssps_set_data(ssps_handle* h, const double* Xminus, const double* Xplus, int32_t n, int32_t V, const double* prior) {
    CKS(bind(h));
    if (!Xminus || !Xplus || !prior) return fail(SSPS_ERR_INVALID, "ssps_set_data: null argument");
    CKS(check_dims(n, V));
    h->n = n; h->V = V; h->has_data = false;
    CKS(upload(h, h->Xminus, Xminus, (size_t)n * V));
    CKS(upload(h, h->Xplus, Xplus, (size_t)n * V));
    return finish_data(h, prior);
}

int ssps_set_timeseries(ssps_handle* h, const double* X, const int32_t* series_len, int32_t n_series, int32_t V,
                        const double* prior) {
    CKS(bind(h));
    if (!X || !series_len || !prior || n_series < 1) return fail(SSPS_ERR_INVALID, "ssps_set_timeseries: bad argument");
    // combine_X (dbn_preprocess.jl:70-87): Xminus = rows 1..T-1, Xplus = rows 2..T of every series
    std::vector<int> row_map;
    int total = 0;
    for (int m = 0; m < n_series; ++m) {
        if (series_len[m] < 1) return fail(SSPS_ERR_INVALID, "series %d has length %d", m, series_len[m]);
        for (int t = 0; t + 1 < series_len[m]; ++t) row_map.push_back(total + t);
        total += series_len[m];
    }
    const int n = (int)row_map.size();
    CKS(check_dims(n, V));
    h->n = n; h->V = V; h->has_data = false;
    DevBuf<double> Xd;
    DevBuf<int> map_d;
    CKS(upload(h, Xd, X, (size_t)total * V));
    CKS(upload(h, map_d, row_map.data(), row_map.size()));
    CK(h->Xminus.alloc((size_t)n * V)); CK(h->Xplus.alloc((size_t)n * V));
    lag_gather_kernel<<<div_up((size_t)n * V, 256), 256, 0, h->stream>>>(Xd.p, total, V, map_d.p, n, h->Xminus.p, h->Xplus.p);
    CKS(launch_ok(h, "lag_gather"));
    CK(cudaStreamSynchronize(h->stream));
    return finish_data(h, prior);
}

int ssps_get_design(ssps_handle* h, double* Z, double* Y) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "no data set");
    const size_t nv = (size_t)h->n * h->V;
    if (Z) CKS(download(h, Z, h->Z.p, nv));
    if (Y) CKS(download(h, Y, h->Y.p, nv));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

int ssps_get_gram(ssps_handle* h, double* G, double* R, double* yty) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "no data set");
    const size_t vv = (size_t)h->V * h->V;
    if (G) CKS(download(h, G, h->G.p, vv));
    if (R) CKS(download(h, R, h->R.p, vv));
    if (yty) CKS(download(h, yty, h->yty.p, (size_t)h->V));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

int ssps_get_derived(ssps_handle* h, double* lambda_min, double* prob_param) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "no data set");
    if (lambda_min) *lambda_min = h->lambda_min;
    if (prob_param) CKS(download(h, prob_param, h->prob_param.p, (size_t)h->V));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

int ssps_get_dims(ssps_handle* h, int32_t* n, int32_t* V, int32_t* n_chains) {
    if (!h) return fail(SSPS_ERR_INVALID, "null handle");
    if (n) *n = h->n;
    if (V) *V = h->V;
    if (n_chains) *n_chains = h->n_chains;
    return SSPS_OK;
}

// ------------------------------------------------------------------------------------------------
// K2
// ------------------------------------------------------------------------------------------------
int ssps_score(ssps_handle* h, int32_t n_sets, const int32_t* child, const int32_t* set_offsets,
               const int32_t* set_members, const double* lambda, double* lml_out, double* logprior_out,
               double* min_pivot_out) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "ssps_score before ssps_set_data");
    if (n_sets < 0 || (n_sets > 0 && (!child || !set_offsets))) return fail(SSPS_ERR_INVALID, "ssps_score: bad argument");
    if (logprior_out && !lambda) return fail(SSPS_ERR_INVALID, "ssps_score: logprior_out needs lambda");
    if (n_sets == 0) return SSPS_OK;
    const int V = h->V;
    int kmax = 0;
    for (int s = 0; s < n_sets; ++s) {
        const int k = set_offsets[s + 1] - set_offsets[s];
        if (k < 0 || k > V) return fail(SSPS_ERR_INVALID, "set %d has %d members", s, k);
        if (child[s] < 0 || child[s] >= V) return fail(SSPS_ERR_INVALID, "set %d: child %d out of range", s, child[s]);
        if (k > 0 && !set_members) return fail(SSPS_ERR_INVALID, "ssps_score: null set_members");
        for (int a = 0; a < k; ++a) {
            const int j = set_members[set_offsets[s] + a];
            if (j < 0 || j >= V) return fail(SSPS_ERR_INVALID, "set %d: parent %d out of range", s, j);
            if (a > 0 && j <= set_members[set_offsets[s] + a - 1])
                return fail(SSPS_ERR_INVALID, "set %d: members must be sorted ascending and unique", s);
        }
        kmax = std::max(kmax, k);
    }
    const int total = set_offsets[n_sets];
    const int kcap = std::max(kmax, 1);
    const int block = 128, grid = div_up((size_t)n_sets, block);
    const size_t nthreads = (size_t)grid * block;
    if (h->is_poly) {       // interaction columns: every set from scratch (csrc/ssps_poly.cuh)
        int m_warp = 0;                                 // widest design the WARP tier will see
        for (int s = 0; s < n_sets; ++s) {
            const int k = set_offsets[s + 1] - set_offsets[s];
            int m = 0;
            const int cls = poly_class(k, h->poly, h->n, m);
            if (cls == PC_REFUSE || (cls != PC_SAT && cls != PC_EMPTY && k > SSPS_POLY_KLIST))
                return fail(SSPS_ERR_UNSUPPORTED, "regression_deg=%d: set %d has %d parents and a design of fewer than n - 1 = %d columns; "
                            "the interaction-column scorer handles at most %d parents for such sets", h->cfg.regression_deg, s, k, h->n - 1, SSPS_POLY_KW);
            if (cls == PC_WARP) m_warp = std::max(m_warp, m);
        }
        const int mcap_p = (m_warp + 31) & ~31;
        const size_t wdoubles = poly_warp_scratch_doubles(h->n, mcap_p);
        int pgrid = std::min(grid, 4 * h->sm_count);
        if (mcap_p) pgrid = std::max(1, std::min(pgrid, (int)((size_t)2 << 30) / (int)std::min<size_t>(4 * wdoubles * 8, (size_t)1 << 30)));   // <= 2 GB of WARP-tier scratch
        const size_t pthreads = (size_t)pgrid * block;
        DevBuf<int> d_child, d_off, d_mem;
        DevBuf<double> d_lam, d_lml, d_lp, arena, warena;
        CKS(upload(h, d_child, child, (size_t)n_sets));
        CKS(upload(h, d_off, set_offsets, (size_t)n_sets + 1));
        CKS(upload(h, d_mem, set_members, (size_t)std::max(total, 1)));
        if (lambda) CKS(upload(h, d_lam, lambda, (size_t)n_sets));
        CK(d_lml.alloc(n_sets)); CK(d_lp.alloc(n_sets));
        CK(arena.alloc(pthreads * (size_t)SSPS_POLY_SCRATCH_DOUBLES));
        CK(warena.alloc(pthreads / 32 * wdoubles));
        poly_score_kernel<<<pgrid, block, 0, h->stream>>>(h->dev, h->poly, n_sets, d_child.p, d_off.p, d_mem.p, lambda ? d_lam.p : nullptr,
                                                          arena.p, warena.p, mcap_p, d_lml.p, logprior_out ? d_lp.p : nullptr);
        CKS(launch_ok(h, "poly_score_kernel"));
        if (lml_out) CKS(download(h, lml_out, d_lml.p, (size_t)n_sets));
        if (logprior_out) CKS(download(h, logprior_out, d_lp.p, (size_t)n_sets));
        CK(cudaStreamSynchronize(h->stream));
        if (min_pivot_out) for (int e = 0; e < n_sets; ++e) min_pivot_out[e] = std::nan("");
        return SSPS_OK;
    }
    DevBuf<int> d_child, d_off, d_mem;
    DevBuf<double> d_lam, d_lml, d_lp, d_minp, arena_d;
    DevBuf<uint16_t> arena_t;
    CKS(upload(h, d_child, child, (size_t)n_sets));
    CKS(upload(h, d_off, set_offsets, (size_t)n_sets + 1));
    CKS(upload(h, d_mem, set_members, (size_t)total));
    if (lambda) CKS(upload(h, d_lam, lambda, (size_t)n_sets));
    CK(d_lml.alloc(n_sets)); CK(d_lp.alloc(n_sets)); CK(d_minp.alloc(n_sets));
    CK(arena_d.alloc(nthreads * (size_t)(SSPS_TRI(kcap) + kcap)));
    CK(arena_t.alloc(nthreads * (size_t)kcap));
    score_kernel<<<grid, block, 0, h->stream>>>(h->dev, n_sets, d_child.p, d_off.p, d_mem.p, lambda ? d_lam.p : nullptr,
                                                arena_d.p, arena_t.p, kcap, d_lml.p, logprior_out ? d_lp.p : nullptr, d_minp.p);
    CKS(launch_ok(h, "score_kernel"));
    if (lml_out) CKS(download(h, lml_out, d_lml.p, (size_t)n_sets));
    if (logprior_out) CKS(download(h, logprior_out, d_lp.p, (size_t)n_sets));
    if (min_pivot_out) CKS(download(h, min_pivot_out, d_minp.p, (size_t)n_sets));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

}  // extern "C"

// ================================================================================================
// chains: allocation, launches, snapshots, change-point log
// ================================================================================================
namespace {

int reserve_snapshots(ssps_handle* h);

// K4 reductions over chains: enough (edge, chain-chunk) threads to fill the GPU; one chunk when V^2 alone does
void chain_chunks(const ssps_handle* h, size_t vv, int& chunks, int& chunk_len) {
    const size_t want = (size_t)h->sm_count * 2048;
    chunks = (int)std::max<size_t>(1, std::min<size_t>((want + vv - 1) / vv, (size_t)std::max(1, h->n_chains / 32)));
    chunks = std::min(chunks, 64);
    if (const char* e = getenv("SSPS_K4_CHUNKS")) chunks = std::max(1, std::min(atoi(e), h->n_chains));     // test hook
    chunk_len = (h->n_chains + chunks - 1) / chunks;
    chunks = (h->n_chains + chunk_len - 1) / chunk_len;
}

int alloc_chains(ssps_handle* h, int n_chains) {
    const int V = h->V;
    const size_t vv = (size_t)V * V;
    h->n_chains = 0;
    h->snaps.clear(); h->snap_free.clear();
    h->nch_pad = ((n_chains + 31) / 32) * 32;
    // bound on |S|: the reference has none; the default is everything the factor arena supports (a lane owns up to SSPS_KU
    // columns of 32 rows), i.e. V itself up to V = 128.  Proposals beyond the bound are rejected (ssps.h: max_indegree).
    int kcap = h->cfg.max_indegree > 0 ? h->cfg.max_indegree : std::min(V, 32 * SSPS_KU);
    h->kcap = std::min(kcap, V);
    if (h->kcap > 32 * SSPS_KU) return fail(SSPS_ERR_UNSUPPORTED, "max_indegree=%d: the per-thread factor arena supports at most %d", h->kcap, 32 * SSPS_KU);
    CK(h->k.alloc((size_t)V * h->nch_pad)); CK(h->lambda.alloc((size_t)V * h->nch_pad));
    CK(h->slots.alloc(vv * h->nch_pad));
    CK(h->counts.alloc(vv * n_chains));
    CK(h->lam_sum.alloc((size_t)V * n_chains)); CK(h->lam_sq.alloc((size_t)V * n_chains));
    cudaStream_t s = h->stream;
    CK(cudaMemsetAsync(h->k.p, 0, h->k.bytes(), s));
    CK(cudaMemsetAsync(h->lambda.p, 0, h->lambda.bytes(), s));
    CK(cudaMemsetAsync(h->slots.p, 0, h->slots.bytes(), s));
    CK(cudaMemsetAsync(h->counts.p, 0, h->counts.bytes(), s));
    CK(cudaMemsetAsync(h->lam_sum.p, 0, h->lam_sum.bytes(), s));
    CK(cudaMemsetAsync(h->lam_sq.p, 0, h->lam_sq.bytes(), s));
    ChainState& st = h->st;
    st.n_chains = n_chains; st.nch_pad = h->nch_pad;
    st.k = h->k.p; st.lambda = h->lambda.p; st.slots = h->slots.p; st.counts = h->counts.p;
    st.lam_sum = h->lam_sum.p; st.lam_sq = h->lam_sq.p;

    // persistent MH grid: as many CTAs as stay resident, never more than there are warp-tasks
    h->mh_smem = SSPS_SMEM_BYTES;
    int occ = 1;
    {
        // Shared memory holds only the SMALL-tier factors; what the carve-out leaves is L1 for the G / R / move
        // tables.  First find the register-limited occupancy with the largest carve-out, then shrink the
        // carve-out to what that many CTAs need (SSPS_CARVEOUT=<percent> overrides, for experiments).
        // Once per device and process: cudaGetDeviceProperties and the attribute calls cost milliseconds (up to 100 ms
        // measured on a fresh context), which a host that creates a handle per data set would pay on every one.
        static std::mutex cfg_mu;
        static std::map<int, int> cfg_occ;
        std::lock_guard<std::mutex> g(cfg_mu);
        auto it = cfg_occ.find(h->device);
        if (it == cfg_occ.end()) {
            int o_min = 1 << 30;
            const void* fns[4] = {(const void*)mh_sweeps_kernel<PhiloxSource, false>, (const void*)mh_sweeps_kernel<PhiloxSource, true>,
                                  (const void*)mh_sweeps_kernel<ReplaySource, false>, (const void*)mh_sweeps_kernel<ReplaySource, true>};
            int smem_per_sm = 0;
            CK(cudaDeviceGetAttribute(&smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, h->device));
            for (const void* f : fns) {
                int o = 0;
                CK(cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
                CK(cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->mh_smem));
                CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, f, SSPS_BLOCK, h->mh_smem));
                o_min = std::min(o_min, o);
            }
            o_min = std::max(1, o_min);
            int pct = (int)std::ceil(100.0 * (double)o_min * (double)(h->mh_smem + 1024) / (double)smem_per_sm) + 2;
            if (const char* e = getenv("SSPS_CARVEOUT")) pct = atoi(e);
            pct = std::max(0, std::min(100, pct));
            for (const void* f : fns) CK(cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            it = cfg_occ.emplace(h->device, o_min).first;
        }
        occ = it->second;
    }
    occ = std::max(1, occ);
    if (const char* e = getenv("SSPS_MAX_OCC")) occ = std::max(1, std::min(occ, atoi(e)));      // experiments: resident CTAs per SM
    const int n_tasks = V * (h->nch_pad / 32);
    const int warps_per_cta = SSPS_BLOCK / 32;
    // room for the scheduler to spread a slab over up to SSPS_MAX_SPLIT warps
    h->mh_grid = std::max(1, std::min(h->sm_count * occ, div_up((size_t)n_tasks * SSPS_MAX_SPLIT, warps_per_cta)));
    const size_t nthreads = (size_t)h->mh_grid * SSPS_BLOCK;
    const size_t kc = (size_t)h->kcap;
    CK(h->big_d.alloc(nthreads * SSPS_ROWLEN(kc)));
    CK(h->scr_d.alloc(nthreads / 32 * SSPS_ROWLEN(kc))); CK(h->scr_w.alloc(nthreads / 32 * kc)); CK(h->scr_r.alloc(nthreads / 32 * kc));
    CK(h->big_w.alloc(nthreads * kc));
    CK(h->big_r.alloc(nthreads * kc));
    const size_t xthreads = h->is_uniform ? std::max(nthreads, (size_t)h->nch_pad) : nthreads;
    CK(h->x_d.alloc(h->is_uniform ? (size_t)h->nch_pad * (SSPS_TRI(kc) + kc) : (h->cfg.verify ? nthreads * (SSPS_TRI(kc) + kc) : 1)));
    CK(h->x_t.alloc(xthreads * 2 * kc));
    if (h->is_uniform) {
        const size_t W = ((size_t)V + 31) / 32;
        CK(h->u_par.alloc((size_t)V * W * h->nch_pad)); CK(h->u_chi.alloc((size_t)V * W * h->nch_pad));
        CK(h->u_lml.alloc((size_t)V * h->nch_pad)); CK(h->u_lamlog.alloc((size_t)V * h->nch_pad));
    }
    CK(h->task_counter.alloc(1));
    CK(h->order.alloc(V)); CK(h->vcost.alloc(V));
    CK(h->task_start.alloc((size_t)V + 1)); CK(h->task_split.alloc(V)); CK(h->task_est.alloc(2 * (size_t)V));
    {
        std::vector<int> ones((size_t)V, 1);
        CK(cudaMemcpyAsync(h->task_split.p, ones.data(), sizeof(int) * (size_t)V, cudaMemcpyHostToDevice, s));
        CK(cudaStreamSynchronize(s));
    }
    CK(h->stats.alloc(ST_N)); CK(h->khist.alloc(64));
    CK(cudaMemsetAsync(h->stats.p, 0, h->stats.bytes(), s));
    CK(cudaMemsetAsync(h->khist.p, 0, h->khist.bytes(), s));
    CK(cudaMemsetAsync(h->vcost.p, 0, h->vcost.bytes(), s));
    h->have_cost = false;

    if (h->is_poly) {
        int occ_p = 1;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_p, (const void*)mh_poly_kernel<PhiloxSource>, 128, 0));
        h->poly_grid = std::max(1, std::min(h->sm_count * std::max(1, occ_p), div_up((size_t)n_tasks, 4)));
        // WARP tier: scratch for the widest design a set within max_indegree can have below saturation (csrc/ssps_poly.cuh);
        // the grid shrinks if that scratch would take more than a quarter of the free memory
        const int m_scored = poly_max_scored_width(std::min(h->kcap, SSPS_POLY_KLIST), h->cfg.regression_deg, h->n);
        h->poly_mcap_p = (m_scored > SSPS_POLY_MMAX || (h->poly.force_warp && m_scored > 0)) ? ((m_scored + 31) & ~31) : 0;
        const size_t wdoubles = poly_warp_scratch_doubles(h->n, h->poly_mcap_p);
        if (h->poly_mcap_p) {
            size_t free_b = 0, total_b = 0;
            CK(cudaMemGetInfo(&free_b, &total_b));
            const size_t per_cta = 4 * wdoubles * sizeof(double);
            h->poly_grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)h->poly_grid, (free_b / 4) / per_cta));
        }
        CK(h->poly_arena.alloc((size_t)h->poly_grid * 128 * SSPS_POLY_SCRATCH_DOUBLES));
        CK(h->poly_warena.alloc((size_t)h->poly_grid * 4 * wdoubles));
        h->kcap = std::min(h->kcap, SSPS_POLY_KLIST);
    }

    h->thinning = 0; h->sweeps_done = 0; h->samples_recorded = 0;
    h->kernel_ms = 0.0; h->mh_launches = 0;
    h->snap_grid = h->cfg.psrf_grid == 0 ? 250 : (h->cfg.psrf_grid < 0 ? 0 : h->cfg.psrf_grid);

    h->cp_parent.clear(); h->cp_lambda.clear(); h->cp_present.clear(); h->cp_lambda_last.clear();
    h->pev_cap = h->lev_cap = 0;
    h->n_chains = n_chains;
    CKS(reserve_snapshots(h));
    h->allocs_in_run = 0;
    return SSPS_OK;
}

int fetch_state_host(ssps_handle* h, int chain, std::vector<int>& indeg, std::vector<int>& members_padded,
                     std::vector<double>& lambda) {
    const int V = h->V;
    DevBuf<int> d_indeg, d_mem;
    DevBuf<double> d_lam;
    CK(d_indeg.alloc(V)); CK(d_mem.alloc((size_t)V * V)); CK(d_lam.alloc(V));
    export_chain_kernel<<<div_up(V, 128), 128, 0, h->stream>>>(h->dev, h->st, chain, d_indeg.p, d_mem.p, d_lam.p);
    CKS(launch_ok(h, "export_chain"));
    indeg.resize(V); members_padded.resize((size_t)V * V); lambda.resize(V);
    CKS(download(h, indeg.data(), d_indeg.p, (size_t)V));
    CKS(download(h, members_padded.data(), d_mem.p, (size_t)V * V));
    CKS(download(h, lambda.data(), d_lam.p, (size_t)V));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

// index 0 of the change-point log = the initial state (state_updates.jl:166-186)
int changepoints_begin(ssps_handle* h) {
    const int V = h->V, C = h->n_chains;
    h->cp_parent.assign(C, {}); h->cp_lambda.assign(C, {});
    h->cp_present.assign((size_t)C * V * V, 0);
    h->cp_lambda_last.assign((size_t)C * V, 0.0);
    std::vector<int> indeg, mem;
    std::vector<double> lam;
    for (int c = 0; c < C; ++c) {
        CKS(fetch_state_host(h, c, indeg, mem, lam));
        for (int i = 0; i < V; ++i) {
            h->cp_lambda[c].push_back({0, i, lam[i]});
            h->cp_lambda_last[(size_t)c * V + i] = lam[i];
            for (int a = 0; a < indeg[i]; ++a) {
                const int j = mem[(size_t)i * V + a];
                h->cp_parent[c].push_back({0, i, j, 1});
                h->cp_present[((size_t)c * V + i) * V + j] = 1;
            }
        }
    }
    return SSPS_OK;
}

int changepoints_alloc(ssps_handle* h, int max_rec_local, int max_sweeps_local) {
    const size_t nmc = (size_t)h->n_chains * h->V;
    const int pcap = 2 * max_sweeps_local, lcap = max_rec_local;
    if (pcap > h->pev_cap) { CK(h->pev.alloc(nmc * pcap)); h->pev_cap = pcap; }
    if (lcap > h->lev_cap) { CK(h->lev_val.alloc(nmc * lcap)); CK(h->lev_idx.alloc(nmc * lcap)); h->lev_cap = lcap; }
    if (h->pev_n.n < nmc) { CK(h->pev_n.alloc(nmc)); CK(h->lev_n.alloc(nmc)); }
    return SSPS_OK;
}

// Turn one launch's toggle log into change points relative to the previously RECORDED state.
int changepoints_drain(ssps_handle* h, int64_t rec_base) {
    const int V = h->V, C = h->n_chains;
    const size_t nmc = (size_t)C * V;
    std::vector<int> pn(nmc), ln(nmc);
    std::vector<uint32_t> pev(nmc * h->pev_cap);
    std::vector<double> lval(nmc * h->lev_cap);
    std::vector<uint16_t> lidx(nmc * h->lev_cap);
    CKS(download(h, pn.data(), h->pev_n.p, nmc));
    CKS(download(h, ln.data(), h->lev_n.p, nmc));
    CKS(download(h, pev.data(), h->pev.p, pev.size()));
    CKS(download(h, lval.data(), h->lev_val.p, lval.size()));
    CKS(download(h, lidx.data(), h->lev_idx.p, lidx.size()));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<std::pair<int, int>> touched;   // (rel sample, parent)
    for (int c = 0; c < C; ++c)
        for (int i = 0; i < V; ++i) {
            const size_t mc = (size_t)c * V + i;
            if (pn[mc] > h->pev_cap || ln[mc] > h->lev_cap) return fail(SSPS_ERR_NOMEM, "change-point log overflow");
            uint8_t* present = &h->cp_present[mc * V];
            // toggles are in time order; those sharing a recorded-sample index collapse to their net effect
            const uint32_t* ev = &pev[mc * h->pev_cap];
            int a = 0;
            while (a < pn[mc]) {
                const int rel = (int)((ev[a] >> 16) & 0x7fffu);
                int b = a;
                touched.clear();
                std::vector<uint8_t> now;   // small: at most 2 * thinning toggles
                for (; b < pn[mc] && (int)((ev[b] >> 16) & 0x7fffu) == rel; ++b) {
                    const int j = (int)(ev[b] & 0xffffu);
                    const uint8_t val = (uint8_t)(ev[b] >> 31);
                    bool found = false;
                    for (size_t t = 0; t < touched.size(); ++t)
                        if (touched[t].second == j) { now[t] = val; found = true; break; }
                    if (!found) { touched.emplace_back(rel, j); now.push_back(val); }
                }
                std::vector<size_t> order(touched.size());
                for (size_t t = 0; t < order.size(); ++t) order[t] = t;
                std::sort(order.begin(), order.end(), [&](size_t x, size_t y) { return touched[x].second < touched[y].second; });
                for (size_t t : order) {
                    const int j = touched[t].second;
                    if (present[j] != now[t]) {
                        h->cp_parent[c].push_back({rec_base + rel, i, j, (int32_t)now[t]});
                        present[j] = now[t];
                    }
                }
                a = b;
            }
            for (int e = 0; e < ln[mc]; ++e) {
                const double v = lval[mc * h->lev_cap + e];
                h->cp_lambda[c].push_back({rec_base + (int64_t)lidx[mc * h->lev_cap + e], i, v});
                h->cp_lambda_last[mc] = v;
            }
        }
    return SSPS_OK;
}

// The snapshot pool: ONE slab sized at ssps_init_chains, sub-allocated in fixed slots.  (Round 1 kept a copy of the
// tallies every snap_grid samples forever and refilled its pool with cudaMalloc -- 2..180 ms at random between the
// launches of ssps_run, 22-39 % of a driver-timed bench step.)
int reserve_snapshots(ssps_handle* h) {
    h->snaps.clear(); h->snap_free.clear();
    h->snap_slots = 0; h->snap_per = 0;
    if (h->snap_grid <= 0 || h->n_chains <= 0) return SSPS_OK;
    const size_t vv = (size_t)h->V * h->V, nc = (size_t)h->n_chains;
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t per = up(vv * nc * sizeof(uint32_t)) + 2 * up((size_t)h->V * nc * sizeof(double));
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    free_b += BlockCache::get().cached_bytes(h->device);
    int want = h->cfg.snapshot_slots > 0 ? h->cfg.snapshot_slots : 32;
    if (const char* e = getenv("SSPS_SNAP_SLOTS")) want = atoi(e);      // experiments
    const size_t room = (free_b / 4) / per;
    int slots = (int)std::min<size_t>((size_t)std::max(want, 4), std::max<size_t>(room, 4));
    CK(h->snap_slab.alloc((size_t)slots * per));
    h->snap_per = per; h->snap_slots = slots;
    for (int e = slots; e-- > 0;) h->snap_free.push_back(e);
    return SSPS_OK;
}

// Which snapshots can still be read?  A PSRF / posterior query at length L with burn-in fraction beta reads the
// tallies at b = floor(beta L) and at the split point b + (L - b) / 2 (evaluate_convergence,
// postprocess_samples.jl:155-180); L is the current length or, when that is not snapshot-aligned, the last aligned
// one (at most 4 grid steps back).  So with beta >= min_burnin a snapshot older than min_burnin (n - 4 g) is dead.
// When the pool is full: recycle the dead ones; if none, drop every other snapshot and double the grid.
int take_snapshot(ssps_handle* h) {
    if (h->snap_grid == 0 || h->snap_slots == 0 || h->samples_recorded % h->snap_grid != 0) return SSPS_OK;
    const double beta = h->cfg.min_burnin > 0.0 ? std::min(h->cfg.min_burnin, 0.95) : 0.5;
    while (h->snap_free.empty()) {
        const int64_t dead_below = (int64_t)floor(beta * (double)std::max<int64_t>(h->samples_recorded - 4 * h->snap_grid, 0));
        std::vector<Snapshot> kept;
        for (auto& s : h->snaps) {
            if (s.n < dead_below) h->snap_free.push_back(s.slot);      // same stream: safe to overwrite in order
            else kept.push_back(s);
        }
        if (h->snap_free.empty()) {
            h->snap_grid *= 2;
            kept.clear();
            for (auto& s : h->snaps) {
                if (s.n % h->snap_grid == 0) kept.push_back(s);
                else h->snap_free.push_back(s.slot);
            }
            if (getenv("SSPS_TRACE"))
                fprintf(stderr, "[ssps trace] snapshot pool full at n = %lld: grid doubled to %lld\n", (long long)h->samples_recorded, (long long)h->snap_grid);
        }
        h->snaps.swap(kept);
        if (h->samples_recorded % h->snap_grid != 0) return SSPS_OK;       // no longer a grid point
    }
    const size_t vv = (size_t)h->V * h->V, nc = (size_t)h->n_chains;
    auto up = [](size_t x) { return (x + 255) / 256 * 256; };
    Snapshot sn;
    sn.slot = h->snap_free.back(); h->snap_free.pop_back();
    unsigned char* base = h->snap_slab.p + (size_t)sn.slot * h->snap_per;
    sn.counts = (uint32_t*)base;
    sn.lam_sum = (double*)(base + up(vv * nc * sizeof(uint32_t)));
    sn.lam_sq = (double*)(base + up(vv * nc * sizeof(uint32_t)) + up((size_t)h->V * nc * sizeof(double)));
    sn.n = h->samples_recorded;
    CK(cudaMemcpyAsync(sn.counts, h->counts.p, vv * nc * sizeof(uint32_t), cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaMemcpyAsync(sn.lam_sum, h->lam_sum.p, (size_t)h->V * nc * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaMemcpyAsync(sn.lam_sq, h->lam_sq.p, (size_t)h->V * nc * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    h->snaps.push_back(sn);
    return SSPS_OK;
}

const Snapshot* find_snapshot(const ssps_handle* h, int64_t n) {
    for (const auto& s : h->snaps)
        if (s.n == n) return &s;
    return nullptr;
}

// A run = a sequence of MH launches on the handle's stream.  Split into begin / one launch / end so that ssps_run_multi can
// keep the GPUs of several handles busy from ONE host thread (every piece only enqueues work; run_end synchronises).
struct RunCtx {
    int64_t n_sweeps = 0, done = 0;
    int thinning = 1;
    bool log_cp = false, replay = false;
    int64_t max_rec = 0, dflt_rec = 0;
    DevBuf<double> replay_dev;
    size_t ev_used = 0;
    std::chrono::steady_clock::time_point t0;
    long long allocs0 = 0;
};

int run_begin(ssps_handle* h, RunCtx& cx, int64_t n_sweeps, int thinning, const double* replay_host) {
    if (!h->has_data || h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "ssps_run before ssps_set_data / ssps_init_chains");
    if (thinning < 1) return fail(SSPS_ERR_INVALID, "thinning must be >= 1");
    if (n_sweeps < 0 || n_sweeps % thinning != 0) return fail(SSPS_ERR_INVALID, "n_sweeps must be a non-negative multiple of thinning");
    if (h->thinning == 0) h->thinning = thinning;
    if (h->thinning != thinning) return fail(SSPS_ERR_INVALID, "thinning changed from %d to %d between runs", h->thinning, thinning);
    cx.n_sweeps = n_sweeps; cx.done = 0; cx.thinning = thinning; cx.ev_used = 0;
    cx.log_cp = h->cfg.record_changepoints != 0;
    cx.replay = replay_host != nullptr;
    cx.max_rec = cx.log_cp ? 4096 : 30000;          // rel indices are 15/16-bit on the device
    cx.allocs0 = g_device_allocs;
    if (cx.log_cp && h->cp_parent.empty()) CKS(changepoints_begin(h));
    if (replay_host) CKS(upload(h, cx.replay_dev, replay_host, (size_t)n_sweeps * h->n_chains * h->V * 6));
    cx.t0 = std::chrono::steady_clock::now();
    return SSPS_OK;
}

// enqueue ONE launch (up to the next snapshot point)
int run_launch(ssps_handle* h, RunCtx& cx) {
    const int V = h->V, C = h->n_chains, thinning = cx.thinning;
    const int64_t dflt_rec = h->snap_grid > 0 ? h->snap_grid : 1024;
    // launch boundary = next snapshot point (or the default batch), capped by the rel-index width
    int64_t rec = std::min<int64_t>((cx.n_sweeps - cx.done) / thinning, std::min(cx.max_rec, dflt_rec));
    if (h->snap_grid > 0) rec = std::min(rec, h->snap_grid - h->samples_recorded % h->snap_grid);
    // no measured task costs yet: a short calibration launch, so that the scheduler can split the heavy vertices for the rest
    // of the batch (the first 100-sweep launch of a fresh handle ran 25 % longer than the later ones)
    static const int calib = getenv("SSPS_CALIB_SWEEPS") ? atoi(getenv("SSPS_CALIB_SWEEPS")) : 8;      // experiments: 0 = no calibration launch
    if (!h->have_cost && !cx.replay && rec > 24 && !h->is_uniform && !h->is_poly && calib > 0) rec = calib;
    const int n_local = (int)(rec * thinning);
    MhParams prm{};
    prm.sweep0 = h->sweeps_done; prm.n_local = n_local; prm.thinning = thinning;
    prm.n_slabs = h->nch_pad / 32;
    prm.n_tasks = V * prm.n_slabs;
    prm.task_counter = h->task_counter.p;
    prm.order = h->order.p; prm.vcost = h->vcost.p; prm.start = h->task_start.p; prm.split = h->task_split.p; prm.V = V;
    prm.big_d = h->big_d.p; prm.big_w = h->big_w.p; prm.big_r = h->big_r.p; prm.scr_d = h->scr_d.p; prm.scr_w = h->scr_w.p; prm.scr_r = h->scr_r.p;
    prm.x_d = h->x_d.p; prm.x_t = h->x_t.p;
    prm.kcap = h->kcap; prm.stride = h->mh_grid * SSPS_BLOCK;
    prm.stats = h->stats.p;
    prm.khist = h->cfg.collect_khist ? h->khist.p : nullptr;
    if (cx.log_cp) {
        CKS(changepoints_alloc(h, (int)rec, n_local));
        prm.pev = h->pev.p; prm.pev_n = h->pev_n.p; prm.pev_cap = h->pev_cap;
        prm.lev_val = h->lev_val.p; prm.lev_idx = h->lev_idx.p; prm.lev_n = h->lev_n.p; prm.lev_cap = h->lev_cap;
    }
    // task schedule: from the cycles every vertex cost in the previous launch (first launch: by in-degree, no split)
    if (!h->have_cost) {
        cost_from_state_kernel<<<V, 128, 0, h->stream>>>(h->st, V, h->vcost.p);
        CKS(launch_ok(h, "cost_from_state"));
    }
    schedule_kernel<<<1, 256, 0, h->stream>>>(h->vcost.p, V, prm.n_slabs, h->mh_grid * (SSPS_BLOCK / 32), h->have_cost ? 1 : 0,
                                              h->task_split.p, h->order.p, h->task_start.p, h->task_est.p);
    CKS(launch_ok(h, "order"));
    CK(cudaMemsetAsync(h->vcost.p, 0, h->vcost.bytes(), h->stream));
    h->have_cost = true;
    CK(cudaMemsetAsync(h->task_counter.p, 0, sizeof(int), h->stream));
    if (cx.ev_used == h->ev_pool.size()) {
        cudaEvent_t a, b;
        CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
        h->ev_pool.emplace_back(a, b);
    }
    CK(cudaEventRecord(h->ev_pool[cx.ev_used].first, h->stream));
    const double* rp = cx.replay ? cx.replay_dev.p + (size_t)cx.done * C * V * 6 : nullptr;
    if (h->is_uniform) {
        UniformScratch U{h->u_par.p, h->u_chi.p, h->u_lml.p, h->u_lamlog.p, (V + 31) / 32, h->nch_pad};
        prm.stride = h->nch_pad;
        const int ugrid = h->nch_pad / 128 + (h->nch_pad % 128 ? 1 : 0);
        if (cx.replay) {
            ReplaySource src{rp, C, V};
            mh_uniform_kernel<ReplaySource><<<ugrid, 128, 0, h->stream>>>(h->dev, h->st, prm, src, U, h->x_d.p, h->x_t.p);
        } else {
            PhiloxSource src{h->seed, h->chain_offset};
            mh_uniform_kernel<PhiloxSource><<<ugrid, 128, 0, h->stream>>>(h->dev, h->st, prm, src, U, h->x_d.p, h->x_t.p);
        }
    } else if (h->is_poly) {
        if (cx.replay) {
            ReplaySource src{rp, C, V};
            mh_poly_kernel<ReplaySource><<<h->poly_grid, 128, 0, h->stream>>>(h->dev, h->poly, h->st, prm, src, h->poly_arena.p, h->poly_warena.p, h->poly_mcap_p);
        } else {
            PhiloxSource src{h->seed, h->chain_offset};
            mh_poly_kernel<PhiloxSource><<<h->poly_grid, 128, 0, h->stream>>>(h->dev, h->poly, h->st, prm, src, h->poly_arena.p, h->poly_warena.p, h->poly_mcap_p);
        }
    } else if (cx.replay) {
        ReplaySource src{rp, C, V};
        if (h->cfg.verify) mh_sweeps_kernel<ReplaySource, true><<<h->mh_grid, SSPS_BLOCK, h->mh_smem, h->stream>>>(h->dev, h->st, prm, src);
        else mh_sweeps_kernel<ReplaySource, false><<<h->mh_grid, SSPS_BLOCK, h->mh_smem, h->stream>>>(h->dev, h->st, prm, src);
    } else {
        PhiloxSource src{h->seed, h->chain_offset};
        if (h->cfg.verify) mh_sweeps_kernel<PhiloxSource, true><<<h->mh_grid, SSPS_BLOCK, h->mh_smem, h->stream>>>(h->dev, h->st, prm, src);
        else mh_sweeps_kernel<PhiloxSource, false><<<h->mh_grid, SSPS_BLOCK, h->mh_smem, h->stream>>>(h->dev, h->st, prm, src);
    }
    CKS(launch_ok(h, "mh_sweeps_kernel"));
    CK(cudaEventRecord(h->ev_pool[cx.ev_used].second, h->stream));
    ++cx.ev_used; ++h->mh_launches;
    const int64_t rec_base = h->samples_recorded;
    h->sweeps_done += n_local; h->samples_recorded += rec; cx.done += n_local;
    if (cx.log_cp) CKS(changepoints_drain(h, rec_base));
    CKS(take_snapshot(h));
    if (cx.ev_used >= 256) {   // resolve timings in bounded batches
        CK(cudaStreamSynchronize(h->stream));
        for (size_t e = 0; e < cx.ev_used; ++e) { float ms = 0; CK(cudaEventElapsedTime(&ms, h->ev_pool[e].first, h->ev_pool[e].second)); h->kernel_ms += ms; }
        cx.ev_used = 0;
    }
    return SSPS_OK;
}

int run_end(ssps_handle* h, RunCtx& cx, int64_t* sweeps_done_out) {
    CK(cudaStreamSynchronize(h->stream));
    if (getenv("SSPS_TRACE")) {
        const double ms = 1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - cx.t0).count();
        float kms = 0; for (size_t e = 0; e < cx.ev_used; ++e) { float m1 = 0; cudaEventElapsedTime(&m1, h->ev_pool[e].first, h->ev_pool[e].second); kms += m1;
            if (atoi(getenv("SSPS_TRACE")) >= 2) fprintf(stderr, "[ssps trace] launch %zu: %.2f ms\n", e, m1); }
        if (ms > 1.5 * kms + 2.0) fprintf(stderr, "[ssps trace] run_sweeps %.1f ms wall, %.1f ms in the MH kernel (%lld sweeps)\n", ms, kms, (long long)cx.done);
    }
    for (size_t e = 0; e < cx.ev_used; ++e) { float ms = 0; CK(cudaEventElapsedTime(&ms, h->ev_pool[e].first, h->ev_pool[e].second)); h->kernel_ms += ms; }
    cx.ev_used = 0;
    if (sweeps_done_out) *sweeps_done_out = cx.done;
    h->allocs_in_run += g_device_allocs - cx.allocs0;
    cx.replay_dev.release();
    // (a proposal that would exceed max_indegree is rejected and counted -- ST_KCAP_REJECT -- never an error after the fact)
    return SSPS_OK;
}

// One batch of sweeps for every micro-chain.  replay_host == nullptr => Philox streams.
int run_sweeps(ssps_handle* h, int64_t n_sweeps, double timeout_s, int thinning, const double* replay_host,
               int64_t* sweeps_done_out) {
    RunCtx cx;
    CKS(run_begin(h, cx, n_sweeps, thinning, replay_host));
    while (cx.done < cx.n_sweeps) {
        if (timeout_s > 0.0) {
            CK(cudaStreamSynchronize(h->stream));
            const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - cx.t0).count();
            if (el >= timeout_s) break;
        }
        CKS(run_launch(h, cx));
    }
    return run_end(h, cx, sweeps_done_out);
}

}  // namespace

extern "C" {

int ssps_init_chains(ssps_handle* h, int32_t n_chains, uint64_t seed, uint32_t chain_id_offset,
                     const int32_t* init_offsets, const int32_t* init_members, const double* init_lambda) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "ssps_init_chains before ssps_set_data");
    if (n_chains < 1) return fail(SSPS_ERR_INVALID, "n_chains must be >= 1");
    const bool explicit_state = init_offsets || init_members || init_lambda;
    if (explicit_state && (!init_offsets || !init_lambda))
        return fail(SSPS_ERR_INVALID, "init_offsets / init_members / init_lambda are all-or-nothing");
    const bool trace = getenv("SSPS_TRACE") != nullptr;
    const auto t_in = std::chrono::steady_clock::now();
    const long long allocs_in = g_device_allocs;
    CKS(alloc_chains(h, n_chains));
    if (trace)
        fprintf(stderr, "[ssps trace] init_chains: buffers %.2f ms (%lld cudaMalloc)\n",
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_in).count(), g_device_allocs - allocs_in);
    h->seed = seed; h->chain_offset = chain_id_offset;
    const int V = h->V;
    const size_t nmc = (size_t)n_chains * V;
    const int blocks = div_up((size_t)V * h->nch_pad, 128);
    if (explicit_state) {
        for (size_t m = 0; m < nmc; ++m) {
            const int k = init_offsets[m + 1] - init_offsets[m];
            if (k < 0 || k > V) return fail(SSPS_ERR_INVALID, "initial parent set %zu has %d members", m, k);
            if (k > 0 && !init_members) return fail(SSPS_ERR_INVALID, "null init_members");
            for (int a = 0; a < k; ++a) {
                const int j = init_members[init_offsets[m] + a];
                if (j < 0 || j >= V || (a > 0 && j <= init_members[init_offsets[m] + a - 1]))
                    return fail(SSPS_ERR_INVALID, "initial parent set %zu must be sorted, unique, in range", m);
            }
            const double l = init_lambda[m];
            if (!(l >= h->lambda_min && l <= h->cfg.lambda_max))
                return fail(SSPS_ERR_INVALID, "initial lambda %g outside [%g, %g]", l, h->lambda_min, h->cfg.lambda_max);
        }
        DevBuf<int> d_off, d_mem;
        DevBuf<double> d_lam;
        CKS(upload(h, d_off, init_offsets, nmc + 1));
        CKS(upload(h, d_mem, init_members, (size_t)init_offsets[nmc]));
        CKS(upload(h, d_lam, init_lambda, nmc));
        import_state_kernel<<<blocks, 128, 0, h->stream>>>(h->dev, h->st, d_off.p, d_mem.p, d_lam.p);
        CKS(launch_ok(h, "import_state"));
        CK(cudaStreamSynchronize(h->stream));       // the uploads go out of scope
    } else {
        init_chains_kernel<<<blocks, 128, 0, h->stream>>>(h->dev, h->st, seed, chain_id_offset);
        CKS(launch_ok(h, "init_chains"));
    }
    // the initial sets must fit the bound the sweeps run under: refuse HERE, not after hours of sampling
    {
        int kmax = 0;
        CK(cudaMemsetAsync(h->task_counter.p, 0, sizeof(int), h->stream));
        max_indegree_kernel<<<blocks, 128, 0, h->stream>>>(h->st, V, h->task_counter.p);
        CKS(launch_ok(h, "max_indegree"));
        CK(cudaMemcpyAsync(&kmax, h->task_counter.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        int bound = h->kcap;
        const char* why = "max_indegree";
        if (h->is_poly && kmax > SSPS_POLY_KW) {
            // interaction columns: sets of more than 64 parents are scored only when their design is saturated (closed form)
            int m = 0;
            if (poly_class(SSPS_POLY_KW + 1, h->cfg.regression_deg, h->n, m) != PC_SAT) { bound = std::min(bound, SSPS_POLY_KW); why = "the interaction-column scorer's limit for designs narrower than n - 1 columns"; }
        }
        if (trace)
            fprintf(stderr, "[ssps trace] init_chains: total %.2f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_in).count());
        if (kmax > bound) {
            h->n_chains = 0;
            return fail(SSPS_ERR_INDEGREE, "an initial parent set has %d members, above %s (%d); lower the prior's in-degree "
                        "(--large-indeg) or raise max_indegree", kmax, why, bound);
        }
    }
    return SSPS_OK;
}

int ssps_run(ssps_handle* h, int64_t n_sweeps, double timeout_s, int32_t thinning, int64_t* sweeps_done) {
    CKS(bind(h));
    return run_sweeps(h, n_sweeps, timeout_s, thinning, nullptr, sweeps_done);
}

int ssps_replay(ssps_handle* h, int64_t n_sweeps, const double* draws, int32_t thinning) {
    CKS(bind(h));
    if (!draws) return fail(SSPS_ERR_INVALID, "ssps_replay: null draws");
    return run_sweeps(h, n_sweeps, 0.0, thinning, draws, nullptr);
}

int ssps_export_draws(ssps_handle* h, int64_t first_sweep, int64_t n_sweeps, double* out) {
    CKS(bind(h));
    if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "no chains");
    if (first_sweep < 0 || n_sweeps < 0 || !out) return fail(SSPS_ERR_INVALID, "ssps_export_draws: bad argument");
    if (n_sweeps == 0) return SSPS_OK;
    const size_t per = (size_t)h->n_chains * h->V;
    if ((double)per * (double)n_sweeps * 6.0 * 8.0 > 8e9) return fail(SSPS_ERR_NOMEM, "ssps_export_draws: more than 8 GB requested");
    const size_t total = per * (size_t)n_sweeps;
    DevBuf<double> d;
    CK(d.alloc(total * 6));
    PhiloxSource src{h->seed, h->chain_offset};
    export_draws_kernel<<<div_up(total, 256), 256, 0, h->stream>>>(src, h->n_chains, h->V, first_sweep, (int)n_sweeps, d.p);
    CKS(launch_ok(h, "export_draws"));
    CKS(download(h, out, d.p, total * 6));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

int ssps_get_state(ssps_handle* h, int32_t chain, int32_t* indeg, int32_t* members, int64_t cap, int64_t* n_members,
                   double* lambda, double* lml) {
    CKS(bind(h));
    if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "no chains");
    if (chain < 0 || chain >= h->n_chains) return fail(SSPS_ERR_INVALID, "chain %d out of range", chain);
    const int V = h->V;
    std::vector<int> kk, mem;
    std::vector<double> lam;
    CKS(fetch_state_host(h, chain, kk, mem, lam));
    int64_t total = 0;
    for (int i = 0; i < V; ++i) total += kk[i];
    if (n_members) *n_members = total;
    if (indeg) memcpy(indeg, kk.data(), sizeof(int) * (size_t)V);
    if (lambda) memcpy(lambda, lam.data(), sizeof(double) * (size_t)V);
    if (members) {
        if (cap < total) return fail(SSPS_ERR_INVALID, "members capacity %lld < %lld", (long long)cap, (long long)total);
        int64_t o = 0;
        for (int i = 0; i < V; ++i)
            for (int a = 0; a < kk[i]; ++a) members[o++] = mem[(size_t)i * V + a];
    }
    if (lml && h->is_poly) {
        std::vector<int32_t> ch(V), off(V + 1, 0), mm;
        for (int i = 0; i < V; ++i) {
            ch[i] = i; off[i + 1] = off[i] + kk[i];
            for (int a = 0; a < kk[i]; ++a) mm.push_back(mem[(size_t)i * V + a]);
        }
        if (mm.empty()) mm.push_back(0);
        return ssps_score(h, V, ch.data(), off.data(), mm.data(), nullptr, lml, nullptr, nullptr);
    }
    if (lml) {
        const int block = 128, grid = div_up((size_t)V, block);
        const size_t nthreads = (size_t)grid * block;
        DevBuf<double> d_lml, arena_d;
        DevBuf<uint16_t> arena_t;
        int kmax = 1;
        for (int i = 0; i < V; ++i) kmax = std::max(kmax, kk[i]);
        CK(d_lml.alloc(V)); CK(arena_d.alloc(nthreads * (size_t)(SSPS_TRI(kmax) + kmax))); CK(arena_t.alloc(nthreads * (size_t)kmax));
        state_lml_kernel<<<grid, block, 0, h->stream>>>(h->dev, h->st, chain, arena_d.p, arena_t.p, kmax, (int)nthreads, d_lml.p);
        CKS(launch_ok(h, "state_lml"));
        CKS(download(h, lml, d_lml.p, (size_t)V));
        CK(cudaStreamSynchronize(h->stream));
    }
    return SSPS_OK;
}

// ------------------------------------------------------------------------------------------------
// tallies / posterior / PSRF (K4)
// ------------------------------------------------------------------------------------------------
int ssps_edge_counts(ssps_handle* h, double* counts_child_major, int64_t* n_samples, int32_t* n_chains) {
    CKS(bind(h));
    if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "no chains");
    const size_t vv = (size_t)h->V * h->V;
    if (counts_child_major) {
        DevBuf<double> out;
        CK(out.alloc(vv));
        int chunks = 1, chunk_len = h->n_chains;
        chain_chunks(h, vv, chunks, chunk_len);
        CK(cudaMemsetAsync(out.p, 0, out.bytes(), h->stream));
        edge_sum_kernel<<<dim3(div_up(vv, 256), chunks), 256, 0, h->stream>>>(h->counts.p, nullptr, h->n_chains, vv, chunk_len, out.p);
        CKS(launch_ok(h, "edge_sum"));
        CKS(download(h, counts_child_major, out.p, vv));
        CK(cudaStreamSynchronize(h->stream));
    }
    if (n_samples) *n_samples = h->samples_recorded;
    if (n_chains) *n_chains = h->n_chains;
    return SSPS_OK;
}

int ssps_chain_edge_counts(ssps_handle* h, uint32_t* counts) {
    CKS(bind(h));
    if (h->n_chains <= 0 || !counts) return fail(SSPS_ERR_INVALID, "no chains / null buffer");
    CKS(download(h, counts, h->counts.p, (size_t)h->V * h->V * h->n_chains));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

int ssps_edge_posterior(ssps_handle* h, double burnin, double* probs_child_major, double* lambda_mean,
                        int64_t* burnin_used, int64_t* n_used) {
    CKS(bind(h));
    if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "no chains");
    if (!(burnin >= 0.0 && burnin < 1.0)) return fail(SSPS_ERR_INVALID, "burnin must be in [0, 1)");
    const int64_t n = h->samples_recorded;
    if (n < 1) return fail(SSPS_ERR_UNAVAILABLE, "no recorded samples yet");
    const int64_t want = (int64_t)floor(burnin * (double)n);   // postprocess_samples.jl:186
    const Snapshot* lo = nullptr;
    for (const auto& s : h->snaps)
        if (s.n <= want && s.n < n) lo = &s;
    const int64_t b = lo ? lo->n : 0;
    const size_t vv = (size_t)h->V * h->V;
    const double denom = (double)(n - b) * (double)h->n_chains;
    if (probs_child_major) {
        DevBuf<double> out;
        CK(out.alloc(vv));
        int chunks = 1, chunk_len = h->n_chains;
        chain_chunks(h, vv, chunks, chunk_len);
        CK(cudaMemsetAsync(out.p, 0, out.bytes(), h->stream));
        edge_sum_kernel<<<dim3(div_up(vv, 256), chunks), 256, 0, h->stream>>>(h->counts.p, lo ? lo->counts : nullptr, h->n_chains, vv, chunk_len, out.p);
        CKS(launch_ok(h, "edge_sum"));
        CKS(download(h, probs_child_major, out.p, vv));
        CK(cudaStreamSynchronize(h->stream));
        for (size_t e = 0; e < vv; ++e) probs_child_major[e] /= denom;
    }
    if (lambda_mean) {
        DevBuf<double> out;
        CK(out.alloc(h->V));
        lam_sum_kernel<<<h->V, 256, 0, h->stream>>>(h->lam_sum.p, lo ? lo->lam_sum : nullptr, h->n_chains, h->V, out.p);
        CKS(launch_ok(h, "lam_sum"));
        CKS(download(h, lambda_mean, out.p, (size_t)h->V));
        CK(cudaStreamSynchronize(h->stream));
        for (int i = 0; i < h->V; ++i) lambda_mean[i] /= denom;
    }
    if (burnin_used) *burnin_used = b;
    if (n_used) *n_used = n;
    return SSPS_OK;
}

int64_t ssps_psrf_sums_len(int32_t V) { return 3 * (int64_t)V * V + 3 * (int64_t)V + 2; }

int ssps_psrf_sums(ssps_handle* h, double burnin, double* sums, int32_t on_device) {
    CKS(bind(h));
    if (h->n_chains <= 0 || !sums) return fail(SSPS_ERR_INVALID, "no chains / null buffer");
    if (!(burnin >= 0.0 && burnin < 1.0)) return fail(SSPS_ERR_INVALID, "burnin must be in [0, 1)");
    if (h->snap_grid <= 0) return fail(SSPS_ERR_UNAVAILABLE, "snapshots disabled (psrf_grid < 0)");
    // largest usable length: len, burn-in index b and split index s must all have snapshots
    // (evaluate_convergence, postprocess_samples.jl:167-168; b = 0 needs none)
    const Snapshot *S_len = nullptr, *S_s = nullptr, *S_b = nullptr;
    int64_t len = 0, b = 0, s = 0;
    std::vector<int64_t> cand;
    cand.push_back(h->samples_recorded);
    for (auto it = h->snaps.rbegin(); it != h->snaps.rend(); ++it)
        if (it->n < h->samples_recorded) cand.push_back(it->n);
    bool ok = false;
    for (int64_t L : cand) {
        b = (int64_t)floor(burnin * (double)L);
        s = b + (L - b) / 2;
        if (s - b < 2 || L - s < 2) continue;
        S_s = find_snapshot(h, s);
        S_b = b == 0 ? nullptr : find_snapshot(h, b);
        S_len = find_snapshot(h, L);
        if (S_s && (b == 0 || S_b) && (S_len || L == h->samples_recorded)) { len = L; ok = true; break; }
    }
    if (!ok) return fail(SSPS_ERR_UNAVAILABLE, "no snapshot-aligned (burn-in, split, length) triple yet: %lld samples, grid %lld",
                         (long long)h->samples_recorded, (long long)h->snap_grid);
    const int V = h->V;
    const size_t vv = (size_t)V * V;
    const size_t total = (size_t)ssps_psrf_sums_len(V);
    double* dst = sums;
    if (!on_device) { CK(h->sums_scratch.alloc(total)); dst = h->sums_scratch.p; }
    const double N1 = (double)(s - b), N2 = (double)(len - s);
    const uint32_t* c_len = S_len ? S_len->counts : h->counts.p;
    const double* ls_len = S_len ? S_len->lam_sum : h->lam_sum.p;
    const double* lq_len = S_len ? S_len->lam_sq : h->lam_sq.p;
    {
        int chunks = 1, chunk_len = h->n_chains;
        chain_chunks(h, vv, chunks, chunk_len);
        double* part = dst;
        if (chunks > 1) { CK(h->psrf_partial.alloc((size_t)chunks * 3 * vv)); part = h->psrf_partial.p; }
        psrf_sums_kernel<<<dim3(div_up(vv, 256), chunks), 256, 0, h->stream>>>(c_len, S_s->counts, S_b ? S_b->counts : nullptr, h->n_chains, vv,
                                                                                N1, N2, chunk_len, part);
        CKS(launch_ok(h, "psrf_sums"));
        if (chunks > 1) {
            psrf_sums_reduce_kernel<<<div_up(3 * vv, 256), 256, 0, h->stream>>>(part, chunks, 3 * vv, dst);
            CKS(launch_ok(h, "psrf_sums_reduce"));
        }
    }
    psrf_lambda_sums_kernel<<<V, 256, 0, h->stream>>>(ls_len, S_s->lam_sum, S_b ? S_b->lam_sum : nullptr, lq_len,
                                                                        S_s->lam_sq, S_b ? S_b->lam_sq : nullptr, h->n_chains, V, N1, N2,
                                                                        dst + 3 * vv);
    CKS(launch_ok(h, "psrf_lambda_sums"));
    // tail: number of half-chains H, and H * n (so that a SUM all-reduce over ranks keeps n = tail[1] / tail[0])
    const double tail[2] = {2.0 * (double)h->n_chains, 2.0 * (double)h->n_chains * N1};
    CK(cudaMemcpyAsync(dst + 3 * vv + 3 * (size_t)V, tail, sizeof tail, cudaMemcpyHostToDevice, h->stream));
    if (!on_device) CKS(download(h, sums, dst, total));
    CK(cudaStreamSynchronize(h->stream));
    return SSPS_OK;
}

int ssps_psrf_finish(ssps_handle* h, const double* sums, int32_t on_device, double* max_edge_psrf,
                     double* edge_psrf_child_major, double* lambda_psrf) {
    CKS(bind(h));
    if (!h->has_data || !sums) return fail(SSPS_ERR_INVALID, "no data / null buffer");
    const int V = h->V;
    const size_t vv = (size_t)V * V;
    const size_t total = (size_t)ssps_psrf_sums_len(V);
    const double* src = sums;
    if (!on_device) {
        CK(h->sums_scratch.alloc(total));
        CK(cudaMemcpyAsync(h->sums_scratch.p, sums, total * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        src = h->sums_scratch.p;
    }
    double tail[2];
    CK(cudaMemcpyAsync(tail, src + 3 * vv + 3 * (size_t)V, sizeof tail, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    const double H = tail[0], n = tail[1] / tail[0];
    if (!(H >= 2.0) || !(n >= 2.0)) return fail(SSPS_ERR_INVALID, "psrf sums carry H=%g n=%g", H, n);
    CK(h->psrf_scratch.alloc(vv + V));
    CK(h->max_bits.alloc(1));
    CK(cudaMemsetAsync(h->max_bits.p, 0, sizeof(unsigned long long), h->stream));
    psrf_finish_kernel<<<div_up(vv, 256), 256, 0, h->stream>>>(src, vv, vv, H, n, h->psrf_scratch.p, h->max_bits.p);
    CKS(launch_ok(h, "psrf_finish(edges)"));
    psrf_finish_kernel<<<div_up((size_t)V, 128), 128, 0, h->stream>>>(src + 3 * vv, (size_t)V, (size_t)V, H, n, h->psrf_scratch.p + vv, nullptr);
    CKS(launch_ok(h, "psrf_finish(lambda)"));
    unsigned long long bits = 0;
    CKS(download(h, &bits, h->max_bits.p, 1));
    if (edge_psrf_child_major) CKS(download(h, edge_psrf_child_major, h->psrf_scratch.p, vv));
    if (lambda_psrf) CKS(download(h, lambda_psrf, h->psrf_scratch.p + vv, (size_t)V));
    CK(cudaStreamSynchronize(h->stream));
    if (max_edge_psrf) {
        double m;
        memcpy(&m, &bits, sizeof m);
        *max_edge_psrf = bits ? m : NAN;   // NaN: every edge indicator was constant (W == 0 everywhere)
    }
    return SSPS_OK;
}

int ssps_psrf(ssps_handle* h, double burnin, double* max_edge_psrf, double* edge_psrf_child_major) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "no data");
    DevBuf<double> sums;
    CK(sums.alloc((size_t)ssps_psrf_sums_len(h->V)));
    CKS(ssps_psrf_sums(h, burnin, sums.p, 1));
    return ssps_psrf_finish(h, sums.p, 1, max_edge_psrf, edge_psrf_child_major, nullptr);
}

// ------------------------------------------------------------------------------------------------
// change-point log
// ------------------------------------------------------------------------------------------------
int ssps_changepoints_size(ssps_handle* h, int32_t chain, int64_t* n_parent_events, int64_t* n_lambda_events) {
    CKS(bind(h));
    if (!h->cfg.record_changepoints) return fail(SSPS_ERR_UNAVAILABLE, "record_changepoints was 0 at ssps_create");
    if (chain < 0 || chain >= h->n_chains) return fail(SSPS_ERR_INVALID, "chain %d out of range", chain);
    if (h->cp_parent.empty()) CKS(changepoints_begin(h));
    if (n_parent_events) *n_parent_events = (int64_t)h->cp_parent[chain].size();
    if (n_lambda_events) *n_lambda_events = (int64_t)h->cp_lambda[chain].size();
    return SSPS_OK;
}

int ssps_changepoints_get(ssps_handle* h, int32_t chain, int64_t* pe_n, int32_t* pe_child, int32_t* pe_parent,
                          int32_t* pe_value, int64_t* le_n, int32_t* le_child, double* le_value) {
    int64_t np = 0, nl = 0;
    CKS(ssps_changepoints_size(h, chain, &np, &nl));
    std::vector<ParentEvent> pe = h->cp_parent[chain];
    std::stable_sort(pe.begin(), pe.end(), [](const ParentEvent& a, const ParentEvent& b) {
        if (a.child != b.child) return a.child < b.child;
        if (a.parent != b.parent) return a.parent < b.parent;
        return a.n < b.n;
    });
    for (int64_t e = 0; e < np; ++e) {
        if (pe_n) pe_n[e] = pe[e].n;
        if (pe_child) pe_child[e] = pe[e].child;
        if (pe_parent) pe_parent[e] = pe[e].parent;
        if (pe_value) pe_value[e] = pe[e].value;
    }
    std::vector<LambdaEvent> le = h->cp_lambda[chain];
    std::stable_sort(le.begin(), le.end(), [](const LambdaEvent& a, const LambdaEvent& b) {
        if (a.child != b.child) return a.child < b.child;
        return a.n < b.n;
    });
    for (int64_t e = 0; e < nl; ++e) {
        if (le_n) le_n[e] = le[e].n;
        if (le_child) le_child[e] = le[e].child;
        if (le_value) le_value[e] = le[e].value;
    }
    return SSPS_OK;
}

// ------------------------------------------------------------------------------------------------
// stats / diagnostics
// ------------------------------------------------------------------------------------------------
int ssps_get_stats(ssps_handle* h, ssps_run_stats* out) {
    CKS(bind(h));
    if (!out) return fail(SSPS_ERR_INVALID, "null out");
    memset(out, 0, sizeof *out);
    out->kernel_launches = h->kernel_launches;
    if (h->n_chains > 0) {
        unsigned long long st[ST_N];
        CK(cudaMemcpyAsync(st, h->stats.p, sizeof st, cudaMemcpyDeviceToHost, h->stream));     // (on the handle's stream: no device-wide sync)
        CK(cudaStreamSynchronize(h->stream));
        out->sweeps_done = h->sweeps_done;
        out->samples_recorded = h->samples_recorded;
        out->n_proposals = (int64_t)st[ST_PROP];
        out->n_accepted = (int64_t)st[ST_ACCP];
        out->n_lambda_accepted = (int64_t)st[ST_ACCL];
        out->sum_k = (int64_t)st[ST_SUMK];
        out->algorithmic_flops = (double)st[ST_F3] / 3.0;   // sum (k^3 + 6k^2 + 6k) / 3 = sum k^3/3 + 2k^2 + 2k
        out->algorithmic_bytes = 8.0 * (double)st[ST_BYTES];
        out->n_tier1 = (int64_t)st[ST_TIER1]; out->n_big_steps = (int64_t)st[ST_BIG_STEPS];
        out->n_lambda_exact = (int64_t)st[ST_LAM_EXACT]; out->n_mismatch = (int64_t)st[ST_MISMATCH];
        out->big_cycles = (int64_t)st[ST_BIG_CYCLES];
        out->kernel_ms = h->kernel_ms;
        out->mh_launches = h->mh_launches;
        out->n_decision_mismatch = (int64_t)st[ST_DEC_MISMATCH];
        out->n_lambda_decision_mismatch = (int64_t)st[ST_LAM_DEC_MISMATCH];
        { uint32_t b = (uint32_t)st[ST_MARGIN_USE]; float f; memcpy(&f, &b, sizeof f); out->margin_use_max = (double)f; }
        { uint32_t b = (uint32_t)st[ST_LAM_MARGIN_USE]; float f; memcpy(&f, &b, sizeof f); out->lambda_margin_use_max = (double)f; }
        out->n_indegree_rejects = (int64_t)st[ST_KCAP_REJECT];
        out->n_mid_steps = (int64_t)st[ST_MID_STEPS];
        out->snapshot_slots = h->snap_slots;
        out->snapshots_live = (int64_t)h->snaps.size();
        out->snapshot_grid = h->snap_grid;
        out->device_allocs_in_run = h->allocs_in_run;
    }
    return SSPS_OK;
}

int ssps_get_raw_stats(ssps_handle* h, int64_t* out, int32_t len) {
    CKS(bind(h));
    if (!out || len < 1 || h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "bad buffer or no chains");
    unsigned long long st[ST_N];
    CK(cudaMemcpyAsync(st, h->stats.p, sizeof st, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int a = 0; a < len; ++a) out[a] = a < ST_N ? (int64_t)st[a] : 0;
    return SSPS_OK;
}

int ssps_get_khist(ssps_handle* h, int64_t* hist, int32_t len) {
    CKS(bind(h));
    if (!hist || len < 1) return fail(SSPS_ERR_INVALID, "bad buffer");
    if (h->n_chains <= 0 || !h->cfg.collect_khist) return fail(SSPS_ERR_UNAVAILABLE, "collect_khist was 0 or no chains");
    unsigned long long hh[64];
    CK(cudaMemcpyAsync(hh, h->khist.p, sizeof hh, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int a = 0; a < len; ++a) hist[a] = a < 64 ? (int64_t)hh[a] : 0;
    return SSPS_OK;
}

int ssps_get_vertex_cost(ssps_handle* h, int64_t* cycles, int32_t* mean_indeg_x1000, int32_t* split) {
    CKS(bind(h));
    if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "no chains");
    const int V = h->V;
    CK(cudaStreamSynchronize(h->stream));
    if (cycles) {
        std::vector<unsigned long long> c(V);
        CK(cudaMemcpy(c.data(), h->vcost.p, sizeof(unsigned long long) * V, cudaMemcpyDeviceToHost));
        for (int i = 0; i < V; ++i) cycles[i] = (int64_t)c[i];
    }
    if (split) CK(cudaMemcpy(split, h->task_split.p, sizeof(int) * V, cudaMemcpyDeviceToHost));
    if (mean_indeg_x1000) {
        std::vector<int> k((size_t)V * h->nch_pad);
        CK(cudaMemcpy(k.data(), h->k.p, sizeof(int) * k.size(), cudaMemcpyDeviceToHost));
        for (int i = 0; i < V; ++i) {
            long long s = 0;
            for (int c = 0; c < h->n_chains; ++c) s += k[(size_t)i * h->nch_pad + c];
            mean_indeg_x1000[i] = (int32_t)(1000 * s / h->n_chains);
        }
    }
    return SSPS_OK;
}

int64_t ssps_bounds_failures(void) {
#ifdef SSPS_BOUNDS
    unsigned long long v = 0;
    if (cudaMemcpyFromSymbol(&v, g_ssps_bounds_fail, sizeof v) != cudaSuccess) { cudaGetLastError(); return -2; }
    return (int64_t)v;
#else
    return -1;
#endif
}

int ssps_release_cached_memory(int32_t device) {
    int prev = 0;
    CK(cudaGetDevice(&prev));
    CK(cudaSetDevice(device));
    CK(cudaDeviceSynchronize());
    BlockCache::get().trim(device);
    CK(cudaSetDevice(prev));
    return SSPS_OK;
}

int ssps_measure_fp64_peak(ssps_handle* h, double* tflops) {
    CKS(bind(h));
    if (!tflops) return fail(SSPS_ERR_INVALID, "null out");
    const int block = 256, grid = h->sm_count * 8, iters = 1 << 16;
    DevBuf<double> out;
    CK(out.alloc((size_t)grid * block));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(a, h->stream));
        dfma_peak_kernel<<<grid, block, 0, h->stream>>>(out.p, iters);
        CK(cudaEventRecord(b, h->stream));
        CKS(launch_ok(h, "dfma_peak"));
        CK(cudaEventSynchronize(b));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, a, b));
        const double flops = 2.0 * 8.0 * (double)iters * (double)grid * block;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    *tflops = best;
    return SSPS_OK;
}

int ssps_measure_gather_peak(ssps_handle* h, int64_t table_bytes, double* l1_gbs, double* l2_gbs, double* smem_gbs) {
    CKS(bind(h));
    if (table_bytes < 8) return fail(SSPS_ERR_INVALID, "ssps_measure_gather_peak: table_bytes < 8");
    const unsigned n_elems = (unsigned)std::min<int64_t>(table_bytes / 8, (int64_t)1 << 28);
    const int block = 256, grid = h->sm_count * 8, iters = 1 << 12;
    DevBuf<double> table, out;
    CK(table.alloc(n_elems)); CK(out.alloc((size_t)grid * block));
    CK(cudaMemsetAsync(table.p, 0, table.bytes(), h->stream));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    double* res[3] = {l1_gbs, l2_gbs, smem_gbs};
    for (int mode = 0; mode < 3; ++mode) {
        if (!res[mode]) continue;
        double best = 0.0;
        for (int rep = 0; rep < 4; ++rep) {
            CK(cudaEventRecord(a, h->stream));
            if (mode == 0) gather_peak_kernel<0><<<grid, block, 0, h->stream>>>(table.p, n_elems, iters, out.p);
            else if (mode == 1) gather_peak_kernel<1><<<grid, block, 0, h->stream>>>(table.p, n_elems, iters, out.p);
            else gather_peak_kernel<2><<<grid, block, 0, h->stream>>>(table.p, n_elems, iters, out.p);
            CK(cudaEventRecord(b, h->stream));
            CKS(launch_ok(h, "gather_peak"));
            CK(cudaEventSynchronize(b));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, a, b));
            const double bytes = 8.0 * 8.0 * (double)iters * (double)grid * block;
            if (rep > 0) best = std::max(best, bytes / (ms * 1e-3) / 1e9);
        }
        *res[mode] = best;
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    return SSPS_OK;
}

}  // extern "C"

// ================================================================================================
// Multi-GPU: chains are sharded over handles (one per GPU, in one process or one process per GPU); the only
// exchange is a SUM all-reduce of additive statistics -- the split-PSRF sums of ssps_psrf_sums and the pooled
// tallies of ssps_edge_posterior -- over an NCCL communicator the LIBRARY owns.  It replaces the
// reference's offline pooling of per-chain JSON files (scripts/postprocess_samples.jl:242-259).
// libnccl.so.2 is resolved at run time (dlopen): single-GPU users need no NCCL at all, and inside a torch process the
// copy torch already loaded is the one that is used.
// ================================================================================================
namespace {

struct NcclId { char internal[128]; };      // == ncclUniqueId (nccl.h)
typedef void* NcclComm;                     // == ncclComm_t
enum { NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MAX = 2 };      // ncclDataType_t / ncclRedOp_t values (nccl.h)

struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // the copy this process already has (torch's), if any
        if (!lib) if (const char* e = getenv("SSPS_NCCL_LIB")) lib = dlopen(e, RTLD_NOW);
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW);
        if (!lib) return;
        api.lib = lib;
        *(void**)&api.GetUniqueId = dlsym(lib, "ncclGetUniqueId");
        *(void**)&api.CommInitRank = dlsym(lib, "ncclCommInitRank");
        *(void**)&api.CommDestroy = dlsym(lib, "ncclCommDestroy");
        *(void**)&api.AllReduce = dlsym(lib, "ncclAllReduce");
        *(void**)&api.GroupStart = dlsym(lib, "ncclGroupStart");
        *(void**)&api.GroupEnd = dlsym(lib, "ncclGroupEnd");
        *(void**)&api.GetErrorString = dlsym(lib, "ncclGetErrorString");
        *(void**)&api.GetVersion = dlsym(lib, "ncclGetVersion");
    });
    if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce || !api.GroupStart || !api.GroupEnd) return nullptr;
    return &api;
}

int need_nccl(NcclApi*& api) {
    api = nccl_api();
    if (!api) return fail(SSPS_ERR_UNAVAILABLE, "libnccl.so.2 could not be loaded (%s); set SSPS_NCCL_LIB or LD_LIBRARY_PATH", dlerror() ? dlerror() : "symbols missing");
    return SSPS_OK;
}

#define NK(api, call)                                                                                          \
    do {                                                                                                       \
        int r_ = (call);                                                                                       \
        if (r_ != 0) return fail(SSPS_ERR_CUDA, "%s: %s", #call, (api)->GetErrorString ? (api)->GetErrorString(r_) : "nccl error"); \
    } while (0)

// un-normalised pooled tallies of the local chains on the device: out[0 .. V*V) = sum over chains of post-burn-in counts,
// out[V*V .. +V) = the same for lambda, out[V*V + V] = (n - b) * local chains
int posterior_sums(ssps_handle* h, double burnin, double* out_dev, int64_t* b_out, int64_t* n_out) {
    if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "no chains");
    if (!(burnin >= 0.0 && burnin < 1.0)) return fail(SSPS_ERR_INVALID, "burnin must be in [0, 1)");
    const int64_t n = h->samples_recorded;
    if (n < 1) return fail(SSPS_ERR_UNAVAILABLE, "no recorded samples yet");
    const int64_t want = (int64_t)floor(burnin * (double)n);   // postprocess_samples.jl:186
    const Snapshot* lo = nullptr;
    for (const auto& sn : h->snaps)
        if (sn.n <= want && sn.n < n && (!lo || sn.n > lo->n)) lo = &sn;
    const int64_t b = lo ? lo->n : 0;
    const size_t vv = (size_t)h->V * h->V;
    int chunks = 1, chunk_len = h->n_chains;
    chain_chunks(h, vv, chunks, chunk_len);
    CK(cudaMemsetAsync(out_dev, 0, (vv + (size_t)h->V + 1) * sizeof(double), h->stream));
    edge_sum_kernel<<<dim3(div_up(vv, 256), chunks), 256, 0, h->stream>>>(h->counts.p, lo ? lo->counts : nullptr, h->n_chains, vv, chunk_len, out_dev);
    CKS(launch_ok(h, "edge_sum"));
    lam_sum_kernel<<<h->V, 256, 0, h->stream>>>(h->lam_sum.p, lo ? lo->lam_sum : nullptr, h->n_chains, h->V, out_dev + vv);
    CKS(launch_ok(h, "lam_sum"));
    const double denom = (double)(n - b) * (double)h->n_chains;
    CK(cudaMemcpyAsync(out_dev + vv + h->V, &denom, sizeof denom, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));       // `denom` is a stack variable
    if (b_out) *b_out = b;
    if (n_out) *n_out = n;
    return SSPS_OK;
}

// SUM all-reduce of `count` doubles staged in every handle's comm_buf (in place); one group call for all local handles
int comm_allreduce(ssps_handle* const* hs, int n, size_t count, int op) {
    NcclApi* api = nullptr;
    bool any = false;
    for (int e = 0; e < n; ++e) any = any || hs[e]->comm != nullptr;
    if (!any) {
        if (n == 1) return SSPS_OK;             // a single handle without a communicator: the local sums are the global ones
        return fail(SSPS_ERR_INVALID, "%d handles but no communicator: call ssps_comm_init_all first", n);
    }
    CKS(need_nccl(api));
    NK(api, api->GroupStart());
    for (int e = 0; e < n; ++e) {
        ssps_handle* h = hs[e];
        if (!h->comm) { api->GroupEnd(); return fail(SSPS_ERR_INVALID, "handle %d has no communicator", e); }
        CK(cudaSetDevice(h->device));
        NK(api, api->AllReduce(h->comm_buf.p, h->comm_buf.p, count, NCCL_FLOAT64, op, h->comm, h->stream));
    }
    NK(api, api->GroupEnd());
    for (int e = 0; e < n; ++e) { CK(cudaSetDevice(hs[e]->device)); CK(cudaStreamSynchronize(hs[e]->stream)); }
    return SSPS_OK;
}

int check_handles(ssps_handle* const* hs, int n) {
    if (!hs || n < 1) return fail(SSPS_ERR_INVALID, "need at least one handle");
    for (int e = 0; e < n; ++e) {
        if (!hs[e]) return fail(SSPS_ERR_INVALID, "handle %d is null", e);
        if (hs[e]->V != hs[0]->V) return fail(SSPS_ERR_INVALID, "handle %d holds a different data set (V=%d vs %d)", e, hs[e]->V, hs[0]->V);
    }
    return SSPS_OK;
}

}  // namespace

extern "C" {

int ssps_comm_unique_id(void* id_out) {
    if (!id_out) return fail(SSPS_ERR_INVALID, "null id buffer");
    NcclApi* api = nullptr;
    CKS(need_nccl(api));
    NcclId id;
    NK(api, api->GetUniqueId(&id));
    memcpy(id_out, &id, sizeof id);
    return SSPS_OK;
}

int ssps_comm_init_rank(ssps_handle* h, int32_t n_ranks, int32_t rank, const void* id) {
    CKS(bind(h));
    if (!id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(SSPS_ERR_INVALID, "ssps_comm_init_rank: bad argument");
    if (h->comm) return fail(SSPS_ERR_INVALID, "the handle already has a communicator");
    NcclApi* api = nullptr;
    CKS(need_nccl(api));
    NcclId nid;
    memcpy(&nid, id, sizeof nid);
    NcclComm c = nullptr;
    NK(api, api->CommInitRank(&c, n_ranks, nid, rank));
    h->comm = c; h->comm_rank = rank; h->comm_size = n_ranks;
    return SSPS_OK;
}

int ssps_comm_init_all(ssps_handle* const* hs, int32_t n) {
    CKS(check_handles(hs, n));
    NcclApi* api = nullptr;
    CKS(need_nccl(api));
    for (int e = 0; e < n; ++e) {
        if (hs[e]->comm) return fail(SSPS_ERR_INVALID, "handle %d already has a communicator", e);
        for (int f = 0; f < e; ++f) if (hs[f]->device == hs[e]->device) return fail(SSPS_ERR_INVALID, "handles %d and %d share device %d", f, e, hs[e]->device);
    }
    NcclId id;
    NK(api, api->GetUniqueId(&id));
    std::vector<NcclComm> comms((size_t)n, nullptr);
    NK(api, api->GroupStart());
    for (int e = 0; e < n; ++e) {
        CK(cudaSetDevice(hs[e]->device));
        NK(api, api->CommInitRank(&comms[e], n, id, e));
    }
    NK(api, api->GroupEnd());
    for (int e = 0; e < n; ++e) { hs[e]->comm = comms[e]; hs[e]->comm_rank = e; hs[e]->comm_size = n; }
    return SSPS_OK;
}

int ssps_comm_destroy(ssps_handle* h) {
    CKS(bind(h));
    if (!h->comm) return SSPS_OK;
    NcclApi* api = nullptr;
    CKS(need_nccl(api));
    CK(cudaStreamSynchronize(h->stream));
    NK(api, api->CommDestroy(h->comm));
    h->comm = nullptr; h->comm_rank = 0; h->comm_size = 1;
    return SSPS_OK;
}

int ssps_comm_info(ssps_handle* h, int32_t* n_ranks, int32_t* rank, int32_t* nccl_version) {
    if (!h) return fail(SSPS_ERR_INVALID, "null handle");
    if (n_ranks) *n_ranks = h->comm ? h->comm_size : 1;
    if (rank) *rank = h->comm ? h->comm_rank : 0;
    if (nccl_version) {
        *nccl_version = 0;
        NcclApi* api = nccl_api();
        if (api && api->GetVersion) api->GetVersion(nccl_version);
    }
    return SSPS_OK;
}

// ssps_run on every handle at once: one host thread keeps all GPUs busy (launches are enqueued round-robin, one
// synchronisation at the end; with a timeout, one check per round of launches)
int ssps_run_multi(ssps_handle* const* hs, int32_t n, int64_t n_sweeps, double timeout_s, int32_t thinning, int64_t* sweeps_done) {
    CKS(check_handles(hs, n));
    std::vector<RunCtx> cx((size_t)n);
    for (int e = 0; e < n; ++e) { CKS(bind(hs[e])); CKS(run_begin(hs[e], cx[e], n_sweeps, thinning, nullptr)); }
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        bool left = false;
        for (int e = 0; e < n; ++e) left = left || cx[e].done < cx[e].n_sweeps;
        if (!left) break;
        if (timeout_s > 0.0) {
            for (int e = 0; e < n; ++e) { CK(cudaSetDevice(hs[e]->device)); CK(cudaStreamSynchronize(hs[e]->stream)); }
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() >= timeout_s) break;
        }
        for (int e = 0; e < n; ++e)
            if (cx[e].done < cx[e].n_sweeps) { CKS(bind(hs[e])); CKS(run_launch(hs[e], cx[e])); }
    }
    int64_t done_min = n_sweeps;
    for (int e = 0; e < n; ++e) {
        int64_t d = 0;
        CKS(bind(hs[e]));
        CKS(run_end(hs[e], cx[e], &d));
        done_min = std::min(done_min, d);
    }
    if (sweeps_done) *sweeps_done = done_min;
    return SSPS_OK;
}

// split-PSRF over the chains of ALL ranks.  hs = the handles this process drives: n = 1 in the process-per-GPU model
// (collective: every rank calls it), all local handles after ssps_comm_init_all.  Outputs as ssps_psrf_finish.
int ssps_psrf_global(ssps_handle* const* hs, int32_t n, double burnin, double* max_edge_psrf, double* edge_psrf_child_major,
                     double* lambda_psrf) {
    CKS(check_handles(hs, n));
    const size_t total = (size_t)ssps_psrf_sums_len(hs[0]->V);
    // every rank must split its chains at the same samples: exchange (L, -L) with MAX first
    int first_err = SSPS_OK;
    std::string first_msg;
    for (int e = 0; e < n; ++e) {
        CKS(bind(hs[e]));
        CK(hs[e]->comm_buf.alloc(total + 2));
        int rc = ssps_psrf_sums(hs[e], burnin, hs[e]->comm_buf.p, 1);
        if (rc != SSPS_OK && first_err == SSPS_OK) { first_err = rc; first_msg = g_err; }
        // length used = tail[1] / tail[0] * 2 + b ... carried explicitly instead: the half length n1 identifies the triple
        double chk[2] = {-1.0, -1.0};
        if (rc == SSPS_OK) {
            double tail[2];
            CK(cudaMemcpyAsync(tail, hs[e]->comm_buf.p + total - 2, sizeof tail, cudaMemcpyDeviceToHost, hs[e]->stream));
            CK(cudaStreamSynchronize(hs[e]->stream));
            chk[0] = tail[1] / tail[0]; chk[1] = -chk[0];
        } else {
            CK(cudaMemsetAsync(hs[e]->comm_buf.p, 0, total * sizeof(double), hs[e]->stream));
            chk[0] = 1e300; chk[1] = 1e300;        // poisons the agreement test on every rank
        }
        CK(cudaMemcpyAsync(hs[e]->comm_buf.p + total, chk, sizeof chk, cudaMemcpyHostToDevice, hs[e]->stream));
        CK(cudaStreamSynchronize(hs[e]->stream));
    }
    // two all-reduces: the check words with MAX, then the sums with SUM (both tiny: latency-bound over NVLink)
    {
        std::vector<DevBuf<double>> tmp((size_t)n);
        bool any = false;
        for (int e = 0; e < n; ++e) any = any || hs[e]->comm != nullptr;
        if (any) {
            NcclApi* api = nullptr;
            CKS(need_nccl(api));
            NK(api, api->GroupStart());
            for (int e = 0; e < n; ++e) {
                if (!hs[e]->comm) { api->GroupEnd(); return fail(SSPS_ERR_INVALID, "handle %d has no communicator", e); }
                CK(cudaSetDevice(hs[e]->device));
                NK(api, api->AllReduce(hs[e]->comm_buf.p + total, hs[e]->comm_buf.p + total, 2, NCCL_FLOAT64, NCCL_MAX, hs[e]->comm, hs[e]->stream));
            }
            NK(api, api->GroupEnd());
        } else if (n > 1) return fail(SSPS_ERR_INVALID, "%d handles but no communicator: call ssps_comm_init_all first", n);
    }
    for (int e = 0; e < n; ++e) {
        double chk[2];
        CK(cudaSetDevice(hs[e]->device));
        CK(cudaMemcpyAsync(chk, hs[e]->comm_buf.p + total, sizeof chk, cudaMemcpyDeviceToHost, hs[e]->stream));
        CK(cudaStreamSynchronize(hs[e]->stream));
        if (chk[0] >= 1e300) {
            if (first_err != SSPS_OK) { g_err = first_msg; return first_err; }
            return fail(SSPS_ERR_UNAVAILABLE, "another rank has no snapshot-aligned (burn-in, split, length) triple yet");
        }
        if (chk[0] != -chk[1])
            return fail(SSPS_ERR_UNAVAILABLE, "ranks would split their chains at different samples (half lengths %g..%g): run the same "
                        "number of sweeps with the same psrf_grid / snapshot_slots on every rank", -chk[1], chk[0]);
    }
    CKS(comm_allreduce(hs, n, total, NCCL_SUM));
    CKS(bind(hs[0]));
    return ssps_psrf_finish(hs[0], hs[0]->comm_buf.p, 1, max_edge_psrf, edge_psrf_child_major, lambda_psrf);
}

// evaluate_summary over the chains of ALL ranks (postprocess_samples.jl:183-191)
int ssps_edge_posterior_global(ssps_handle* const* hs, int32_t n, double burnin, double* probs_child_major, double* lambda_mean,
                               int64_t* burnin_used, int64_t* n_used, int64_t* n_chains_total) {
    CKS(check_handles(hs, n));
    const int V = hs[0]->V;
    const size_t vv = (size_t)V * V, total = vv + (size_t)V + 1;
    int64_t b0 = 0, n0 = 0;
    for (int e = 0; e < n; ++e) {
        int64_t b = 0, nn = 0;
        CKS(bind(hs[e]));
        CK(hs[e]->comm_buf.alloc((size_t)ssps_psrf_sums_len(V) + 2));
        CKS(posterior_sums(hs[e], burnin, hs[e]->comm_buf.p, &b, &nn));
        if (e == 0) { b0 = b; n0 = nn; }
        else if (b != b0 || nn != n0) return fail(SSPS_ERR_INVALID, "handles %d and 0 are at different samples (%lld/%lld vs %lld/%lld)", e, (long long)b, (long long)nn, (long long)b0, (long long)n0);
    }
    CKS(comm_allreduce(hs, n, total, NCCL_SUM));
    CKS(bind(hs[0]));
    std::vector<double> host(total);
    CKS(download(hs[0], host.data(), hs[0]->comm_buf.p, total));
    CK(cudaStreamSynchronize(hs[0]->stream));
    const double denom = host[vv + V];
    if (probs_child_major) for (size_t e = 0; e < vv; ++e) probs_child_major[e] = host[e] / denom;
    if (lambda_mean) for (int i = 0; i < V; ++i) lambda_mean[i] = host[vv + i] / denom;
    if (burnin_used) *burnin_used = b0;
    if (n_used) *n_used = n0;
    if (n_chains_total) *n_chains_total = (int64_t)llround(denom / (double)(n0 - b0));
    return SSPS_OK;
}

}  // extern "C"

// ================================================================================================
// K5: convergence statistics from the change-point logs (csrc/ssps_neff.cuh)
// ================================================================================================
namespace {

template <bool BINARY>
int run_neff(ssps_handle* h0, int Q, int C, const std::vector<long long>& off, const std::vector<int>& pos, const std::vector<double>& val,
             int b, int s, int len, double* mean_out, double* psrf_out, double* neff_out) {
    DevBuf<long long> d_off;
    DevBuf<int> d_pos;
    DevBuf<double> d_val, d_mean, d_psrf, d_neff;
    CKS(upload(h0, d_off, off.data(), off.size()));
    CKS(upload(h0, d_pos, pos.data(), std::max<size_t>(pos.size(), 1)));
    CKS(upload(h0, d_val, val.data(), std::max<size_t>(val.size(), 1)));
    CK(d_mean.alloc(Q)); CK(d_psrf.alloc(Q)); CK(d_neff.alloc(Q));
    NeffLog lg{d_off.p, d_pos.p, d_val.p, C};
    neff_kernel<BINARY><<<div_up((size_t)Q * 32, 128), 128, 0, h0->stream>>>(lg, Q, b, s, len, d_mean.p, d_psrf.p, d_neff.p);
    CKS(launch_ok(h0, "neff_kernel"));
    if (mean_out) CKS(download(h0, mean_out, d_mean.p, (size_t)Q));
    if (psrf_out) CKS(download(h0, psrf_out, d_psrf.p, (size_t)Q));
    if (neff_out) CKS(download(h0, neff_out, d_neff.p, (size_t)Q));
    CK(cudaStreamSynchronize(h0->stream));
    return SSPS_OK;
}

}  // namespace

extern "C" {

int ssps_convergence(ssps_handle* const* hs, int32_t n, double burnin, int64_t length, double* edge_mean, double* edge_psrf,
                     double* edge_neff, double* lambda_mean, double* lambda_psrf, double* lambda_neff, int64_t* burnin_index,
                     int64_t* split_index) {
    CKS(check_handles(hs, n));
    if (!(burnin >= 0.0 && burnin < 1.0)) return fail(SSPS_ERR_INVALID, "burnin must be in [0, 1)");
    const int V = hs[0]->V;
    int C = 0;
    for (int e = 0; e < n; ++e) {
        ssps_handle* h = hs[e];
        if (!h->cfg.record_changepoints) return fail(SSPS_ERR_UNAVAILABLE, "ssps_convergence needs the change-point log (record_changepoints = 1)");
        if (h->n_chains <= 0) return fail(SSPS_ERR_INVALID, "handle %d has no chains", e);
        if (h->samples_recorded != hs[0]->samples_recorded) return fail(SSPS_ERR_INVALID, "handles 0 and %d recorded %lld and %lld samples", e, (long long)hs[0]->samples_recorded, (long long)h->samples_recorded);
        CKS(bind(h));
        if (h->cp_parent.empty()) CKS(changepoints_begin(h));
        C += h->n_chains;
    }
    const int64_t len = length > 0 ? length : hs[0]->samples_recorded;
    if (len > hs[0]->samples_recorded) return fail(SSPS_ERR_INVALID, "stop point %lld beyond the %lld recorded samples", (long long)len, (long long)hs[0]->samples_recorded);
    const int64_t b = (int64_t)floor(burnin * (double)len);            // postprocess_samples.jl:167
    const int64_t s = b + (len - b) / 2;                               // :168
    if (s - b < 3 || len - s < 3) return fail(SSPS_ERR_UNAVAILABLE, "%lld samples are too few for split-chain statistics", (long long)len);
    if (len > 2000000000LL) return fail(SSPS_ERR_UNSUPPORTED, "stop point beyond 2^31 samples");
    if (burnin_index) *burnin_index = b;
    if (split_index) *split_index = s;
    CKS(bind(hs[0]));
    // ---- edge indicators: Q = V * V sequences per chain ----
    if (edge_mean || edge_psrf || edge_neff) {
        const size_t Q = (size_t)V * V;
        std::vector<long long> off(Q * C + 1, 0);
        int c0 = 0;
        for (int e = 0; e < n; ++e) {
            for (int c = 0; c < hs[e]->n_chains; ++c)
                for (const ParentEvent& ev : hs[e]->cp_parent[c]) ++off[((size_t)ev.child * V + ev.parent) * C + (c0 + c) + 1];
            c0 += hs[e]->n_chains;
        }
        for (size_t i = 0; i < Q * C; ++i) off[i + 1] += off[i];
        std::vector<int> pos((size_t)off[Q * C]);
        std::vector<double> val((size_t)off[Q * C]);
        std::vector<long long> fill(off.begin(), off.end() - 1);
        c0 = 0;
        for (int e = 0; e < n; ++e) {
            for (int c = 0; c < hs[e]->n_chains; ++c)
                for (const ParentEvent& ev : hs[e]->cp_parent[c]) {      // time order inside every (child, parent): kept by the stable fill
                    const long long at = fill[((size_t)ev.child * V + ev.parent) * C + (c0 + c)]++;
                    pos[(size_t)at] = (int)ev.n; val[(size_t)at] = (double)ev.value;
                }
            c0 += hs[e]->n_chains;
        }
        CKS(run_neff<true>(hs[0], (int)Q, C, off, pos, val, (int)b, (int)s, (int)len, edge_mean, edge_psrf, edge_neff));
    }
    // ---- lambda_i: Q = V real-valued sequences per chain ----
    if (lambda_mean || lambda_psrf || lambda_neff) {
        const size_t Q = (size_t)V;
        std::vector<long long> off(Q * C + 1, 0);
        int c0 = 0;
        for (int e = 0; e < n; ++e) {
            for (int c = 0; c < hs[e]->n_chains; ++c)
                for (const LambdaEvent& ev : hs[e]->cp_lambda[c]) ++off[(size_t)ev.child * C + (c0 + c) + 1];
            c0 += hs[e]->n_chains;
        }
        for (size_t i = 0; i < Q * C; ++i) off[i + 1] += off[i];
        std::vector<int> pos((size_t)off[Q * C]);
        std::vector<double> val((size_t)off[Q * C]);
        std::vector<long long> fill(off.begin(), off.end() - 1);
        c0 = 0;
        for (int e = 0; e < n; ++e) {
            for (int c = 0; c < hs[e]->n_chains; ++c)
                for (const LambdaEvent& ev : hs[e]->cp_lambda[c]) {
                    const long long at = fill[(size_t)ev.child * C + (c0 + c)]++;
                    pos[(size_t)at] = (int)ev.n; val[(size_t)at] = ev.value;
                }
            c0 += hs[e]->n_chains;
        }
        CKS(run_neff<false>(hs[0], (int)Q, C, off, pos, val, (int)b, (int)s, (int)len, lambda_mean, lambda_psrf, lambda_neff));
    }
    return SSPS_OK;
}

}  // extern "C"

// ================================================================================================
// K2 without allocation (ssps_score_static).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (graph capture, serving loops).  The per-thread factor scratch of
// the scorer lives in ONE file-scope device array of a fixed size; everything else is the caller's device memory.
// ================================================================================================
namespace {

// Held by every GPU that loads the library until the process exits.  Shared by all calls on a device: calls that
// overlap on the GPU race on it (ssps.h).  Keeps the last call's contents; score_set writes every element before it
// reads it (row a of the factor, w(a), the ids), so nothing has to start at zero.
static __device__ __align__(256) unsigned char g_score_static_scratch[SSPS_SCORE_STATIC_SCRATCH_BYTES];

// In host code the name above is a host-side stand-in, not a device address: the address the kernel gets comes from
// cudaGetSymbolAddress, once per device, with that device current.  Fixed table: no host allocation either.
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kStaticMaxDevices] = {};

int static_scratch_ptr(int device, unsigned char** out) {
    if (device < 0 || device >= kStaticMaxDevices) return fail(SSPS_ERR_UNSUPPORTED, "ssps_score_static: device %d beyond the table of %d", device, kStaticMaxDevices);
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[device]) {
        int cur = -1;
        CK(cudaGetDevice(&cur));
        if (cur != device) return fail(SSPS_ERR_CUDA, "ssps_score_static: device %d is current, the handle lives on %d", cur, device);
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_score_static_scratch);
        if (e != cudaSuccess || !p) return fail(SSPS_ERR_CUDA, "cudaGetSymbolAddress(g_score_static_scratch): %s", cudaGetErrorString(e));
        g_static_ptr[device] = p;
    }
    *out = (unsigned char*)g_static_ptr[device];
    return SSPS_OK;
}

// bytes of scratch one thread needs for sets of at most kcap parents: factor rows + w (fp64), ids (uint16)
inline size_t static_bytes_per_thread(int kcap) { return (size_t)(SSPS_TRI(kcap) + kcap) * sizeof(double) + (size_t)kcap * sizeof(uint16_t); }

// A fixed number of threads walks the sets grid-stride; thread t owns column t of the interleaved scratch (element e at
// [e * stride + t], the layout of score_kernel's arena).  The sets come from device memory the host has not seen, so
// the checks ssps_score makes on the host are made here: an invalid set gets NaN, the scratch is never indexed by it.
__global__ void score_static_kernel(SspsDev P, int n_sets, int kcap, const int* __restrict__ child, const int* __restrict__ offsets,
                                    const int* __restrict__ members, const double* __restrict__ lambda,
                                    double* __restrict__ arena_d, uint16_t* __restrict__ arena_t,
                                    double* __restrict__ lml_out, double* __restrict__ lp_out, double* __restrict__ minp_out) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int stride = gridDim.x * blockDim.x;
    Scratch sc{arena_d + tid, arena_t + tid, stride, kcap};
    const double nan = __longlong_as_double(0x7ff8000000000000LL);
    for (int s = tid; s < n_sets; s += stride) {
        const int i = child[s];
        const long long o0 = offsets[s], o1 = offsets[s + 1];
        const long long kk = o1 - o0;
        bool ok = i >= 0 && i < P.V && o0 >= 0 && kk >= 0 && kk <= (long long)kcap;
        const int k = ok ? (int)kk : 0;
        const int* mem = members + (ok ? o0 : 0);
        for (int a = 0; a < k; ++a) {
            const int j = mem[a];
            if (j < 0 || j >= P.V || (a > 0 && j <= mem[a - 1])) { ok = false; break; }
            sc.T(a) = (uint16_t)j;
        }
        if (!ok) {
            if (lml_out) lml_out[s] = nan;
            if (minp_out) minp_out[s] = nan;
            if (lp_out) lp_out[s] = nan;
            continue;
        }
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
}

}  // namespace

extern "C" {

int ssps_score_static(ssps_handle* h, int32_t n_sets, int32_t max_parents, const int32_t* d_child, const int32_t* d_set_offsets,
                      const int32_t* d_set_members, const double* d_lambda, double* d_lml_out, double* d_logprior_out,
                      double* d_min_pivot_out) {
    CKS(bind(h));
    if (!h->has_data) return fail(SSPS_ERR_INVALID, "ssps_score_static before ssps_set_data");
    if (h->is_poly) return fail(SSPS_ERR_UNSUPPORTED, "ssps_score_static: regression_deg=%d scores every set from scratch and needs far more scratch "
                                "than the fixed array of %u bytes; use ssps_score", h->cfg.regression_deg, (unsigned)SSPS_SCORE_STATIC_SCRATCH_BYTES);
    if (n_sets < 0 || n_sets > SSPS_SCORE_STATIC_MAX_SETS)
        return fail(SSPS_ERR_INVALID, "ssps_score_static: n_sets=%d outside [0, %d]", n_sets, (int)SSPS_SCORE_STATIC_MAX_SETS);
    if (max_parents < 0 || max_parents > SSPS_SCORE_STATIC_MAX_PARENTS)
        return fail(SSPS_ERR_INVALID, "ssps_score_static: max_parents=%d outside [0, %d] (the fixed scratch array of %u bytes is sized for that; "
                    "use ssps_score for larger sets)", max_parents, (int)SSPS_SCORE_STATIC_MAX_PARENTS, (unsigned)SSPS_SCORE_STATIC_SCRATCH_BYTES);
    if (n_sets > 0 && (!d_child || !d_set_offsets || (max_parents > 0 && !d_set_members))) return fail(SSPS_ERR_INVALID, "ssps_score_static: null input");
    if (d_logprior_out && !d_lambda) return fail(SSPS_ERR_INVALID, "ssps_score_static: d_logprior_out needs d_lambda");
    unsigned char* base = nullptr;
    CKS(static_scratch_ptr(h->device, &base));          // also on an empty batch: the warm-up call before a graph capture
    if (n_sets == 0) return SSPS_OK;
    const int kcap = std::max(max_parents, 1);
    const int block = 128;
    const size_t per_thread = static_bytes_per_thread(kcap);
    const size_t fit = (size_t)SSPS_SCORE_STATIC_SCRATCH_BYTES / per_thread / block;         // CTAs the array has room for
    if (fit < 1) return fail(SSPS_ERR_UNSUPPORTED, "ssps_score_static: scratch array too small for one CTA at max_parents=%d", max_parents);
    const int grid = (int)std::min<size_t>(std::min<size_t>(fit, (size_t)div_up((size_t)n_sets, block)), (size_t)h->sm_count * 8);
    const size_t nthreads = (size_t)grid * block;
    double* arena_d = (double*)base;                                                          // 256-byte aligned
    uint16_t* arena_t = (uint16_t*)(base + nthreads * (size_t)(SSPS_TRI(kcap) + kcap) * sizeof(double));
    score_static_kernel<<<grid, block, 0, h->stream>>>(h->dev, n_sets, kcap, d_child, d_set_offsets, d_set_members, d_lambda, arena_d, arena_t,
                                                       d_lml_out, d_logprior_out, d_min_pivot_out);
    return launch_ok(h, "score_static_kernel");
}

}  // extern "C"
