"""ctypes front-end of the CPU ORACLE (oracle/ssps_oracle.c).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under ssps_b200/ imports this module.
PARITY UNPINNED: see the header of ssps_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libssps_oracle.so")

SCORER_CHOL = 0
SCORER_QR = 1
FLAG_LML_CACHE = 1
FLAG_STALE_PRIOR = 2
FLAG_UNIFORM = 4

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_lp = C.POINTER(C.c_int64)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ssps_oracle.c")
    hdr = os.path.join(_HERE, "ssps_oracle.h")
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.exists(p) and os.path.getmtime(p) > os.path.getmtime(_LIB_PATH) for p in (src, hdr))
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_problem_create.restype = C.c_void_p
        L.orc_problem_create.argtypes = [_dp, _dp, C.c_int, C.c_int, _dp, C.c_double, C.c_double,
                                         C.c_double, C.c_int]
        L.orc_problem_destroy.argtypes = [C.c_void_p]
        for name in ("G", "R", "Z", "Y", "yty", "prob_param"):
            f = getattr(L, "orc_problem_" + name)
            f.restype = _dp
            f.argtypes = [C.c_void_p]
        L.orc_problem_lambda_min.restype = C.c_double
        L.orc_problem_lambda_min.argtypes = [C.c_void_p]
        L.orc_problem_lml.restype = C.c_double
        L.orc_problem_lml.argtypes = [C.c_void_p, C.c_int, _ip, C.c_int, C.c_int]
        L.orc_problem_logprior.restype = C.c_double
        L.orc_problem_logprior.argtypes = [C.c_void_p, C.c_int, _ip, C.c_int, C.c_double, C.c_int]
        L.orc_lml_qr.restype = C.c_double
        L.orc_lml_qr.argtypes = [_dp, _dp, C.c_int, _ip, C.c_int, C.c_int, _ip]
        L.orc_lml_chol.restype = C.c_double
        L.orc_lml_chol.argtypes = [_dp, _dp, C.c_double, C.c_int, C.c_int, _ip, C.c_int, _dp, _ip]
        L.orc_lml_hill.restype = C.c_double
        L.orc_lml_hill.argtypes = [_dp, _dp, C.c_int, _ip, C.c_int]
        L.orc_logprior_ref.restype = C.c_double
        L.orc_logprior_ref.argtypes = [_ip, C.c_int, C.c_int, _dp, C.c_double]
        L.orc_logprior_closed.restype = C.c_double
        L.orc_logprior_closed.argtypes = [_ip, C.c_int, C.c_int, _dp, C.c_double]
        L.orc_move_probs.argtypes = [C.c_int, C.c_int, C.c_double, _dp]
        L.orc_ith_nonparent.restype = C.c_int
        L.orc_ith_nonparent.argtypes = [C.c_int, _ip, C.c_int]
        L.orc_searchsorted_exclude.restype = C.c_int
        L.orc_searchsorted_exclude.argtypes = [_ip, C.c_int, C.c_int]
        L.orc_n_design_cols.restype = C.c_int
        L.orc_n_design_cols.argtypes = [C.c_int, C.c_int]
        L.orc_construct_B.argtypes = [_dp, C.c_int, _ip, C.c_int, C.c_int, _dp]
        L.orc_standardize.argtypes = [_dp, C.c_int, C.c_int, _dp]
        L.orc_gram.argtypes = [_dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp]
        L.orc_philox4x32_10.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.orc_step_draws.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, _dp]
        L.orc_chain_create.restype = C.c_void_p
        L.orc_chain_create.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_chain_destroy.argtypes = [C.c_void_p]
        L.orc_chain_init.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32]
        L.orc_chain_set_state.argtypes = [C.c_void_p, _ip, _ip, _dp]
        L.orc_chain_sweep.argtypes = [C.c_void_p, _dp, _ip]
        L.orc_chain_run.argtypes = [C.c_void_p, C.c_int64]
        L.orc_chain_sweeps_done.restype = C.c_int64
        L.orc_chain_sweeps_done.argtypes = [C.c_void_p]
        L.orc_chain_indegree.restype = C.c_int
        L.orc_chain_indegree.argtypes = [C.c_void_p, C.c_int]
        L.orc_chain_parents.argtypes = [C.c_void_p, C.c_int, _ip]
        L.orc_chain_lambda.restype = C.c_double
        L.orc_chain_lambda.argtypes = [C.c_void_p, C.c_int]
        L.orc_chain_lml.restype = C.c_double
        L.orc_chain_lml.argtypes = [C.c_void_p, C.c_int]
        for name in ("n_proposals", "n_accepted", "cache_hits", "recorded_n"):
            f = getattr(L, "orc_chain_" + name)
            f.restype = C.c_int64
            f.argtypes = [C.c_void_p]
        L.orc_chain_k_hist.argtypes = [C.c_void_p, _lp, C.c_int]
        L.orc_chain_recorder_enable.argtypes = [C.c_void_p]
        L.orc_chain_record.argtypes = [C.c_void_p]
        L.orc_chain_n_parent_events.restype = C.c_int64
        L.orc_chain_n_parent_events.argtypes = [C.c_void_p, C.c_int]
        L.orc_chain_parent_events.argtypes = [C.c_void_p, C.c_int, _lp, _ip, _ip]
        L.orc_chain_n_lambda_events.restype = C.c_int64
        L.orc_chain_n_lambda_events.argtypes = [C.c_void_p, C.c_int]
        L.orc_chain_lambda_events.argtypes = [C.c_void_p, C.c_int, _lp, _dp]
        L.orc_chain_edge_counts.argtypes = [C.c_void_p, _lp]
        L.orc_psrf_from_counts.restype = C.c_double
        L.orc_psrf_from_counts.argtypes = [_lp, _lp, C.c_int, C.c_int64, C.c_int64]
        L.orc_psrf_binary_dense.restype = C.c_double
        L.orc_psrf_binary_dense.argtypes = [C.POINTER(C.POINTER(C.c_uint8)), C.c_int, C.c_int64, C.c_double]
        L.orc_neff_binary_dense.restype = C.c_double
        L.orc_neff_binary_dense.argtypes = [C.POINTER(C.POINTER(C.c_uint8)), C.c_int, C.c_int64, C.c_double]
        L.orc_psrf_real_dense.restype = C.c_double
        L.orc_psrf_real_dense.argtypes = [C.POINTER(_dp), C.c_int, C.c_int64, C.c_double]
        L.orc_run_chains_mt.restype = C.c_double
        L.orc_run_chains_mt.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64,
                                        C.c_uint64, _lp, _lp]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _l(a):
    return a.ctypes.data_as(_lp)


def _fcol(a):
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def _ints(S):
    return np.ascontiguousarray(np.asarray(S, dtype=np.int32).reshape(-1))


# --------------------------------------------------------------------------------------------
# stateless helpers
# --------------------------------------------------------------------------------------------
def standardize(X):
    X = _fcol(X)
    n, V = X.shape
    Z = np.empty_like(X, order="F")
    lib().orc_standardize(_d(X), n, V, _d(Z))
    return Z


def gram(Z, Y):
    Z = _fcol(Z); Y = _fcol(Y)
    n, V = Z.shape
    G = np.empty((V, V), order="F"); R = np.empty((V, V), order="F"); yty = np.empty(V)
    lib().orc_gram(_d(Z), _d(Y), n, V, _d(G), _d(R), _d(yty))
    return G, R, yty


def construct_B(Xminus, S, deg):
    Xminus = _fcol(Xminus)
    n = Xminus.shape[0]
    S = _ints(S)
    w = lib().orc_n_design_cols(len(S), deg)
    B = np.empty((n, max(w, 0)), order="F")
    if w > 0:
        lib().orc_construct_B(_d(Xminus), n, _i(S), len(S), deg, _d(B))
    return B


def lml_qr(Xminus, y, S, deg=1, return_rank=False):
    Xminus = _fcol(Xminus)
    y = np.ascontiguousarray(y, dtype=np.float64)
    S = _ints(S)
    rank = C.c_int(0)
    v = lib().orc_lml_qr(_d(Xminus), _d(y), Xminus.shape[0], _i(S), len(S), deg, C.byref(rank))
    return (v, rank.value) if return_rank else v


def lml_chol(G, Ri, yty, n, S, return_diag=False):
    G = _fcol(G)
    Ri = np.ascontiguousarray(Ri, dtype=np.float64)
    S = _ints(S)
    minr = C.c_double(0.0); skipped = C.c_int(0)
    v = lib().orc_lml_chol(_d(G), _d(Ri), float(yty), int(n), G.shape[0], _i(S), len(S),
                           C.byref(minr), C.byref(skipped))
    return (v, minr.value, skipped.value) if return_diag else v


def lml_hill(Xminus, y, S):
    Xminus = _fcol(Xminus)
    y = np.ascontiguousarray(y, dtype=np.float64)
    S = _ints(S)
    return lib().orc_lml_hill(_d(Xminus), _d(y), Xminus.shape[0], _i(S), len(S))


def logprior_ref(S, V, conf_col, lam):
    S = _ints(S); cc = np.ascontiguousarray(conf_col, dtype=np.float64)
    return lib().orc_logprior_ref(_i(S), len(S), V, _d(cc), float(lam))


def logprior_closed(S, V, conf_col, lam):
    S = _ints(S); cc = np.ascontiguousarray(conf_col, dtype=np.float64)
    return lib().orc_logprior_closed(_i(S), len(S), V, _d(cc), float(lam))


def move_probs(k, V, prob_param):
    out = np.empty(3)
    lib().orc_move_probs(int(k), int(V), float(prob_param), _d(out))
    return out


def ith_nonparent(i, parents_1based):
    S = _ints(parents_1based)
    return lib().orc_ith_nonparent(int(i), _i(S), len(S))


def searchsorted_exclude(parents_1based, x):
    S = _ints(parents_1based)
    return lib().orc_searchsorted_exclude(_i(S), len(S), int(x))


def philox(ctr, key):
    c = (C.c_uint32 * 4)(*ctr); k = (C.c_uint32 * 2)(*key); o = (C.c_uint32 * 4)()
    lib().orc_philox4x32_10(c, k, o)
    return list(o)


def step_draws(seed, chain, vertex, sweep):
    out = np.empty(6)
    lib().orc_step_draws(int(seed), int(chain), int(vertex), int(sweep), _d(out))
    return out


# --------------------------------------------------------------------------------------------
class Problem:
    """Data + derived parameters (SSPS.jl:155-169)."""

    def __init__(self, Xminus, Xplus, prior, large_indeg=20.0, lambda_max=15.0, lambda_prop_std=0.25,
                 regression_deg=1):
        self.Xminus = _fcol(Xminus); self.Xplus = _fcol(Xplus); self.prior = _fcol(prior)
        self.n, self.V = self.Xminus.shape
        assert self.Xplus.shape == (self.n, self.V) and self.prior.shape == (self.V, self.V)
        self.regression_deg = regression_deg
        self.lambda_max = lambda_max; self.lambda_prop_std = lambda_prop_std; self.large_indeg = large_indeg
        self._h = lib().orc_problem_create(_d(self.Xminus), _d(self.Xplus), self.n, self.V, _d(self.prior),
                                           float(large_indeg), float(lambda_max), float(lambda_prop_std),
                                           int(regression_deg))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_problem_destroy(self._h); self._h = None

    def _mat(self, name, shape):
        p = getattr(lib(), "orc_problem_" + name)(self._h)
        return np.ctypeslib.as_array(p, shape=(int(np.prod(shape)),)).reshape(shape, order="F").copy(order="F")

    @property
    def G(self): return self._mat("G", (self.V, self.V))
    @property
    def R(self): return self._mat("R", (self.V, self.V))
    @property
    def Z(self): return self._mat("Z", (self.n, self.V))
    @property
    def Y(self): return self._mat("Y", (self.n, self.V))
    @property
    def yty(self): return self._mat("yty", (self.V,))
    @property
    def prob_param(self): return self._mat("prob_param", (self.V,))
    @property
    def lambda_min(self): return lib().orc_problem_lambda_min(self._h)

    def lml(self, child, S, scorer=SCORER_QR):
        S = _ints(S)
        return lib().orc_problem_lml(self._h, int(child), _i(S), len(S), scorer)

    def logprior(self, child, S, lam, form=0):
        """form 0: reference two-branch formula; 1: closed form; 2: grouped form (kernel order)."""
        S = _ints(S)
        return lib().orc_problem_logprior(self._h, int(child), _i(S), len(S), float(lam), form)

    def run_chains_mt(self, n_chains, n_threads, n_sweeps, seed=1, scorer=SCORER_QR,
                      flags=FLAG_LML_CACHE, want_counts=True):
        counts = np.zeros((self.V, self.V), dtype=np.int64) if want_counts else None
        nprop = C.c_int64(0)
        secs = lib().orc_run_chains_mt(self._h, scorer, flags, n_chains, n_threads, int(n_sweeps), int(seed),
                                       _l(counts) if want_counts else None, C.byref(nprop))
        return secs, nprop.value, counts


class Chain:
    def __init__(self, problem: Problem, scorer=SCORER_CHOL, flags=0):
        self.P = problem
        self._h = lib().orc_chain_create(problem._h, scorer, flags)
        self.V = problem.V

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_chain_destroy(self._h); self._h = None

    def init(self, seed, chain_id):
        lib().orc_chain_init(self._h, int(seed), int(chain_id))

    def set_state(self, parent_sets, lambdas):
        offs = np.zeros(self.V + 1, dtype=np.int32)
        for i, ps in enumerate(parent_sets):
            offs[i + 1] = offs[i] + len(ps)
        mem = _ints([j for ps in parent_sets for j in sorted(ps)]) if offs[-1] else np.zeros(1, np.int32)
        lam = np.ascontiguousarray(lambdas, dtype=np.float64)
        lib().orc_chain_set_state(self._h, _i(offs), _i(mem), _d(lam))

    def sweep(self, draws=None):
        acc = np.zeros((self.V, 3), dtype=np.int32)
        if draws is not None:
            draws = np.ascontiguousarray(draws, dtype=np.float64).reshape(self.V, 6)
            lib().orc_chain_sweep(self._h, _d(draws), _i(acc))
        else:
            lib().orc_chain_sweep(self._h, None, _i(acc))
        return acc

    def run(self, n_sweeps):
        lib().orc_chain_run(self._h, int(n_sweeps))

    def parents(self, child):
        k = lib().orc_chain_indegree(self._h, child)
        out = np.zeros(max(k, 1), dtype=np.int32)
        lib().orc_chain_parents(self._h, child, _i(out))
        return out[:k].tolist()

    def parent_sets(self):
        return [self.parents(i) for i in range(self.V)]

    def lambdas(self):
        return np.array([lib().orc_chain_lambda(self._h, i) for i in range(self.V)])

    def lmls(self):
        return np.array([lib().orc_chain_lml(self._h, i) for i in range(self.V)])

    @property
    def n_proposals(self): return lib().orc_chain_n_proposals(self._h)
    @property
    def n_accepted(self): return lib().orc_chain_n_accepted(self._h)
    @property
    def cache_hits(self): return lib().orc_chain_cache_hits(self._h)
    @property
    def recorded_n(self): return lib().orc_chain_recorded_n(self._h)

    def k_hist(self):
        h = np.zeros(self.V + 1, dtype=np.int64)
        lib().orc_chain_k_hist(self._h, _l(h), len(h))
        return h

    def recorder_enable(self):
        lib().orc_chain_recorder_enable(self._h)

    def record(self):
        lib().orc_chain_record(self._h)

    def parent_events(self, child):
        m = lib().orc_chain_n_parent_events(self._h, child)
        n = np.zeros(max(m, 1), np.int64); p = np.zeros(max(m, 1), np.int32); v = np.zeros(max(m, 1), np.int32)
        lib().orc_chain_parent_events(self._h, child, _l(n), _i(p), _i(v))
        return n[:m], p[:m], v[:m]

    def lambda_events(self, child):
        m = lib().orc_chain_n_lambda_events(self._h, child)
        n = np.zeros(max(m, 1), np.int64); v = np.zeros(max(m, 1), np.float64)
        lib().orc_chain_lambda_events(self._h, child, _l(n), _d(v))
        return n[:m], v[:m]

    def edge_counts(self):
        out = np.zeros((self.V, self.V), dtype=np.int64)
        lib().orc_chain_edge_counts(self._h, _l(out))
        return out

    def results_dict(self, t_elapsed=0.0):
        """The reference's per-chain JSON object (state_updates.jl:167-179, SSPS.jl:223-230); 1-based keys."""
        ps = []
        for i in range(self.V):
            n, p, v = self.parent_events(i)
            d = {}
            for a in range(len(n)):
                d.setdefault(str(int(p[a]) + 1), []).append([int(n[a]), int(v[a])])
            ps.append(d)
        lam = []
        for i in range(self.V):
            n, v = self.lambda_events(i)
            lam.append([[int(n[a]), float(v[a])] for a in range(len(n))])
        return {"parent_sets": ps, "lambda": lam, "n": int(self.recorded_n), "t_elapsed": float(t_elapsed)}


def psrf_from_counts(c1, c2, N1, N2):
    c1 = np.ascontiguousarray(c1, dtype=np.int64); c2 = np.ascontiguousarray(c2, dtype=np.int64)
    return lib().orc_psrf_from_counts(_l(c1), _l(c2), len(c1), int(N1), int(N2))


def _u8_ptrs(seqs):
    arrs = [np.ascontiguousarray(s, dtype=np.uint8) for s in seqs]
    ptrs = (C.POINTER(C.c_uint8) * len(arrs))(*[a.ctypes.data_as(C.POINTER(C.c_uint8)) for a in arrs])
    return arrs, ptrs


def psrf_binary_dense(seqs, burnin=0.5):
    arrs, ptrs = _u8_ptrs(seqs)
    return lib().orc_psrf_binary_dense(ptrs, len(arrs), len(arrs[0]), float(burnin))


def neff_binary_dense(seqs, burnin=0.5):
    arrs, ptrs = _u8_ptrs(seqs)
    return lib().orc_neff_binary_dense(ptrs, len(arrs), len(arrs[0]), float(burnin))


def psrf_real_dense(seqs, burnin=0.5):
    arrs = [np.ascontiguousarray(s, dtype=np.float64) for s in seqs]
    ptrs = (_dp * len(arrs))(*[_d(a) for a in arrs])
    return lib().orc_psrf_real_dense(ptrs, len(arrs), len(arrs[0]), float(burnin))
