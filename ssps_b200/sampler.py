"""Host-side handle over the C ABI (include/ssps.h): one `Sampler` = one ssps_handle = one GPU.

Method names follow the reference's functions for this path so that tests read like the
reference's own code:

    Sampler.set_data          <- combine_X / standardize_X / vectorize_X  (SSPS/src/dbn_preprocess.jl:70-108)
    Sampler.log_marg_lik      <- log_marg_lik / compute_lml                (SSPS/src/dbn_distributions.jl:235-264)
    Sampler.conf_parents_logpdf <- logpdf(::ConfParentPrior, ...)          (SSPS/src/dbn_distributions.jl:96-153)
    Sampler.generate          <- Gen.generate(vertex_lambda_dbn_model)     (SSPS/src/mcmc_inference.jl:40)
    Sampler.update_loop       <- vertex_lambda_update_loop x n sweeps      (SSPS/src/state_updates.jl:35-51)
    Sampler.results           <- update_results_storediff                  (SSPS/src/state_updates.jl:161-227)

Indices are 0-based here (the 1-based shift happens only where reference file formats are
written: ssps_b200.results_json).  numpy arrays cross the boundary as plain pointers.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import Config, RunStats, check

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return None if a is None else a.ctypes.data_as(_ip)


def _l(a):
    return None if a is None else a.ctypes.data_as(_lp)


def _fcol(a) -> np.ndarray:
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def sets_to_csr(sets: Sequence[Sequence[int]]):
    offs = np.zeros(len(sets) + 1, dtype=np.int32)
    for s, members in enumerate(sets):
        offs[s + 1] = offs[s] + len(members)
    mem = np.zeros(max(int(offs[-1]), 1), dtype=np.int32)
    pos = 0
    for members in sets:
        m = len(members)
        mem[pos:pos + m] = np.asarray(members, dtype=np.int32)
        pos += m
    return offs, mem


def release_cached_memory(device: int = 0) -> None:
    """ssps_release_cached_memory: hand the device blocks kept from destroyed handles back to the driver."""
    check(_lib.load().ssps_release_cached_memory(int(device)))


class Sampler:
    """One GPU's worth of SSPS chains behind libssps_b200.so."""

    def __init__(self, device: int = -1, regression_deg: int = 1, large_indeg: float = 20.0,
                 lambda_max: float = 15.0, lambda_prop_std: float = 0.25, max_indegree: int = 0,
                 psrf_grid: int = 0, record_changepoints: bool = False, collect_khist: bool = False,
                 verify: bool = False, min_burnin: float = 0.0, snapshot_slots: int = 0, proposal: str = "sparse"):
        self._L = _lib.load()
        cfg = Config()
        cfg.struct_size = C.sizeof(Config)
        cfg.device = device
        cfg.regression_deg = regression_deg
        cfg.max_indegree = max_indegree
        cfg.large_indeg = large_indeg
        cfg.lambda_max = lambda_max
        cfg.lambda_prop_std = lambda_prop_std
        cfg.psrf_grid = psrf_grid
        cfg.record_changepoints = int(record_changepoints)
        cfg.collect_khist = int(collect_khist)
        cfg.verify = int(verify)
        cfg.min_burnin = float(min_burnin)
        cfg.snapshot_slots = int(snapshot_slots)
        if proposal not in ("sparse", "uniform"):
            raise ValueError("proposal must be 'sparse' or 'uniform' (SSPS.jl:87-88)")
        cfg.proposal = 1 if proposal == "uniform" else 0
        self._h = C.c_void_p()
        check(self._L.ssps_create(C.byref(cfg), C.byref(self._h)))
        self.cfg = cfg
        self.n = self.V = 0
        self.n_chains = 0

    # ------------------------------------------------------------------ lifetime
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._L.ssps_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_stream(self, cuda_stream: int) -> None:
        """cuda_stream: a raw cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream)."""
        check(self._L.ssps_set_stream(self._h, C.c_void_p(cuda_stream)))

    # ------------------------------------------------------------------ K1
    def set_data(self, Xminus, Xplus, prior) -> None:
        Xm, Xp, pr = _fcol(Xminus), _fcol(Xplus), _fcol(prior)
        n, V = Xm.shape
        if Xp.shape != (n, V) or pr.shape != (V, V):
            raise ValueError("Xminus %s, Xplus %s, prior %s do not agree" % (Xm.shape, Xp.shape, pr.shape))
        check(self._L.ssps_set_data(self._h, _d(Xm), _d(Xp), n, V, _d(pr)))
        self.n, self.V, self.n_chains = n, V, 0

    def set_timeseries(self, series: Sequence[np.ndarray], prior) -> None:
        """series: list of (T_m x V) arrays, rows already sorted by timestep (ingest_timeseries!)."""
        X = _fcol(np.concatenate([np.asarray(s, dtype=np.float64) for s in series], axis=0))
        lens = np.asarray([np.asarray(s).shape[0] for s in series], dtype=np.int32)
        pr = _fcol(prior)
        V = X.shape[1]
        if pr.shape != (V, V):
            raise ValueError("prior %s does not match V=%d" % (pr.shape, V))
        check(self._L.ssps_set_timeseries(self._h, _d(X), _i(lens), len(lens), V, _d(pr)))
        self.n, self.V, self.n_chains = int(lens.sum() - len(lens)), V, 0

    def design(self):
        Z = np.empty((self.n, self.V), order="F")
        Y = np.empty((self.n, self.V), order="F")
        check(self._L.ssps_get_design(self._h, _d(Z), _d(Y)))
        return Z, Y

    def gram(self):
        G = np.empty((self.V, self.V), order="F")
        R = np.empty((self.V, self.V), order="F")
        yty = np.empty(self.V)
        check(self._L.ssps_get_gram(self._h, _d(G), _d(R), _d(yty)))
        return G, R, yty

    def derived(self):
        lmin = C.c_double(0.0)
        pp = np.empty(self.V)
        check(self._L.ssps_get_derived(self._h, C.byref(lmin), _d(pp)))
        return lmin.value, pp

    # ------------------------------------------------------------------ K2
    def score(self, children: Sequence[int], sets: Sequence[Sequence[int]], lambdas=None, want_min_pivot=False):
        """Batched (child, S) -> (lml, logprior[, min pivot ratio]); logprior is None without lambdas."""
        ns = len(sets)
        ch = np.ascontiguousarray(children, dtype=np.int32)
        offs, mem = sets_to_csr(sets)
        lam = None if lambdas is None else np.ascontiguousarray(lambdas, dtype=np.float64)
        lml = np.empty(ns)
        lp = None if lam is None else np.empty(ns)
        mp = np.empty(ns) if want_min_pivot else None
        check(self._L.ssps_score(self._h, ns, _i(ch), _i(offs), _i(mem), _d(lam), _d(lml), _d(lp), _d(mp)))
        return (lml, lp, mp) if want_min_pivot else (lml, lp)

    def score_static(self, n_sets: int, max_parents: int, d_child: int, d_set_offsets: int, d_set_members: int,
                     d_lambda: Optional[int] = None, d_lml_out: Optional[int] = None, d_logprior_out: Optional[int] = None,
                     d_min_pivot_out: Optional[int] = None) -> None:
        """ssps_score_static: the scores of score() from DEVICE buffers of the caller (addresses as ints; int32 CSR, fp64
        outputs), the factor scratch from a fixed array inside the library: no allocation anywhere in the call.  Enqueued on
        the sampler's stream and NOT synchronised; at most 64 parents per set (include/ssps.h).  No counterpart in the
        reference (added for callers that cannot call cudaMalloc)."""
        check(self._L.ssps_score_static(self._h, n_sets, max_parents, d_child, d_set_offsets, d_set_members, d_lambda, d_lml_out,
                                        d_logprior_out, d_min_pivot_out))

    def log_marg_lik(self, ind: int, parent_inds: Sequence[int]) -> float:
        return float(self.score([ind], [sorted(parent_inds)])[0][0])

    def conf_parents_logpdf(self, parents: Sequence[int], idx: int, lambda_: float) -> float:
        return float(self.score([idx], [sorted(parents)], [lambda_])[1][0])

    # ------------------------------------------------------------------ K3
    def generate(self, n_chains: int, seed: int = 0, chain_id_offset: int = 0,
                 init_parent_sets: Optional[Sequence[Sequence[Sequence[int]]]] = None, init_lambda=None) -> None:
        """Initial trace of every chain; explicit state = resume (recover_state!, SSPS.jl:24-43)."""
        if init_parent_sets is not None:
            flat = [sorted(ps) for chain in init_parent_sets for ps in chain]
            if len(flat) != n_chains * self.V:
                raise ValueError("init_parent_sets must be n_chains x V lists")
            offs, mem = sets_to_csr(flat)
            lam = np.ascontiguousarray(init_lambda, dtype=np.float64).reshape(n_chains * self.V)
            check(self._L.ssps_init_chains(self._h, n_chains, seed, chain_id_offset, _i(offs), _i(mem), _d(lam)))
        else:
            check(self._L.ssps_init_chains(self._h, n_chains, seed, chain_id_offset, None, None, None))
        self.n_chains = n_chains

    def update_loop(self, n_sweeps: int, thinning: int = 1, timeout: float = 0.0) -> int:
        done = C.c_int64(0)
        check(self._L.ssps_run(self._h, int(n_sweeps), float(timeout), int(thinning), C.byref(done)))
        return done.value

    def replay(self, draws: np.ndarray, thinning: int = 1) -> None:
        """draws: (n_sweeps, n_chains, V, 6) = (z_lambda, u_lambda, u_action, u_add, u_rem, u_accept)."""
        d = np.ascontiguousarray(draws, dtype=np.float64)
        if d.ndim != 4 or d.shape[1:] != (self.n_chains, self.V, 6):
            raise ValueError("draws must be (n_sweeps, %d, %d, 6)" % (self.n_chains, self.V))
        check(self._L.ssps_replay(self._h, d.shape[0], _d(d), int(thinning)))

    def export_draws(self, first_sweep: int, n_sweeps: int) -> np.ndarray:
        """The (z, u_lam, u_act, u_add, u_rem, u_acc) numbers update_loop uses for sweeps
        first_sweep+1 .. first_sweep+n_sweeps, shaped like replay()'s input."""
        out = np.empty((n_sweeps, self.n_chains, self.V, 6))
        check(self._L.ssps_export_draws(self._h, int(first_sweep), int(n_sweeps), _d(out)))
        return out

    def state(self, chain: int, want_lml: bool = True):
        V = self.V
        indeg = np.zeros(V, dtype=np.int32)
        mem = np.zeros(V * V, dtype=np.int32)
        tot = C.c_int64(0)
        lam = np.zeros(V)
        lml = np.zeros(V) if want_lml else None
        check(self._L.ssps_get_state(self._h, chain, _i(indeg), _i(mem), mem.size, C.byref(tot), _d(lam), _d(lml)))
        sets, o = [], 0
        for i in range(V):
            sets.append(mem[o:o + indeg[i]].tolist())
            o += int(indeg[i])
        return sets, lam, lml

    # ------------------------------------------------------------------ K4
    def edge_counts(self):
        cnt = np.zeros((self.V, self.V))
        ns, nc = C.c_int64(0), C.c_int32(0)
        check(self._L.ssps_edge_counts(self._h, _d(cnt), C.byref(ns), C.byref(nc)))
        return cnt, ns.value, nc.value

    def chain_edge_counts(self) -> np.ndarray:
        cnt = np.zeros((self.n_chains, self.V, self.V), dtype=np.uint32)
        check(self._L.ssps_chain_edge_counts(self._h, cnt.ctypes.data_as(C.POINTER(C.c_uint32))))
        return cnt

    def edge_posterior(self, burnin: float = 0.5):
        probs = np.zeros((self.V, self.V))
        lam = np.zeros(self.V)
        b, n = C.c_int64(0), C.c_int64(0)
        check(self._L.ssps_edge_posterior(self._h, float(burnin), _d(probs), _d(lam), C.byref(b), C.byref(n)))
        return probs, lam, b.value, n.value

    def psrf_sums_len(self) -> int:
        return int(self._L.ssps_psrf_sums_len(self.V))

    def psrf_sums(self, burnin: float = 0.5) -> np.ndarray:
        sums = np.zeros(self.psrf_sums_len())
        check(self._L.ssps_psrf_sums(self._h, float(burnin), sums.ctypes.data_as(C.c_void_p), 0))
        return sums

    def psrf_sums_device(self, burnin: float, device_ptr: int) -> None:
        check(self._L.ssps_psrf_sums(self._h, float(burnin), C.c_void_p(device_ptr), 1))

    def psrf_finish(self, sums, device_ptr: Optional[int] = None):
        mx = C.c_double(0.0)
        edge = np.zeros((self.V, self.V))
        lam = np.zeros(self.V)
        if device_ptr is not None:
            check(self._L.ssps_psrf_finish(self._h, C.c_void_p(device_ptr), 1, C.byref(mx), _d(edge), _d(lam)))
        else:
            s = np.ascontiguousarray(sums, dtype=np.float64)
            check(self._L.ssps_psrf_finish(self._h, s.ctypes.data_as(C.c_void_p), 0, C.byref(mx), _d(edge), _d(lam)))
        return mx.value, edge, lam

    def psrf(self, burnin: float = 0.5):
        mx = C.c_double(0.0)
        edge = np.zeros((self.V, self.V))
        check(self._L.ssps_psrf(self._h, float(burnin), C.byref(mx), _d(edge)))
        return mx.value, edge

    def convergence(self, burnin: float = 0.5, length: int = 0, others: Sequence["Sampler"] = ()):
        """ssps_convergence: evaluate_convergence + evaluate_summary of postprocess_samples.jl on the device, from the
        change-point logs of this sampler's chains (and those of `others`, pooled).  Returns a dict with V x V arrays
        edge_mean / edge_psrf / edge_neff, length-V arrays lambda_mean / lambda_psrf / lambda_neff, and the exact
        burn-in and split indices."""
        hs = [self] + list(others)
        arr = (C.c_void_p * len(hs))(*[x._h for x in hs])
        V = self.V
        out = {k: np.zeros((V, V)) for k in ("edge_mean", "edge_psrf", "edge_neff")}
        out.update({k: np.zeros(V) for k in ("lambda_mean", "lambda_psrf", "lambda_neff")})
        b, sp = C.c_int64(0), C.c_int64(0)
        check(self._L.ssps_convergence(arr, len(hs), float(burnin), int(length), _d(out["edge_mean"]), _d(out["edge_psrf"]),
                                       _d(out["edge_neff"]), _d(out["lambda_mean"]), _d(out["lambda_psrf"]), _d(out["lambda_neff"]),
                                       C.byref(b), C.byref(sp)))
        out["burnin_index"], out["split_index"] = b.value, sp.value
        return out

    # ------------------------------------------------------------------ change points
    def changepoints(self, chain: int):
        npar, nlam = C.c_int64(0), C.c_int64(0)
        check(self._L.ssps_changepoints_size(self._h, chain, C.byref(npar), C.byref(nlam)))
        a, b = max(npar.value, 1), max(nlam.value, 1)
        pn = np.zeros(a, np.int64); pc = np.zeros(a, np.int32); pp = np.zeros(a, np.int32); pv = np.zeros(a, np.int32)
        ln = np.zeros(b, np.int64); lc = np.zeros(b, np.int32); lv = np.zeros(b)
        check(self._L.ssps_changepoints_get(self._h, chain, _l(pn), _i(pc), _i(pp), _i(pv), _l(ln), _i(lc), _d(lv)))
        m, q = npar.value, nlam.value
        return (pn[:m], pc[:m], pp[:m], pv[:m]), (ln[:q], lc[:q], lv[:q])

    def results(self, chain: int, t_elapsed: float = 0.0) -> dict:
        """The reference's per-chain results object (update_results_storediff + SSPS.jl:223-230);
        parent keys are 1-based strings, exactly what JSON.print writes for the DefaultDict."""
        (pn, pc, pp, pv), (ln, lc, lv) = self.changepoints(chain)
        ps = [dict() for _ in range(self.V)]
        for e in range(len(pn)):
            ps[int(pc[e])].setdefault(str(int(pp[e]) + 1), []).append([int(pn[e]), int(pv[e])])
        lam = [[] for _ in range(self.V)]
        for e in range(len(ln)):
            lam[int(lc[e])].append([int(ln[e]), float(lv[e])])
        return {"parent_sets": ps, "lambda": lam, "n": int(self.stats().samples_recorded), "t_elapsed": float(t_elapsed)}

    # ------------------------------------------------------------------ diagnostics
    def stats(self) -> RunStats:
        st = RunStats()
        check(self._L.ssps_get_stats(self._h, C.byref(st)))
        return st

    def khist(self, length: int = 64) -> np.ndarray:
        h = np.zeros(length, dtype=np.int64)
        check(self._L.ssps_get_khist(self._h, _l(h), length))
        return h

    def raw_stats(self, length: int = 32) -> np.ndarray:
        h = np.zeros(length, dtype=np.int64)
        check(self._L.ssps_get_raw_stats(self._h, _l(h), length))
        return h

    def vertex_cost(self):
        """(warp cycles per vertex in the last MH launch, mean in-degree per vertex now, warps per slab in that launch)."""
        c = np.zeros(self.V, dtype=np.int64)
        k = np.zeros(self.V, dtype=np.int32)
        sp = np.zeros(self.V, dtype=np.int32)
        check(self._L.ssps_get_vertex_cost(self._h, _l(c), k.ctypes.data_as(C.POINTER(C.c_int32)), sp.ctypes.data_as(C.POINTER(C.c_int32))))
        return c, k / 1000.0, sp

    def measure_fp64_peak(self) -> float:
        t = C.c_double(0.0)
        check(self._L.ssps_measure_fp64_peak(self._h, C.byref(t)))
        return t.value

    def measure_gather_peak(self, table_bytes: int):
        """Useful GB/s of random 8-byte gathers from a table of `table_bytes`: (read-only L1 path, L2 only, shared memory)."""
        a, b, c = C.c_double(0.0), C.c_double(0.0), C.c_double(0.0)
        check(self._L.ssps_measure_gather_peak(self._h, int(table_bytes), C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value
