"""ctypes binding of libssps_b200.so -- one declaration per entry point of include/ssps.h.

The library is loaded lazily and loudly: if the shared object is missing or a symbol is absent the
import of the product path fails; nothing here falls back to a CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os

from . import _build

OK = 0
ABI_VERSION = 4
ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_NOMEM, ERR_INDEGREE, ERR_UNAVAILABLE = -1, -2, -3, -4, -5, -6

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lp = C.POINTER(C.c_int64)
_up = C.POINTER(C.c_uint32)
_hp = C.c_void_p


class Config(C.Structure):
    """ssps_config (include/ssps.h)."""
    _fields_ = [
        ("struct_size", C.c_int32), ("device", C.c_int32), ("regression_deg", C.c_int32),
        ("max_indegree", C.c_int32), ("large_indeg", C.c_double), ("lambda_max", C.c_double),
        ("lambda_prop_std", C.c_double), ("psrf_grid", C.c_int32), ("record_changepoints", C.c_int32),
        ("collect_khist", C.c_int32), ("verify", C.c_int32),
        ("min_burnin", C.c_double), ("snapshot_slots", C.c_int32), ("proposal", C.c_int32),
    ]


class RunStats(C.Structure):
    """ssps_run_stats (include/ssps.h)."""
    _fields_ = [
        ("sweeps_done", C.c_int64), ("samples_recorded", C.c_int64), ("n_proposals", C.c_int64),
        ("n_accepted", C.c_int64), ("n_lambda_accepted", C.c_int64), ("sum_k", C.c_int64),
        ("algorithmic_flops", C.c_double), ("algorithmic_bytes", C.c_double), ("kernel_ms", C.c_double),
        ("kernel_launches", C.c_int64), ("mh_launches", C.c_int64),
        ("n_tier1", C.c_int64), ("n_lambda_exact", C.c_int64), ("n_big_steps", C.c_int64), ("big_cycles", C.c_int64), ("n_mismatch", C.c_int64),
        ("n_decision_mismatch", C.c_int64), ("n_lambda_decision_mismatch", C.c_int64),
        ("margin_use_max", C.c_double), ("lambda_margin_use_max", C.c_double),
        ("n_indegree_rejects", C.c_int64), ("n_mid_steps", C.c_int64),
        ("snapshot_slots", C.c_int64), ("snapshots_live", C.c_int64), ("snapshot_grid", C.c_int64),
        ("device_allocs_in_run", C.c_int64),
    ]


# name -> (restype, argtypes); the single source of truth for tests/test_abi.py as well
SIGNATURES = {
    "ssps_last_error": (C.c_char_p, []),
    "ssps_abi_version": (C.c_int, []),
    "ssps_create": (C.c_int, [C.POINTER(Config), C.POINTER(_hp)]),
    "ssps_destroy": (None, [_hp]),
    "ssps_set_stream": (C.c_int, [_hp, C.c_void_p]),
    "ssps_set_data": (C.c_int, [_hp, _dp, _dp, C.c_int32, C.c_int32, _dp]),
    "ssps_set_timeseries": (C.c_int, [_hp, _dp, _ip, C.c_int32, C.c_int32, _dp]),
    "ssps_get_design": (C.c_int, [_hp, _dp, _dp]),
    "ssps_get_gram": (C.c_int, [_hp, _dp, _dp, _dp]),
    "ssps_get_derived": (C.c_int, [_hp, _dp, _dp]),
    "ssps_get_dims": (C.c_int, [_hp, _ip, _ip, _ip]),
    "ssps_score": (C.c_int, [_hp, C.c_int32, _ip, _ip, _ip, _dp, _dp, _dp, _dp]),
    "ssps_init_chains": (C.c_int, [_hp, C.c_int32, C.c_uint64, C.c_uint32, _ip, _ip, _dp]),
    "ssps_run": (C.c_int, [_hp, C.c_int64, C.c_double, C.c_int32, _lp]),
    "ssps_replay": (C.c_int, [_hp, C.c_int64, _dp, C.c_int32]),
    "ssps_export_draws": (C.c_int, [_hp, C.c_int64, C.c_int64, _dp]),
    "ssps_get_state": (C.c_int, [_hp, C.c_int32, _ip, _ip, C.c_int64, _lp, _dp, _dp]),
    "ssps_edge_counts": (C.c_int, [_hp, _dp, _lp, _ip]),
    "ssps_chain_edge_counts": (C.c_int, [_hp, _up]),
    "ssps_edge_posterior": (C.c_int, [_hp, C.c_double, _dp, _dp, _lp, _lp]),
    "ssps_psrf_sums_len": (C.c_int64, [C.c_int32]),
    "ssps_psrf_sums": (C.c_int, [_hp, C.c_double, C.c_void_p, C.c_int32]),
    "ssps_psrf_finish": (C.c_int, [_hp, C.c_void_p, C.c_int32, _dp, _dp, _dp]),
    "ssps_psrf": (C.c_int, [_hp, C.c_double, _dp, _dp]),
    "ssps_changepoints_size": (C.c_int, [_hp, C.c_int32, _lp, _lp]),
    "ssps_changepoints_get": (C.c_int, [_hp, C.c_int32, _lp, _ip, _ip, _ip, _lp, _ip, _dp]),
    "ssps_comm_unique_id": (C.c_int, [C.c_void_p]),
    "ssps_comm_init_rank": (C.c_int, [_hp, C.c_int32, C.c_int32, C.c_void_p]),
    "ssps_comm_init_all": (C.c_int, [C.POINTER(_hp), C.c_int32]),
    "ssps_comm_destroy": (C.c_int, [_hp]),
    "ssps_comm_info": (C.c_int, [_hp, _ip, _ip, _ip]),
    "ssps_run_multi": (C.c_int, [C.POINTER(_hp), C.c_int32, C.c_int64, C.c_double, C.c_int32, _lp]),
    "ssps_psrf_global": (C.c_int, [C.POINTER(_hp), C.c_int32, C.c_double, _dp, _dp, _dp]),
    "ssps_edge_posterior_global": (C.c_int, [C.POINTER(_hp), C.c_int32, C.c_double, _dp, _dp, _lp, _lp, _lp]),
    "ssps_convergence": (C.c_int, [C.POINTER(_hp), C.c_int32, C.c_double, C.c_int64, _dp, _dp, _dp, _dp, _dp, _dp, _lp, _lp]),
    "ssps_get_stats": (C.c_int, [_hp, C.POINTER(RunStats)]),
    "ssps_get_khist": (C.c_int, [_hp, _lp, C.c_int32]),
    "ssps_get_raw_stats": (C.c_int, [_hp, _lp, C.c_int32]),
    "ssps_get_vertex_cost": (C.c_int, [_hp, _lp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "ssps_measure_fp64_peak": (C.c_int, [_hp, _dp]),
    "ssps_measure_gather_peak": (C.c_int, [_hp, C.c_int64, _dp, _dp, _dp]),
    "ssps_release_cached_memory": (C.c_int, [C.c_int32]),
    "ssps_bounds_failures": (C.c_int64, []),
    # device pointers (plain addresses): the call allocates nothing and stages nothing
    "ssps_score_static": (C.c_int, [_hp, C.c_int32, C.c_int32] + [C.c_void_p] * 7),
}

_lib = None


class SspsError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__("libssps_b200: %s (status %d)" % (message, code))
        self.code = code


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True) -> C.CDLL:
    """dlopen the in-tree library and bind every symbol of include/ssps.h."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path):
        if not build_if_missing:
            raise ImportError("libssps_b200.so is not built (python -m ssps_b200._build); no CPU fallback exists")
        _build.build()
    L = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        f = getattr(L, name)          # AttributeError if the symbol is missing: fail loudly
        f.restype = res
        f.argtypes = args
    if L.ssps_abi_version() != ABI_VERSION:
        raise ImportError("libssps_b200.so ABI version %d, binding expects %d" % (L.ssps_abi_version(), ABI_VERSION))
    _lib = L
    return L


def check(status: int) -> None:
    if status != OK:
        msg = load().ssps_last_error()
        raise SspsError(status, msg.decode("utf-8", "replace") if msg else "unknown error")
