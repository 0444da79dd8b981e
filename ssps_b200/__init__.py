"""ssps_b200 -- B200-native parent-set scoring / Metropolis-Hastings hot path of SSPS.

Layout (only what the path needs):
    csrc/            sm_100a CUDA kernels (ssps_kernels.cuh) + the C ABI (ssps_capi.cu) -> libssps_b200.so
    _lib.py          ctypes binding of include/ssps.h
    sampler.py       host-side handle mirroring the reference's function names for this path
    preprocess.py    reference input-file ingest (dbn_preprocess.jl)
    ssps_wrapper.py  CLI twin of SSPS/ssps_wrapper.jl (same flags, same JSON)
    postprocess.py   pooled edge posterior / PSRF summary in postprocess_samples.jl's format
    multi_gpu.py     chains sharded over GPUs (one process per GPU, or one process for all); the PSRF / posterior
                     all-reduce runs on the library's own NCCL communicator
    simulate.py      restatement of scripts/simulate_data.jl (synthetic inputs)

Importing the package never touches oracle/ (test infrastructure) and has no CPU fallback.
"""
from ._lib import SspsError, load, lib_path  # noqa: F401
from .sampler import Sampler, release_cached_memory  # noqa: F401
from .multi_gpu import MultiGpuSampler, ShardedSampler  # noqa: F401

__all__ = ["Sampler", "MultiGpuSampler", "ShardedSampler", "SspsError", "load", "lib_path", "release_cached_memory"]
