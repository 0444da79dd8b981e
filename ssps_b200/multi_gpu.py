"""Chains sharded over the GPUs of one box.  No data-path collective.

The posterior factorises over chains (and vertices), so each GPU runs its own contiguous block of chain ids on its
own full copy of Z/G/R.  The only exchange is a SUM all-reduce of additive sufficient statistics -- the
split-PSRF sums and the pooled tallies -- over an NCCL communicator owned by libssps_b200.so itself
(ssps_comm_* / ssps_psrf_global / ssps_edge_posterior_global, include/ssps.h): a `ccall` host gets the same calls,
torch is not on this path.  The reference pools chains offline by reading their JSON files
(scripts/postprocess_samples.jl:242-259).

Two host models, as in include/ssps.h:
    ShardedSampler   one process per GPU (torchrun, mpirun, ...): collective calls, every rank makes them
    MultiGpuSampler  one process drives all GPUs (what `ssps_wrapper --gpus N` uses): ssps_run_multi keeps them busy
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import check
from .sampler import Sampler, _d

COMM_ID_BYTES = 128


def shard_chains(n_chains_total: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous block (first chain id, count) of `rank`; the first n % world ranks hold one extra chain."""
    if not (0 <= rank < world_size):
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    base, extra = divmod(n_chains_total, world_size)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def allreduce_sum_(tensor):
    """In-place SUM all-reduce over torch's default process group (no-op when not initialised).  Kept for hosts that
    prefer their own communicator (ssps_psrf_sums / ssps_psrf_finish with device pointers); the samplers below do not use it."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
    return tensor


def comm_unique_id() -> bytes:
    """ssps_comm_unique_id: 128 opaque bytes rank 0 hands to the other ranks (any host channel will do)."""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    check(_lib.load().ssps_comm_unique_id(buf))
    return buf.raw


def _handles(samplers: Sequence[Sampler]):
    arr = (C.c_void_p * len(samplers))(*[s._h for s in samplers])
    return arr, len(samplers)


def _psrf_global(samplers: Sequence[Sampler], burnin: float):
    L = _lib.load()
    V = samplers[0].V
    arr, n = _handles(samplers)
    mx = C.c_double(0.0)
    edge, lam = np.zeros((V, V)), np.zeros(V)
    check(L.ssps_psrf_global(arr, n, float(burnin), C.byref(mx), _d(edge), _d(lam)))
    return mx.value, edge, lam


def _edge_posterior_global(samplers: Sequence[Sampler], burnin: float):
    L = _lib.load()
    V = samplers[0].V
    arr, n = _handles(samplers)
    probs, lam = np.zeros((V, V)), np.zeros(V)
    b, nn, nc = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    check(L.ssps_edge_posterior_global(arr, n, float(burnin), _d(probs), _d(lam), C.byref(b), C.byref(nn), C.byref(nc)))
    return probs, lam, b.value, nn.value, nc.value


class ShardedSampler:
    """One process per GPU.  Construct on every rank; the communicator is bootstrapped through torch.distributed
    (whatever backend: the 128-byte id is broadcast as an object) unless `comm_id`, `rank` and `world` are passed."""

    def __init__(self, n_chains_total: int, seed: int = 0, device: Optional[int] = None, rank: Optional[int] = None,
                 world: Optional[int] = None, comm_id: Optional[bytes] = None, stream: Optional[int] = None,
                 **sampler_kwargs):
        if rank is None or world is None:
            import torch.distributed as dist
            ok = dist.is_available() and dist.is_initialized()
            rank = dist.get_rank() if ok else 0
            world = dist.get_world_size() if ok else 1
        self.rank, self.world = int(rank), int(world)
        self.first, self.count = shard_chains(n_chains_total, self.world, self.rank)
        self.n_chains_total, self.seed = n_chains_total, seed
        if device is None:
            import torch
            device = torch.cuda.current_device()
        self.device = int(device)
        self.sampler = Sampler(device=self.device, **sampler_kwargs)
        if stream is not None:
            self.sampler.set_stream(stream)
        if self.world > 1:
            if comm_id is None:
                import torch.distributed as dist
                box = [comm_unique_id() if self.rank == 0 else None]
                dist.broadcast_object_list(box, src=0)
                comm_id = box[0]
            buf = C.create_string_buffer(comm_id, COMM_ID_BYTES)
            check(_lib.load().ssps_comm_init_rank(self.sampler._h, self.world, self.rank, buf))

    def set_data(self, Xminus, Xplus, prior):
        self.sampler.set_data(Xminus, Xplus, prior)

    def generate(self):
        # chain ids key the Philox streams: ranks use disjoint blocks, so an N-GPU run draws exactly the
        # chains a 1-GPU run of n_chains_total would
        self.sampler.generate(self.count, seed=self.seed, chain_id_offset=self.first)

    def update_loop(self, n_sweeps: int, thinning: int = 1, timeout: float = 0.0) -> int:
        return self.sampler.update_loop(n_sweeps, thinning=thinning, timeout=timeout)

    def psrf(self, burnin: float = 0.5):
        """(max edge split-PSRF over ALL ranks' chains, V x V edge PSRFs, lambda PSRFs) -- collective."""
        return _psrf_global([self.sampler], burnin)

    def edge_posterior(self, burnin: float = 0.5):
        """Pooled post-burn-in edge frequencies and lambda means over all ranks (evaluate_summary) -- collective."""
        probs, lam, _, _, _ = _edge_posterior_global([self.sampler], burnin)
        return probs, lam

    def comm_info(self):
        nr, rk, ver = C.c_int32(0), C.c_int32(0), C.c_int32(0)
        check(_lib.load().ssps_comm_info(self.sampler._h, C.byref(nr), C.byref(rk), C.byref(ver)))
        return nr.value, rk.value, ver.value

    def close(self):
        if self.sampler._h:
            _lib.load().ssps_comm_destroy(self.sampler._h)
        self.sampler.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class MultiGpuSampler:
    """One process, one Sampler per device, driven together (ssps_comm_init_all / ssps_run_multi)."""

    def __init__(self, n_chains_total: int, devices: Sequence[int], seed: int = 0, **sampler_kwargs):
        self.devices = [int(d) for d in devices]
        if len(set(self.devices)) != len(self.devices) or not self.devices:
            raise ValueError("devices must be distinct and non-empty")
        self.world = len(self.devices)
        self.n_chains_total, self.seed = n_chains_total, seed
        self.samplers = [Sampler(device=d, **sampler_kwargs) for d in self.devices]
        self.shards = [shard_chains(n_chains_total, self.world, r) for r in range(self.world)]
        if any(c < 1 for _, c in self.shards):
            raise ValueError("%d chains cannot be spread over %d GPUs" % (n_chains_total, self.world))
        if self.world > 1:
            arr, n = _handles(self.samplers)
            check(_lib.load().ssps_comm_init_all(arr, n))

    @property
    def V(self):
        return self.samplers[0].V

    def set_data(self, Xminus, Xplus, prior):
        for s in self.samplers:
            s.set_data(Xminus, Xplus, prior)

    def set_timeseries(self, series, prior):
        for s in self.samplers:
            s.set_timeseries(series, prior)

    def generate(self, init_parent_sets=None, init_lambda=None):
        for s, (first, count) in zip(self.samplers, self.shards):
            if init_parent_sets is not None:
                s.generate(count, seed=self.seed, chain_id_offset=first,
                           init_parent_sets=init_parent_sets[first:first + count], init_lambda=init_lambda[first:first + count])
            else:
                s.generate(count, seed=self.seed, chain_id_offset=first)

    def update_loop(self, n_sweeps: int, thinning: int = 1, timeout: float = 0.0) -> int:
        arr, n = _handles(self.samplers)
        done = C.c_int64(0)
        check(_lib.load().ssps_run_multi(arr, n, int(n_sweeps), float(timeout), int(thinning), C.byref(done)))
        return done.value

    def psrf(self, burnin: float = 0.5):
        return _psrf_global(self.samplers, burnin)

    def edge_posterior(self, burnin: float = 0.5):
        return _edge_posterior_global(self.samplers, burnin)

    def convergence(self, burnin: float = 0.5, length: int = 0):
        """Pooled over all GPUs' chains (the logs live on the host of this one process; computed on the first device)."""
        return self.samplers[0].convergence(burnin, length, others=self.samplers[1:])

    @property
    def n_chains(self):
        return self.n_chains_total

    @property
    def record_changepoints(self):
        return bool(self.samplers[0].cfg.record_changepoints)

    def samples_recorded(self) -> int:
        return int(self.samplers[0].stats().samples_recorded)

    def locate(self, chain: int) -> Tuple[Sampler, int]:
        """(sampler, local index) of global chain id `chain`."""
        for s, (first, count) in zip(self.samplers, self.shards):
            if first <= chain < first + count:
                return s, chain - first
        raise IndexError(chain)

    def results(self, chain: int, t_elapsed: float = 0.0) -> dict:
        s, c = self.locate(chain)
        return s.results(c, t_elapsed=t_elapsed)

    def stats(self):
        return [s.stats() for s in self.samplers]

    def close(self):
        L = _lib.load()
        for s in self.samplers:
            if s._h:
                L.ssps_comm_destroy(s._h)
            s.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
