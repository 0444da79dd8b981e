"""Chains sharded over the GPUs of one box: one process per GPU, no data-path collective.

The posterior factorises over chains (and vertices), so each rank runs its own contiguous block
of chain ids on its own full copy of Z/G/R.  The only exchange is the reduction of
additive sufficient statistics -- edge tallies and the split-PSRF sums of ssps_psrf_sums -- with
one torch.distributed all_reduce(SUM) (NCCL over NVLink on GPUs; gloo in the CPU tests).
The reference pools chains offline by reading their JSON files (postprocess_samples.jl:242-259).
"""
from __future__ import annotations

from typing import Tuple


def shard_chains(n_chains_total: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous block (first chain id, count) of `rank`; the first n % world ranks hold one extra chain."""
    if not (0 <= rank < world_size):
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    base, extra = divmod(n_chains_total, world_size)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


def allreduce_sum_(tensor):
    """In-place SUM all-reduce over the default process group (no-op when not initialised)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
    return tensor


class ShardedSampler:
    """A `Sampler` per rank plus the cross-rank reductions.  Construct after init_process_group."""

    def __init__(self, n_chains_total: int, seed: int = 0, **sampler_kwargs):
        import torch
        import torch.distributed as dist
        from .sampler import Sampler
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.first, self.count = shard_chains(n_chains_total, self.world, self.rank)
        self.n_chains_total, self.seed = n_chains_total, seed
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.sampler = Sampler(device=self.device.index, **sampler_kwargs)
        self.sampler.set_stream(torch.cuda.current_stream().cuda_stream)
        self._sums = None

    def set_data(self, Xminus, Xplus, prior):
        self.sampler.set_data(Xminus, Xplus, prior)

    def generate(self):
        # chain ids key the Philox streams: ranks use disjoint blocks, so an N-GPU run draws exactly the
        # chains a 1-GPU run of n_chains_total would
        self.sampler.generate(self.count, seed=self.seed, chain_id_offset=self.first)

    def update_loop(self, n_sweeps: int, thinning: int = 1, timeout: float = 0.0) -> int:
        return self.sampler.update_loop(n_sweeps, thinning=thinning, timeout=timeout)

    def psrf(self, burnin: float = 0.5):
        """max edge split-PSRF over ALL ranks' chains (+ the V x V matrix and lambda PSRFs)."""
        import torch
        if self._sums is None:
            self._sums = torch.zeros(self.sampler.psrf_sums_len(), dtype=torch.float64, device=self.device)
        self.sampler.psrf_sums_device(burnin, self._sums.data_ptr())
        allreduce_sum_(self._sums)
        return self.sampler.psrf_finish(None, device_ptr=self._sums.data_ptr())

    def edge_posterior(self, burnin: float = 0.5):
        """Pooled post-burn-in edge frequencies over all ranks (evaluate_summary)."""
        import torch
        probs, lam, b, n = self.sampler.edge_posterior(burnin)
        w = float(self.count * (n - b))
        t = torch.from_numpy(probs * w).to(self.device)
        tl = torch.from_numpy(lam * w).to(self.device)
        tw = torch.tensor([w], dtype=torch.float64, device=self.device)
        for x in (t, tl, tw):
            allreduce_sum_(x)
        return (t / tw).cpu().numpy(), (tl / tw).cpu().numpy()

    def close(self):
        self.sampler.close()
