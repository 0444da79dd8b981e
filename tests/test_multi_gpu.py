"""N > 1 host logic on CPU: world_size-2 gloo processes shard chains and SUM-all-reduce the PSRF
sufficient statistics exactly as the ranks of a multi-GPU run do with NCCL; the result must equal
the oracle's PSRF over the pooled chains."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_chains_partitions_exactly():
    from ssps_b200.multi_gpu import shard_chains
    for total in (1, 7, 8, 512, 1023):
        for world in (1, 2, 4, 8):
            blocks = [shard_chains(total, world, r) for r in range(world)]
            assert sum(c for _, c in blocks) == total
            nxt = 0
            for first, count in blocks:
                assert first == nxt
                nxt += count
            assert max(c for _, c in blocks) - min(c for _, c in blocks) <= 1
    with pytest.raises(ValueError):
        shard_chains(8, 2, 2)


def _local_sums(c_b, c_s, c_len, N1, N2):
    """numpy statement of psrf_sums_kernel for ONE edge over the local chains (+ the [H, H*n] tail)."""
    m1 = (c_s - c_b) / N1
    m2 = (c_len - c_s) / N2
    sm = (m1 + m2).sum()
    sm2 = (m1 * m1 + m2 * m2).sum()
    sv = (m1 * (1 - m1) * N1 / (N1 - 1) + m2 * (1 - m2) * N2 / (N2 - 1)).sum()
    H = 2.0 * len(c_b)
    return np.array([sm, sm2, sv, H, H * N1])


def _finish(s):
    """numpy statement of psrf_finish_kernel."""
    sm, sm2, sv, H, Hn = s
    n = Hn / H
    mbar = sm / H
    B = n * max(sm2 - H * mbar * mbar, 0.0) / H
    W = sv / H
    return np.sqrt(((n - 1) * W / n + B / n) / W)


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from ssps_b200.multi_gpu import allreduce_sum_, shard_chains
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(1234)                     # same data on every rank
    n_chains, L = 7, 200
    seqs = (rng.random((n_chains, L)) < rng.uniform(0.2, 0.6, (n_chains, 1))).astype(np.int64)
    b, s = L // 2, L // 2 + L // 4
    first, count = shard_chains(n_chains, world, rank)
    mine = seqs[first:first + count]
    cum = np.cumsum(mine, axis=1)
    sums = _local_sums(cum[:, b - 1].astype(float), cum[:, s - 1].astype(float), cum[:, L - 1].astype(float), s - b, L - s)
    t = torch.from_numpy(sums)
    allreduce_sum_(t)
    q.put((rank, t.numpy().copy(), first, count))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_psrf_reduction_matches_pooled_oracle(oracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got.sort(key=lambda g: g[0])
    assert np.array_equal(got[0][1], got[1][1])                      # both ranks hold the same reduced statistics
    assert (got[0][2], got[0][3], got[1][2], got[1][3]) == (0, 4, 4, 3)
    rng = np.random.default_rng(1234)
    n_chains, L = 7, 200
    seqs = (rng.random((n_chains, L)) < rng.uniform(0.2, 0.6, (n_chains, 1))).astype(np.uint8)
    want = oracle.psrf_binary_dense([seqs[c] for c in range(n_chains)], 0.5)
    assert got[0][1][3] == 14.0 and got[0][1][4] / got[0][1][3] == 50.0   # H half-chains, n survives the SUM
    assert _finish(got[0][1]) == pytest.approx(want, rel=1e-12)
