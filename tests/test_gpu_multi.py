"""Multi-GPU through the library's own NCCL communicator (include/ssps.h "Multi-GPU"): N GPUs x C chains must give the
tallies, the posterior and the PSRF of 1 GPU x N C chains.  Skipped below 2 GPUs (gpurun --gpus 2)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from conftest import make_synth  # noqa: E402


def _n_gpus():
    import torch
    return torch.cuda.device_count()


KW = dict(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=50)


def _single_gpu_reference(xm, xp, prior, n_chains, n_sweeps, seed):
    from ssps_b200 import Sampler
    with Sampler(device=0, **KW) as s:
        s.set_data(xm, xp, prior)
        s.generate(n_chains, seed=seed)
        s.update_loop(n_sweeps)
        probs, lam, b, n = s.edge_posterior(0.5)
        mx, edge, lam_psrf = s.psrf_finish(s.psrf_sums(0.5))
        counts = s.chain_edge_counts()
    return probs, lam, mx, edge, lam_psrf, counts


def test_global_calls_without_a_communicator_equal_the_local_ones():
    """n = 1 and no communicator: ssps_psrf_global / ssps_edge_posterior_global are the local statistics (1 GPU is enough)."""
    from ssps_b200 import MultiGpuSampler
    ds, xm, xp = make_synth(30, seed=3)
    probs, lam, mx, edge, lam_psrf, _ = _single_gpu_reference(xm, xp, ds.prior_graph, 24, 400, 9)
    with MultiGpuSampler(24, [0], seed=9, **KW) as m:
        m.set_data(xm, xp, ds.prior_graph)
        m.generate()
        assert m.update_loop(400) == 400
        mx2, edge2, lam2 = m.psrf(0.5)
        p2, l2, b, n, nc = m.edge_posterior(0.5)
    assert (b, n, nc) == (200, 400, 24)
    assert np.array_equal(p2, probs) and np.array_equal(l2, lam)
    assert mx2 == mx and np.array_equal(edge2, edge, equal_nan=True) and np.array_equal(lam2, lam_psrf)


@pytest.mark.skipif("_n_gpus() < 2")
def test_one_process_two_gpus_equals_one_gpu():
    from ssps_b200 import MultiGpuSampler
    ds, xm, xp = make_synth(30, seed=3)
    C, n_sweeps, seed = 40, 400, 9                  # 2 x 20 chains
    probs, lam, mx, edge, lam_psrf, counts = _single_gpu_reference(xm, xp, ds.prior_graph, C, n_sweeps, seed)
    with MultiGpuSampler(C, [0, 1], seed=seed, **KW) as m:
        m.set_data(xm, xp, ds.prior_graph)
        m.generate()
        assert m.update_loop(n_sweeps) == n_sweeps
        got = np.concatenate([s.chain_edge_counts() for s in m.samplers], axis=0)
        mx2, edge2, lam2 = m.psrf(0.5)
        p2, l2, b, n, nc = m.edge_posterior(0.5)
    assert np.array_equal(got, counts)              # the same chains, bit for bit (Philox keyed by global chain id)
    assert (b, n, nc) == (n_sweeps // 2, n_sweeps, C)
    np.testing.assert_allclose(p2, probs, rtol=0, atol=1e-15)      # integer tallies: exact up to the final division
    np.testing.assert_allclose(l2, lam, rtol=1e-13)
    assert abs(mx2 - mx) <= 1e-12 * mx
    ok = np.isfinite(edge)
    assert np.array_equal(ok, np.isfinite(edge2))
    np.testing.assert_allclose(edge2[ok], edge[ok], rtol=1e-12)
    np.testing.assert_allclose(lam2, lam_psrf, rtol=1e-10)


def _rank_main(rank, world, port, C, n_sweeps, seed, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from conftest import make_synth as mk
    from ssps_b200 import ShardedSampler
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)   # bootstrap only
    ds, xm, xp = mk(30, seed=3)
    with ShardedSampler(C, seed=seed, device=rank, **KW) as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate()
        s.update_loop(n_sweeps)
        mx, edge, lam_psrf = s.psrf(0.5)
        probs, lam = s.edge_posterior(0.5)
        info = s.comm_info()
    q.put((rank, mx, edge, probs, info))
    dist.destroy_process_group()


@pytest.mark.skipif("_n_gpus() < 2")
def test_one_process_per_gpu_equals_one_gpu():
    import torch.multiprocessing as mp
    ds, xm, xp = make_synth(30, seed=3)
    C, n_sweeps, seed = 40, 400, 9
    probs, lam, mx, edge, lam_psrf, _ = _single_gpu_reference(xm, xp, ds.prior_graph, C, n_sweeps, seed)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, C, n_sweeps, seed, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, mx2, edge2, p2, info in out:
        assert info[0] == 2 and info[1] == rank and info[2] >= 22000
        assert abs(mx2 - mx) <= 1e-12 * mx
        np.testing.assert_allclose(p2, probs, rtol=0, atol=1e-15)
        ok = np.isfinite(edge)
        np.testing.assert_allclose(edge2[ok], edge[ok], rtol=1e-12)
