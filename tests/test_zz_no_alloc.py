"""ssps_score_static: K2 for callers that cannot call cudaMalloc (graph capture, serving loops).  No counterpart in the
reference.  The entry point takes DEVICE pointers and draws its per-thread
factor scratch from one fixed `static __device__` array inside the library; it must return what ssps_score returns.

Collected last on purpose (file name).  Passes alone in a fresh process and after the
other GPU tests (it creates its own Sampler and its own device buffers; the shared scratch array needs no clearing)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from conftest import make_synth  # noqa: E402

MAX_PARENTS = 64            # SSPS_SCORE_STATIC_MAX_PARENTS (include/ssps.h)


def _csr(sets):
    offs = np.zeros(len(sets) + 1, dtype=np.int32)
    offs[1:] = np.cumsum([len(s) for s in sets])
    mem = np.asarray([j for s in sets for j in s] or [0], dtype=np.int32)
    return offs, mem


def _static_scores(s, torch, children, sets, lam, max_parents):
    """one call of ssps_score_static through the C ABI; the device buffers are the test's (torch is only the allocator)"""
    dev = torch.device("cuda", 0)
    offs, mem = _csr(sets)
    d_ch = torch.from_numpy(np.ascontiguousarray(children, dtype=np.int32)).to(dev)
    d_off = torch.from_numpy(offs).to(dev)
    d_mem = torch.from_numpy(mem).to(dev)
    d_lam = torch.from_numpy(np.ascontiguousarray(lam, dtype=np.float64)).to(dev)
    out = torch.full((3, len(sets)), -12345.0, dtype=torch.float64, device=dev)        # a value no score takes
    torch.cuda.synchronize()                    # the sampler has its own stream: inputs must have landed
    s.score_static(len(sets), max_parents, d_ch.data_ptr(), d_off.data_ptr(), d_mem.data_ptr(), d_lam.data_ptr(),
                   out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr())
    torch.cuda.synchronize()                    # the call does not synchronise
    res = out.cpu().numpy()
    return res[0], res[1], res[2]


def _check(a, b, what):
    # the same device routine in the same operation order (fmad off): identical bits expected; the assertion uses the K2
    # tolerance of the parity tests (1e-10) tightened to 1e-13, and reports bit equality
    assert np.array_equal(np.isnan(a), np.isnan(b)), what
    np.testing.assert_allclose(a, b, rtol=1e-13, atol=0.0, equal_nan=True, err_msg=what)
    return bool(np.array_equal(a, b, equal_nan=True))


def test_score_static_matches_score_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    from ssps_b200 import Sampler
    from ssps_b200._lib import SspsError
    V = 100
    ds, xm, xp = make_synth(V, T=12, M=4, seed=77)              # n = 44 < 64: the large sets are rank deficient (pivot skipping)
    rng = np.random.default_rng(5)
    with Sampler(device=0, large_indeg=15.0, lambda_prop_std=3.0) as s:
        s.set_data(xm, xp, ds.prior_graph)
        # call 1: 5000 sets of 0..64 parents (more sets than scratch columns: the grid-stride walk reuses them)
        n1 = 5000
        sets1 = [sorted(rng.choice(V, size=int(rng.integers(0, MAX_PARENTS + 1)), replace=False).tolist()) for _ in range(n1)]
        sets1[0], sets1[1] = [], sorted(rng.choice(V, size=MAX_PARENTS, replace=False).tolist())
        ch1 = rng.integers(0, V, n1)
        lam1 = rng.uniform(0.5, 14.0, n1)
        want1 = s.score(ch1, sets1, lam1, want_min_pivot=True)
        got1 = _static_scores(s, torch, ch1, sets1, lam1, MAX_PARENTS)
        # call 2, straight after, smaller bound (another scratch layout over the array's stale contents)
        n2 = 70000
        sets2 = [sorted(rng.choice(V, size=int(rng.integers(0, 9)), replace=False).tolist()) for _ in range(n2)]
        ch2 = rng.integers(0, V, n2)
        lam2 = rng.uniform(0.5, 14.0, n2)
        got2 = _static_scores(s, torch, ch2, sets2, lam2, 8)
        want2 = s.score(ch2, sets2, lam2, want_min_pivot=True)
        bits = []
        for got, want, tag in ((got1, want1, "call 1"), (got2, want2, "call 2")):
            for g, w, name in zip(got, want, ("lml", "logprior", "min pivot")):
                assert not np.any(g == -12345.0), "%s: %s has entries the kernel never wrote" % (tag, name)
                bits.append(_check(g, w, "%s: %s" % (tag, name)))
        assert np.isfinite(got1[0]).all() and np.isfinite(got2[0]).all()
        print("ssps_score_static vs ssps_score: bit-identical = %s" % bits)

        # sets the device must refuse (the host never sees them): too many members for the stated bound, unsorted, out of range
        bad_sets = [[1, 2, 3], [5, 4], [3, 3], [0, V], [2, 7], []]
        bad_ch = np.array([0, 1, 2, 3, V, 4])
        lml, lp, mp = _static_scores(s, torch, bad_ch, bad_sets, np.full(6, 2.0), 2)
        assert np.isnan(lml[:5]).all() and np.isnan(lp[:5]).all() and np.isnan(mp[:5]).all()
        assert lml[5] == s.score([4], [[]], [2.0])[0][0]

        # sizes beyond the fixed array: refused with a message, nothing launched
        with pytest.raises(SspsError, match="max_parents=65"):
            s.score_static(1, MAX_PARENTS + 1, 1, 1, 1)
        with pytest.raises(SspsError, match="n_sets"):
            s.score_static((1 << 24) + 1, 4, 1, 1, 1)
        s.score_static(0, 4, None, None, None)              # the empty batch (warm-up before a capture) is legal


def test_score_static_refuses_interaction_designs():
    from ssps_b200 import Sampler
    from ssps_b200._lib import SspsError
    ds, xm, xp = make_synth(20, seed=3)
    with Sampler(device=0, regression_deg=2) as s:
        s.set_data(xm, xp, ds.prior_graph)
        with pytest.raises(SspsError, match="regression_deg=2"):
            s.score_static(1, 4, 1, 1, 1)
