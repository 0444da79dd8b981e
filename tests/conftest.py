import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def c1():
    """Config C1: the reference's shipped example (run_ssps/example_*), committed as a fixture."""
    d = np.load(os.path.join(ROOT, "tests", "golden", "example_c1.npz"))
    return {k: d[k] for k in d.files}


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def c1_problem(c1, oracle):
    return oracle.Problem(c1["Xminus"], c1["Xplus"], c1["prior"], large_indeg=20.0, lambda_max=15.0,
                          lambda_prop_std=3.0)


def make_synth(V, T=8, M=4, seed=20200701):
    from ssps_b200.simulate import simulate
    ds = simulate(V, T, M, seed=seed)
    xm, xp = ds.lagged()
    return ds, xm, xp


@pytest.fixture(scope="session")
def synth_v1000():
    """V = 1000, T = 100, M = 4 (n = 396): the shape of BASELINE.json's config C5 (and of the V = 1000 replay test)."""
    return make_synth(1000, T=100, M=4)


def pytest_sessionfinish(session, exitstatus):
    """A library built with -DSSPS_BOUNDS (tools/bounds_run.sh) counts failed index checks on the device: report them."""
    try:
        from ssps_b200 import _lib
        if _lib._lib is None:
            return
        n = _lib._lib.ssps_bounds_failures()
    except Exception:
        return
    if n >= 0:
        print("\n[ssps bounds check] failed index checks on the device: %d" % n)
        if n > 0:
            session.exitstatus = 1
