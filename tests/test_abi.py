"""The C-ABI library loads on a CPU-only box, exports every symbol include/ssps.h declares, the
ctypes binding covers all of them, and the product path fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "ssps.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ssps_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    from ssps_b200 import _lib
    L = _lib.load()
    names = _declared()
    assert len(names) >= 29
    raw = ctypes.CDLL(_lib.lib_path())
    for n in names:
        assert hasattr(raw, n), "library does not export %s" % n
        assert n in _lib.SIGNATURES, "binding misses %s" % n
    assert set(_lib.SIGNATURES) == set(names)
    assert L.ssps_abi_version() == 4
    assert L.ssps_psrf_sums_len(10) == 3 * 100 + 30 + 2


def test_struct_layouts_match_header():
    from ssps_b200 import _lib
    assert ctypes.sizeof(_lib.Config) == 72
    assert ctypes.sizeof(_lib.RunStats) == 88 + 5 * 8 + 10 * 8


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ssps_b200 import Sampler, SspsError
    with pytest.raises(SspsError) as e:
        Sampler()
    assert e.value.code == -2      # SSPS_ERR_CUDA


def test_create_validates_config():
    from ssps_b200 import _lib
    L = _lib.load()
    cfg = _lib.Config()
    h = ctypes.c_void_p()
    cfg.struct_size = 12
    assert L.ssps_create(ctypes.byref(cfg), ctypes.byref(h)) == _lib.ERR_INVALID
    assert b"struct_size" in L.ssps_last_error()
    cfg.struct_size = ctypes.sizeof(_lib.Config)
    cfg.regression_deg = 1
    cfg.large_indeg, cfg.lambda_max, cfg.lambda_prop_std = -1.0, 15.0, 0.25
    assert L.ssps_create(ctypes.byref(cfg), ctypes.byref(h)) == _lib.ERR_INVALID
    assert L.ssps_create(None, None) == _lib.ERR_INVALID


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "ssps_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert not re.search(r"#\s*include[^\n]*oracle", src), f      # comments may cite it; code may not use it
                assert "libssps_oracle" not in src and "orc_" not in re.sub(r"//[^\n]*|/\*.*?\*/", "", src, flags=re.S), f
