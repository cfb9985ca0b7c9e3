"""CPU-only: the C-ABI library loads and exports every symbol include/flivo.h declares; no compute calls."""
import ctypes
import os
import re


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()
    hdr = open(os.path.join(os.path.dirname(pkg.HERE), "include", "flivo.h")).read()
    declared = sorted(set(re.findall(r"\b(flivo_[a-z0-9_]+)\s*\(", hdr)))
    assert declared == sorted(pkg.SYMBOLS)
    for name in declared:
        assert hasattr(L, name), name
    assert b"sm_100a" in L.flivo_version()


def test_struct_sizes_match_header(pkg):
    assert ctypes.sizeof(pkg.State) == 36 * 8
    assert ctypes.sizeof(pkg.Options) == 24
    assert ctypes.sizeof(pkg.Cfg) == 8 + 8 + 23 * 8 + 8
    assert ctypes.sizeof(pkg.Report) == 16 + 8 + 23 * 8 + 8
    assert ctypes.sizeof(pkg.Stats) == 7 * 8 + 16 * 4 * 3


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the product must fail loudly rather than compute on the host."""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("GPU present")
    import pytest
    with pytest.raises(pkg.FlivoError):
        pkg.Flivo()


def test_repack_v1_is_pure(pkg):
    import numpy as np
    from oracle import pyoracle as orc
    rng = np.random.default_rng(0)
    A = rng.normal(size=(12, 12)); H = A @ A.T; z = rng.normal(size=12)
    HPH = np.zeros(81); HPL = np.zeros(9)
    rc = pkg.lib().flivo_repack_v1(H.ctypes.data, z.ctypes.data, 0.001, HPH.ctypes.data, HPL.ctypes.data)
    assert rc == 0
    oH, oL = orc.repack_v1(H, z, 0.001)
    assert np.array_equal(HPH.reshape(9, 9), oH) and np.array_equal(HPL, oL)
