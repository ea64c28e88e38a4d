"""CPU: the C-ABI libraries load, export every symbol include/lbm_b200.h declares, and refuse to
compute without a GPU (no CPU fallback).  No compute calls here."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "lbm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lbm_[a-z_0-9]+)\s*\(", text)))


@pytest.fixture(scope="module", autouse=True)
def _built():
    from latticeboltzmann_b200 import build
    build.build()


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_library_exports_every_declared_symbol(prec):
    from latticeboltzmann_b200 import _capi
    lib = _capi.load(prec)
    declared = _declared()
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib.dll, name), f"{name} declared in lbm_b200.h but not exported by {lib.path}"
    assert sorted(_capi.SYMBOLS) == declared
    assert lib.dll.lbm_sizeof_real() == (8 if prec == "dp" else 4)
    assert prec.encode() in lib.dll.lbm_version()


def test_scene_reference_is_initialize_arrays():
    from latticeboltzmann_b200 import _capi
    sc = _capi.load("dp").dll.lbm_scene_reference()
    assert (sc.rho0, sc.channel_walls, sc.bar_i0, sc.bar_i1, sc.bar_j0, sc.bar_j1, sc.disc_r) == (0.1, 1, 20, 220, 100, 105, -1)


def test_reynolds_from_column_host_order():
    """The serial accumulation of ComputeReynolds (LB.c:529-546) is pure host arithmetic."""
    from latticeboltzmann_b200 import _capi
    for prec in ("dp", "sp"):
        lib = _capi.load(prec)
        t = lib.dtype
        uy = (np.arange(1, 11) / 7).astype(t)
        fl = np.array([0, 1, 1, 1, 0, 1, 1, 1, 1, 0], dtype=np.int8)
        tot = t(0)
        for v, k in zip(uy, fl):
            if k:
                tot = t(tot + v)
        nu = t(t(1) / t(3)) * t(t(0.7) - t(t(1) / t(2)))
        want = t(t(t(tot / t(7)) * t(10)) / nu)
        got = lib.dll.lbm_reynolds_from_column(uy.ctypes.data_as(ctypes.c_void_p), fl.ctypes.data_as(ctypes.c_void_p), 10, t(0.7))
        assert t(got) == want


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from latticeboltzmann_b200 import Lattice, LbmError
    with pytest.raises(LbmError):
        Lattice(32, 32)


def test_c_driver_links_against_the_abi():
    exe = os.path.join(ROOT, "driver", "lbm_b200_dp")
    assert os.path.exists(exe)
    out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "liblbm_b200_dp.so" in out and "not found" not in out.split("liblbm_b200_dp.so")[1].splitlines()[0]


def test_product_never_touches_the_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    for base in ("latticeboltzmann_b200", "include", "driver"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".c", "Makefile")):
                    assert "oracle" not in open(os.path.join(dp, f), errors="ignore").read().lower(), os.path.join(dp, f)
