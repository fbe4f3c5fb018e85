"""CPU checks of the drop-in boundary: the shared library loads, exports every symbol include/gpirt_b200.h declares,
fails loudly without a device, and the host-only bookkeeping entry points are bit-exact against the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import gpirt_b200
from gpirt_b200 import _lib
from oracle import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "gpirt_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(gpirt_[A-Za-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 40
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/gpirt_b200.h but not exported"
    assert set(_lib.SYMBOLS) <= set(names)


def test_no_cpu_fallback():
    if _lib.lib().gpirt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(gpirt_b200.GpirtError) as e:
        gpirt_b200.K([0.0], [0.0], [1.0, 1.0, 1.0])
    assert e.value.code == _lib.ERR_NO_DEVICE
    with pytest.raises(gpirt_b200.GpirtError) as e:
        gpirt_b200.Sampler(10, 5, 3)
    assert e.value.code == _lib.ERR_NO_DEVICE


def test_product_never_touches_the_oracle():
    """no import, include, link or path reference to oracle/ anywhere in the product tree"""
    pat = re.compile(r"(from|import)\s+oracle|#include\s+[\"<][^\n]*orc_|oracle/|liborc|\borc\.")
    for base, _, files in os.walk(os.path.join(ROOT, "gpirt_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                src = open(os.path.join(base, f)).read()
                assert not pat.search(src), f"{f} references the oracle"


def test_argument_validation_before_any_device_work():
    L = _lib.lib()
    h = C.c_void_p()
    for field, value, code in (("horizon", 2, _lib.ERR_UNSUPPORTED), ("constant_IRF", 1, _lib.ERR_UNSUPPORTED), ("C", 22, _lib.ERR_CATEGORIES),
                               ("C", 1, _lib.ERR_CATEGORIES), ("n", 0, _lib.ERR_SHAPE), ("p_beta", 4, _lib.ERR_SHAPE), ("world", 0, _lib.ERR_INVALID)):
        cfg = _lib.Config(n=10, m_total=5, horizon=1, C=3, constant_IRF=0, p_beta=3, theta_os=1.0, theta_ls=10.0, kernel=0, device=0,
                          seed=1, chain=0, rank=0, world=1)
        setattr(cfg, field, value)
        assert L.gpirt_create(C.byref(cfg), C.byref(h)) == code, field
        assert not h.value and L.gpirt_last_error()
    assert L.gpirt_sweep(None, 1) == _lib.ERR_INVALID
    with pytest.raises(gpirt_b200.GpirtError):
        gpirt_b200.gpirtMCMC(np.ones((4, 3)), 2, 0, vote_codes={"yea": [1]})
    # recover_fstar: a thresholds cube must carry one slice per session
    n, m, Ccat = 6, 3, 3
    thr = gpirt_b200.default_thresholds(m, Ccat)
    with pytest.raises(ValueError, match="thresholds"):
        gpirt_b200.recover_fstar(1, np.zeros((n, m, 2)), np.ones((n, m, 2)), np.zeros((n, 2)), np.zeros((2, m, 2)), thr[:, :, None])


def test_threshold_bookkeeping_is_bit_exact():
    rng = np.random.default_rng(0)
    for Ccat in (2, 3, 5, 7, 21):
        d = rng.standard_normal(Ccat - 1)
        t = gpirt_b200.delta_to_threshold(d)
        to = orc.delta_to_threshold(d)
        assert np.array_equal(t, to) and np.isneginf(t[0]) and np.isposinf(t[-1])
        assert np.array_equal(gpirt_b200.threshold_to_delta(t), orc.threshold_to_delta(to))


def test_default_thresholds_are_qnorm_quantiles():
    for Ccat in (2, 5, 7):
        t = gpirt_b200.default_thresholds(3, Ccat)
        assert np.allclose(orc.pnorm(t[0, 1:Ccat]), np.arange(1, Ccat) / Ccat, atol=1e-15)
        assert np.isneginf(t[:, 0]).all() and np.isposinf(t[:, Ccat]).all()


def test_store_index_arithmetic_of_the_mcmc_wrapper():
    # n_samples = int(sample/THIN); stored when iter >= burn and iter % THIN == 0 (src/gpirtMCMC.cpp:48,334-335)
    for sample, burn, thin in ((10, 10, 5), (6, 4, 2), (5, 0, 1)):
        n_samples = int(sample / thin)
        idx = [(it - burn) // thin for it in range(sample + burn) if it >= burn and it % thin == 0]
        assert idx == list(range(n_samples))


def test_r_shim_builds_against_stand_in_rcpp_and_links_the_c_abi():
    """integration/r-package/src/*.cpp (the bodies a maintainer swaps into the R package) compile against the stand-in headers and resolve
    gpirt_mcmc / gpirt_recover_fstar from libgpirt_b200.so; running them needs a GPU (tests/test_gpu_entry.py)."""
    import ctypes as C
    import os
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(["make", "-C", os.path.join(root, "tests", "rshim")], check=True, capture_output=True)
    L = C.CDLL(os.path.join(root, "tests", "rshim", "_build", "librshim_test.so"))
    for sym in ("shim_gpirtMCMC", "shim_recover_fstar", "shim_last_error", "shim_normals_used"):
        getattr(L, sym)


def test_memory_estimate_reproduces_the_reference_arithmetic_and_warning():
    """src/gpirtMCMC.cpp:47-82: MB of the kept draws, total over what is stored, warning above 10 000 MB. Host arithmetic: runs without a device."""
    from gpirt_b200.api import memory_estimate

    a = _lib.McmcArgs()
    a.n, a.m, a.horizon, a.C = 2000, 5000, 1, 5
    a.sample_iterations, a.burn_iterations, a.THIN, a.store_f, a.store_fstar = 1000, 500, 4, 0, 0
    num, text = memory_estimate(a)
    mb = 1024.0 * 1024.0
    assert num["n_samples"] == 250
    assert num["theta_mb"] == 250 * 2000 * 8.0 / mb and num["beta_mb"] == 250 * 3 * 5000 * 8.0 / mb
    assert num["threshold_mb"] == 250 * 5000 * 6 * 8.0 / mb  # the reference's C is thresholds.n_cols = categories + 1
    assert num["total_mb"] == num["theta_mb"] + num["beta_mb"] + num["threshold_mb"] and num["warn"] == 0
    assert "=== MEMORY ESTIMATE ===" in text and "Samples to store: 250 (thinned from 1000)" in text and "DISABLED - will skip" in text and "WARNING" not in text
    a.store_f = 1  # 250 draws of 2000 x 5000 doubles: 19 GB
    num, text = memory_estimate(a)
    assert num["warn"] == 1 and num["total_mb"] > 10000 and "WARNING: Estimated memory usage exceeds 10 GB!" in text and "F samples:         19073.5 MB (ENABLED)" in text
