"""gpirt_chol_device_noalloc (include/gpirt_b200.h; no counterpart in the reference): the
Cholesky factorisation for callers that cannot allocate device memory. Called through the C-ABI with device pointers, checked against the
existing entry point gpirt_chol on the same matrices, on calls in a row that share the entry point's static scratch.

Collected last on purpose: a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _matrix(n, seed):
    import gpirt_b200

    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(-3, 3, n))
    return gpirt_b200.K(x, x, np.array([0.5, 0.3, 0.0])) + np.eye(n) * 0.5  # distinct theta, large jitter: the well-conditioned fixture of test_gpu_parity


def _device_matrix(A, ld):
    """column-major n x n with leading dimension ld on cuda:0, as a (n, ld) row-major torch tensor (row j = column j of A)"""
    import torch

    n = A.shape[0]
    buf = np.zeros((n, ld))
    buf[:, :n] = A.T
    return torch.from_numpy(buf).to("cuda:0")


def test_no_alloc_cholesky_equals_gpirt_chol_on_calls_in_a_row():
    import torch

    import gpirt_b200
    from gpirt_b200 import _lib

    L = _lib.lib()
    fn = L.gpirt_chol_device_noalloc
    fn.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    cases = [(333, 336), (130, 144), (333, 336), (2048, 2048), (60, 64)]  # ragged tiles, a smaller task list after a larger one, the maximum, one tile
    mats = [_matrix(n, 10 + k) for k, (n, _) in enumerate(cases)]
    want = [gpirt_b200.chol(A) for A in mats]  # the existing entry point
    dev = [_device_matrix(A, ld) for A, (_, ld) in zip(mats, cases)]
    info = torch.full((len(cases),), 7, dtype=torch.int32, device="cuda:0")  # garbage on entry: the call clears its word
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    free = []
    for k, (n, ld) in enumerate(cases):  # back to back on one stream, no synchronisation in between: every call re-initialises the shared scratch
        _lib.check(fn(0, C.c_void_p(dev[k].data_ptr()), ld, n, C.c_void_p(info.data_ptr() + 4 * k), C.c_void_p(stream)))
        if k < 2:
            torch.cuda.synchronize()
            free.append(torch.cuda.mem_get_info(0)[0])
    torch.cuda.synchronize()
    free.append(torch.cuda.mem_get_info(0)[0])
    assert info.cpu().tolist() == [0] * len(cases)
    for k, (n, ld) in enumerate(cases):
        got = dev[k].cpu().numpy()[:, :n].T
        assert np.all(np.triu(got, 1) == 0.0)
        # the tolerances of test_chol_well_conditioned_elementwise (north_star: 1e-10 element-wise on a well-conditioned fixture, backward 1e-14)
        assert np.max(np.abs(got - want[k])) / np.max(np.abs(want[k])) < 1e-10, f"call {k} (n = {n})"
        assert np.max(np.abs(got @ got.T - mats[k])) / np.max(np.abs(mats[k])) < 1e-14
    # no device memory was taken by the calls after the first (the first one loads the kernels' module)
    assert free[1] == free[0] and free[2] == free[0], free


def test_no_alloc_cholesky_rejects_what_its_static_scratch_cannot_hold():
    import torch

    from gpirt_b200 import _lib

    L = _lib.lib()
    fn = L.gpirt_chol_device_noalloc
    fn.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    buf = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    info = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    rc = fn(0, C.c_void_p(buf.data_ptr()), 2050, 2049, C.c_void_p(info.data_ptr()), None)
    assert rc == _lib.ERR_SHAPE and b"2048" in L.gpirt_last_error()
    rc = fn(0, C.c_void_p(buf.data_ptr()), 3, 3, C.c_void_p(info.data_ptr()), None)  # odd leading dimension
    assert rc == _lib.ERR_SHAPE
    rc = fn(0, None, 4, 4, C.c_void_p(info.data_ptr()), None)
    assert rc == _lib.ERR_INVALID
    # not positive definite: reported through the caller's info word, as gpirt_chol reports it through its return code
    A = np.eye(70)
    A[65, 65] = -1.0
    d = _device_matrix(A, 80)
    _lib.check(fn(0, C.c_void_p(d.data_ptr()), 80, 70, C.c_void_p(info.data_ptr()), None))
    torch.cuda.synchronize()
    assert int(info.cpu()[0]) == 66
