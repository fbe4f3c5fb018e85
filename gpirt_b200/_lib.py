"""ctypes binding of libgpirt_b200.so (the C-ABI in include/gpirt_b200.h).

There is no fallback: if the shared library is missing this module raises at import of the
symbols, and every compute call fails with GPIRT_ERR_NO_DEVICE when no CUDA device is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgpirt_b200.so")
NGRID = 1001

OK, ERR_INVALID, ERR_SHAPE, ERR_CATEGORIES, ERR_NOT_PD, ERR_CUDA, ERR_UNSUPPORTED, ERR_STREAM_OVERRUN, ERR_NO_DEVICE, ERR_ESS_STALLED = range(10)
BLOCK = {"f": 0, "fstar": 1, "theta": 2, "beta": 3, "thr": 4}
XCHG = {"fstar": 0, "beta": 1, "thr": 2, "theta": 3, "item_ll": 4, "theta_part": 5, "items_packed": 6}


class GpirtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gpirt_b200 error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("n", C.c_int32), ("m_total", C.c_int32), ("horizon", C.c_int32), ("C", C.c_int32),
                ("constant_IRF", C.c_int32), ("p_beta", C.c_int32), ("theta_os", C.c_double), ("theta_ls", C.c_double),
                ("kernel", C.c_int32), ("device", C.c_int32), ("seed", C.c_uint64), ("chain", C.c_uint32),
                ("rank", C.c_int32), ("world", C.c_int32)]


class Streams(C.Structure):
    _fields_ = [("z", C.c_void_p), ("z_stride", C.c_int64), ("u", C.c_void_p), ("u_stride", C.c_int64), ("nu", C.c_void_p)]


PROGRESS_FN = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.c_void_p)


class McmcArgs(C.Structure):
    _fields_ = [("y", C.c_void_p), ("n", C.c_int32), ("m", C.c_int32), ("horizon", C.c_int32),
                ("theta", C.c_void_p),
                ("sample_iterations", C.c_int32), ("burn_iterations", C.c_int32), ("THIN", C.c_int32),
                ("beta_prior_means", C.c_void_p), ("beta_prior_sds", C.c_void_p),
                ("theta_prior_means", C.c_void_p), ("theta_prior_sds", C.c_void_p),
                ("theta_os", C.c_double), ("theta_ls", C.c_double), ("KERNEL", C.c_char_p),
                ("thresholds", C.c_void_p), ("C", C.c_int32),
                ("constant_IRF", C.c_int32), ("store_f", C.c_int32), ("store_fstar", C.c_int32),
                ("beta_init", C.c_void_p), ("f_init", C.c_void_p),
                ("seed", C.c_uint64), ("chain", C.c_uint32), ("device", C.c_int32),
                ("theta_draws", C.c_void_p), ("beta_draws", C.c_void_p), ("threshold_draws", C.c_void_p),
                ("f_draws", C.c_void_p), ("fstar_draws", C.c_void_p), ("ll_draws", C.c_void_p),
                ("progress", PROGRESS_FN), ("user", C.c_void_p), ("f_init_z", C.c_void_p),
                ("n_devices", C.c_int32), ("devices", C.c_void_p)]


class MemEstimate(C.Structure):
    _fields_ = [("n_samples", C.c_int32), ("theta_mb", C.c_double), ("beta_mb", C.c_double), ("f_mb", C.c_double), ("fstar_mb", C.c_double),
                ("threshold_mb", C.c_double), ("total_mb", C.c_double), ("warn", C.c_int32)]


class RecoverArgs(C.Structure):
    _fields_ = [("seed", C.c_int32), ("f", C.c_void_p), ("y", C.c_void_p), ("n", C.c_int32), ("m", C.c_int32),
                ("horizon", C.c_int32), ("theta", C.c_void_p), ("beta", C.c_void_p), ("thresholds", C.c_void_p),
                ("C", C.c_int32), ("beta_prior_means", C.c_void_p), ("beta_prior_sds", C.c_void_p),
                ("constant_IRF", C.c_int32), ("device", C.c_int32), ("fstar", C.c_void_p)]


# every symbol include/gpirt_b200.h declares (tests check the library exports each one)
SYMBOLS = [
    "gpirt_last_error", "gpirt_device_count", "gpirt_K", "gpirt_K_time", "gpirt_delta_to_threshold",
    "gpirt_threshold_to_delta", "gpirt_chol", "gpirt_chol_device_noalloc", "gpirt_double_solve", "gpirt_create", "gpirt_destroy", "gpirt_release_cached_memory", "gpirt_set_stream",
    "gpirt_set_data_f64", "gpirt_set_data_i8", "gpirt_set_priors", "gpirt_set_state", "gpirt_get_state", "gpirt_get_L",
    "gpirt_store_open", "gpirt_store_push", "gpirt_store_count", "gpirt_store_flush", "gpirt_store_flush_slice", "gpirt_store_close", "gpirt_store_sample_bytes",
    "gpirt_set_iter", "gpirt_init_f_from_z", "gpirt_panel_create", "gpirt_panel_destroy", "gpirt_panel_session", "gpirt_panel_get_L_time",
    "gpirt_panel_draw_theta", "gpirt_panel_sweep", "gpirt_panel_total_ll", "gpirt_panel_sync", "gpirt_set_solver", "gpirt_set_overlap", "gpirt_set_chol_ctas", "gpirt_set_screen", "gpirt_debug_theta_threads", "gpirt_get_exact_evals", "gpirt_get_diag", "gpirt_sync", "gpirt_update_cholesky_cache", "gpirt_draw_f", "gpirt_draw_fstar",
    "gpirt_draw_theta", "gpirt_gather_f", "gpirt_draw_beta", "gpirt_draw_threshold", "gpirt_ll_bar", "gpirt_sweep",
    "gpirt_sweep_phase_items", "gpirt_sweep_phase_theta", "gpirt_sweep_phase_close", "gpirt_theta_items_begin", "gpirt_theta_items_partials",
    "gpirt_theta_items_decide", "gpirt_items_pack", "gpirt_exchange_pull", "gpirt_chain_export", "gpirt_chain_attach", "gpirt_chain_detach", "gpirt_exchange_buffer",
    "gpirt_total_ll", "gpirt_init_draw", "gpirt_mcmc", "gpirt_memory_estimate", "gpirt_mcmc_chains", "gpirt_copy_data", "gpirt_set_mean_rows", "gpirt_interp1", "gpirt_ipc_export", "gpirt_ipc_attach", "gpirt_ipc_detach", "gpirt_peer_wait", "gpirt_recover_fstar", "gpirt_profile_enable", "gpirt_profile_read",
    "gpirt_profile_reset", "gpirt_profile_timeline", "gpirt_debug_dist_plan", "gpirt_debug_uniforms", "gpirt_debug_normals", "gpirt_launch_count", "gpirt_debug_probit", "gpirt_debug_phi_tab", "gpirt_debug_cell32", "gpirt_debug_fp64_peak", "gpirt_debug_chol_trace",
]

_lib = None


def lib():
    """Load the shared library (built by __graft_entry__.build() / make -C gpirt_b200/csrc)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C gpirt_b200/csrc` "
                              "(gpirt_b200 has no CPU or pure-Python path)")
        L = C.CDLL(LIB_PATH)
        L.gpirt_last_error.restype = C.c_char_p
        L.gpirt_launch_count.restype = C.c_longlong
        L.gpirt_store_sample_bytes.restype = C.c_int64
        for s in SYMBOLS:
            getattr(L, s)  # AttributeError if the build is stale
        _lib = L
    return _lib


def check(rc):
    if rc != OK:
        raise GpirtError(rc, lib().gpirt_last_error().decode())
