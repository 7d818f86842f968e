"""ctypes binding of libshopt_b200.so (include/shopt_b200.h).

The shared library is the product; this module only declares its C ABI.  There is no fallback of any kind:
if the library has not been built, importing the compute functions raises, and if no CUDA device is
present every compute call returns SHOPT_E_NODEVICE, which `check()` turns into an exception.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libshopt_b200.so")

OK, E_NODEVICE, E_BADARG, E_CUDA, E_NOMEM, E_UNSUPPORTED = 0, -1, -2, -3, -4, -5
FIT_CONVERGED, FIT_MAXITER, FIT_LS_FAILED, FIT_NONFINITE, FIT_THREW = 0, 1, 2, 3, -1
PCA_OK, PCA_FAILED, PCA_NOCONV = 0, 1, 2

_ERR = {E_NODEVICE: "no CUDA device", E_BADARG: "bad argument", E_CUDA: "CUDA error",
        E_NOMEM: "device allocation failed", E_UNSUPPORTED: "size not supported by the kernels"}


class ShoptError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        super().__init__(f"libshopt_b200: {_ERR.get(code, 'error')} ({code}) {detail}".strip())


dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int32)
lp = C.POINTER(C.c_int64)
vp = C.c_void_p


class Params(C.Structure):
    """struct shopt_params (shopt.yml keys on the hot path)."""
    _fields_ = [("n_a", C.c_int), ("g_tol_model", C.c_double), ("g_tol_learned", C.c_double),
                ("max_iter", C.c_int), ("s_lo", C.c_double), ("s_hi", C.c_double), ("pca_terms", C.c_int),
                ("pratio", C.c_double), ("lanczos", C.c_int), ("unity_sum", C.c_int), ("degree", C.c_int),
                ("iterations", C.c_int), ("training_ratio", C.c_double), ("median_constraint", C.c_int),
                ("sigma_clip", C.c_int), ("mode", C.c_int), ("chisq_fill", C.c_double), ("profile", C.c_int)]


class StageIO(C.Structure):
    """struct shopt_stage_io."""
    _fields_ = [("stamps", vp), ("r", C.c_int), ("c", C.c_int), ("n_local", C.c_int64), ("u", vp), ("v", vp),
                ("x0_model", vp), ("x0_learned", vp), ("n_global", C.c_int64), ("offset", C.c_int64), ("perm", vp),
                ("model", vp), ("learned", vp), ("pgf", vp), ("flags", vp), ("pca_k", vp), ("poly", vp),
                ("n_per_round", vp), ("counts", C.c_int64 * 4), ("weights", vp), ("x0_pixels", vp),
                ("field", C.c_double * 30), ("field_rank", C.c_int32),
                ("polyfit_rank", C.c_int32), ("polyfit_info", C.c_int32), ("stamps_f32", C.c_int32), ("perm_dev", vp),
                ("model_status", vp), ("model_iters", vp), ("learned_status", vp), ("learned_iters", vp)]


class SummaryFull(C.Structure):
    """struct shopt_summary_full."""
    _fields_ = [("poly", dp), ("r", C.c_int), ("c", C.c_int), ("degree", C.c_int), ("sC", dp), ("g1C", dp), ("g2C", dp),
                ("n_stars", C.c_int64), ("u", dp), ("v", dp), ("s_model", dp), ("g1_model", dp), ("g2_model", dp),
                ("s_data", dp), ("g1_data", dp), ("g2_data", dp), ("stars", dp), ("pgf", dp),
                ("n_train", C.c_int64), ("train_u", dp), ("train_v", dp),
                ("n_valid", C.c_int64), ("valid_u", dp), ("valid_v", dp), ("valid_stars", dp),
                ("n_flag_snr", C.c_int64), ("n_flag_blacklist", C.c_int64), ("n_flag_failed", C.c_int64),
                ("flag_snr", dp), ("flag_blacklist", dp), ("flag_failed", dp)]


ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p)
HOST_ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int, C.c_void_p)
MODE_PCA, MODE_SMOOTH, MODE_CHISQ = 0, 1, 2
PROFILE_GAUSSIAN, PROFILE_KOLMOGOROV = 0, 1
FLAG_BLACKLISTED, FLAG_PCA_FAILED, FLAG_LEARNED_OOB, FLAG_TRAINING, FLAG_REJECTED = 1, 2, 4, 8, 16
POLY_RANK_DEFICIENT, POLY_NONFINITE, POLY_ILL_CONDITIONED = 1, 2, 4
UNIQUE_ID_BYTES = 128

# name -> (restype, argtypes); every symbol include/shopt_b200.h declares
SIGNATURES = {
    "shopt_b200_version": (C.c_int, []),
    "shopt_device_count": (C.c_int, []),
    "shopt_set_device": (C.c_int, [C.c_int]),
    "shopt_last_error": (C.c_char_p, []),
    "shopt_last_timing": (C.c_int, [dp]),
    "shopt_release_buffers": (C.c_int, []),
    "shopt_launch_count": (C.c_int64, []),
    "shopt_bench_dfma": (C.c_int, [dp]),
    "shopt_comm_unique_id": (C.c_int, [vp]),
    "shopt_comm_init": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(vp)]),
    "shopt_comm_destroy": (C.c_int, [vp]),
    "shopt_comm_rank": (C.c_int, [vp]),
    "shopt_comm_world": (C.c_int, [vp]),
    "shopt_comm_allreduce": (C.c_int, [vp, C.c_int64, C.c_int, vp, vp]),
    "shopt_fit_analytic_batch": (C.c_int, [dp, C.c_int, C.c_int, C.c_int64, C.c_int, dp, C.c_double, C.c_int,
                                           dp, dp, dp, ip, ip]),
    "shopt_fit_analytic_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, C.c_int, vp, C.c_double, C.c_int,
                                               vp, vp, vp, vp, vp, vp]),
    "shopt_fit_profile_batch": (C.c_int, [C.c_int, dp, C.c_int, C.c_int, C.c_int64, C.c_int, dp, C.c_double, C.c_int,
                                          dp, dp, dp, ip, ip]),
    "shopt_fit_profile_batch_dev": (C.c_int, [C.c_int, vp, C.c_int, C.c_int, C.c_int64, C.c_int, vp, C.c_double, C.c_int,
                                              vp, vp, vp, vp, vp, vp]),
    "shopt_profile_cost_grad_batch": (C.c_int, [C.c_int, dp, C.c_int, C.c_int, C.c_int64, C.c_int, dp, dp, dp]),
    "shopt_profile_cost_grad_batch_dev": (C.c_int, [C.c_int, vp, C.c_int, C.c_int, C.c_int64, C.c_int, vp, vp, vp, vp]),
    "shopt_analytic_cost_grad_batch": (C.c_int, [dp, C.c_int, C.c_int, C.c_int64, C.c_int, dp, dp, dp]),
    "shopt_analytic_cost_grad_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, C.c_int, vp, vp, vp, vp]),
    "shopt_s_out_of_bounds": (C.c_int64, [dp, C.c_int64, C.c_double, C.c_double, lp]),
    "shopt_pca_batch": (C.c_int, [dp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_double, C.c_int, C.c_int, dp, ip, ip]),
    "shopt_pca_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_double, C.c_int, C.c_int,
                                      vp, vp, vp, vp]),
    "shopt_smooth_batch": (C.c_int, [dp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, dp]),
    "shopt_smooth_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, vp, vp]),
    "shopt_smooth_mode_batch": (C.c_int, [dp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, dp]),
    "shopt_smooth_mode_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, vp, vp]),
    "shopt_chisq_batch": (C.c_int, [dp, dp, dp, C.c_double, C.c_int, C.c_int, C.c_int64, C.c_int, dp]),
    "shopt_chisq_batch_dev": (C.c_int, [vp, vp, vp, C.c_double, C.c_int, C.c_int, C.c_int64, C.c_int, vp, vp]),
    "shopt_ingest_batch": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, dp]),
    "shopt_ingest_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, vp, vp]),
    "shopt_snr_cut": (C.c_int64, [C.POINTER(C.c_float), C.c_int64, C.c_double, lp, dp]),
    "shopt_summary_write": (C.c_int, [C.c_char_p, dp, C.c_int, C.c_int, C.c_int, dp, dp, dp]),
    "shopt_summary_write_full": (C.c_int, [C.c_char_p, C.POINTER(SummaryFull)]),
    "shopt_summary_read": (C.c_int, [C.c_char_p, dp, C.c_int64, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                     C.POINTER(C.c_int), dp, dp, dp]),
    "shopt_field_interp": (C.c_int, [dp, dp, C.POINTER(dp), C.c_int, C.c_int64, dp]),
    "shopt_field_eval": (C.c_int, [dp, dp, dp, C.c_int64, dp]),
    "shopt_polyfit": (C.c_int, [dp, C.c_int, C.c_int, C.c_int64, dp, dp, lp, C.c_int64, C.c_int, C.c_int, C.c_int,
                                C.c_int, dp, dp, lp, lp, lp]),
    "shopt_polyfit_acc_len": (C.c_int64, [C.c_int, C.c_int, C.c_int]),
    "shopt_polyfit_centre_dev": (C.c_int, [vp, vp, C.c_int64, vp, vp]),
    "shopt_polyfit_normal_dev": (C.c_int, [vp, C.c_int, C.c_int, vp, vp, vp, C.c_int64, C.c_int, vp, vp, vp]),
    "shopt_polyfit_solve_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int64, vp, vp, vp]),
    "shopt_polyfit_factor_host": (C.c_int, [dp, C.c_int, dp, C.c_int64, dp, C.c_int, dp, ip]),
    "shopt_polyfit_mse_dev": (C.c_int, [vp, C.c_int, C.c_int, vp, vp, vp, C.c_int64, C.c_int, vp, vp, vp]),
    "shopt_polyfit_mse_noalloc_dev": (C.c_int, [vp, C.c_int, C.c_int, vp, vp, vp, C.c_int64, C.c_int, vp, vp, vp]),
    "shopt_polyfit_mse_noalloc_max_stars": (C.c_int64, [C.c_int, C.c_int, C.c_int]),
    "shopt_worst_10_percent": (C.c_int64, [dp, C.c_int64, C.c_int, lp]),
    "shopt_psf_at_batch": (C.c_int, [dp, C.c_int, C.c_int, C.c_int, dp, dp, C.c_int64, dp]),
    "shopt_psf_at_batch_dev": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int64, vp, vp]),
    "shopt_default_params": (None, [C.POINTER(Params)]),
    "shopt_training_split": (C.c_int64, [ip, C.c_int64, lp, C.c_double, lp]),
    "shopt_training_split_sharded_host": (C.c_int64, [ip, C.c_int64, C.c_int64, C.c_int64, lp, C.c_double, vp, vp, lp,
                                                      C.POINTER(C.c_uint32), lp]),
    "shopt_select_worst_sharded_host": (C.c_int64, [dp, C.POINTER(C.c_uint32), C.c_int64, C.c_int64, vp, vp,
                                                    C.POINTER(C.c_uint8)]),
    "shopt_median_sharded_host": (C.c_int, [dp, C.c_int64, C.c_int64, vp, vp, dp]),
    "shopt_stage_run_dev": (C.c_int, [C.POINTER(Params), C.POINTER(StageIO), C.c_void_p, C.c_void_p, C.c_void_p]),
    "shopt_stage_run": (C.c_int, [C.POINTER(Params), C.POINTER(StageIO)]),
    "shopt_stage_run_comm": (C.c_int, [C.POINTER(Params), C.POINTER(StageIO), C.c_void_p, C.c_void_p]),
}

_LIB = None


def lib():
    """Load libshopt_b200.so; raises if it has not been built (python __graft_entry__.py / make -C shopt_b200/csrc)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ShoptError(E_NODEVICE, f"{LIB_PATH} is missing: build it with `make -C shopt_b200/csrc`; "
                                         "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(code):
    if code != OK:
        detail = lib().shopt_last_error()
        raise ShoptError(code, detail.decode() if detail else "")
    return code
