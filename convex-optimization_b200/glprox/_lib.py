"""ctypes binding of libglprox.so (C ABI declared in include/glprox.h).

The library is the product: if it is missing or cannot be loaded this module raises -- there is no
Python / NumPy / torch fallback for the numerics.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libglprox.so")

GLP_F32, GLP_F64 = 0, 1
GLP_PROXGD, GLP_FPROXGD, GLP_FGD, GLP_GD, GLP_SGD = 0, 1, 2, 3, 4
STEP_TYPES = {"line_search": 0, "fixed": 1, "diminishing": 2, "diminishing2": 3, "bb": 4}
BACKENDS = {"auto": 0, "stream": 1, "persistent": 2}

GLP_OK = 0
GLP_ERR_ARG, GLP_ERR_CUDA, GLP_ERR_NO_DEVICE, GLP_ERR_NCCL = -1, -2, -3, -4
GLP_ERR_STATE, GLP_ERR_DEVICE_FAULT, GLP_ERR_UNBOUND = -5, -6, -7


class GlpOpts(C.Structure):
    _fields_ = [("maxit", C.c_int32), ("step_type", C.c_int32), ("stable_len_threshold", C.c_int32),
                ("maxit_line_search_iter", C.c_int32), ("thres", C.c_double), ("alpha0", C.c_double),
                ("ftol", C.c_double), ("line_search_attenuation_coeffi", C.c_double),
                ("backend", C.c_int32), ("poll_interval", C.c_int32), ("delta", C.c_double)]


class GlpResult(C.Structure):
    _fields_ = [("status", C.c_int32), ("iters", C.c_int32), ("done", C.c_int32),
                ("backend_used", C.c_int32), ("fval", C.c_double), ("device_ms", C.c_double),
                ("sparsity", C.c_double), ("n_fwd", C.c_int64), ("n_tpass", C.c_int64),
                ("n_launches", C.c_int64), ("stage_end", C.c_int32 * 3), ("reserved", C.c_int32)]


class GlpStepParams(C.Structure):
    _fields_ = [("algo", C.c_int32), ("reserved", C.c_int32), ("theta", C.c_double), ("t", C.c_double),
                ("mu", C.c_double), ("thres", C.c_double)]


class GlpProfile(C.Structure):
    _fields_ = [("fwd_launches", C.c_int64), ("tpass_launches", C.c_int64), ("fwd_ms_sum", C.c_double),
                ("fwd_ms_min", C.c_double), ("tpass_ms_sum", C.c_double), ("tpass_ms_min", C.c_double),
                ("bytes_per_launch", C.c_double)]


class GlpAdmmOpts(C.Structure):
    _fields_ = [("maxit", C.c_int32), ("converge_len", C.c_int32), ("thres", C.c_double), ("tau", C.c_double),
                ("rho", C.c_double)]


class GlpAdmmPrimalOpts(C.Structure):
    _fields_ = [("maxit", C.c_int32), ("converge_len", C.c_int32), ("step_type", C.c_int32), ("pad", C.c_int32),
                ("thres", C.c_double), ("tau", C.c_double), ("rho", C.c_double), ("eta_0", C.c_double)]


class GlpMetrics(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("sumsq_x", "sumsq_ref", "sumsq_diff", "reg", "max_abs", "count", "rr",
                                          "objective", "sparsity", "err_exact", "err_ref")]


class GlproxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libglprox error %d: %s" % (code, msg))
        self.code = code


# name -> (restype, argtypes); every symbol include/glprox.h declares
_VP, _I, _I64, _D = C.c_void_p, C.c_int, C.c_int64, C.c_double
SIGNATURES = {
    "glp_version": (_I, []),
    "glp_last_error": (C.c_char_p, []),
    "glp_device_count": (_I, []),
    "glp_default_opts": (_I, [_I, C.POINTER(GlpOpts)]),
    "glp_create": (_I, [C.POINTER(_VP), _I, _I, _I, _I, _I, _VP, _I64, _VP, _VP, _VP]),
    "glp_destroy": (_I, [_VP]),
    "glp_fwd": (_I, [_VP, _VP, _I, _VP, _VP]),
    "glp_tpass_prox": (_I, [_VP, C.POINTER(GlpStepParams), _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "glp_solve": (_I, [_VP, _I, _VP, _D, C.POINTER(GlpOpts), _VP, C.POINTER(GlpResult), _VP, _I]),
    "glp_solve_begin": (_I, [_VP, _I, _VP, _D, C.POINTER(GlpOpts)]),
    "glp_solve_step": (_I, [_VP, _I]),
    "glp_solve_poll": (_I, [_VP, C.POINTER(GlpResult)]),
    "glp_solve_end": (_I, [_VP, _VP, C.POINTER(GlpResult), _VP, _I]),
    "glp_debug_read": (_I, [_VP, _I, _VP]),
    "glp_set_profiling": (_I, [_VP, _I]),
    "glp_get_profile": (_I, [_VP, C.POINTER(GlpProfile)]),
    "glp_solve_host": (_I, [_I, _I, _I, _I, _I, _I, _VP, _I64, _VP, _VP, _D, C.POINTER(GlpOpts), _VP,
                            C.POINTER(GlpResult), _VP, _I]),
    "glp_comm_export": (_I, [_VP, _I, _VP, C.POINTER(_VP)]),
    "glp_comm_attach": (_I, [_VP, _I, _I, _I64, _VP, _I]),
    "glp_group_solve": (_I, [C.POINTER(_VP), _I, _I, C.POINTER(_VP), _D, C.POINTER(GlpOpts), C.POINTER(_VP),
                             C.POINTER(GlpResult), _VP, _I]),
    "glp_gemm_tf32": (_I, [_I, _VP, _I, _I64, _I64, _I, _VP, _VP, _I64, _VP, _VP, _I64, _VP, _I64, _I]),
    "glp_set_tensor_mode": (_I, [_VP, _I]),
    "glp_tf32_split": (_I, [_VP, _VP, _I64, _I64, _I64, _VP, _VP, _I64, _I]),
    "glp_nccl_unique_id": (_I, [_VP]),
    "glp_nccl_comm_create": (_I, [C.POINTER(_VP), _VP, _I, _I, _I]),
    "glp_nccl_comm_destroy": (_I, [_VP]),
    "glp_comm_attach_nccl": (_I, [_VP, _VP, _I, _I, _I64]),
    "glp_admm_default_opts": (_I, [C.POINTER(GlpAdmmOpts)]),
    "glp_admm_dual": (_I, [_VP, _VP, _D, C.POINTER(GlpAdmmOpts), _VP, C.POINTER(GlpResult), _VP, _I]),
    "glp_admm_primal_default_opts": (_I, [C.POINTER(GlpAdmmPrimalOpts)]),
    "glp_admm_primal": (_I, [_VP, _VP, _D, C.POINTER(GlpAdmmPrimalOpts), _VP, C.POINTER(GlpResult), _VP, _I]),
    "glp_alm_dual": (_I, [_VP, _VP, _D, C.POINTER(GlpAdmmOpts), _VP, C.POINTER(GlpResult), _VP, _I]),
    "glp_metrics_eval": (_I, [_VP, _VP, _VP, _D, _D, C.POINTER(GlpMetrics)]),
    "glp_metrics_sums_noalloc": (_I, [_I, _VP, _I, _VP, _VP, _I64, _I, _D, _VP]),
    "glp_randn": (_I, [_VP, _I, _I64, _I64, _I64, C.c_uint64, _I64, _I64, _VP]),
    "glp_rhs_exact": (_I, [_VP, _I, _I64, _I64, _VP, _I, _VP, _I64, _VP, _VP]),
    "glp_decide": (_I, [_I, _VP, _VP, _VP, _D, _VP, _VP, _VP, _I]),
    "glp_solve_sparsity_hist": (_I, [_VP, _VP, _I]),
}

_lib = None


def load():
    """Load libglprox.so (once).  Raises ImportError with build instructions when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s not found: the CUDA library is the only implementation of this path (no CPU fallback). "
            "Build it with `python __graft_entry__.py build` (nvcc, sm_100a)." % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == ABI mismatch with include/glprox.h
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != GLP_OK:
        msg = load().glp_last_error()
        raise GlproxError(rc, msg.decode("utf-8", "replace") if msg else "")
    return rc
