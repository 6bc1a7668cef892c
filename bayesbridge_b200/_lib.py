"""ctypes binding of libbayesbridge_b200.so (include/bayesbridge_b200.h).

The shared library is the product; this module only declares its signatures.  There is no
CPU fallback: if the library cannot be loaded the import error says how to build it, and every
compute entry point fails when no CUDA device is usable.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbayesbridge_b200.so")
LIB_PATH = os.environ.get("BB_LIB_PATH", LIB_PATH)   # development: try another build of the same library

dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int)
up = C.POINTER(C.c_uint)

BB_TRI, BB_STB = 0, 1
MEM_HOST, MEM_DEVICE = 0, 1
ST_TAU, ST_SIG2, ST_OMEGA, ST_U, ST_BETA, ST_LAMBDA, ST_ALPHA, ST_ALL = 1, 2, 4, 8, 16, 32, 64, 127
OPT_WIDE_GAUSSIAN = 1
OUT_BETA, OUT_U, OUT_OMEGA, OUT_SHAPE, OUT_LAMBDA, OUT_SIG2, OUT_TAU, OUT_ALPHA = 1, 2, 4, 8, 16, 32, 64, 128

# every symbol include/bayesbridge_b200.h declares
EXPORTS = [
    "bridge_regression", "bridge_reg_stable", "bb_set_seed", "bb_set_device", "bb_set_verbose", "bb_set_interrupt", "bb_scalar_status",
    "bb_last_error", "bb_device_count", "bb_model_create", "bb_model_create_factors", "bb_gram", "bb_gram_noalloc",
    "bb_model_create_gram", "bb_model_resid_partial", "bb_model_set_r0", "bb_model_free", "bb_model_clone", "bb_model_status", "bb_model_get",
    "bb_model_precompute_ms", "bb_chains_create", "bb_chains_free", "bb_chains_set", "bb_chains_get", "bb_chains_alpha",
    "bb_chains_option", "bb_set_wide_gaussian", "bb_chains_reserve", "bb_chains_run", "bb_chains_store_state", "bb_chains_fetch", "bb_chains_summary", "bb_chains_summary_begin", "bb_chains_summary_add", "bb_chains_summary_end", "bb_launch_count", "bb_measure_fp64_fma",
    "bb_bridge_batched", "bb_bridge_batched_multi", "bb_bridge_batched_summary",
    "rtexpon_rate_left", "rtexpon_rate_both", "rtexpon_rate", "rtnorm_left", "rtnorm_both", "rtnorm", "rrtgamma_rate", "retstable_LD",
]


class Hyper(C.Structure):
    """bb_hyper (BridgeWrapper.cpp:89-100)."""
    _fields_ = [("sig2_shape", C.c_double), ("sig2_scale", C.c_double), ("nu_shape", C.c_double), ("nu_rate", C.c_double),
                ("alpha_a", C.c_double), ("alpha_b", C.c_double), ("true_sig2", C.c_double), ("true_tau", C.c_double),
                ("true_alpha", C.c_double), ("betaburn", C.c_int)]


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C bayesbridge_b200/csrc` (needs nvcc; there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.bb_last_error.restype = C.c_char_p
    L.bb_device_count.restype = C.c_int
    L.bb_launch_count.restype = C.c_longlong
    L.bb_set_seed.argtypes = [C.c_ulonglong]
    L.bb_set_device.argtypes = [C.c_int]
    L.bb_set_verbose.argtypes = [C.c_int]
    L.bridge_regression.argtypes = [dp] * 18 + [ip] * 4 + [dp, ip, ip, ip]
    L.bridge_regression.restype = None
    L.bridge_reg_stable.argtypes = [dp] * 16 + [ip] * 4 + [dp, ip]
    L.bridge_reg_stable.restype = None
    L.bb_model_create.restype = vp
    L.bb_model_create.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_int]
    L.bb_model_create_factors.restype = vp
    L.bb_model_create_factors.argtypes = [vp, vp, C.c_int, C.c_int, dp, dp, dp, dp, C.c_int, C.c_int]
    L.bb_gram.argtypes = [vp, vp, C.c_int, C.c_int, vp, C.c_int]
    L.bb_gram_noalloc.argtypes = [vp, vp, C.c_int, C.c_int, vp, C.c_int]
    L.bb_model_create_gram.restype = vp
    L.bb_model_create_gram.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    L.bb_model_resid_partial.argtypes = [vp, vp, vp, C.c_int, vp]
    L.bb_model_set_r0.argtypes = [vp, C.c_double, C.c_double]
    L.bb_model_free.argtypes = [vp]
    L.bb_model_free.restype = None
    L.bb_model_get.argtypes = [vp] + [dp] * 6
    L.bb_model_precompute_ms.restype = C.c_double
    L.bb_model_precompute_ms.argtypes = [vp]
    L.bb_chains_create.restype = vp
    L.bb_chains_create.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_ulonglong, C.c_uint]
    L.bb_chains_free.argtypes = [vp]
    L.bb_chains_free.restype = None
    L.bb_chains_set.argtypes = [vp] + [dp] * 5
    L.bb_chains_get.argtypes = [vp] + [dp] * 6 + [up]
    L.bb_chains_alpha.argtypes = [vp, dp, dp]
    L.bb_chains_option.argtypes = [vp, C.c_int, C.c_int]
    L.bb_set_wide_gaussian.argtypes = [C.c_int]
    L.bb_set_wide_gaussian.restype = None
    L.bb_chains_reserve.argtypes = [vp, C.c_longlong, C.c_int]
    L.bb_chains_run.argtypes = [vp, C.POINTER(Hyper), C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_longlong, dp]
    L.bb_chains_store_state.argtypes = [vp, C.c_longlong]
    L.bb_chains_fetch.argtypes = [vp, C.c_int, C.c_longlong, C.c_longlong, vp]
    L.bb_chains_summary.argtypes = [vp, C.c_int, C.c_longlong, C.c_longlong, C.c_double, vp, vp]
    L.bb_chains_summary_begin.argtypes = [vp, C.c_int]
    L.bb_chains_summary_add.argtypes = [vp, C.c_longlong, C.c_longlong]
    L.bb_chains_summary_end.argtypes = [vp, C.c_double, vp, vp]
    L.bb_measure_fp64_fma.argtypes = [C.c_int, dp]
    L.bb_bridge_batched.argtypes = [C.c_int, C.c_int, dp, dp, C.c_int, C.c_int, C.POINTER(Hyper), C.c_int, C.c_ulonglong,
                                    C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int] + [dp] * 7 + [up, dp]
    L.bb_bridge_batched_multi.argtypes = [C.c_int, C.c_int, dp, dp, C.c_int, C.c_int, C.POINTER(Hyper), C.c_int, C.c_ulonglong,
                                          C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, ip] + [dp] * 7 + [up, dp]
    L.bb_set_interrupt.argtypes = [C.c_void_p]
    L.bb_set_interrupt.restype = None
    L.bb_scalar_status.restype = C.c_uint
    L.bb_model_clone.restype = vp
    L.bb_model_clone.argtypes = [vp, C.c_int]
    L.bb_model_status.argtypes = [vp]
    L.bb_bridge_batched_summary.argtypes = [C.c_int, C.c_int, dp, dp, C.c_int, C.c_int, C.POINTER(Hyper), C.c_int, C.c_ulonglong,
                                            C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int] + [dp] * 4 + [up, dp]
    for name, nd in (("rtexpon_rate_left", 3), ("rtexpon_rate_both", 4), ("rtexpon_rate", 4), ("rtnorm_left", 4),
                     ("rtnorm_both", 5), ("rtnorm", 5), ("rrtgamma_rate", 4), ("retstable_LD", 4)):
        fn = getattr(L, name)
        fn.argtypes = [dp] * nd + [ip]
        fn.restype = None
    _lib = L
    return L


def last_error():
    return load().bb_last_error().decode()
