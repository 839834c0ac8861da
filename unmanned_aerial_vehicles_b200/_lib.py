"""ctypes binding of libgpk.so (C ABI declared in include/gpk.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``csrc/Makefile``.  There is no CPU
fallback: if the shared object is missing, every compute entry point raises ``GpkUnavailable``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GPK_LIB", os.path.join(_HERE, "libgpk.so"))

# every symbol include/gpk.h declares (tests check the .so exports exactly these)
SYMBOLS = (
    "gpk_abi_version", "gpk_pad", "gpk_lml_out_len", "gpk_grad_workspace_doubles", "gpk_gram",
    "gpk_potrf_inv", "gpk_potrs", "gpk_lauum", "gpk_lml_grad", "gpk_lml_eval", "gpk_cross_gram",
    "gpk_predict_mean", "gpk_predict_std", "gpk_predict_workspace_doubles", "gpk_predict_mean_workspace_doubles", "gpk_predict_mean_fused", "gpk_predict_mean_fused_static", "gpk_predict", "gpk_predict_stacked", "gpk_residual_metrics_out_len", "gpk_residual_metrics_workspace_doubles", "gpk_residual_metrics", "gpk_extract_lower", "gpk_pad_copy", "gpk_gemm_nt",
    "gpk_dmma_probe", "gpk_launch_count", "gpk_profile_begin", "gpk_profile_end", "gpk_profile_launch", "gpk_leaf_phase_cycles", "gpk_gemv", "gpk_trimv", "gpk_fix_diag", "gpk_logdiag",
    "gpk_grad_block_workspace_doubles", "gpk_grad_padded_dim", "gpk_grad_block", "gpk_grad_block_sum",
)


class GpkUnavailable(RuntimeError):
    pass


class GpkError(RuntimeError):
    pass


_lib = None
_p = C.c_void_p
_i = C.c_int
_d = C.c_double
_dp = C.POINTER(C.c_double)


MAX_STACKED = 16   # GPK_MAX_STACKED in include/gpk.h
STATIC_SCRATCH_DOUBLES = 524288           # GPK_STATIC_SCRATCH_DOUBLES
ERR_STATIC_SCRATCH_TOO_SMALL = -2000      # GPK_ERR_STATIC_SCRATCH_TOO_SMALL


class StackedModel(C.Structure):
    """GpkStackedModel of include/gpk.h: one model's arguments of gpk_predict_stacked"""
    _fields_ = [("Xq_h", _p), ("D", _i), ("X_d", _p), ("N", _i), ("inv_ls_h", _p), ("c", _d), ("noise", _d),
                ("alpha_d", _p), ("T", _i), ("Mlow_d", _p), ("NP", _i), ("y_mean_d", _p), ("y_std_d", _p),
                ("mean_h", _p), ("std_h", _p), ("neg_h", _p), ("ws_d", _p), ("ws_doubles", C.c_size_t)]


def _declare(lib):
    lib.gpk_abi_version.restype = _i
    lib.gpk_pad.argtypes = [_i]
    lib.gpk_lml_out_len.argtypes = [_i, _i]
    lib.gpk_grad_workspace_doubles.argtypes = [_i, _i]
    lib.gpk_grad_workspace_doubles.restype = C.c_size_t
    lib.gpk_gram.argtypes = [_p, _i, _i, _dp, _d, _d, _d, _p, _i, _i, _p]
    lib.gpk_potrf_inv.argtypes = [_p, _p, _p, _i, _p, _p]
    lib.gpk_potrs.argtypes = [_p, _p, _i, _p, _p, _p, _i, _p]
    lib.gpk_lauum.argtypes = [_p, _p, _i, _p]
    lib.gpk_lml_grad.argtypes = [_p, _i, _i, _dp, _d, _d, _p, _p, _p, _i, _p, _i, _i, _p, _p, _p]
    lib.gpk_lml_eval.argtypes = [_p, _i, _i, _dp, _d, _d, _d, _p, _i, _p, _p, _p, _p, _p, _i, _i, _p, _p, _p, _p]
    lib.gpk_cross_gram.argtypes = [_p, _i, _p, _i, _i, _dp, _d, _p, _i, _i, _p]
    lib.gpk_predict_mean.argtypes = [_p, _i, _i, _i, _p, _i, _p, _p, _p, _p]
    lib.gpk_predict_std.argtypes = [_p, _i, _i, _p, _i, _d, _p, _i, _p, _p, _p, _p]
    lib.gpk_predict_workspace_doubles.argtypes = [_i, _i, _i, _i]
    lib.gpk_predict_workspace_doubles.restype = C.c_size_t
    lib.gpk_predict_mean_workspace_doubles.argtypes = [_i, _i, _i, _i]
    lib.gpk_predict_mean_workspace_doubles.restype = C.c_size_t
    lib.gpk_predict_mean_fused.argtypes = [_p, _i, _i, _p, _i, _i, _dp, _d, _p, _i, _p, _p, _p, _p, _p]
    lib.gpk_predict_mean_fused_static.argtypes = [_p, _i, _i, _p, _i, _i, _dp, _d, _p, _i, _p, _p, _p, _p]
    lib.gpk_predict.argtypes = [_p, _i, _i, _p, _i, _dp, _d, _d, _p, _i, _p, _i, _p, _p, _p, _p, _p, _p, C.c_size_t, _p]
    lib.gpk_predict_stacked.argtypes = [C.POINTER(StackedModel), _i, _i, _p]
    lib.gpk_residual_metrics_workspace_doubles.argtypes = [_i]
    lib.gpk_residual_metrics_workspace_doubles.restype = C.c_size_t
    lib.gpk_residual_metrics.argtypes = [_p, _p, _i, _i, _p, _p, _p]
    lib.gpk_extract_lower.argtypes = [_p, _i, _p, _i, _p]
    lib.gpk_pad_copy.argtypes = [_p, _i, _i, _p, _i, _i, _p]
    lib.gpk_gemm_nt.argtypes = [_p, _i, _p, _i, _p, _i, _i, _i, _i, _d, _i, _i, _i, _i, _p]
    lib.gpk_dmma_probe.argtypes = [_i, _i, _p, _dp, _p]
    lib.gpk_gemv.argtypes = [_p, _i, _i, _i, _p, _i, _p, _d, _d, _i, _p]
    lib.gpk_trimv.argtypes = [_p, _i, _p, _p, _i, _i, _p]
    lib.gpk_fix_diag.argtypes = [_p, _i, _i, _i, _d, _d, _p]
    lib.gpk_logdiag.argtypes = [_p, _i, _i, _p, _p]
    lib.gpk_grad_block_workspace_doubles.argtypes = [_i, _i, _i]
    lib.gpk_grad_block_workspace_doubles.restype = C.c_size_t
    lib.gpk_grad_padded_dim.argtypes = [_i]
    lib.gpk_grad_block.argtypes = [_p, _i, C.c_longlong, _p, _i, C.c_longlong, _i, _dp, _d, _p, _p, _i, _p, _i, _i, _i,
                                   _p, _i, _p]
    lib.gpk_grad_block_sum.argtypes = [_p, _i, _i, _i, _p, _p]
    lib.gpk_leaf_phase_cycles.argtypes = [_p]
    lib.gpk_launch_count.restype = C.c_ulonglong
    lib.gpk_profile_launch.argtypes = [_i, _dp, _dp, C.POINTER(C.c_int)]
    lib.gpk_profile_end.argtypes = [_dp, _dp, C.POINTER(C.c_int)]
    for name in SYMBOLS:
        fn = getattr(lib, name)
        if name not in ("gpk_grad_workspace_doubles", "gpk_predict_workspace_doubles", "gpk_predict_mean_workspace_doubles",
                        "gpk_residual_metrics_workspace_doubles", "gpk_grad_block_workspace_doubles",
                        "gpk_launch_count"):
            fn.restype = _i


def load():
    """Load libgpk.so (once).  Raises GpkUnavailable if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GpkUnavailable(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C unmanned_aerial_vehicles_b200/csrc`. There is no CPU fallback."
            )
        lib = C.CDLL(LIB_PATH)
        _declare(lib)
        if lib.gpk_abi_version() != 1:
            raise GpkUnavailable("libgpk.so ABI version mismatch; rebuild")
        _lib = lib
    return _lib


def check(rc: int, what: str):
    if rc != 0:
        raise GpkError(f"{what} failed with code {rc}" + (" (CUDA error %d)" % (-rc - 1000) if rc <= -1000 else ""))


def darr(values):
    """host double array for the *_h arguments"""
    vals = [float(v) for v in values]
    return (C.c_double * len(vals))(*vals)
