"""ctypes binding of libpmx.so (include/pmx.h).  No fallback: a missing library is an error.

``lib()`` loads the in-tree shared object (building it first if the sources are newer and nvcc
is present).  Every device entry point takes raw device pointers and a ``cudaStream_t``; the
``Engine`` class passes ``tensor.data_ptr()`` and ``torch.cuda.current_stream().cuda_stream``.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpmx.so")

OK, E_INVALID, E_CUDA, E_VALUE, E_KEY, E_CAPACITY, E_JSON, E_LIMIT = 0, -1, -2, -3, -4, -5, -6, -7
QUANTILE_BINS = 2048
QUANTILE_PASSES = 6
QUANTILE_GATHER_ROW = 1024
FUSED_MAX_WINDOW_STEPS = 513     # pmx_fused_align_rolling_rate: window_steps - 1 <= 512
INT64_MAX = (1 << 63) - 1


class PmxError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libpmx error {code}: {message}")
        self.code = code
        self.message = message


class LabelsResult(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("n_strings", C.c_int64), ("strings_len", C.c_int64)]


class TokenizeResult(C.Structure):
    _fields_ = [("n_samples", C.c_int64), ("n_series", C.c_int64), ("extra_len", C.c_int64),
                ("result_type", C.c_int32)]


_p = C.c_void_p
_i64 = C.c_int64
_u64 = C.c_uint64
_int = C.c_int

# name -> (restype, argtypes); mirrors include/pmx.h one to one
SIGNATURES = {
    "pmx_version": (_int, []),
    "pmx_last_error": (C.c_char_p, []),
    "pmx_ctx_create": (_int, [_int, C.POINTER(_p)]),
    "pmx_ctx_destroy": (None, [_p]),
    "pmx_ctx_scratch_bytes": (C.c_size_t, [_p]),
    "pmx_tokenize": (_int, [_p, C.c_size_t, _p, _p, _i64, _p, _p, _p, _i64, _p, C.c_size_t,
                            C.POINTER(TokenizeResult)]),
    "pmx_tokenize_mt": (_int, [_p, C.c_size_t, _p, _p, _i64, _p, _p, _p, _i64, _p, C.c_size_t, _int,
                               C.POINTER(TokenizeResult)]),
    "pmx_labels_intern": (_int, [_p, C.c_size_t, _p, _p, _i64, _p, _p, _p, _i64, _p, C.c_size_t, _p, _i64,
                                 C.POINTER(LabelsResult)]),
    "pmx_tokenize_device": (_int, [_p, _p, _u64, _p, _p, _i64, _p, _p, _p, _i64, _p, _p]),
    "pmx_parse_f64": (_int, [_p, _p, _u64, _p, _i64, _p, _p, _p]),
    "pmx_parse_ts_ms": (_int, [_p, _p, _u64, _p, _i64, _p, _p, _p]),
    "pmx_parse_samples": (_int, [_p, _p, _u64, _p, _p, _i64, _p, _p, _p, _p]),
    "pmx_align": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _p, _p]),
    "pmx_bucket": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p]),
    "pmx_rolling": (_int, [_p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p, _p]),
    "pmx_rolling_ragged": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _p, _p, _p]),
    "pmx_rate": (_int, [_p, _p, _i64, _i64, _i64, _i64, _int, _p, _p, _p, _p]),
    "pmx_rate_ragged": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _int, _p, _p, _p, _p]),
    "pmx_fused_align_rolling_rate": (_int, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _i64, _i64,
                                            _p, _p, _p, _p, _p, _p]),
    "pmx_group_index": (_int, [_p, _i64, _i64, _p, _p]),
    "pmx_group_reduce": (_int, [_p, _p, _i64, _i64, _p, _p, _i64, _int, _p, _p, _p, _p, _p, _p]),
    "pmx_group_reduce_noalloc": (_int, [_p, _i64, _i64, _p, _p, _i64, _p, _p, _p, _p, _p, _p]),
    "pmx_group_pack": (_int, [_p, _p, _p, _p, _i64, _p, _p]),
    "pmx_group_combine": (_int, [_p, _int, _i64, _p, _p, _p, _p, _p, _p]),
    "pmx_pivot_gather": (_int, [_p, _p, _p, _i64, _p, _p]),
    "pmx_ratio_pct": (_int, [_p, _p, _i64, _p, _i64, _i64, _p, _i64, _p, _p]),
    "pmx_where_gt": (_int, [_p, _p, _i64, C.c_double, _p, _p]),
    "pmx_slab_tail_lengths": (_int, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p]),
    "pmx_slab_tail_offsets": (_int, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _p]),
    "pmx_slab_tail_gather": (_int, [_p, _p, _p, _p, _p, _i64, _i64, _p, _p, _p]),
    "pmx_slab_append": (_int, [_p, _p, _p, _p, _i64, _p, _p, _p, _p, _i64, _i64, _p, _p, _p]),
    "pmx_csr_copy": (_int, [_p, _p, _p, _p, _p, _i64, _p, _p, _p]),
    "pmx_quantile_hist": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _i64, _i64, _int, _p, _p, _p]),
    "pmx_quantile_hist_lp": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _i64, _i64, _int, _int, _p, _p, _p]),
    "pmx_quantile_list_pass": (_int, [_i64, _i64, _i64]),
    "pmx_quantile_init": (_int, [_p, _p, _p, _p, _i64, _p, _i64, _p, _p]),
    "pmx_quantile_tail": (_int, [_p, _p, _i64, _i64, _p, _i64, _i64, _int, _p, _p]),
    "pmx_quantile_tail_filter": (_int, [_p, _i64, _i64, _int, _p, _p, _p]),
    "pmx_quantile_tail_merge": (_int, [_p, _p, _int, _i64, _i64, _int, _p, _p, _p]),
    "pmx_quantile_pick": (_int, [_p, _p, _i64, _p, _i64, _int, _p, _p]),
    "pmx_quantile_finish": (_int, [_p, _p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "pmx_quantile_select": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "pmx_peer_put": (_int, [_p, _p, _int, _int, C.c_uint64, C.c_uint64, _p]),
    "pmx_bsel_widths": (_int, [_p, _p, _p]),
    "pmx_bsel_begin": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _p, _p, _i64, _p, _p]),
    "pmx_bsel_begin_csr": (_int, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _i64, _p, _p]),
    "pmx_bsel_sweep": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _p, _i64, _p, _int, _p, _p]),
    "pmx_bsel_hist": (_int, [_p, _i64, _i64, _i64, _p, _i64, _p, _int, _p, _p, _p, _p, _p, _p, _p]),
    "pmx_bsel_rows": (_int, [_p, _i64, _i64, _i64, _i64, _p, _int, _p, _p]),
    "pmx_bsel_finish": (_int, [_p, _i64, _i64, _i64, _i64, _p, _int, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "pmx_summary_quantiles": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _p, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "pmx_summary_sweep": (_int, [_p, _p, _i64, _i64, _p, _i64, _p, _p, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "pmx_summary_brackets_csr": (_int, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p, _i64, _p]),
}

_lib = None


def lib():
    """The loaded library; raises if it cannot be found or built (there is no CPU path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        from . import build as _build
        try:
            _build.build()
        except Exception as e:  # noqa: BLE001
            raise ImportError(
                f"libpmx.so is missing at {LIB_PATH} and could not be built ({e}); run "
                "`python -m k8s_rocm_metrics_dashboard_b200.build` where nvcc is available. "
                "This engine has no CPU fallback.") from e
    l = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(l, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = l
    return l


def check(rc):
    if rc != OK:
        raise PmxError(rc, lib().pmx_last_error().decode("utf-8", "replace"))
