"""ctypes binding of ``libdlwp_b200.so`` (C ABI in ``include/dlwp_b200.h``).

There is no CPU fallback: if the library has not been built, or no sm_100 device is present,
the calls raise.  Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or
``python dlwp-benchmark_b200/build.py``.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdlwp_b200.so")

ABI_VERSION = 3
MODEL_FNO, MODEL_AFNO = 0, 1
AFNO_FILTER_AFNO2D, AFNO_FILTER_FNO2D = 0, 1
PREC_FP32, PREC_FP32_TC, PREC_BF16 = 0, 1, 2
FFT_NORM_BACKWARD, FFT_NORM_FORWARD, FFT_NORM_ORTHO = 0, 1, 2
STATIC_SCRATCH_BYTES, STATIC_MAX_ROWS = 16 << 20, 64      # DLWP_STATIC_SCRATCH_BYTES / DLWP_STATIC_MAX_ROWS


class ModelDesc(C.Structure):
    """Mirror of ``dlwp_model_desc`` (field order and types must match the header)."""
    _fields_ = [
        ("kind", C.c_int32), ("precision", C.c_int32), ("max_batch", C.c_int32),
        ("height", C.c_int32), ("width", C.c_int32),
        ("constant_channels", C.c_int32), ("prescribed_channels", C.c_int32),
        ("prognostic_channels", C.c_int32), ("context_size", C.c_int32),
        ("fft_norm", C.c_int32), ("n_rows", C.c_int32), ("n_cols", C.c_int32),
        ("rows_in", C.POINTER(C.c_int32)), ("rows_out", C.POINTER(C.c_int32)),
        ("hidden_channels", C.c_int32), ("lifting_channels", C.c_int32),
        ("projection_channels", C.c_int32), ("n_layers", C.c_int32),
        ("embed_dim", C.c_int32), ("depth", C.c_int32), ("num_blocks", C.c_int32),
        ("mlp_hidden", C.c_int32), ("patch_h", C.c_int32), ("patch_w", C.c_int32),
        ("use_pos_embed", C.c_int32), ("hidden_size_factor", C.c_int32),
        ("sparsity_threshold", C.c_float), ("layernorm_eps", C.c_float),
        ("afno_filter", C.c_int32),
    ]


class FrbWeights(C.Structure):
    """Mirror of ``dlwp_frb_weights`` (FourierResidualBlock parameters, device pointers)."""
    _fields_ = [(n, C.c_void_p) for n in (
        "f1_w1", "f1_w2", "conv1_w", "conv1_b", "f2_w1", "f2_w2", "conv2_w", "conv2_b", "short_w", "short_b",
        "norm1_w", "norm1_b", "norm2_w", "norm2_b")] + [("n_groups", C.c_int32), ("eps", C.c_float)]


# name -> (restype, argtypes); every symbol the header declares
_P = C.c_void_p
SYMBOLS = {
    "dlwp_abi_version": (C.c_int, []),
    "dlwp_last_error": (C.c_char_p, []),
    "dlwp_plan_create": (C.c_int, [C.POINTER(ModelDesc), C.c_int, C.POINTER(_P)]),
    "dlwp_plan_destroy": (C.c_int, [_P]),
    "dlwp_plan_workspace_bytes": (C.c_int, [_P, C.POINTER(C.c_size_t)]),
    "dlwp_plan_uses_tensor_cores": (C.c_int, [_P, C.POINTER(C.c_int)]),
    "dlwp_plan_check_guards": (C.c_int, [_P, _P]),
    "dlwp_plan_debug_read": (C.c_int, [_P, C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "dlwp_plan_set_tensor": (C.c_int, [_P, C.c_char_p, _P, C.c_int, C.POINTER(C.c_int64), _P]),
    "dlwp_plan_commit_weights": (C.c_int, [_P, _P]),
    "dlwp_step": (C.c_int, [_P, _P, _P, C.c_int, _P]),
    "dlwp_rollout": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, _P, _P]),
    "dlwp_rollout_host": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, _P, _P]),
    "dlwp_metrics_accumulate": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "dlwp_denormalize_split": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dlwp_spectral_conv2d": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, _P]),
    "dlwp_spectral_conv2d_static": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                              C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, _P]),
    "dlwp_fourier_residual_block_workspace": (C.c_int, [C.c_int] * 7 + [C.POINTER(C.c_size_t)]),
    "dlwp_fourier_residual_block": (C.c_int, [_P, C.POINTER(FrbWeights), _P] + [C.c_int] * 7 + [_P, C.c_size_t, _P]),
    "dlwp_linear_forward": (C.c_int, [_P] * 6 + [C.c_int] * 4 + [_P]),
    "dlwp_linear_backward": (C.c_int, [_P] * 9 + [C.c_int] * 4 + [_P]),
    "dlwp_spectral_train_workspace": (C.c_int, [C.c_int] * 7 + [C.POINTER(C.c_size_t)]),
    "dlwp_spectral_forward": (C.c_int, [_P] * 3 + [C.c_int] * 7 + [C.POINTER(C.c_int32)] * 2 + [C.c_int, _P, C.c_size_t, _P]),
    "dlwp_spectral_backward": (C.c_int, [_P] * 5 + [C.c_int] * 7 + [C.POINTER(C.c_int32)] * 2 + [C.c_int, _P, C.c_size_t, _P]),
    "dlwp_layernorm_forward": (C.c_int, [_P] * 4 + [C.c_int, C.c_int, C.c_float, _P]),
    "dlwp_layernorm_backward": (C.c_int, [_P] * 6 + [C.c_int, C.c_int, C.c_float, _P]),
    "dlwp_afno_filter_workspace": (C.c_int, [C.c_int] * 7 + [C.POINTER(C.c_size_t)]),
    "dlwp_afno_filter_forward": (C.c_int, [_P] * 7 + [C.c_int] * 6 + [C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_float, _P, C.c_size_t, _P]),
    "dlwp_afno_filter_backward": (C.c_int, [_P] * 11 + [C.c_int] * 6 + [C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_float, _P, C.c_size_t,
                                            _P]),
    "dlwp_nchw_to_nhwc": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, _P]),
    "dlwp_nhwc_to_nchw": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, _P]),
    "dlwp_launch_count": (C.c_int64, []),
    "dlwp_selftest_umma": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, _P]),
    "dlwp_selftest_umma_bmn": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dlwp_profile_begin": (C.c_int, [_P]),
    "dlwp_profile_end": (C.c_int, [_P]),
    "dlwp_profile_count": (C.c_int, []),
    "dlwp_profile_entry": (C.c_int, [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
}

_lib = None
_lock = threading.Lock()


class NativeError(RuntimeError):
    pass


def lib():
    """Load the shared library once; raise (never fall back) when it is missing."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.isfile(LIB_PATH):
                    raise NativeError(
                        f"{LIB_PATH} is missing: build the CUDA extension first (__graft_entry__.build()). "
                        "dlwp_benchmark_b200 has no CPU or PyTorch fallback.")
                handle = C.CDLL(LIB_PATH)
                for name, (res, args) in SYMBOLS.items():
                    fn = getattr(handle, name)
                    fn.restype, fn.argtypes = res, args
                if handle.dlwp_abi_version() != ABI_VERSION:
                    raise NativeError(f"ABI mismatch: library {handle.dlwp_abi_version()}, binding {ABI_VERSION}")
                _lib = handle
    return _lib


def check(status: int, what: str = ""):
    if status != 0:
        msg = lib().dlwp_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'dlwp call'} failed ({status}): {msg}")


def launch_count() -> int:
    return int(lib().dlwp_launch_count())


def int32_array(values):
    arr = (C.c_int32 * max(len(values), 1))(*values)
    return arr


class kernel_profile:
    """``with kernel_profile(stream_ptr) as prof: ...`` -> ``prof.entries`` = {kernel name:
    (total device ms, launches)} for every kernel the library launched inside the block."""

    def __init__(self, stream: int):
        self._stream = C.c_void_p(stream)
        self.entries = {}

    def __enter__(self):
        check(lib().dlwp_profile_begin(self._stream), "dlwp_profile_begin")
        return self

    def __exit__(self, *exc):
        check(lib().dlwp_profile_end(self._stream), "dlwp_profile_end")
        name, ms, n = C.c_char_p(), C.c_double(), C.c_int64()
        for i in range(lib().dlwp_profile_count()):
            check(lib().dlwp_profile_entry(i, C.byref(name), C.byref(ms), C.byref(n)))
            self.entries[name.value.decode()] = (float(ms.value), int(n.value))
        return False
