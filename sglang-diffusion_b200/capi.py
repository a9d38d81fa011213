"""ctypes binding of the C ABI in include/b200_w8a8.h (libb200_w8a8.so).

This is the lowest Python layer: device pointers and sizes in, status codes out.  It
exists so that the parity tests can exercise the library exactly through the boundary a
non-Python host (cgo / JNI / FFI) would bind, without the torch dispatcher in between.
There is no fallback of any kind: a missing .so raises ImportError at import time.
"""
from __future__ import annotations

import ctypes
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libb200_w8a8.so")

DTYPE_BF16, DTYPE_F16, DTYPE_F32, DTYPE_S32 = 0, 1, 2, 3
OK, ERR_ARG, ERR_ARCH, ERR_CUDA = 0, -1, -2, -3

# every symbol include/b200_w8a8.h declares (tests check the .so exports all of them)
SYMBOLS = [
    "b2_abi_version", "b2_last_error", "b2_device_info",
    "b2_quant_int8_rows", "b2_quant_fp8_rows", "b2_quant_fp8_group128",
    "b2_gemm_s8", "b2_gemm_e4m3", "b2_gemm_e4m3_block128",
    "b2_linear_w8a8_int8_workspace", "b2_linear_w8a8_int8",
    "b2_linear_w8a8_fp8_workspace", "b2_linear_w8a8_fp8",
    "b2_launch_count", "b2_launch_count_reset", "b2_set_gemm_config", "b2_last_gemm_config",
    "b2_set_profile_buffer", "b2_set_weight_prefetch", "b2_get_weight_prefetch",
    "b2_allreduce_symm",
    "b2_quant_int8_rows_sum", "b2_quant_fp8_tensor", "b2_silu_mul_quant", "b2_rmsnorm_quant",
    "b2_quant_fp8_group128_v2", "b2_bmm_e4m3_workspace", "b2_bmm_e4m3",
    "b2_set_allreduce_timeout_ms", "b2_get_allreduce_timeout_ms",
    "b2_set_fused_decode", "b2_get_fused_decode",
    "b2_gemm_s8_push_symm", "b2_gemm_e4m3_push_symm", "b2_allreduce_symm_gathered",
    "b2_gemm_s8_scatter_symm", "b2_gemm_e4m3_scatter_symm", "b2_allreduce_symm_gathered_bcast",
    "b2_allreduce_symm_gathered_rmsnorm_quant", "b2_reduce_scatter_symm_rmsnorm_quant_allgather",
    "b2_linear_w8a8_int8_static_max_bytes", "b2_linear_w8a8_int8_static",
    "b2_moe_max_padded", "b2_moe_align_block128", "b2_quant_int8_rows_gather", "b2_gemm_s8_grouped", "b2_moe_combine",
]

if not os.path.isfile(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python sglang-diffusion_b200/build.py` "
        "(nvcc, sm_100a).  b200_w8a8 has no CPU or PyTorch fallback."
    )

lib = ctypes.CDLL(LIB_PATH)

_vp, _i, _i64, _f, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t

lib.b2_abi_version.restype = _i
lib.b2_last_error.restype = ctypes.c_char_p
lib.b2_device_info.argtypes = [ctypes.POINTER(_i)] * 3
lib.b2_device_info.restype = _i
lib.b2_quant_int8_rows.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _i, _i, _vp]
lib.b2_quant_fp8_rows.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _i, _i, _vp]
lib.b2_quant_fp8_group128.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _i64, _i64, _i, _i, _f, _f, _f, _vp]
_gemm = [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i, _vp, _i64, _i, _i, _i, _i, _vp]
lib.b2_gemm_s8.argtypes = _gemm
lib.b2_gemm_e4m3.argtypes = _gemm
lib.b2_gemm_e4m3_block128.argtypes = [_vp, _i64, _vp, _i64, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _i, _vp, _i64,
                                      _i, _i, _i, _i, _vp]
lib.b2_linear_w8a8_int8_workspace.argtypes = [_i, _i]
lib.b2_linear_w8a8_int8_workspace.restype = _sz
lib.b2_linear_w8a8_fp8_workspace.argtypes = [_i, _i]
lib.b2_linear_w8a8_fp8_workspace.restype = _sz
_lin = [_vp, _i, _i64, _vp, _i64, _vp, _vp, _i, _vp, _i64, _i, _i, _i, _vp, _sz, _vp]
lib.b2_linear_w8a8_int8.argtypes = _lin
lib.b2_linear_w8a8_fp8.argtypes = _lin
for _n in ("b2_quant_int8_rows", "b2_quant_fp8_rows", "b2_quant_fp8_group128", "b2_gemm_s8", "b2_gemm_e4m3",
           "b2_gemm_e4m3_block128", "b2_linear_w8a8_int8", "b2_linear_w8a8_fp8"):
    getattr(lib, _n).restype = _i
lib.b2_launch_count.restype = ctypes.c_uint64
lib.b2_launch_count_reset.restype = None
lib.b2_set_gemm_config.argtypes = [_i]
lib.b2_set_gemm_config.restype = None
lib.b2_last_gemm_config.restype = ctypes.c_char_p
lib.b2_set_profile_buffer.argtypes = [_vp]
lib.b2_set_profile_buffer.restype = None
lib.b2_set_weight_prefetch.argtypes = [_i]
lib.b2_set_weight_prefetch.restype = None
lib.b2_get_weight_prefetch.restype = _i
lib.b2_allreduce_symm.argtypes = [_vp, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _vp]
lib.b2_allreduce_symm.restype = _i
lib.b2_quant_int8_rows_sum.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _vp, _i, _i, _vp]
lib.b2_quant_fp8_tensor.argtypes = [_vp, _i, _vp, _vp, _vp, _i64, _i64, _i, _vp]
lib.b2_silu_mul_quant.argtypes = [_vp, _i, _i64, _i, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _i64, _i64, _i, _vp]
lib.b2_rmsnorm_quant.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _f, _i, _vp, _i64, _vp, _i64, _i64, _vp, _i64, _i64, _i, _vp]
lib.b2_quant_fp8_group128_v2.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _i64, _i64, _i64, _i, _vp]
lib.b2_bmm_e4m3_workspace.argtypes = [_i, _i]
lib.b2_bmm_e4m3_workspace.restype = _sz
lib.b2_bmm_e4m3.argtypes = [_vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _vp, _i64, _i64, _i, _i, _i, _i, _i, _vp, _sz, _vp]
for _n in ("b2_quant_int8_rows_sum", "b2_quant_fp8_tensor", "b2_silu_mul_quant", "b2_rmsnorm_quant",
           "b2_quant_fp8_group128_v2", "b2_bmm_e4m3"):
    getattr(lib, _n).restype = _i
lib.b2_set_allreduce_timeout_ms.argtypes = [ctypes.c_longlong]
lib.b2_set_allreduce_timeout_ms.restype = None
lib.b2_get_allreduce_timeout_ms.restype = ctypes.c_longlong
_push = [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i, _vp, _vp, _i64, _i, _i, _i, _i, _i, _i, _vp]
lib.b2_gemm_s8_push_symm.argtypes = _push
lib.b2_gemm_e4m3_push_symm.argtypes = _push
lib.b2_gemm_s8_scatter_symm.argtypes = _push   # same layout: the multicast pointer becomes the host array of peer buffers
lib.b2_gemm_e4m3_scatter_symm.argtypes = _push
lib.b2_allreduce_symm_gathered.argtypes = [_vp, _vp, _vp, _i64, _i64, _i, _i, _i, _i, _vp]
lib.b2_allreduce_symm_gathered_rmsnorm_quant.argtypes = [_vp, _vp, _vp, _i64, _vp, _f, _i, _vp, _i64, _vp, _i64, _i64, _vp,
                                                         _i64, _i, _i, _i64, _i, _i, _i, _i, _vp]
lib.b2_allreduce_symm_gathered_rmsnorm_quant.restype = _i
lib.b2_reduce_scatter_symm_rmsnorm_quant_allgather.argtypes = [_vp, _vp, _vp, _i64, _vp, _f, _i, _vp, _i64, _vp, _i64, _i, _i,
                                                               _i64, _i, _i, _i, _i, _vp]
lib.b2_reduce_scatter_symm_rmsnorm_quant_allgather.restype = _i
lib.b2_allreduce_symm_gathered_bcast.argtypes = [_vp, _vp, _vp, _vp, _i64, _i64, _i, _i, _i, _i, _vp]
for _n in ("b2_gemm_s8_push_symm", "b2_gemm_e4m3_push_symm", "b2_allreduce_symm_gathered",
           "b2_gemm_s8_scatter_symm", "b2_gemm_e4m3_scatter_symm", "b2_allreduce_symm_gathered_bcast"):
    getattr(lib, _n).restype = _i
lib.b2_linear_w8a8_int8_static_max_bytes.argtypes = []
lib.b2_linear_w8a8_int8_static_max_bytes.restype = _sz
lib.b2_linear_w8a8_int8_static.argtypes = [_vp, _i, _i64, _vp, _i64, _vp, _vp, _i, _vp, _i64, _i, _i, _i, _vp]
lib.b2_linear_w8a8_int8_static.restype = _i
lib.b2_moe_max_padded.argtypes = [_i64, _i]
lib.b2_moe_max_padded.restype = _i64
lib.b2_moe_align_block128.argtypes = [_vp, _i, _i, _vp, _vp, _vp, _vp, _i64, _vp]
lib.b2_quant_int8_rows_gather.argtypes = [_vp, _i, _i64, _vp, _i, _i, _vp, _i64, _vp, _i64, _i, _vp]
lib.b2_gemm_s8_grouped.argtypes = [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _i, _vp, _i64, _i, _i64, _i, _i, _vp]
lib.b2_moe_combine.argtypes = [_vp, _i64, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _vp]
for _n in ("b2_moe_align_block128", "b2_quant_int8_rows_gather", "b2_gemm_s8_grouped", "b2_moe_combine"):
    getattr(lib, _n).restype = _i
QMODE_INT8_ROWS, QMODE_FP8_ROWS, QMODE_FP8_GROUP128, QMODE_FP8_GROUP128_V2 = 0, 1, 2, 3


class B2Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200_w8a8 error {code}: {msg}")
        self.code = code


def check(status: int) -> None:
    if status != OK:
        raise B2Error(status, lib.b2_last_error().decode())


def last_error() -> str:
    return lib.b2_last_error().decode()


def device_info():
    a, b, c = _i(), _i(), _i()
    check(lib.b2_device_info(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
    return a.value, b.value, c.value


def launch_count() -> int:
    return int(lib.b2_launch_count())


def reset_launch_count() -> None:
    lib.b2_launch_count_reset()


def set_gemm_config(config_id: int) -> None:
    lib.b2_set_gemm_config(int(config_id))


def last_gemm_config() -> str:
    return lib.b2_last_gemm_config().decode()


def set_weight_prefetch(enabled: bool) -> None:
    """b2_set_weight_prefetch: promise that W operands are static (not written by the preceding kernel), so that
    decode-shape GEMMs launch programmatically and fetch weights under the tail of the preceding kernel."""
    lib.b2_set_weight_prefetch(1 if enabled else 0)


lib.b2_set_fused_decode.argtypes = [_i]
lib.b2_set_fused_decode.restype = None
lib.b2_get_fused_decode.restype = _i


class fused_decode:
    """`with capi.fused_decode(): ...` -- M <= 32 linears run as ONE kernel (b2_set_fused_decode; off by default)."""

    def __init__(self, enabled: bool = True):
        self._on = 1 if enabled else 0

    def __enter__(self):
        self._prev = int(lib.b2_get_fused_decode())
        lib.b2_set_fused_decode(self._on)
        return self

    def __exit__(self, *exc):
        lib.b2_set_fused_decode(self._prev)
        return False


class static_weights:
    """`with capi.static_weights(): ...` -- scope of the promise above (restores the previous setting)."""

    def __enter__(self):
        self._prev = int(lib.b2_get_weight_prefetch())
        lib.b2_set_weight_prefetch(1)
        return self

    def __exit__(self, *exc):
        lib.b2_set_weight_prefetch(self._prev)
        return False
