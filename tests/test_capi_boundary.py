"""The drop-in boundary on a CPU-only box: libb200_w8a8.so loads, exports exactly the symbols
include/b200_w8a8.h declares, reports errors through status codes (no compute without a GPU),
and the torch shim registers the reference's op schemas verbatim."""
from __future__ import annotations

import ctypes
import os
import re
import subprocess

import pytest
import torch

from conftest import ROOT

import sglang_diffusion_b200 as b2

HEADER = os.path.join(ROOT, "include", "b200_w8a8.h")


def declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_all_exported():
    syms = declared_symbols()
    assert len(syms) >= 17
    lib = ctypes.CDLL(b2.capi.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), f"{s} is declared in include/b200_w8a8.h but not exported"
    assert sorted(b2.capi.SYMBOLS) == syms, "capi.SYMBOLS out of sync with the header"


def test_only_c_abi_is_exported():
    out = subprocess.check_output(["nm", "-D", "--defined-only", b2.capi.LIB_PATH], text=True)
    names = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert names and all(n.startswith("b2_") for n in names), names


def test_abi_version_and_launch_counter():
    assert b2.capi.lib.b2_abi_version() == 1
    b2.capi.reset_launch_count()
    assert b2.capi.launch_count() == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_compute_fails_loudly_without_gpu():
    buf = (ctypes.c_uint8 * 4096)()
    p = ctypes.addressof(buf)
    st = b2.capi.lib.b2_quant_int8_rows(p, b2.capi.DTYPE_BF16, 64, p, 64, p, 1, 64, None)
    assert st == b2.capi.ERR_ARCH
    assert "no CPU fallback" in b2.capi.last_error()
    st = b2.capi.lib.b2_gemm_s8(p, 64, p, 64, p, p, None, 0, p, 64, b2.capi.DTYPE_BF16, 1, 64, 64, None)
    assert st == b2.capi.ERR_ARCH
    with pytest.raises(b2.capi.B2Error):
        b2.capi.check(st)


SCHEMAS = {
    # sgl-kernel/csrc/common_extension.cc:129-158
    "int8_scaled_mm": "sgl_kernel::int8_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType out_dtype, Tensor? bias) -> Tensor",
    "fp8_scaled_mm": "sgl_kernel::fp8_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType out_dtype, Tensor? bias) -> Tensor",
    "fp8_blockwise_scaled_mm": "sgl_kernel::fp8_blockwise_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType out_dtype) -> Tensor",
    "sgl_per_token_quant_fp8": "sgl_kernel::sgl_per_token_quant_fp8(Tensor input, Tensor! output_q, Tensor! output_s) -> ()",
    "sgl_per_token_group_quant_8bit": "sgl_kernel::sgl_per_token_group_quant_8bit(Tensor input, Tensor! output_q, Tensor! output_s, int group_size, float eps, float fp8_min, float fp8_max, bool scale_ue8m0) -> ()",
    # sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:419-447
    "per_token_quant_int8_cpu": "sgl_kernel::per_token_quant_int8_cpu(Tensor A) -> (Tensor, Tensor)",
    "int8_scaled_mm_cpu": "sgl_kernel::int8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales1, Tensor scales2, Tensor? bias, ScalarType out_dtype, bool is_vnni) -> Tensor",
    "int8_scaled_mm_with_quant": "sgl_kernel::int8_scaled_mm_with_quant(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, ScalarType out_dtype, bool is_vnni) -> Tensor",
}


@pytest.mark.parametrize("name", sorted(SCHEMAS))
def test_reference_schemas_registered_verbatim(name):
    op = getattr(torch.ops.sgl_kernel, name).default
    assert str(op._schema) == str(torch._C.parse_schema(SCHEMAS[name]))
    assert torch._C._dispatch_has_kernel_for_dispatch_key(f"sgl_kernel::{name}", "CUDA")
    # no CPU kernel is registered: the product has no CPU path
    assert not torch._C._dispatch_has_kernel_for_dispatch_key(f"sgl_kernel::{name}", "CPU")


def test_cpu_tensors_are_rejected_not_computed():
    a = torch.zeros(4, 64, dtype=torch.int8)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.sgl_kernel.int8_scaled_mm(a, a.t(), torch.ones(4), torch.ones(4), torch.bfloat16, None)


def test_product_package_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import/link/execute it."""
    pkg = os.path.join(ROOT, "sglang-diffusion_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "restate" not in txt and "sgl_cpu_oracle" not in txt and "orc_" not in txt, f
    out = subprocess.check_output(["ldd", b2.capi.LIB_PATH], text=True)
    assert "oracle" not in out


def test_static_weight_flag_is_scoped_and_thread_local():
    """b2_set_weight_prefetch / capi.static_weights(): default off, restored on exit, nested scopes keep it on."""
    from sglang_diffusion_b200 import capi

    assert capi.lib.b2_get_weight_prefetch() == 0
    with capi.static_weights():
        assert capi.lib.b2_get_weight_prefetch() == 1
        with capi.static_weights():
            assert capi.lib.b2_get_weight_prefetch() == 1
        assert capi.lib.b2_get_weight_prefetch() == 1
    assert capi.lib.b2_get_weight_prefetch() == 0
    import threading

    seen = []
    with capi.static_weights():
        t = threading.Thread(target=lambda: seen.append(capi.lib.b2_get_weight_prefetch()))
        t.start()
        t.join()
    assert seen == [0]


def test_allreduce_symm_rejects_bad_arguments_without_touching_the_gpu():
    """Argument validation of b2_allreduce_symm happens before any CUDA call (status code + b2_last_error)."""
    import ctypes

    from sglang_diffusion_b200 import capi

    buf = (ctypes.c_void_p * 8)()
    p = ctypes.cast(buf, ctypes.c_void_p)
    f = capi.lib.b2_allreduce_symm
    assert f(None, p, None, p, 1024, capi.DTYPE_BF16, 0, 2, 0, 64, None) == -1  # B2_ERR_ARG
    assert "null pointer" in capi.last_error()
    assert f(p, p, None, p, 1024, capi.DTYPE_F32, 0, 2, 0, 64, None) != 0 and "bfloat16 or float16" in capi.last_error()
    assert f(p, p, None, p, 1024, capi.DTYPE_BF16, 0, 1, 0, 64, None) != 0 and "world size" in capi.last_error()
    assert f(p, p, None, p, 1024, capi.DTYPE_BF16, 9, 8, 0, 64, None) != 0
    assert f(p, p, None, p, 1001, capi.DTYPE_BF16, 0, 2, 0, 64, None) != 0 and "multiple of 8" in capi.last_error()
    assert f(p, p, None, p, 1024, capi.DTYPE_BF16, 0, 2, 2, 64, None) != 0 and "multicast" in capi.last_error()
    assert f(p, p, None, None, 1024, capi.DTYPE_BF16, 0, 2, 0, 64, None) != 0 and "in-place" in capi.last_error()
