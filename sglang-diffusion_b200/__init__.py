"""sglang-diffusion_b200 -- B200-native (sm_100a) W8A8 quantized-linear hot path.

Drop-in for the `sgl_kernel` W8A8 operators of analytics-zoo/sglang-diffusion:
importing this package loads

  libb200_w8a8.so      hand-written CUDA kernels behind the C ABI of include/b200_w8a8.h
  b200_w8a8_torch.so   TORCH_LIBRARY_FRAGMENT(sgl_kernel) registration of the same ops

so that `torch.ops.sgl_kernel.{int8_scaled_mm, fp8_scaled_mm, fp8_blockwise_scaled_mm,
sgl_per_token_quant_fp8, sgl_per_token_group_quant_8bit}` (schemas of
sgl-kernel/csrc/common_extension.cc:129-158) and the CPU-path names
`{per_token_quant_int8_cpu, int8_scaled_mm_cpu, int8_scaled_mm_with_quant}`
(sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:419-447) dispatch to it for CUDA tensors.

The directory name carries a hyphen (it is the reference's name); import it as
`sglang_diffusion_b200` (the alias package next to it) or through importlib.
There is no CPU / eager-PyTorch fallback: missing binaries raise ImportError here and a
non-sm_100 device raises NotImplementedError at call time.
"""
from __future__ import annotations

import os

import torch

from . import capi  # raises ImportError when libb200_w8a8.so is missing

_PKG = os.path.dirname(os.path.abspath(__file__))
TORCH_SO = os.path.join(_PKG, "b200_w8a8_torch.so")

if not os.path.isfile(TORCH_SO):
    raise ImportError(
        f"{TORCH_SO} is missing: build it with `python sglang-diffusion_b200/build.py`. "
        "b200_w8a8 has no CPU or PyTorch fallback."
    )
torch.ops.load_library(TORCH_SO)

from . import ops  # noqa: E402  (registers the fake/meta kernels)
from .ops import (  # noqa: E402,F401
    fp8_blockwise_scaled_mm,
    fp8_scaled_mm,
    int8_scaled_mm,
    int8_scaled_mm_with_quant,
    per_token_group_quant_fp8,
    per_token_quant_int8,
    per_token_quant_int8_cpu,
    sgl_per_token_group_quant_8bit,
    sgl_per_token_quant_fp8,
    sglang_per_token_group_quant_fp8,
    sglang_per_token_quant_fp8,
    bmm_fp8,
    rmsnorm_quant,
    scaled_fp8_quant,
    sgl_per_tensor_quant_fp8,
    silu_and_mul_quant,
    static_quant_fp8,
)
from .linear import (  # noqa: E402,F401
    W8A8Fp8LinearMethod,
    W8A8Int8LinearMethod,
    W8A8BlockFp8LinearMethod,
    QuantLinear,
    apply_fp8_linear,
    cutlass_w8a8_block_fp8_linear_with_fallback,
    install_into_sglang,
    shard_rows,
)

from .host_api import HostW8A8Int8Linear  # noqa: E402,F401
from . import moe  # noqa: E402,F401
from .moe import W8A8Int8MoEMethod, fused_experts  # noqa: E402,F401

__all__ = [
    "HostW8A8Int8Linear", "moe", "W8A8Int8MoEMethod", "fused_experts",
    "capi", "ops", "int8_scaled_mm", "fp8_scaled_mm", "fp8_blockwise_scaled_mm", "sgl_per_token_quant_fp8",
    "sgl_per_token_group_quant_8bit", "per_token_quant_int8", "per_token_quant_int8_cpu",
    "int8_scaled_mm_with_quant", "sglang_per_token_quant_fp8", "sglang_per_token_group_quant_fp8",
    "per_token_group_quant_fp8", "bmm_fp8", "rmsnorm_quant", "scaled_fp8_quant", "sgl_per_tensor_quant_fp8",
    "silu_and_mul_quant", "static_quant_fp8", "W8A8Int8LinearMethod", "W8A8Fp8LinearMethod", "W8A8BlockFp8LinearMethod",
    "QuantLinear", "apply_fp8_linear", "cutlass_w8a8_block_fp8_linear_with_fallback", "install_into_sglang", "shard_rows",
]
