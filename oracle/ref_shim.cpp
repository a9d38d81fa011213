// TEST INFRASTRUCTURE ONLY -- part of the oracle, never of the product path.
//
// Registration shim for the reference's own CPU implementation of the W8A8
// linear path.  The reference sources are compiled where they lie
// (/root/reference/sgl-kernel/csrc/cpu/{gemm,gemm_int8,gemm_fp8}.cpp, see
// oracle/build_ref.py); nothing from them is copied here.  This file only
// declares the five entry points (prototypes as in
// sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:138-183) and exposes them to
// Python under the namespace `sgl_kernel_ref`, so that they can live in the same
// process as the product's `sgl_kernel` CUDA ops without a double m.def().
#include <ATen/ATen.h>
#include <torch/library.h>

#include <optional>
#include <tuple>
#include <vector>

at::Tensor convert_weight_packed(at::Tensor& weight);
std::tuple<at::Tensor, at::Tensor> per_token_quant_int8_cpu(at::Tensor& A);
at::Tensor int8_scaled_mm_cpu(
    at::Tensor& mat1,
    at::Tensor& mat2,
    at::Tensor& scales1,
    at::Tensor& scales2,
    const std::optional<at::Tensor>& bias,
    at::ScalarType out_dtype,
    bool is_vnni);
at::Tensor fp8_scaled_mm_cpu(
    at::Tensor& mat1,
    at::Tensor& mat2,
    at::Tensor& scales2,
    std::vector<int64_t> block_size,
    const std::optional<at::Tensor>& bias,
    at::ScalarType out_dtype,
    bool is_vnni);
at::Tensor int8_scaled_mm_with_quant(
    at::Tensor& mat1,
    at::Tensor& mat2,
    at::Tensor& scales2,
    const std::optional<at::Tensor>& bias,
    at::ScalarType out_dtype,
    bool is_vnni);

TORCH_LIBRARY_FRAGMENT(sgl_kernel_ref, m) {
  m.def("convert_weight_packed(Tensor weight) -> Tensor");
  m.impl("convert_weight_packed", c10::kCPU, &convert_weight_packed);
  m.def("per_token_quant_int8_cpu(Tensor A) -> (Tensor, Tensor)");
  m.impl("per_token_quant_int8_cpu", c10::kCPU, &per_token_quant_int8_cpu);
  m.def(
      "int8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales1, Tensor scales2, Tensor? bias, "
      "ScalarType out_dtype, bool is_vnni) -> Tensor");
  m.impl("int8_scaled_mm_cpu", c10::kCPU, &int8_scaled_mm_cpu);
  m.def(
      "fp8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales2, int[] block_size, Tensor? bias, "
      "ScalarType out_dtype, bool is_vnni) -> Tensor");
  m.impl("fp8_scaled_mm_cpu", c10::kCPU, &fp8_scaled_mm_cpu);
  m.def(
      "int8_scaled_mm_with_quant(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, ScalarType "
      "out_dtype, bool is_vnni) -> Tensor");
  m.impl("int8_scaled_mm_with_quant", c10::kCPU, &int8_scaled_mm_with_quant);
}
