// torch.ops.sgl_kernel.* registration shim over the C ABI of libb200_w8a8.so.
//
// Schemas are the reference's, verbatim (sgl-kernel/csrc/common_extension.cc:129-158 for the
// CUDA ops; sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:419-447 for the three CPU-path op
// names, which are given CUDA kernels here so that the CPU oracle and the GPU path share
// call sites).  Argument checks and messages follow int8_gemm_kernel.cu:670-695,
// fp8_gemm_kernel.cu:1457-1487, fp8_blockwise_gemm_kernel.cu:374-403,
// per_token_quant_fp8.cu:242-249 and per_token_group_quant_8bit.cu:125-131.
// This file contains no arithmetic: every op validates, allocates the output with the torch
// caching allocator (graph-safe) and calls one b2_* function on the current stream.
#include <ATen/ATen.h>
#include <ATen/core/dispatch/Dispatcher.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include <cmath>
#include <optional>
#include <tuple>

#include "b200_w8a8.h"

namespace {

inline void* cur_stream(const at::Tensor& t) {
  return static_cast<void*>(c10::cuda::getCurrentCUDAStream(t.device().index()).stream());
}

inline void check_status(int st, const char* op) {
  if (st == B2_OK) return;
  if (st == B2_ERR_ARCH) {
    TORCH_CHECK_NOT_IMPLEMENTED(false, op, ": ", b2_last_error());
  }
  TORCH_CHECK(false, op, ": ", b2_last_error());
}

inline int act_dtype(at::ScalarType t, const char* what) {
  if (t == at::kBFloat16) return B2_DTYPE_BF16;
  if (t == at::kHalf) return B2_DTYPE_F16;
  TORCH_CHECK(false, what, " must be Half or BFloat16");
  return -1;
}

inline int bias_dtype_of(const at::Tensor& b) {
  if (b.scalar_type() == at::kBFloat16) return B2_DTYPE_BF16;
  if (b.scalar_type() == at::kHalf) return B2_DTYPE_F16;
  if (b.scalar_type() == at::kFloat) return B2_DTYPE_F32;
  TORCH_CHECK(false, "bias must be Half, BFloat16 or Float32");
  return -1;
}

void check_scaled_mm_common(const at::Tensor& mat_a, const at::Tensor& mat_b, const at::Tensor& scales_a,
                            const at::Tensor& scales_b, at::ScalarType out_dtype,
                            const std::optional<at::Tensor>& bias) {
  TORCH_CHECK(mat_a.is_cuda(), "mat_a must be a CUDA tensor");
  TORCH_CHECK(mat_b.is_cuda(), "mat_b must be a CUDA tensor");
  TORCH_CHECK(mat_a.dim() == 2, "mat_a must be a 2D tensor");
  TORCH_CHECK(mat_b.dim() == 2, "mat_b must be a 2D tensor");
  TORCH_CHECK(mat_a.stride(1) == 1, "mat_a must be a row major tensor");
  TORCH_CHECK(mat_b.stride(0) == 1, "mat_b must be a column major tensor");
  TORCH_CHECK(mat_a.size(1) == mat_b.size(0), "mat_a and mat_b shapes cannot be multiplied");
  TORCH_CHECK(mat_a.size(1) % 16 == 0, "mat_a.size(1) must be multiple of 16 for memory alignment");
  TORCH_CHECK(mat_b.size(0) % 16 == 0, "mat_b.size(0) must be multiple of 16 for memory alignment");
  TORCH_CHECK(mat_b.size(1) % 8 == 0, "mat_b.size(1) must be multiple of 8 for memory alignment");
  TORCH_CHECK(out_dtype == at::kHalf || out_dtype == at::kBFloat16, "out_dtype must be Half or BFloat16");
  TORCH_CHECK(scales_a.numel() == mat_a.size(0), "size of scales_a is not matched");
  TORCH_CHECK(scales_b.numel() == mat_b.size(1), "size of scales_b is not matched");
  TORCH_CHECK(scales_a.is_contiguous(), "scales_a must be contiguous");
  TORCH_CHECK(scales_b.is_contiguous(), "scales_b msut be contiguous");
  TORCH_CHECK(scales_a.scalar_type() == at::kFloat, "scales_a must be Float32");
  TORCH_CHECK(scales_b.scalar_type() == at::kFloat, "scales_b must be Float32");
  if (bias) {
    TORCH_CHECK(bias->numel() == mat_b.size(1), "size of bias is not matched");
    TORCH_CHECK(bias->is_contiguous(), "bias must be contiguous");
    TORCH_CHECK(bias->scalar_type() == out_dtype, "bias dtype must match output dtype");
  }
}

// mat_b is logical [K,N] with stride(0)==1: the same memory as the [N,K] row-major weight.
inline int64_t ldw_of(const at::Tensor& mat_b) { return mat_b.size(1) > 1 ? mat_b.stride(1) : mat_b.size(0); }

at::Tensor int8_scaled_mm(const at::Tensor& mat_a, const at::Tensor& mat_b, const at::Tensor& scales_a,
                          const at::Tensor& scales_b, at::ScalarType out_dtype,
                          const std::optional<at::Tensor>& bias) {
  check_scaled_mm_common(mat_a, mat_b, scales_a, scales_b, out_dtype, bias);
  TORCH_CHECK(mat_a.scalar_type() == at::kChar, "mat_a must be Int8");
  TORCH_CHECK(mat_b.scalar_type() == at::kChar, "mat_b must be Int8");
  c10::cuda::CUDAGuard guard(mat_a.device());
  const int64_t M = mat_a.size(0), K = mat_a.size(1), N = mat_b.size(1);
  at::Tensor out = at::empty({M, N}, mat_a.options().dtype(out_dtype));
  const int st = b2_gemm_s8(mat_a.data_ptr<int8_t>(), M > 1 ? mat_a.stride(0) : K, mat_b.data_ptr<int8_t>(),
                            ldw_of(mat_b), scales_a.data_ptr<float>(), scales_b.data_ptr<float>(),
                            bias ? bias->data_ptr() : nullptr, bias ? bias_dtype_of(*bias) : B2_DTYPE_BF16,
                            out.data_ptr(), N, act_dtype(out_dtype, "out_dtype"), (int)M, (int)N, (int)K,
                            cur_stream(mat_a));
  check_status(st, "int8_scaled_mm");
  return out;
}

at::Tensor fp8_scaled_mm(const at::Tensor& mat_a, const at::Tensor& mat_b, const at::Tensor& scales_a,
                         const at::Tensor& scales_b, at::ScalarType out_dtype,
                         const std::optional<at::Tensor>& bias) {
  check_scaled_mm_common(mat_a, mat_b, scales_a, scales_b, out_dtype, bias);
  TORCH_CHECK(mat_a.scalar_type() == at::kFloat8_e4m3fn, "mat_a must be Float8_e4m3fn");
  TORCH_CHECK(mat_b.scalar_type() == at::kFloat8_e4m3fn, "mat_b must be Float8_e4m3fn");
  c10::cuda::CUDAGuard guard(mat_a.device());
  const int64_t M = mat_a.size(0), K = mat_a.size(1), N = mat_b.size(1);
  at::Tensor out = at::empty({M, N}, mat_a.options().dtype(out_dtype));
  const int st = b2_gemm_e4m3(static_cast<const uint8_t*>(mat_a.data_ptr()), M > 1 ? mat_a.stride(0) : K,
                              static_cast<const uint8_t*>(mat_b.data_ptr()), ldw_of(mat_b),
                              scales_a.data_ptr<float>(), scales_b.data_ptr<float>(),
                              bias ? bias->data_ptr() : nullptr, bias ? bias_dtype_of(*bias) : B2_DTYPE_BF16,
                              out.data_ptr(), N, act_dtype(out_dtype, "out_dtype"), (int)M, (int)N, (int)K,
                              cur_stream(mat_a));
  check_status(st, "fp8_scaled_mm");
  return out;
}

at::Tensor fp8_blockwise_scaled_mm(const at::Tensor& mat_a, const at::Tensor& mat_b, const at::Tensor& scales_a,
                                   const at::Tensor& scales_b, at::ScalarType out_dtype) {
  TORCH_CHECK(mat_a.is_cuda(), "mat_a must be a CUDA tensor");
  TORCH_CHECK(mat_b.is_cuda(), "mat_b must be a CUDA tensor");
  TORCH_CHECK(mat_a.dim() == 2, "mat_a must be a 2D tensor");
  TORCH_CHECK(mat_b.dim() == 2, "mat_b must be a 2D tensor");
  TORCH_CHECK(mat_a.stride(1) == 1, "mat_a must be a row major tensor");
  TORCH_CHECK(mat_b.stride(0) == 1, "mat_b must be a column major tensor");
  TORCH_CHECK(mat_a.size(1) == mat_b.size(0), "mat_a and mat_b shapes cannot be multiplied");
  TORCH_CHECK(mat_a.size(1) % 16 == 0, "mat_a must be multiple of 16 bytes for memory alignment");
  TORCH_CHECK(mat_a.scalar_type() == at::kFloat8_e4m3fn, "mat_a must be Float8_e4m3fn");
  TORCH_CHECK(mat_b.scalar_type() == at::kFloat8_e4m3fn, "mat_b must be Float8_e4m3fn");
  TORCH_CHECK(out_dtype == at::kHalf || out_dtype == at::kBFloat16, "out_dtype must be Half or BFloat16");
  TORCH_CHECK(scales_a.dim() == 2 && scales_b.dim() == 2, "scales must be 2D tensors");
  TORCH_CHECK(mat_a.size(0) == scales_a.size(0), "size of scales_a is not matched");
  // the reference's floor-division checks let N = 192 through and then read one scale out of bounds; the kernel
  // walks ceil(N/128) blocks, so both extents must sit on the 128 grid
  TORCH_CHECK(mat_a.size(1) % 128 == 0, "mat_a.size(1) must be a multiple of 128 for 128x128 block scales");
  TORCH_CHECK(mat_b.size(1) % 128 == 0, "mat_b.size(1) must be a multiple of 128 for 128x128 block scales");
  TORCH_CHECK(mat_a.size(1) / 128 == scales_a.size(1), "size of scales_a is not matched");
  TORCH_CHECK(mat_b.size(0) / 128 == scales_b.size(0), "size of scales_b is not matched");
  TORCH_CHECK(mat_b.size(1) / 128 == scales_b.size(1), "size of scales_b is not matched");
  TORCH_CHECK(scales_a.scalar_type() == at::kFloat, "scales_a must be Float32");
  TORCH_CHECK(scales_b.scalar_type() == at::kFloat, "scales_b must be Float32");
  c10::cuda::CUDAGuard guard(mat_a.device());
  const int64_t M = mat_a.size(0), K = mat_a.size(1), N = mat_b.size(1);
  TORCH_CHECK((N * 2) % 16 == 0, "out must be multiple of 16 bytes for memory alignment");
  at::Tensor out = at::empty({M, N}, mat_a.options().dtype(out_dtype));
  // scales are consumed through their strides: the reference's M-major / K-major layouts
  // (fp8_blockwise_gemm_kernel.cu:396-403) need no copy, and neither does anything else.
  const int st = b2_gemm_e4m3_block128(
      static_cast<const uint8_t*>(mat_a.data_ptr()), M > 1 ? mat_a.stride(0) : K,
      static_cast<const uint8_t*>(mat_b.data_ptr()), ldw_of(mat_b), scales_a.data_ptr<float>(), scales_a.stride(0),
      scales_a.stride(1), scales_b.data_ptr<float>(), /*n stride*/ scales_b.stride(1), /*k stride*/ scales_b.stride(0),
      nullptr, B2_DTYPE_BF16, out.data_ptr(), N, act_dtype(out_dtype, "out_dtype"), (int)M, (int)N, (int)K,
      cur_stream(mat_a));
  check_status(st, "fp8_blockwise_scaled_mm");
  return out;
}

void sgl_per_token_quant_fp8(at::Tensor input, at::Tensor output_q, at::Tensor output_s) {
  TORCH_CHECK(input.is_cuda() && output_q.is_cuda() && output_s.is_cuda(), "tensors must be CUDA tensors");
  TORCH_CHECK(input.is_contiguous() && output_q.is_contiguous() && output_s.is_contiguous(),
              "tensors must be contiguous");
  TORCH_CHECK(input.dim() == 2, "input must be a 2D tensor");
  const int64_t M = input.size(0), K = input.size(1);
  TORCH_CHECK(K % 4 == 0, "Hidden dimension must be divisible by 4, but got ", K);
  TORCH_CHECK(output_q.scalar_type() == at::kFloat8_e4m3fn, "output_q must be Float8_e4m3fn");
  TORCH_CHECK(output_s.scalar_type() == at::kFloat && output_s.numel() == M, "output_s must be Float32 [M] / [M,1]");
  TORCH_CHECK(output_q.numel() == input.numel(), "output_q shape mismatch");
  c10::cuda::CUDAGuard guard(input.device());
  const int st = b2_quant_fp8_rows(input.data_ptr(), act_dtype(input.scalar_type(), "input"), K,
                                   static_cast<uint8_t*>(output_q.data_ptr()), K, output_s.data_ptr<float>(), (int)M,
                                   (int)K, cur_stream(input));
  check_status(st, "sgl_per_token_quant_fp8");
}

void sgl_per_token_group_quant_8bit(at::Tensor input, at::Tensor output_q, at::Tensor output_s, int64_t group_size,
                                    double eps, double min_8bit, double max_8bit, bool scale_ue8m0) {
  TORCH_CHECK(input.is_cuda() && output_q.is_cuda() && output_s.is_cuda(), "tensors must be CUDA tensors");
  TORCH_CHECK(input.is_contiguous() && output_q.is_contiguous(), "input and output_q must be contiguous");
  TORCH_CHECK(input.numel() % group_size == 0, "input.numel() must be divisible by group_size");
  TORCH_CHECK(output_s.dim() == 2, "output_s must be a 2D tensor");
  TORCH_CHECK_NOT_IMPLEMENTED(group_size == 128, "sgl_per_token_group_quant_8bit: only group_size 128 is built");
  TORCH_CHECK_NOT_IMPLEMENTED(!scale_ue8m0, "sgl_per_token_group_quant_8bit: UE8M0 scales are not built");
  TORCH_CHECK_NOT_IMPLEMENTED(output_q.scalar_type() == at::kFloat8_e4m3fn,
                              "sgl_per_token_group_quant_8bit: only e4m3 output is built");
  TORCH_CHECK(output_s.scalar_type() == at::kFloat, "output_s must be Float32");
  const int64_t K = input.size(-1);
  const int64_t M = input.numel() / K;
  TORCH_CHECK(output_s.size(0) == M && output_s.size(1) == K / group_size, "output_s shape mismatch");
  c10::cuda::CUDAGuard guard(input.device());
  const int st = b2_quant_fp8_group128(input.data_ptr(), act_dtype(input.scalar_type(), "input"), K,
                                       static_cast<uint8_t*>(output_q.data_ptr()), K, output_s.data_ptr<float>(),
                                       output_s.stride(0), output_s.stride(1), (int)M, (int)K, (float)eps,
                                       (float)min_8bit, (float)max_8bit, cur_stream(input));
  check_status(st, "sgl_per_token_group_quant_8bit");
}

// sgl_per_token_group_quant_8bit_v2 (common_extension.cc:149-152; per_token_group_quant_8bit_v2.cu:396-447):
// input [M, K] -- or [M, 2K] with fuse_silu_and_mul, or [E, T, 2K] + masked_m[E] -- into e4m3 [.., K] and fp32 scales
// (any strides: row-major, column-major, TMA-aligned column-major).
void sgl_per_token_group_quant_8bit_v2(at::Tensor input, at::Tensor output_q, at::Tensor output_s, int64_t group_size,
                                       double eps, double min_8bit, double max_8bit, bool scale_ue8m0,
                                       bool fuse_silu_and_mul, const std::optional<at::Tensor>& masked_m) {
  TORCH_CHECK(input.is_cuda() && output_q.is_cuda() && output_s.is_cuda(), "tensors must be CUDA tensors");
  TORCH_CHECK(input.is_contiguous() && output_q.is_contiguous(), "input and output_q must be contiguous");
  TORCH_CHECK(input.numel() > 0, "input must not be empty");
  TORCH_CHECK(std::abs(1e-10 - eps) < 1e-13, "eps must be 1e-10");
  TORCH_CHECK(input.numel() % group_size == 0, "input.numel() must be divisible by group_size");
  TORCH_CHECK_NOT_IMPLEMENTED(group_size == 128, "sgl_per_token_group_quant_8bit_v2: only group_size 128 is built");
  TORCH_CHECK_NOT_IMPLEMENTED(!scale_ue8m0, "sgl_per_token_group_quant_8bit_v2: UE8M0 scales are not built");
  TORCH_CHECK_NOT_IMPLEMENTED(output_q.scalar_type() == at::kFloat8_e4m3fn,
                              "sgl_per_token_group_quant_8bit_v2: only e4m3 output is built");
  TORCH_CHECK(min_8bit == -448.0 && max_8bit == 448.0, "fp8_min / fp8_max must be -448 / 448");
  TORCH_CHECK(output_s.scalar_type() == at::kFloat, "output_s must be Float32");
  const bool masked = masked_m.has_value();
  TORCH_CHECK(!masked || fuse_silu_and_mul, "masked_m needs fuse_silu_and_mul");
  TORCH_CHECK(output_s.dim() == (masked ? 3 : 2), "output_s has the wrong rank");
  const int64_t K = output_q.size(-1);
  TORCH_CHECK(input.size(-1) == (fuse_silu_and_mul ? 2 * K : K), "input / output_q hidden sizes do not match");
  const int64_t T = output_q.size(-2);
  const int64_t M = output_q.numel() / K;
  TORCH_CHECK(output_s.size(-2) == T && output_s.size(-1) == K / group_size, "output_s shape mismatch");
  if (masked) {
    TORCH_CHECK(masked_m->is_cuda() && masked_m->scalar_type() == at::kInt && masked_m->is_contiguous() &&
                    masked_m->numel() == input.size(0), "masked_m must be an int32 CUDA tensor [num_experts]");
  }
  c10::cuda::CUDAGuard guard(input.device());
  int st;
  if (fuse_silu_and_mul) {
    st = b2_silu_mul_quant(input.data_ptr(), act_dtype(input.scalar_type(), "input"), 2 * K, B2_QMODE_FP8_GROUP128_V2,
                           static_cast<uint8_t*>(output_q.data_ptr()), K, output_s.data_ptr<float>(), output_s.stride(-2),
                           output_s.stride(-1), masked ? output_s.stride(0) : 0, nullptr, 0,
                           masked ? masked_m->data_ptr<int32_t>() : nullptr, T, M, (int)K, cur_stream(input));
  } else {
    st = b2_quant_fp8_group128_v2(input.data_ptr(), act_dtype(input.scalar_type(), "input"), K,
                                  static_cast<uint8_t*>(output_q.data_ptr()), K, output_s.data_ptr<float>(),
                                  output_s.stride(-2), output_s.stride(-1), M, (int)K, cur_stream(input));
  }
  check_status(st, "sgl_per_token_group_quant_8bit_v2");
}

// sgl_per_tensor_quant_fp8 (common_extension.cc:154-155; per_tensor_quant_fp8.cu:98-123)
void sgl_per_tensor_quant_fp8(at::Tensor input, at::Tensor output_q, at::Tensor output_s, bool is_static) {
  TORCH_CHECK(input.is_cuda() && output_q.is_cuda() && output_s.is_cuda(), "tensors must be CUDA tensors");
  TORCH_CHECK(input.is_contiguous() && output_q.is_contiguous() && output_s.is_contiguous(),
              "tensors must be contiguous");
  TORCH_CHECK(output_q.scalar_type() == at::kFloat8_e4m3fn, "output_q must be Float8_e4m3fn");
  TORCH_CHECK(output_s.scalar_type() == at::kFloat && output_s.numel() == 1, "output_s must be one Float32");
  TORCH_CHECK(output_q.numel() >= input.numel(), "output_q is smaller than input");
  c10::cuda::CUDAGuard guard(input.device());
  const int st = b2_quant_fp8_tensor(input.data_ptr(), act_dtype(input.scalar_type(), "input"),
                                     static_cast<uint8_t*>(output_q.data_ptr()), output_s.data_ptr<float>(), nullptr,
                                     input.numel(), 0, is_static ? 1 : 0, cur_stream(input));
  check_status(st, "sgl_per_tensor_quant_fp8");
}

// bmm_fp8 (common_extension.cc:397-401; bmm_fp8.cu): A [b,m,k] e4m3 row-major, B [b,k,n] e4m3 "column major"
// (stride(-2) == 1, i.e. the memory of [b,n,k]), D [b,m,n] bf16 / f16, per-tensor scales.  The cuBLASLt handle of the
// reference signature is not used (no cuBLAS on this path); the workspace holds the broadcast scale vectors.
void bmm_fp8(at::Tensor A, at::Tensor B, at::Tensor D, at::Tensor A_scale, at::Tensor B_scale,
             at::Tensor workspace_buffer, int64_t /*cublas_handle*/) {
  TORCH_CHECK(A.is_cuda(), "A must be a CUDA tensor");
  TORCH_CHECK(B.is_cuda(), "B must be a CUDA tensor");
  TORCH_CHECK(D.is_cuda(), "D must be a CUDA tensor");
  TORCH_CHECK(A.dim() == 3, "Expected 3D tensor for A");
  TORCH_CHECK(B.dim() == 3, "Expected 3D tensor for B");
  TORCH_CHECK(D.dim() == 3, "Expected 3D tensor for D");
  TORCH_CHECK(A.size(0) == B.size(0) && A.size(0) == D.size(0), "Batch sizes must match");
  TORCH_CHECK(A.size(2) == B.size(1), "Incompatible matrix sizes");
  TORCH_CHECK(A.size(1) == D.size(1) && B.size(2) == D.size(2), "Result tensor has incorrect shape");
  TORCH_CHECK(A.scalar_type() == at::kFloat8_e4m3fn && B.scalar_type() == at::kFloat8_e4m3fn,
              "bmm_fp8: only e4m3 x e4m3 is built");
  TORCH_CHECK(A.stride(2) == 1 && D.stride(2) == 1, "A and D must be row major");
  TORCH_CHECK(B.stride(1) == 1, "B must be column major");
  TORCH_CHECK(A_scale.scalar_type() == at::kFloat && B_scale.scalar_type() == at::kFloat && A_scale.numel() == 1 &&
                  B_scale.numel() == 1, "A_scale / B_scale must be single Float32 values");
  c10::cuda::CUDAGuard guard(A.device());
  const int64_t b = A.size(0), m = A.size(1), k = A.size(2), n = B.size(2);
  const int st = b2_bmm_e4m3(static_cast<const uint8_t*>(A.data_ptr()), A.stride(1), A.stride(0),
                             static_cast<const uint8_t*>(B.data_ptr()), n > 1 ? B.stride(2) : k, B.stride(0),
                             A_scale.data_ptr<float>(), B_scale.data_ptr<float>(), D.data_ptr(), D.stride(1), D.stride(0),
                             act_dtype(D.scalar_type(), "D"), (int)b, (int)m, (int)n, (int)k, workspace_buffer.data_ptr(),
                             (size_t)workspace_buffer.numel() * workspace_buffer.element_size(), cur_stream(A));
  check_status(st, "bmm_fp8");
}

// ---- CPU-path op names with CUDA kernels (torch_extension_cpu.cpp:419-447) ------------------

// Returns SIGNED int8 codes (= the CPU op's uint8 code - 128) and fp32 scales [M].
std::tuple<at::Tensor, at::Tensor> per_token_quant_int8_cuda(at::Tensor& A) {
  TORCH_CHECK(A.is_cuda(), "A must be a CUDA tensor");
  TORCH_CHECK(A.dim() == 2, "per_token_quant_int8: expect A to be 2D");
  TORCH_CHECK(A.stride(1) == 1, "per_token_quant_int8: expect last dim contiguous");
  const auto st_ = A.scalar_type();
  TORCH_CHECK(st_ == at::kBFloat16 || st_ == at::kHalf, "per_token_quant_int8: expect A to be bfloat16 or half.");
  c10::cuda::CUDAGuard guard(A.device());
  const int64_t M = A.size(0), K = A.size(1);
  at::Tensor Aq = at::empty({M, K}, A.options().dtype(at::kChar));
  at::Tensor As = at::empty({M}, A.options().dtype(at::kFloat));
  const int st = b2_quant_int8_rows(A.data_ptr(), act_dtype(st_, "A"), M > 1 ? A.stride(0) : K, Aq.data_ptr<int8_t>(),
                                    K, As.data_ptr<float>(), (int)M, (int)K, cur_stream(A));
  check_status(st, "per_token_quant_int8");
  return std::make_tuple(Aq, As);
}

// mat1 int8 [M,K] (signed codes), mat2 int8 [N,K] UNPACKED (is_vnni must describe a CPU
// packing and is ignored on CUDA), scales1 [M], scales2 [N], bias fp32/bf16 [N].
at::Tensor int8_scaled_mm_cpuname_cuda(at::Tensor& mat1, at::Tensor& mat2, at::Tensor& scales1, at::Tensor& scales2,
                                       const std::optional<at::Tensor>& bias, at::ScalarType out_dtype, bool) {
  TORCH_CHECK(mat1.is_cuda() && mat2.is_cuda(), "inputs must be CUDA tensors");
  TORCH_CHECK(mat1.dim() == 2 && mat2.dim() == 2, "int8_scaled_mm: expect 2D inputs");
  TORCH_CHECK(mat1.stride(1) == 1 && mat2.is_contiguous(), "int8_scaled_mm: expect contiguous inputs");
  TORCH_CHECK(mat1.scalar_type() == at::kChar, "int8_scaled_mm (CUDA): expect mat1 to be int8 (signed codes).");
  TORCH_CHECK(mat2.scalar_type() == at::kChar, "int8_scaled_mm: expect mat2 to be int8.");
  TORCH_CHECK(scales1.scalar_type() == at::kFloat && scales2.scalar_type() == at::kFloat,
              "int8_scaled_mm: expect scales to be float32.");
  const int64_t M = mat1.size(0), K = mat1.size(1), N = mat2.size(0);
  TORCH_CHECK(mat2.size(1) == K, "int8_scaled_mm: mat2 must be the unpacked [N, K] weight on CUDA");
  TORCH_CHECK(scales1.numel() == M && scales2.numel() == N, "int8_scaled_mm: scale sizes do not match");
  if (bias) TORCH_CHECK(bias->numel() == N && bias->is_contiguous(), "int8_scaled_mm: bias size mismatch");
  c10::cuda::CUDAGuard guard(mat1.device());
  at::Tensor s1 = scales1.contiguous(), s2 = scales2.contiguous();
  at::Tensor out = at::empty({M, N}, mat1.options().dtype(out_dtype));
  const int st = b2_gemm_s8(mat1.data_ptr<int8_t>(), M > 1 ? mat1.stride(0) : K, mat2.data_ptr<int8_t>(), K,
                            s1.data_ptr<float>(), s2.data_ptr<float>(), bias ? bias->data_ptr() : nullptr,
                            bias ? bias_dtype_of(*bias) : B2_DTYPE_F32, out.data_ptr(), N,
                            act_dtype(out_dtype, "out_dtype"), (int)M, (int)N, (int)K, cur_stream(mat1));
  check_status(st, "int8_scaled_mm");
  return out;
}

at::Tensor int8_scaled_mm_with_quant_cuda(at::Tensor& mat1, at::Tensor& mat2, at::Tensor& scales2,
                                          const std::optional<at::Tensor>& bias, at::ScalarType out_dtype, bool) {
  TORCH_CHECK(mat1.is_cuda() && mat2.is_cuda(), "inputs must be CUDA tensors");
  TORCH_CHECK(mat1.dim() == 2 && mat2.dim() == 2, "int8_scaled_mm_with_quant: expect 2D inputs");
  TORCH_CHECK(mat1.stride(1) == 1 && mat2.is_contiguous(), "int8_scaled_mm_with_quant: expect contiguous inputs");
  const auto st_ = mat1.scalar_type();
  TORCH_CHECK(st_ == at::kBFloat16 || st_ == at::kHalf, "int8_scaled_mm_with_quant: expect A to be bfloat16 or half.");
  TORCH_CHECK(st_ == out_dtype, "int8_scaled_mm_with_quant: expect A has same dtype with out_dtype.");
  TORCH_CHECK(mat2.scalar_type() == at::kChar, "int8_scaled_mm_with_quant: expect mat2 to be int8.");
  TORCH_CHECK(scales2.scalar_type() == at::kFloat, "int8_scaled_mm_with_quant: expect scales to be float32.");
  const int64_t M = mat1.size(0), K = mat1.size(1), N = mat2.size(0);
  TORCH_CHECK(mat2.size(1) == K, "int8_scaled_mm_with_quant: mat2 must be the unpacked [N, K] weight on CUDA");
  TORCH_CHECK(scales2.numel() == N, "int8_scaled_mm_with_quant: scales2 size mismatch");
  if (bias) TORCH_CHECK(bias->numel() == N && bias->is_contiguous(), "int8_scaled_mm_with_quant: bias size mismatch");
  c10::cuda::CUDAGuard guard(mat1.device());
  at::Tensor s2 = scales2.contiguous();
  at::Tensor out = at::empty({M, N}, mat1.options().dtype(out_dtype));
  const size_t ws_bytes = b2_linear_w8a8_int8_workspace((int)M, (int)K);
  at::Tensor ws = at::empty({(int64_t)ws_bytes}, mat1.options().dtype(at::kByte));
  const int st = b2_linear_w8a8_int8(mat1.data_ptr(), act_dtype(st_, "mat1"), M > 1 ? mat1.stride(0) : K,
                                     mat2.data_ptr<int8_t>(), K, s2.data_ptr<float>(),
                                     bias ? bias->data_ptr() : nullptr, bias ? bias_dtype_of(*bias) : B2_DTYPE_F32,
                                     out.data_ptr(), N, (int)M, (int)N, (int)K, ws.data_ptr(), ws_bytes,
                                     cur_stream(mat1));
  check_status(st, "int8_scaled_mm_with_quant");
  return out;
}

bool has_schema(const char* name) {
  return c10::Dispatcher::singleton().findSchema({name, ""}).has_value();
}

}  // namespace

// Declare only what is not declared yet: when the reference's own sgl_kernel extension is
// loaded in the same process its schemas already exist and a second def() would abort.
TORCH_LIBRARY_FRAGMENT(sgl_kernel, m) {
  if (!has_schema("sgl_kernel::int8_scaled_mm"))
    m.def(
        "int8_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType out_dtype, Tensor? "
        "bias) -> Tensor");
  if (!has_schema("sgl_kernel::fp8_scaled_mm"))
    m.def(
        "fp8_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType out_dtype, Tensor? "
        "bias) -> Tensor");
  if (!has_schema("sgl_kernel::fp8_blockwise_scaled_mm"))
    m.def(
        "fp8_blockwise_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType out_dtype) -> "
        "Tensor");
  if (!has_schema("sgl_kernel::sgl_per_token_quant_fp8"))
    m.def("sgl_per_token_quant_fp8(Tensor input, Tensor! output_q, Tensor! output_s) -> ()");
  if (!has_schema("sgl_kernel::sgl_per_token_group_quant_8bit"))
    m.def(
        "sgl_per_token_group_quant_8bit(Tensor input, Tensor! output_q, Tensor! output_s, int group_size,"
        " float eps, float fp8_min, float fp8_max, bool scale_ue8m0) -> ()");
  if (!has_schema("sgl_kernel::sgl_per_token_group_quant_8bit_v2"))
    m.def(
        "sgl_per_token_group_quant_8bit_v2(Tensor input, Tensor! output_q, Tensor! output_s, int group_size,"
        " float eps, float fp8_min, float fp8_max, bool scale_ue8m0, bool fuse_silu_and_mul, Tensor? masked_m) -> ()");
  if (!has_schema("sgl_kernel::sgl_per_tensor_quant_fp8"))
    m.def("sgl_per_tensor_quant_fp8(Tensor input, Tensor! output_q, Tensor! output_s, bool is_static) -> ()");
  if (!has_schema("sgl_kernel::bmm_fp8"))
    m.def(
        "bmm_fp8(Tensor A, Tensor B, Tensor! D, Tensor A_scale, Tensor B_scale, Tensor workspace_buffer, "
        "int cublas_handle) -> ()",
        {at::Tag::needs_fixed_stride_order});
  if (!has_schema("sgl_kernel::per_token_quant_int8_cpu"))
    m.def("per_token_quant_int8_cpu(Tensor A) -> (Tensor, Tensor)");
  if (!has_schema("sgl_kernel::int8_scaled_mm_cpu"))
    m.def(
        "int8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales1, Tensor scales2, Tensor? bias, ScalarType "
        "out_dtype, bool is_vnni) -> Tensor");
  if (!has_schema("sgl_kernel::int8_scaled_mm_with_quant"))
    m.def(
        "int8_scaled_mm_with_quant(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, ScalarType out_dtype, bool "
        "is_vnni) -> Tensor");
}

TORCH_LIBRARY_IMPL(sgl_kernel, CUDA, m) {
  m.impl("int8_scaled_mm", &int8_scaled_mm);
  m.impl("fp8_scaled_mm", &fp8_scaled_mm);
  m.impl("fp8_blockwise_scaled_mm", &fp8_blockwise_scaled_mm);
  m.impl("sgl_per_token_quant_fp8", &sgl_per_token_quant_fp8);
  m.impl("sgl_per_token_group_quant_8bit", &sgl_per_token_group_quant_8bit);
  m.impl("sgl_per_token_group_quant_8bit_v2", &sgl_per_token_group_quant_8bit_v2);
  m.impl("sgl_per_tensor_quant_fp8", &sgl_per_tensor_quant_fp8);
  m.impl("bmm_fp8", &bmm_fp8);
  m.impl("per_token_quant_int8_cpu", &per_token_quant_int8_cuda);
  m.impl("int8_scaled_mm_cpu", &int8_scaled_mm_cpuname_cuda);
  m.impl("int8_scaled_mm_with_quant", &int8_scaled_mm_with_quant_cuda);
}
