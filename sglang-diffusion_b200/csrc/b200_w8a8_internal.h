// Internal declarations shared by the .cu translation units of libb200_w8a8.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200_w8a8.h"

namespace b2 {

// quant.cu
cudaError_t quant_rows_int8(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s, int M,
                            int K, cudaStream_t st);
cudaError_t quant_rows_fp8(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M,
                           int K, cudaStream_t st);
cudaError_t quant_fp8_group128(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                               int64_t s_row_stride, int64_t s_grp_stride, int M, int K, float eps, float fmin,
                               float fmax, cudaStream_t st);

cudaError_t quant_rows_int8_sum(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s,
                                void* row_sum, int M, int K, cudaStream_t st);

// quant_fused.cu
struct FusedQuantArgs {
  const void* x;      // [M, K] (NONE / RMSNORM) or [M, 2K] (SILU_MUL: gate | up)
  int64_t ldx;
  void* residual;     // RMSNORM: optional in/out residual rows
  int64_t ldr;
  const void* weight; // RMSNORM: [K], dtype of x
  float eps;          // RMSNORM epsilon
  void* y;            // optional: the produced row in x's dtype (what the unfused op would have stored)
  int64_t ldy;
  uint8_t* q;
  int64_t ldq;
  float* s;
  int64_t s_row_stride, s_grp_stride, s_exp_stride;
  const int32_t* masked_m;  // optional [E]: valid tokens per expert of an [E, T, ...] layout
  int64_t tokens_per_expert;
  int64_t M;          // rows (E * T for the masked layout)
  int K;
  float q_eps, fmin, fmax;  // group modes
  const int32_t* row_map;   // optional gather: output row r quantizes input row row_map[r] / row_div
  int row_div;
  int row_limit;            // ids >= row_limit are padding rows (all-zero codes)
  // prod 3 (gathered sum + rmsnorm): x is this rank's gather buffer (allreduce_symm.cu, gathered form): 2 sets
  // `gather_set_stride` ELEMENTS apart of `gather_world` slots of M x K elements (ldx == K); the kernel takes the
  // barrier and advances the set counter exactly as allreduce_gathered_kernel does
  int gather_world, gather_rank;
  int64_t gather_set_stride;
  uint32_t* const* gather_pads;
  unsigned long long gather_timeout_ns;
  int gather_max_blocks;    // grid bound = barrier slots of the signal pad; the kernel loops over rows beyond it
  // all-gather form (row modes only): codes go to row out_row0 + r of EVERY rank's buffer out_peers[p] (row stride ldq
  // bytes) and the scales to the fp32 array out_scale_offset bytes into it; a second barrier closes the kernel.
  // nullptr: outputs are local (q, s above).
  uint8_t* const* out_peers;
  int64_t out_row0, out_scale_offset;
};
// prod: 0 none, 1 silu_and_mul, 2 rmsnorm, 3 gathered sum (row-parallel partials of `gather_world` ranks) + rmsnorm;  mode: 0 int8 rows, 1 fp8 rows, 2 fp8 1x128 (v1), 3 fp8 1x128 (v2)
cudaError_t fused_quant_launch(int prod, int mode, int x_dtype, const FusedQuantArgs& a, cudaStream_t st);
cudaError_t quant_fp8_tensor(const void* x, int x_dtype, uint8_t* q, float* s, float* s_repeat, int64_t n,
                             int64_t rows, int is_static, cudaStream_t st);

// moe.cu
cudaError_t moe_align128(const int32_t* topk_ids, int n, int E, int32_t* sorted_ids, int32_t* tile_expert,
                         int32_t* inv_pos, int32_t* num_post_pad, int max_padded, cudaStream_t st);
cudaError_t moe_combine(const void* y, int64_t ldy, const int32_t* inv_pos, const float* w, void* out, int64_t ldo,
                        int tokens, int topk, int H, int dtype, cudaStream_t st);

// gemm_sm100.cu
struct GemmArgs {
  const void* A;   // [M,K] 8-bit, K contiguous
  int64_t lda;
  const void* W;   // [N,K] 8-bit, K contiguous
  int64_t ldw;
  const float* As; // [M] per-token scales (per-tensor kernels) or SFA (block-wise)
  const float* Bs; // [N] per-channel scales or SFB
  const void* bias;
  int bias_dtype;
  void* D;
  int64_t ldd;
  int out_dtype;
  int M, N, K;
  // block-wise (128x128) only
  int64_t sfa_m_stride, sfa_k_stride, sfb_n_stride, sfb_k_stride;
  // grouped (MoE) mode: W is [num_experts * N, K], Bs is [num_experts * N]; tile_expert[b] = expert of rows
  // [128 b, 128 b + 128) of A (expert-sorted, padded per expert), -1 = no tokens
  const int32_t* tile_expert;
  int num_experts;
  // decode-shape row-parallel layers: push the output through the NVSwitch multicast mapping into every rank's gather
  // buffer instead of D (gemm_splitk_sm100.cu SplitParams::mc_*); nullptr = off
  void* mc_slot;
  long long mc_set_bytes;
  const uint32_t* mc_counter;
  // prefill row-parallel layers: scatter epilogue of the persistent kernel (gemm_sm100.cu, RS).  rs_peer_bufs is a HOST
  // array of the `rs_world` ranks' gather buffers (peer-mapped device addresses); D is unused
  void* const* rs_peer_bufs;
  int rs_world, rs_rank;
  long long rs_set_bytes;
  const uint32_t* rs_counter;
  // decode shapes (M <= 32): bf16 / f16 activations quantized inside the split-K kernel (A and As are then unused)
  const void* x;
  int64_t ldx;
  int x_dtype;
  // batched mode (bmm_fp8): `batch` independent problems, strides in BYTES for A / W and in ELEMENTS for D; As / Bs
  // are shared by every batch entry
  int batch;
  int64_t a_batch_stride, w_batch_stride, d_batch_stride;
  int scalar_scales;  // As / Bs are single per-tensor values (persistent kernel only)
};

enum GemmFlavor { GEMM_S8 = 0, GEMM_E4M3 = 1, GEMM_E4M3_BLOCK128 = 2 };

// Returns B2_OK or a B2_ERR_* code; `err` receives a message on failure.
int gemm_launch(GemmFlavor flavor, const GemmArgs& a, int forced_config, const char** used_config,
                cudaStream_t st, char* err, size_t err_len);

// capi.cu
void count_launch(int n = 1);
void* profile_buffer();  // device buffer set by b2_set_profile_buffer (nullptr = profiling off)
int device_sm_count();
int fail_msg(int code, const char* fmt, ...);  // sets b2_last_error() of the calling thread, returns `code`
bool weight_prefetch_enabled();  // b2_set_weight_prefetch: W operands are static, GEMMs may launch programmatically

}  // namespace b2
