/* b200_w8a8 -- C ABI of the B200-native (sm_100a) W8A8 quantized-linear library.
 *
 * This is the drop-in boundary under the `torch.ops.sgl_kernel.*` operators of the
 * reference (analytics-zoo/sglang-diffusion): plain pointers, sizes and a CUDA
 * stream; no torch types.  Every entry point names the reference interface it
 * replaces (paths relative to the reference root).  The torch registration shim
 * (sglang-diffusion_b200/csrc/torch_ops.cpp) and the ctypes binding
 * (sglang-diffusion_b200/capi.py) are both thin callers of these functions.
 *
 * Conventions
 *  - All data pointers are DEVICE pointers valid on the current CUDA device; the
 *    call enqueues work on `stream` and returns without synchronising.  Calls are
 *    CUDA-graph capturable (no allocation, no sync, no host round trip).
 *  - Return value: 0 on success, a negative B2_ERR_* code otherwise;
 *    b2_last_error() returns a thread-local human-readable message.
 *  - Weights `W` are the checkpoint layout [N, K], K contiguous (the reference
 *    passes the same memory as mat_b = weight.t(), i.e. logical [K,N] with
 *    stride(0) == 1; `ldw` = mat_b.stride(1)).
 *  - `ld*` are row strides in ELEMENTS.
 *  - There is no CPU fallback: on a device that is not sm_100 every compute entry
 *    returns B2_ERR_ARCH.
 */
#ifndef B200_W8A8_H_
#define B200_W8A8_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 1

/* element types (activations in, outputs, bias) */
#define B2_DTYPE_BF16 0
#define B2_DTYPE_F16 1
#define B2_DTYPE_F32 2
#define B2_DTYPE_S32 3 /* output only: raw int32 accumulators of b2_gemm_s8 (verification) */

/* status codes */
#define B2_OK 0
#define B2_ERR_ARG (-1)   /* shape / alignment / dtype violation (message says which) */
#define B2_ERR_ARCH (-2)  /* device is not sm_100 (B200) */
#define B2_ERR_CUDA (-3)  /* a CUDA runtime / driver call failed */

typedef void* b2_stream_t; /* cudaStream_t */

int b2_abi_version(void);
const char* b2_last_error(void);

/* SM count and compute capability of the current device (cached per device). */
int b2_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- activation quantizers ------------------------------------------------ */

/* Per-token symmetric int8 quantization with the arithmetic of the reference CPU op
 * per_token_quant_int8_cpu (sgl-kernel/csrc/cpu/gemm_int8.cpp:382-410,
 * vec.h:190-228): amax = max(max|x|,1e-7), s = amax/127, q = RNE(x*(127/amax)).
 * Also the GPU replacement for the Triton per_token_quant_int8
 * (python/sglang/srt/layers/quantization/int8_kernel.py:28-89).
 * q is SIGNED int8 (= reference u8 code - 128).  x_dtype: BF16 or F16. */
int b2_quant_int8_rows(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s, int M, int K,
                       b2_stream_t stream);

/* Per-token e4m3 quantization: sgl_per_token_quant_fp8
 * (sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:242-314; kernel :17-121).
 * s = max|x|/448, q = e4m3_rn_satfinite(clamp(x*(1/s), +-448)); all-zero row -> s=0, q=0. */
int b2_quant_fp8_rows(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M, int K,
                      b2_stream_t stream);

/* 1x128 group e4m3 quantization with fp32 scales: sgl_per_token_group_quant_8bit
 * (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:116-217; kernel :19-114)
 * with group_size = 128, scale_ue8m0 = false.  Scale of (row m, group g) is written to
 * s[m*s_row_stride + g*s_grp_stride]: (1, M) gives the column-major / M-major layout the
 * block-wise GEMM consumes, (K/128, 1) the row-major one. */
int b2_quant_fp8_group128(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                          int64_t s_row_stride, int64_t s_grp_stride, int M, int K, float eps, float fp8_min,
                          float fp8_max, b2_stream_t stream);

/* per_token_quant_int8(x, cal_sum=True) of python/sglang/srt/layers/quantization/int8_kernel.py:46-89 (QoQ): the
 * int8 row quantizer above plus row_sum[m] = sum_k x[m,k] (fp32 accumulation, stored in x's dtype).
 * Requires K % 8 == 0 and 16-byte aligned rows. */
int b2_quant_int8_rows_sum(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s, void* row_sum,
                           int M, int K, b2_stream_t stream);

/* sgl_per_tensor_quant_fp8 (sgl-kernel/csrc/gemm/per_tensor_quant_fp8.cu:98-123) and static_quant_fp8
 * (python/sglang/srt/layers/quantization/fp8_kernel.py:649-700): q = e4m3(clamp(x * (1/s), +-448)) over `numel`
 * contiguous elements.  is_static == 0: s[0] (zeroed by the caller, fp8_kernel.py:1671) first receives max|x|/448
 * (two launches); is_static != 0: s[0] is read.  s_repeat (optional, [rows]) receives s once per token row -- the
 * repeat_scale=True form that feeds the per-token GEMM epilogue (fp8_utils.py:1232-1239). */
int b2_quant_fp8_tensor(const void* x, int x_dtype, uint8_t* q, float* s, float* s_repeat, int64_t numel, int64_t rows,
                        int is_static, b2_stream_t stream);

/* ---- producer-fused quantizers ---------------------------- */

#define B2_QMODE_INT8_ROWS 0       /* per-token int8, arithmetic of b2_quant_int8_rows; s[m] */
#define B2_QMODE_FP8_ROWS 1        /* per-token e4m3, arithmetic of b2_quant_fp8_rows; s[m] */
#define B2_QMODE_FP8_GROUP128 2    /* 1x128 e4m3, arithmetic of b2_quant_fp8_group128 (s = amax/448, q = x/s) */
#define B2_QMODE_FP8_GROUP128_V2 3 /* 1x128 e4m3, arithmetic of per_token_group_quant_8bit_v2.cu (s = amax*(1/448),
                                      q = x*(448/amax)) */

/* silu_and_mul fused into the quantizer: sgl_per_token_group_quant_8bit_v2(fuse_silu_and_mul=true)
 * (sgl-kernel/csrc/gemm/per_token_group_quant_8bit_v2.cu:396-447; kernel :240-394).  x is [M, 2K] (gate | up);
 * y = T(silu(gate)) * up in x's dtype T, then quantized with `qmode` (INT8_ROWS, FP8_ROWS or FP8_GROUP128_V2).
 * Scale of (row m, group g) at s[e*s_exp_stride + t*s_row_stride + g*s_grp_stride] where (e, t) = divmod(m,
 * tokens_per_expert) under the masked layout and (0, m) otherwise.  masked_m (optional, [E] int32): x is
 * [E, tokens_per_expert, 2K] and only the first masked_m[e] tokens of expert e are processed.  y (optional) receives
 * the activation rows themselves ([M, K], dtype T).  K % 8 == 0 (K % 128 == 0 for groups), K <= 32768. */
int b2_silu_mul_quant(const void* x, int x_dtype, int64_t ldx, int qmode, uint8_t* q, int64_t ldq, float* s,
                      int64_t s_row_stride, int64_t s_grp_stride, int64_t s_exp_stride, void* y, int64_t ldy,
                      const int32_t* masked_m, int64_t tokens_per_expert, int64_t M, int K, b2_stream_t stream);

/* (fused-add-)RMSNorm fused into the quantizer: sgl_kernel fused_add_rmsnorm / rmsnorm
 * (sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu; flashinfer norm.cuh FusedAddRMSNormKernel) followed by
 * the activation quantizer of the next linear, in one pass.  residual (optional, in/out, [M,K]): s = x + residual is
 * written back to it, as the reference op does; y = T((s * rsqrt(mean(s^2) + eps)) * weight) is quantized with
 * `qmode` (INT8_ROWS, FP8_ROWS or FP8_GROUP128); y_out (optional) also receives y.  Same shape limits as above. */
int b2_rmsnorm_quant(const void* x, int x_dtype, int64_t ldx, void* residual, int64_t ldr, const void* weight,
                     float eps, int qmode, uint8_t* q, int64_t ldq, float* s, int64_t s_row_stride,
                     int64_t s_grp_stride, void* y_out, int64_t ldy, int64_t M, int K, b2_stream_t stream);

/* sgl_per_token_group_quant_8bit_v2 without fusion (same file, fuse_silu_and_mul = false), fp32 scales, G = 128. */
int b2_quant_fp8_group128_v2(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                             int64_t s_row_stride, int64_t s_grp_stride, int64_t M, int K, b2_stream_t stream);

/* ---- scaled GEMMs ---------------------------------------------------------- */

/* int8_scaled_mm (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747 -- which has NO
 * sm_100 path) with the bit-exact epilogue of the reference CPU op int8_scaled_mm_cpu
 * (sgl-kernel/csrc/cpu/gemm_int8.cpp:422-478; scale_C :20-60):
 *     D[m,n] = out_t( fma( f32(acc)*As[m], Bs[n], bias[n] ) )   (no bias: (f32(acc)*As)*Bs)
 * acc = exact s8 x s8 -> s32 dot over K.  out_dtype BF16/F16, or S32 for the raw
 * accumulators (As/Bs/bias ignored).  bias may be NULL; bias_dtype BF16/F16/F32.
 * Requires K % 16 == 0, N % 8 == 0 (as the reference, :677-679), 16-byte aligned bases. */
int b2_gemm_s8(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As, const float* Bs,
               const void* bias, int bias_dtype, void* D, int64_t ldd, int out_dtype, int M, int N, int K,
               b2_stream_t stream);

/* fp8_scaled_mm (sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1450-1532; sm100 kernel
 * :796-1170):  D = out_t( As[m] * (Bs[n] * f32(A.W^T)) [+ bias] ), e4m3 x e4m3, f32 accumulate.
 * out_dtype BF16/F16/F32. */
int b2_gemm_e4m3(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* As, const float* Bs,
                 const void* bias, int bias_dtype, void* D, int64_t ldd, int out_dtype, int M, int N, int K,
                 b2_stream_t stream);

/* fp8_blockwise_scaled_mm (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:368-465):
 *     D[m,n] = out_t( sum_kb SFA[m,kb] * SFB[n/128,kb] * f32(A_kb . W_kb^T) [+ bias] )
 * SFA element (m,kb) at sfa[m*sfa_m_stride + kb*sfa_k_stride] (reference: M-major => (1, M));
 * SFB element (nb,kb) at sfb[nb*sfb_n_stride + kb*sfb_k_stride].  The reference op has no
 * bias (fp8_utils.py:546-547 adds it in a second kernel); here it may be fused (NULL = none). */
int b2_gemm_e4m3_block128(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* sfa,
                          int64_t sfa_m_stride, int64_t sfa_k_stride, const float* sfb, int64_t sfb_n_stride,
                          int64_t sfb_k_stride, const void* bias, int bias_dtype, void* D, int64_t ldd,
                          int out_dtype, int M, int N, int K, b2_stream_t stream);

/* bmm_fp8 (sgl-kernel/csrc/gemm/bmm_fp8.cu; flashinfer bmm_fp8_internal_cublaslt): for every batch b,
 *     D[b] = out_t( a_scale * b_scale * f32(A[b] . W[b]^T) ),  A [batch, M, K] e4m3, W [batch, N, K] e4m3 (the
 * reference passes B = W.transpose(-2,-1), "column major"), per-tensor fp32 device scalars a_scale / b_scale.
 * ONE launch: the persistent tcgen05 GEMM in batched mode (3-D tensor maps clip every tile to its own batch entry,
 * tiles enumerate batch-major, the epilogue reads the two scalars).  Batch strides: A / W in bytes, D in elements.
 * The workspace arguments mirror the reference's cuBLASLt workspace and are unused (b2_bmm_e4m3_workspace() == 0). */
size_t b2_bmm_e4m3_workspace(int M, int N);
int b2_bmm_e4m3(const uint8_t* A, int64_t lda, int64_t a_batch_stride, const uint8_t* W, int64_t ldw,
                int64_t w_batch_stride, const float* a_scale, const float* b_scale, void* D, int64_t ldd,
                int64_t d_batch_stride, int out_dtype, int batch, int M, int N, int K, void* workspace,
                size_t workspace_bytes, b2_stream_t stream);

/* ---- W8A8-int8 mixture of experts --------------------------- */

/* moe_align_block_size with 128-row blocks (sgl-kernel/csrc/moe/moe_align_kernel.cu; CPU twin
 * sgl-kernel/csrc/cpu/moe.cpp moe_align_block_size): groups the n = tokens * topk routed rows by expert, every
 * expert's run padded to a multiple of 128 rows so that one UMMA tile never mixes experts.
 *   sorted_ids[max_padded]       flat index token*topk + k of every row, padding rows hold n
 *   tile_expert[max_padded/128]  expert of every 128-row block, -1 for blocks past num_post_pad
 *   inv_pos[n]                   row of flat index i (-1 for an expert id outside [0, E))
 *   num_post_pad[1]              rows in use
 * max_padded must be a multiple of 128 and >= b2_moe_max_padded(n, E).  E <= 4096. */
int64_t b2_moe_max_padded(int64_t n, int num_experts);
int b2_moe_align_block128(const int32_t* topk_ids, int n, int num_experts, int32_t* sorted_ids, int32_t* tile_expert,
                          int32_t* inv_pos, int32_t* num_post_pad, int64_t max_padded, b2_stream_t stream);

/* Stage 0 / 2 of fused_experts_int8_kernel_impl (moe_int8.cpp:620-626): per-token int8 quantization (arithmetic of
 * b2_quant_int8_rows) of row row_map[r] / row_div of x into row r of q, for r < P; ids >= row_limit are padding
 * rows and become all-zero codes.  K % 8 == 0, K <= 32768. */
int b2_quant_int8_rows_gather(const void* x, int x_dtype, int64_t ldx, const int32_t* row_map, int row_div,
                              int row_limit, int8_t* q, int64_t ldq, float* s, int64_t P, int K, b2_stream_t stream);

/* Grouped int8 GEMM over expert-sorted rows (stage 1 / 2 of moe_int8.cpp, the tinygemm / brgemm blocks of
 * :640-760): for every 128-row block b with e = tile_expert[b] >= 0,
 *     D[rows of b, :] = out_t( (f32(acc) * As[row]) * Bs[e*N + n] ),  acc = A[rows] . W[e*N + n, :]^T  (exact int32)
 * W is the stacked expert weight [num_experts * N, K] (the checkpoint layout [E, N, K]), Bs [num_experts * N].
 * The arithmetic of one row is that of b2_gemm_s8 with the expert's weight: bit-identical to the dense op. */
int b2_gemm_s8_grouped(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As, const float* Bs,
                       const int32_t* tile_expert, int num_experts, void* D, int64_t ldd, int out_dtype, int64_t P,
                       int N, int K, b2_stream_t stream);

/* Stage 3: out[t, :] = out_t( sum_k topk_weights[t,k] * f32(y[inv_pos[t*topk + k], :]) ), k ascending, fp32
 * accumulation (moe_int8.cpp:830-846).  H % 8 == 0. */
int b2_moe_combine(const void* y, int64_t ldy, const int32_t* inv_pos, const float* topk_weights, void* out,
                   int64_t ldo, int tokens, int topk, int H, int dtype, b2_stream_t stream);

/* ---- fused linear ----------------------------------------------------------- */

/* int8_scaled_mm_with_quant (sgl-kernel/csrc/cpu/gemm_int8.cpp:481-547):
 * bf16/f16 activations in, per-token int8 quantization, int8 GEMM, scaled + biased output
 * of the same dtype.  `workspace` must hold b2_linear_w8a8_int8_workspace(M, K) bytes
 * (quantized activations + row scales; stays L2-resident between the two kernels). */
size_t b2_linear_w8a8_int8_workspace(int M, int K);
int b2_linear_w8a8_int8(const void* x, int x_dtype, int64_t ldx, const int8_t* W, int64_t ldw, const float* Bs,
                        const void* bias, int bias_dtype, void* D, int64_t ldd, int M, int N, int K,
                        void* workspace, size_t workspace_bytes, b2_stream_t stream);

/* Allocation-free form of b2_linear_w8a8_int8.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Same arithmetic and result, bit for bit; the codes + row scales between the quantizer and the GEMM live in
 * a file-scope `static __device__` array of B2_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library
 * holds it until the process exits) instead of a caller workspace.  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine); the caller still passes x, W, Bs, bias and D.
 * Limit: b2_linear_w8a8_int8_workspace(M, K) <= B2_STATIC_SCRATCH_BYTES (4 MiB: M*K <= ~4.19e6, e.g. M = 128 x
 * K = 28672 or M = 256 x K = 14336); larger problems are rejected with B2_ERR_ARG and a message -- use the workspace
 * form.  ALL calls on a device share the one array, which keeps the last call's contents (nothing in it needs to start
 * at zero: every byte the GEMM reads is written by the quantizer of the same call): calls that overlap on the GPU
 * -- two streams, or two threads on one device -- race on it; serialise them on one stream. */
#define B2_STATIC_SCRATCH_BYTES (4u << 20)
size_t b2_linear_w8a8_int8_static_max_bytes(void);
int b2_linear_w8a8_int8_static(const void* x, int x_dtype, int64_t ldx, const int8_t* W, int64_t ldw, const float* Bs,
                               const void* bias, int bias_dtype, void* D, int64_t ldd, int M, int N, int K,
                               b2_stream_t stream);

/* One-kernel form of the two entries above for M <= 32 (calling thread; default 0 = off, or B2_FUSED_DECODE=1 in the
 * environment): the cluster split-K decode GEMM quantizes its own K slice of the activations into resident
 * shared-memory A tiles (row maxima exchanged over DSMEM) -- no quantizer launch, no codes round trip, workspace
 * unused; results bit-identical.  Measured SLOWER than the two programmatically chained launches on every decode
 * shape, hence off by default. */
void b2_set_fused_decode(int enabled);
int b2_get_fused_decode(void);

/* Same composition for per-token fp8 (W8A8Fp8LinearMethod.apply ->
 * apply_fp8_linear, python/sglang/srt/layers/quantization/fp8_utils.py:1194-1276). */
size_t b2_linear_w8a8_fp8_workspace(int M, int K);
int b2_linear_w8a8_fp8(const void* x, int x_dtype, int64_t ldx, const uint8_t* W, int64_t ldw, const float* Bs,
                       const void* bias, int bias_dtype, void* D, int64_t ldd, int M, int N, int K,
                       void* workspace, size_t workspace_bytes, b2_stream_t stream);

/* ---- introspection (tests / bench) ----------------------------------------- */

/* Number of kernels launched by this library on the calling thread since the last reset. */
uint64_t b2_launch_count(void);
void b2_launch_count_reset(void);

/* Force a GEMM tile configuration (0 = automatic).  For benchmarks and parity sweeps. */
void b2_set_gemm_config(int config_id);
/* Name of the tile configuration the last GEMM call on this thread used. */
const char* b2_last_gemm_config(void);

/* Sum of the row-parallel partial outputs over the tensor-parallel ranks of one NVSwitch domain, through peer memory
 * (replaces the NCCL all-reduce of python/sglang/srt/layers/linear.py:1486-1490 for decode-sized messages; the
 * reference's own small-message path is sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:204-260).
 *   peer_bufs_dev    device array of `world` pointers: the SAME symmetric buffer as mapped on every rank (entry
 *                    `rank` is the local one).  Every rank's GEMM has written its [numel] partials there, earlier in
 *                    `stream`.  The two-shot path leaves the sums in the buffer as well.
 *   signal_pads_dev  device array of `world` pointers to u32 signal pads of >= 8 KiB each, zeroed once at allocation
 *                    and mapped like the buffers.  Words [0, world * max_blocks) hold the barrier numbers the peers
 *                    post, words [1024, 1024 + max_blocks) this rank's per-CTA barrier counters (device-resident so
 *                    that CUDA-graph replays advance them), word 1536 the time-out record; nothing is reset between
 *                    calls, so the pads are OWNED by this library: do not use the same symmetric-memory handle for
 *                    barrier() / put_signal() / wait_signal() of another library.
 *   multicast_ptr    NVSwitch multicast mapping of the buffer, or NULL (then only the one-shot path is available)
 *   out              private [numel] output of this rank (16-byte aligned), dtype bf16 / f16 like the buffers; NULL =
 *                    in place: the sums stay in the symmetric buffer only (two-shot path; saves the copy for
 *                    prefill-sized messages when the consumer is the next kernel of the same stream)
 *   algo             0 auto (two-shot when multicast_ptr), 1 one-shot peer loads (fp32 sum in rank order, bit-identical
 *                    on every rank), 2 two-shot multimem.ld_reduce / multimem.st (the switch adds in fp32)
 * The host side only has to allocate and exchange the mappings (torch.distributed._symmetric_memory in
 * sglang_diffusion_b200/comm.py); every rank must make the same call.  No NCCL on this path. */
int b2_allreduce_symm(void* const* peer_bufs_dev, uint32_t* const* signal_pads_dev, void* multicast_ptr, void* out,
                      int64_t numel, int dtype, int rank, int world, int algo, int max_blocks, b2_stream_t stream);

/* GEMM-epilogue-fused all-reduce for decode-shape row-parallel layers (SURVEY 8f rank 2;
 * python/sglang/srt/layers/linear.py:1467-1496, sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:204-260).
 * Step 1, b2_gemm_{s8,e4m3}_push_symm: the scaled GEMM of b2_gemm_s8 / b2_gemm_e4m3 (M <= 128), whose epilogue stores
 * this rank's [M,N] partial output with multimem.st through `multicast_ptr` -- the NVSwitch multicast mapping of the
 * symmetric buffer -- into slot `rank` of EVERY rank's buffer: the partials leave for the peers tile by tile while the
 * GEMM is still running, no separate copy or reduce-scatter step.  The buffer holds two sets of `world` slots of
 * M*N out_dtype elements each (>= 2*world*M*N*2 bytes); the set in use alternates per call (counter in word 1544 of the
 * local signal pad, advanced on the device, so CUDA-graph replays stay consistent).
 * Step 2, b2_allreduce_symm_gathered: ONE barrier ("every rank's GEMM has completed") and a sum over the `world` slots
 * of LOCAL memory in rank order (fp32, one rounding): bit-identical on every rank and equal to algo 1 of
 * b2_allreduce_symm.  Every rank must make the same sequence of calls; the signal pads are the ones of
 * b2_allreduce_symm (same ownership rule).  bias: added by the GEMM (give it on one rank only, linear.py:1480).
 * set_stride_bytes: distance of the two sets inside the gather buffer, the same on every call that shares the buffer
 * (half of it; a multiple of 16; 0 = world * M * N * 2, only safe when every call has the same size). */
int b2_gemm_s8_push_symm(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As, const float* Bs,
                         const void* bias, int bias_dtype, void* multicast_ptr, uint32_t* local_signal_pad,
                         int64_t set_stride_bytes, int rank, int world, int out_dtype, int M, int N, int K,
                         b2_stream_t stream);
int b2_gemm_e4m3_push_symm(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* As,
                           const float* Bs, const void* bias, int bias_dtype, void* multicast_ptr,
                           uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world, int out_dtype,
                           int M, int N, int K, b2_stream_t stream);
int b2_allreduce_symm_gathered(const void* local_buf, uint32_t* const* signal_pads_dev, void* out, int64_t numel,
                               int64_t set_stride_bytes, int dtype, int rank, int world, int max_blocks,
                               b2_stream_t stream);

/* Step 2 fused with what follows a row-parallel layer in a decoder block: all-reduce + fused_add_rmsnorm + the next
 * layer's activation quantizer in ONE launch (the reference's fused all-reduce + RMSNorm hook,
 * python/sglang/srt/distributed/parallel_state.py:648-698 GroupCoordinator.fused_allreduce_rmsnorm, called from
 * python/sglang/srt/layers/layernorm.py forward_with_allreduce_fusion; norm arithmetic of
 * sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu).  After b2_gemm_*_push_symm:
 *   t      = out_t( sum over ranks of the pushed partials )      -- fp32 in rank order, one rounding (== step 2)
 *   sum    = f32(t) + f32(residual);  residual <- out_t(sum)      (residual may be NULL: plain rmsnorm of t)
 *   y      = out_t( (sum * rsqrt(mean(sum^2) + eps)) * weight )   (written to y_out when not NULL)
 *   q, s   = quantizer `qmode` of y (B2_QMODE_INT8_ROWS / FP8_ROWS / FP8_GROUP128), layouts of b2_rmsnorm_quant
 * Equal bit for bit to b2_allreduce_symm_gathered followed by b2_rmsnorm_quant.  One CTA (and one cross-GPU barrier
 * slot) per row, at most min(max_blocks, 1024 / world) CTAs; with more rows the CTAs loop. */
int b2_allreduce_symm_gathered_rmsnorm_quant(const void* local_buf, uint32_t* const* signal_pads_dev, void* residual,
                                             int64_t ldr, const void* weight, float eps, int qmode, uint8_t* q,
                                             int64_t ldq, float* s, int64_t s_row_stride, int64_t s_grp_stride,
                                             void* y_out, int64_t ldy, int M, int N, int64_t set_stride_bytes,
                                             int dtype, int rank, int world, int max_blocks, b2_stream_t stream);

/* Prefill form (M >= 32 * world): GEMM with a SCATTER epilogue + the same gathered sum = a fused reduce-scatter.
 * The sum of a row-parallel layer is only needed row-sharded when the next operation is token-wise (sequence-parallel
 * norm / quantizer, or an all-gather later): each rank then keeps rows [rank * M/world, (rank+1) * M/world).  The
 * persistent GEMM stores every 32-row output box straight into slot `rank` of the OWNING rank's gather buffer (TMA
 * store over NVLink to the peer mapping, tile by tile under the main loop of the next tile), so after ONE barrier
 * b2_allreduce_symm_gathered(local_buf, pads, out, (M/world)*N, set_stride_bytes, ...) sums `world` slots of local memory
 * into out[M/world, N] -- fp32 in rank order, one rounding: equal bit for bit to rows of algo 1 of b2_allreduce_symm.
 * NVLink traffic per rank: (world-1)/world * M*N*2 bytes out, the same in; no separate pass over the partials.
 * peer_bufs_host: HOST array of the `world` ranks' gather buffers as mapped into this process (symmetric memory
 * buffer_ptrs); each holds 2 sets set_stride_bytes apart of `world` slots of (M/world)*N 16-bit elements.
 * Needs M % (32 * world) == 0, N % 8 == 0.  Replaces: RowParallelLinear.forward(skip_all_reduce=True)
 * (python/sglang/srt/layers/linear.py:1467-1490) followed by LayerCommunicator._tp_reduce_scatter
 * (python/sglang/srt/layers/communicator.py:542-560 -> GroupCoordinator.reduce_scatter_tensor,
 * python/sglang/srt/distributed/parallel_state.py:758-762). */
int b2_gemm_s8_scatter_symm(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As,
                            const float* Bs, const void* bias, int bias_dtype, void* const* peer_bufs_host,
                            uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world, int out_dtype,
                            int M, int N, int K, b2_stream_t stream);
int b2_gemm_e4m3_scatter_symm(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* As,
                              const float* Bs, const void* bias, int bias_dtype, void* const* peer_bufs_host,
                              uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world,
                              int out_dtype, int M, int N, int K, b2_stream_t stream);

/* Prefill, sequence-parallel norm ("reduce-scatter + fused next-layer norm for large M"): after
 * b2_gemm_*_scatter_symm every rank sums ITS M/world rows, adds its rows of the residual (residual_local [M/world, N],
 * updated in place -- the reference splits the residual the same way, communicator.py:556-559), applies RMSNorm and
 * the next layer's per-token quantizer, and stores codes and scales straight into rows [rank*M/world, ...) of EVERY
 * rank's activation buffer (act_bufs_dev: DEVICE array of the `world` peer mappings of one symmetric buffer; codes
 * [M, N] at offset 0, fp32 scales [M] at scale_offset_bytes): the all-gather moves 1-byte codes instead of 2-byte
 * sums, the norm is computed once per row instead of on every rank, and the summed rows never reach HBM.  A second
 * barrier closes the kernel: when it completes, this rank's buffer holds all M rows -- the A operand and scales of the
 * next column-parallel b2_gemm_s8 / b2_gemm_e4m3.  NVLink bytes per rank: (world-1)/world * M*N*2 in the GEMM's
 * scatter + (world-1) * (M/world) * (N + 4) here, against 2*(world-1)/world * M*N*2 of an all-reduce.
 * Codes, scales and residual rows equal, bit for bit, b2_allreduce_symm_gathered + b2_rmsnorm_quant on the same rows.
 * qmode: B2_QMODE_INT8_ROWS or B2_QMODE_FP8_ROWS; M % world == 0, N % 16 == 0. */
int b2_reduce_scatter_symm_rmsnorm_quant_allgather(const void* local_buf, uint32_t* const* signal_pads_dev,
                                                   void* residual_local, int64_t ldr, const void* weight, float eps,
                                                   int qmode, void* const* act_bufs_dev, int64_t scale_offset_bytes,
                                                   void* y_local, int64_t ldy, int M, int N, int64_t set_stride_bytes,
                                                   int dtype, int rank, int world, int max_blocks, b2_stream_t stream);

/* All-reduce form of the same (callers that need the full sum on every rank -- RowParallelLinear.forward with
 * reduce_results, linear.py:1486-1490): after b2_gemm_*_scatter_symm, b2_allreduce_symm_gathered_bcast sums this rank's
 * rows (numel_slice = (M/world)*N elements) and stores them into rows [rank*M/world, ...) of every rank's copy of a
 * SECOND symmetric buffer of >= M*N elements (full_bufs_dev: DEVICE array of the `world` peer mappings; plain peer
 * stores, no multicast mapping needed); after a second barrier this rank's copy holds the whole [M,N] sum, which is
 * copied to `out`, or left in place when out == NULL (valid until the next call that uses the buffer).  Bit-identical
 * on every rank and equal to algo 1 of b2_allreduce_symm. */
int b2_allreduce_symm_gathered_bcast(const void* local_buf, uint32_t* const* signal_pads_dev,
                                     void* const* full_bufs_dev, void* out, int64_t numel_slice,
                                     int64_t set_stride_bytes, int dtype, int rank, int world, int max_blocks,
                                     b2_stream_t stream);

/* Waits of b2_allreduce_symm on a peer.  0 (default): wait for ever -- a late peer (first-iteration skew, weight load,
 * debugger) is not an error, hangs are the host watchdog's business, as with the reference's custom all-reduce.
 * ms > 0: a wait gives up after that long (%globaltimer), stores the barrier number in word 1536 of this rank's
 * signal pad (0 = never timed out) and lets the kernel finish; the result of that call is undefined and the host is
 * expected to read the word and tear the group down.  Never traps: the CUDA context stays usable.  Process-wide. */
void b2_set_allreduce_timeout_ms(long long ms);
long long b2_get_allreduce_timeout_ms(void);

/* Static-weight declaration (calling thread; default 0).  When non-zero the caller promises that the W operand of
 * the following GEMM calls is NOT written by the kernel that precedes the GEMM in its stream (inference weights:
 * loaded once).  Decode-shape GEMMs (M <= 128) are then launched with programmatic dependent launch: their CTAs
 * start while the preceding kernel (normally this library's quantizer, which releases its dependents at once) is
 * still running, request their first ring of WEIGHT tiles, and only then wait for that kernel before touching the
 * activations, the scales or the output.  The row quantizers (M <= 128) are launched programmatically as well and wait
 * for their own predecessor before they touch memory, so a whole quantizer -> GEMM -> quantizer -> GEMM chain overlaps.
 * Results are identical; only the launch overlap changes.  The linear
 * methods of the host package (weights are module parameters) turn it on around their calls. */
void b2_set_weight_prefetch(int enabled);
int b2_get_weight_prefetch(void);

/* Developer hook: when non-NULL, the per-token GEMM kernels write per-CTA wait-cycle counters
 * (8 x uint64 per CTA: MMA-waits-for-operands, MMA-waits-for-accumulator, producer-waits-for-stage,
 * epilogue-waits-for-accumulator, CTA lifetime, 3 reserved) into this device buffer
 * (>= 8 * 8 * grid bytes).  NULL (default) disables it. */
void b2_set_profile_buffer(void* device_buffer);

#ifdef __cplusplus
}
#endif
#endif /* B200_W8A8_H_ */
