// C-ABI entry points of libb200_w8a8.so (see include/b200_w8a8.h for the contract
// and the reference interface each function replaces).  Argument validation mirrors
// the reference ops' TORCH_CHECKs (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:670-695,
// fp8_gemm_kernel.cu:1465-1487, fp8_blockwise_gemm_kernel.cu:374-403) but reports
// through status codes + b2_last_error() instead of exceptions.
#include <cstdlib>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "b200_w8a8_internal.h"

namespace b2 {

static thread_local char g_err[512] = "";
static thread_local uint64_t g_launches = 0;
static thread_local int g_forced_config = 0;
static thread_local const char* g_last_config = "";

static thread_local void* g_prof = nullptr;
static thread_local int g_weight_prefetch = 0;

void count_launch(int n) { g_launches += static_cast<uint64_t>(n); }
void* profile_buffer() { return g_prof; }
bool weight_prefetch_enabled() { return g_weight_prefetch != 0; }

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

// the same, for the other translation units of the library
int fail_msg(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

struct DevInfo {
  int sms = 0, major = 0, minor = 0;
  bool valid = false;
};
static DevInfo g_dev[64];

static const DevInfo* dev_info() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  DevInfo& d = g_dev[dev];
  if (!d.valid) {
    if (cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return nullptr;
    cudaDeviceGetAttribute(&d.major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&d.minor, cudaDevAttrComputeCapabilityMinor, dev);
    d.valid = true;
  }
  return &d;
}

int device_sm_count() {
  const DevInfo* d = dev_info();
  const int sms = d ? d->sms : 148;
  // developer experiment (tools/power_probe.py): run the persistent GEMMs on fewer SMs
  static const int cap = [] {
    const char* e = getenv("B2_GEMM_MAX_SMS");
    return e ? atoi(e) : 0;
  }();
  return (cap >= 2 && cap < sms) ? (cap & ~1) : sms;
}

// No CPU path and no other-architecture path: the kernels are sm_100a only.
static int check_arch() {
  const DevInfo* d = dev_info();
  if (!d) {
    cudaGetLastError();
    return fail(B2_ERR_ARCH, "b200_w8a8: no usable CUDA device (this library has no CPU fallback)");
  }
  if (d->major != 10 || d->minor != 0)
    return fail(B2_ERR_ARCH, "b200_w8a8: kernels are built for sm_100a (B200) only, device is sm_%d%d", d->major,
                d->minor);
  return B2_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int check_act_dtype(int dt, const char* what) {
  if (dt != B2_DTYPE_BF16 && dt != B2_DTYPE_F16) return fail(B2_ERR_ARG, "%s must be bfloat16 or float16", what);
  return B2_OK;
}

static int check_cuda(cudaError_t e, const char* what) {
  if (e != cudaSuccess) return fail(B2_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  return B2_OK;
}

static int check_gemm_common(const void* A, int64_t lda, const void* W, int64_t ldw, const void* D, int64_t ldd,
                             int out_bytes, int M, int N, int K) {
  if (M < 0 || N <= 0 || K <= 0) return fail(B2_ERR_ARG, "bad GEMM shape M=%d N=%d K=%d", M, N, K);
  if (!A || !W || !D) return fail(B2_ERR_ARG, "null matrix pointer");
  if (K % 16 != 0) return fail(B2_ERR_ARG, "mat_a/mat_b must be multiple of 16 bytes for memory alignment (K=%d)", K);
  if ((static_cast<int64_t>(N) * 2) % 16 != 0)
    return fail(B2_ERR_ARG, "out must be multiple of 16 bytes for memory alignment (N=%d must be a multiple of 8)", N);
  if (lda < K || ldw < K || ldd < N) return fail(B2_ERR_ARG, "row stride smaller than row length");
  if (lda % 16 != 0 || ldw % 16 != 0) return fail(B2_ERR_ARG, "row strides of A and W must be multiples of 16 bytes");
  if ((ldd * out_bytes) % 16 != 0) return fail(B2_ERR_ARG, "row stride of D must be a multiple of 16 bytes");
  if (!aligned16(A) || !aligned16(W) || !aligned16(D)) return fail(B2_ERR_ARG, "A, W and D must be 16-byte aligned");
  return B2_OK;
}

}  // namespace b2

using namespace b2;

extern "C" {

int b2_abi_version(void) { return B2_ABI_VERSION; }

const char* b2_last_error(void) { return g_err; }

int b2_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  const DevInfo* d = dev_info();
  if (!d) {
    cudaGetLastError();
    return fail(B2_ERR_CUDA, "no CUDA device");
  }
  if (sm_count) *sm_count = d->sms;
  if (cc_major) *cc_major = d->major;
  if (cc_minor) *cc_minor = d->minor;
  return B2_OK;
}

uint64_t b2_launch_count(void) { return g_launches; }
void b2_launch_count_reset(void) { g_launches = 0; }
void b2_set_gemm_config(int config_id) { g_forced_config = config_id; }
const char* b2_last_gemm_config(void) { return g_last_config; }
void b2_set_profile_buffer(void* device_buffer) { g_prof = device_buffer; }
void b2_set_weight_prefetch(int enabled) { g_weight_prefetch = enabled; }
void b2_set_fused_decode(int enabled);
int b2_get_fused_decode(void);
int b2_get_weight_prefetch(void) { return g_weight_prefetch; }

int b2_quant_int8_rows(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s, int M, int K,
                       b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(x_dtype, "per_token_quant_int8: input")) return r;
  if (M < 0 || K <= 0 || !x || !q || !s || ldx < K || ldq < K)
    return fail(B2_ERR_ARG, "per_token_quant_int8: bad arguments (M=%d K=%d)", M, K);
  if (M == 0) return B2_OK;
  if (int r = check_cuda(quant_rows_int8(x, x_dtype, ldx, q, ldq, s, M, K, static_cast<cudaStream_t>(stream)),
                         "per_token_quant_int8 launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_quant_fp8_rows(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M, int K,
                      b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(x_dtype, "sgl_per_token_quant_fp8: input")) return r;
  if (M < 0 || K <= 0 || !x || !q || !s || ldx < K || ldq < K)
    return fail(B2_ERR_ARG, "sgl_per_token_quant_fp8: bad arguments (M=%d K=%d)", M, K);
  if (M == 0) return B2_OK;
  if (int r = check_cuda(quant_rows_fp8(x, x_dtype, ldx, q, ldq, s, M, K, static_cast<cudaStream_t>(stream)),
                         "sgl_per_token_quant_fp8 launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_quant_fp8_group128(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                          int64_t s_row_stride, int64_t s_grp_stride, int M, int K, float eps, float fp8_min,
                          float fp8_max, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(x_dtype, "sgl_per_token_group_quant_8bit: input")) return r;
  if (M < 0 || K <= 0 || !x || !q || !s) return fail(B2_ERR_ARG, "sgl_per_token_group_quant_8bit: bad arguments");
  if (K % 128 != 0) return fail(B2_ERR_ARG, "sgl_per_token_group_quant_8bit: K=%d must be a multiple of group_size 128", K);
  if (ldx % 8 != 0 || ldq % 8 != 0 || !aligned16(x) || (reinterpret_cast<uintptr_t>(q) & 7u))
    return fail(B2_ERR_ARG, "sgl_per_token_group_quant_8bit: input rows must be 16-byte aligned, output rows 8-byte aligned");
  if (M == 0) return B2_OK;
  if (int r = check_cuda(quant_fp8_group128(x, x_dtype, ldx, q, ldq, s, s_row_stride, s_grp_stride, M, K, eps,
                                            fp8_min, fp8_max, static_cast<cudaStream_t>(stream)),
                         "sgl_per_token_group_quant_8bit launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_quant_int8_rows_sum(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s, void* row_sum,
                           int M, int K, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(x_dtype, "per_token_quant_int8: input")) return r;
  if (M < 0 || K <= 0 || !x || !q || !s || !row_sum || ldx < K || ldq < K)
    return fail(B2_ERR_ARG, "per_token_quant_int8(cal_sum): bad arguments (M=%d K=%d)", M, K);
  if (K % 8 != 0 || ldx % 8 != 0 || ldq % 8 != 0 || !aligned16(x) || (reinterpret_cast<uintptr_t>(q) & 7u))
    return fail(B2_ERR_ARG, "per_token_quant_int8(cal_sum): K, ldx and ldq must be multiples of 8 and rows 16-byte aligned");
  if (M == 0) return B2_OK;
  if (int r = check_cuda(quant_rows_int8_sum(x, x_dtype, ldx, q, ldq, s, row_sum, M, K, static_cast<cudaStream_t>(stream)),
                         "per_token_quant_int8(cal_sum) launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_quant_fp8_tensor(const void* x, int x_dtype, uint8_t* q, float* s, float* s_repeat, int64_t numel, int64_t rows,
                        int is_static, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(x_dtype, "sgl_per_tensor_quant_fp8: input")) return r;
  if (numel < 0 || rows < 0 || !x || !q || !s) return fail(B2_ERR_ARG, "sgl_per_tensor_quant_fp8: bad arguments");
  if (!aligned16(x) || !aligned16(q)) return fail(B2_ERR_ARG, "sgl_per_tensor_quant_fp8: input and output must be 16-byte aligned");
  if (numel == 0) return B2_OK;
  if (int r = check_cuda(quant_fp8_tensor(x, x_dtype, q, s, s_repeat, numel, s_repeat ? rows : 0, is_static,
                                          static_cast<cudaStream_t>(stream)),
                         "sgl_per_tensor_quant_fp8 launch"))
    return r;
  count_launch();  // the absmax pass of the dynamic form counted itself
  return B2_OK;
}

static int check_fused_common(const char* op, const void* x, int x_dtype, int64_t ldx, int64_t row_elems, const void* q,
                              int64_t ldq, const void* s, int qmode, int64_t M, int K) {
  if (int r = check_arch()) return r;
  if (x_dtype != B2_DTYPE_BF16 && x_dtype != B2_DTYPE_F16) return fail(B2_ERR_ARG, "%s: input must be bfloat16 or float16", op);
  if (M < 0 || K <= 0 || !x || !q || !s) return fail(B2_ERR_ARG, "%s: bad arguments", op);
  if (K % 8 != 0 || K > 32768) return fail(B2_ERR_ARG, "%s: K=%d must be a multiple of 8 and <= 32768", op, K);
  if ((qmode == B2_QMODE_FP8_GROUP128 || qmode == B2_QMODE_FP8_GROUP128_V2) && K % 128 != 0)
    return fail(B2_ERR_ARG, "%s: K=%d must be a multiple of group_size 128", op, K);
  if (ldx < row_elems || ldx % 8 != 0 || ldq < K || ldq % 8 != 0 || !aligned16(x) || (reinterpret_cast<uintptr_t>(q) & 7u))
    return fail(B2_ERR_ARG, "%s: input rows must be 16-byte aligned, output rows 8-byte aligned", op);
  return B2_OK;
}

int b2_silu_mul_quant(const void* x, int x_dtype, int64_t ldx, int qmode, uint8_t* q, int64_t ldq, float* s,
                      int64_t s_row_stride, int64_t s_grp_stride, int64_t s_exp_stride, void* y, int64_t ldy,
                      const int32_t* masked_m, int64_t tokens_per_expert, int64_t M, int K, b2_stream_t stream) {
  const char* op = "silu_and_mul + quant";
  if (int r = check_fused_common(op, x, x_dtype, ldx, 2ll * K, q, ldq, s, qmode, M, K)) return r;
  if (qmode != B2_QMODE_INT8_ROWS && qmode != B2_QMODE_FP8_ROWS && qmode != B2_QMODE_FP8_GROUP128_V2)
    return fail(B2_ERR_ARG, "%s: qmode must be INT8_ROWS, FP8_ROWS or FP8_GROUP128_V2", op);
  if (masked_m && tokens_per_expert <= 0) return fail(B2_ERR_ARG, "%s: masked layout needs tokens_per_expert > 0", op);
  if (y && (ldy < K || ldy % 8 != 0 || !aligned16(y))) return fail(B2_ERR_ARG, "%s: y rows must be 16-byte aligned", op);
  if (M == 0) return B2_OK;
  FusedQuantArgs a{};
  a.x = x; a.ldx = ldx; a.y = y; a.ldy = ldy; a.q = q; a.ldq = ldq; a.s = s;
  a.s_row_stride = s_row_stride; a.s_grp_stride = s_grp_stride; a.s_exp_stride = masked_m ? s_exp_stride : 0;
  a.masked_m = masked_m; a.tokens_per_expert = masked_m ? tokens_per_expert : 1; a.M = M; a.K = K;
  a.q_eps = 1e-10f; a.fmin = -448.f; a.fmax = 448.f;
  if (int r = check_cuda(fused_quant_launch(1, qmode, x_dtype, a, static_cast<cudaStream_t>(stream)), "silu_and_mul + quant launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_rmsnorm_quant(const void* x, int x_dtype, int64_t ldx, void* residual, int64_t ldr, const void* weight,
                     float eps, int qmode, uint8_t* q, int64_t ldq, float* s, int64_t s_row_stride,
                     int64_t s_grp_stride, void* y_out, int64_t ldy, int64_t M, int K, b2_stream_t stream) {
  const char* op = "rmsnorm + quant";
  if (int r = check_fused_common(op, x, x_dtype, ldx, K, q, ldq, s, qmode, M, K)) return r;
  if (qmode != B2_QMODE_INT8_ROWS && qmode != B2_QMODE_FP8_ROWS && qmode != B2_QMODE_FP8_GROUP128)
    return fail(B2_ERR_ARG, "%s: qmode must be INT8_ROWS, FP8_ROWS or FP8_GROUP128", op);
  if (!weight || !aligned16(weight)) return fail(B2_ERR_ARG, "%s: weight must be a 16-byte aligned [K] vector", op);
  if (residual && (ldr < K || ldr % 8 != 0 || !aligned16(residual))) return fail(B2_ERR_ARG, "%s: residual rows must be 16-byte aligned", op);
  if (y_out && (ldy < K || ldy % 8 != 0 || !aligned16(y_out))) return fail(B2_ERR_ARG, "%s: y rows must be 16-byte aligned", op);
  if (M == 0) return B2_OK;
  FusedQuantArgs a{};
  a.x = x; a.ldx = ldx; a.residual = residual; a.ldr = ldr; a.weight = weight; a.eps = eps; a.y = y_out; a.ldy = ldy;
  a.q = q; a.ldq = ldq; a.s = s; a.s_row_stride = s_row_stride; a.s_grp_stride = s_grp_stride; a.s_exp_stride = 0;
  a.masked_m = nullptr; a.tokens_per_expert = 1; a.M = M; a.K = K; a.q_eps = 1e-10f; a.fmin = -448.f; a.fmax = 448.f;
  if (int r = check_cuda(fused_quant_launch(2, qmode, x_dtype, a, static_cast<cudaStream_t>(stream)), "rmsnorm + quant launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_quant_fp8_group128_v2(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                             int64_t s_row_stride, int64_t s_grp_stride, int64_t M, int K, b2_stream_t stream) {
  const char* op = "sgl_per_token_group_quant_8bit_v2";
  if (int r = check_fused_common(op, x, x_dtype, ldx, K, q, ldq, s, B2_QMODE_FP8_GROUP128_V2, M, K)) return r;
  if (M == 0) return B2_OK;
  FusedQuantArgs a{};
  a.x = x; a.ldx = ldx; a.q = q; a.ldq = ldq; a.s = s; a.s_row_stride = s_row_stride; a.s_grp_stride = s_grp_stride;
  a.tokens_per_expert = 1; a.M = M; a.K = K; a.q_eps = 1e-10f; a.fmin = -448.f; a.fmax = 448.f;
  if (int r = check_cuda(fused_quant_launch(0, B2_QMODE_FP8_GROUP128_V2, x_dtype, a, static_cast<cudaStream_t>(stream)),
                         "sgl_per_token_group_quant_8bit_v2 launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_gemm_s8(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As, const float* Bs,
               const void* bias, int bias_dtype, void* D, int64_t ldd, int out_dtype, int M, int N, int K,
               b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16 && out_dtype != B2_DTYPE_S32)
    return fail(B2_ERR_ARG, "int8_scaled_mm: out_dtype must be Half or BFloat16");
  const int ob = out_dtype == B2_DTYPE_S32 ? 4 : 2;
  if (int r = check_gemm_common(A, lda, W, ldw, D, ldd, ob, M, N, K)) return r;
  if (out_dtype != B2_DTYPE_S32) {
    if (!As || !Bs) return fail(B2_ERR_ARG, "int8_scaled_mm: scales_a / scales_b must not be null");
    if (!aligned16(Bs)) return fail(B2_ERR_ARG, "int8_scaled_mm: scales_b must be 16-byte aligned");
    if (bias && !aligned16(bias)) return fail(B2_ERR_ARG, "int8_scaled_mm: bias must be 16-byte aligned");
    if (bias && bias_dtype != B2_DTYPE_BF16 && bias_dtype != B2_DTYPE_F16 && bias_dtype != B2_DTYPE_F32)
      return fail(B2_ERR_ARG, "int8_scaled_mm: unsupported bias dtype");
  }
  if (M == 0) return B2_OK;
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = As; a.Bs = Bs;
  a.bias = out_dtype == B2_DTYPE_S32 ? nullptr : bias;
  a.bias_dtype = bias_dtype; a.D = D; a.ldd = ldd; a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K;
  char err[256];
  int r = gemm_launch(GEMM_S8, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "int8_scaled_mm: %s", err);
  return B2_OK;
}

int b2_gemm_e4m3(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* As, const float* Bs,
                 const void* bias, int bias_dtype, void* D, int64_t ldd, int out_dtype, int M, int N, int K,
                 b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16 && out_dtype != B2_DTYPE_F32)
    return fail(B2_ERR_ARG, "fp8_scaled_mm: out_dtype must be Half or BFloat16");
  const int ob = out_dtype == B2_DTYPE_F32 ? 4 : 2;
  if (int r = check_gemm_common(A, lda, W, ldw, D, ldd, ob, M, N, K)) return r;
  if (!As || !Bs) return fail(B2_ERR_ARG, "fp8_scaled_mm: scales_a / scales_b must not be null");
  if (!aligned16(Bs)) return fail(B2_ERR_ARG, "fp8_scaled_mm: scales_b must be 16-byte aligned");
  if (bias && !aligned16(bias)) return fail(B2_ERR_ARG, "fp8_scaled_mm: bias must be 16-byte aligned");
  if (bias && bias_dtype != B2_DTYPE_BF16 && bias_dtype != B2_DTYPE_F16 && bias_dtype != B2_DTYPE_F32)
    return fail(B2_ERR_ARG, "fp8_scaled_mm: unsupported bias dtype");
  if (M == 0) return B2_OK;
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = As; a.Bs = Bs; a.bias = bias; a.bias_dtype = bias_dtype;
  a.D = D; a.ldd = ldd; a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K;
  char err[256];
  int r = gemm_launch(GEMM_E4M3, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "fp8_scaled_mm: %s", err);
  return B2_OK;
}

extern "C" int b2_symm_gather_counter_word(void);

static int gemm_push_symm(GemmFlavor flavor, const char* op, const void* A, int64_t lda, const void* W, int64_t ldw,
                          const float* As, const float* Bs, const void* bias, int bias_dtype, void* multicast_ptr,
                          uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world, int out_dtype,
                          int M, int N, int K, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16) return fail(B2_ERR_ARG, "%s: out_dtype must be Half or BFloat16", op);
  if (!multicast_ptr || !local_signal_pad || world < 2 || world > 8 || rank < 0 || rank >= world)
    return fail(B2_ERR_ARG, "%s: multicast pointer / signal pad / rank / world (2..8) invalid", op);
  if (M > 128) return fail(B2_ERR_ARG, "%s: the push epilogue is built for decode shapes (M <= 128)", op);
  if (int r = check_gemm_common(A, lda, W, ldw, multicast_ptr, N, 2, M, N, K)) return r;
  if (!As || !Bs || !aligned16(Bs)) return fail(B2_ERR_ARG, "%s: scales must not be null (Bs 16-byte aligned)", op);
  if (bias && !aligned16(bias)) return fail(B2_ERR_ARG, "%s: bias must be 16-byte aligned", op);
  if ((static_cast<int64_t>(M) * N) % 8 != 0) return fail(B2_ERR_ARG, "%s: M*N must be a multiple of 8", op);
  if (M == 0) return B2_OK;
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = As; a.Bs = Bs; a.bias = bias; a.bias_dtype = bias_dtype;
  a.D = multicast_ptr; a.ldd = N; a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K;
  const long long slot = static_cast<long long>(M) * N * 2;
  a.mc_slot = static_cast<uint8_t*>(multicast_ptr) + slot * rank;
  if (set_stride_bytes == 0) set_stride_bytes = slot * world;
  if (set_stride_bytes % 16 != 0 || set_stride_bytes < slot * world)
    return fail(B2_ERR_ARG, "%s: set stride must be a multiple of 16 and hold world slots", op);
  a.mc_set_bytes = set_stride_bytes;
  a.mc_counter = local_signal_pad + b2_symm_gather_counter_word();
  char err[256];
  int r = gemm_launch(flavor, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "%s: %s", op, err);
  return B2_OK;
}

int b2_gemm_s8_push_symm(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As, const float* Bs,
                         const void* bias, int bias_dtype, void* multicast_ptr, uint32_t* local_signal_pad,
                         int64_t set_stride_bytes, int rank, int world, int out_dtype, int M, int N, int K,
                         b2_stream_t stream) {
  return gemm_push_symm(GEMM_S8, "int8_scaled_mm (push)", A, lda, W, ldw, As, Bs, bias, bias_dtype, multicast_ptr,
                        local_signal_pad, set_stride_bytes, rank, world, out_dtype, M, N, K, stream);
}

int b2_gemm_e4m3_push_symm(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* As,
                           const float* Bs, const void* bias, int bias_dtype, void* multicast_ptr,
                           uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world, int out_dtype,
                           int M, int N, int K, b2_stream_t stream) {
  return gemm_push_symm(GEMM_E4M3, "fp8_scaled_mm (push)", A, lda, W, ldw, As, Bs, bias, bias_dtype, multicast_ptr,
                        local_signal_pad, set_stride_bytes, rank, world, out_dtype, M, N, K, stream);
}

static int gemm_scatter_symm(GemmFlavor flavor, const char* op, const void* A, int64_t lda, const void* W, int64_t ldw,
                             const float* As, const float* Bs, const void* bias, int bias_dtype,
                             void* const* peer_bufs_host, uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank,
                             int world, int out_dtype, int M, int N, int K, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16) return fail(B2_ERR_ARG, "%s: out_dtype must be Half or BFloat16", op);
  if (!peer_bufs_host || !local_signal_pad || world < 2 || world > 8 || rank < 0 || rank >= world)
    return fail(B2_ERR_ARG, "%s: peer buffers / signal pad / rank / world (2..8) invalid", op);
  for (int r = 0; r < world; ++r)
    if (!peer_bufs_host[r] || !aligned16(peer_bufs_host[r])) return fail(B2_ERR_ARG, "%s: gather buffer of rank %d is null or misaligned", op, r);
  if (M <= 0 || M % (32 * world) != 0) return fail(B2_ERR_ARG, "%s: M=%d must be a positive multiple of 32 * world", op, M);
  if (int r = check_gemm_common(A, lda, W, ldw, peer_bufs_host[rank], N, 2, M, N, K)) return r;
  if (!As || !Bs || !aligned16(Bs)) return fail(B2_ERR_ARG, "%s: scales must not be null (Bs 16-byte aligned)", op);
  if (bias && !aligned16(bias)) return fail(B2_ERR_ARG, "%s: bias must be 16-byte aligned", op);
  if (bias && bias_dtype != B2_DTYPE_BF16 && bias_dtype != B2_DTYPE_F16 && bias_dtype != B2_DTYPE_F32)
    return fail(B2_ERR_ARG, "%s: unsupported bias dtype", op);
  if (N % 8 != 0) return fail(B2_ERR_ARG, "%s: N=%d must be a multiple of 8", op, N);
  const long long slot = static_cast<long long>(M / world) * N * 2;
  if (set_stride_bytes == 0) set_stride_bytes = slot * world;
  if (set_stride_bytes % 16 != 0 || set_stride_bytes < slot * world)
    return fail(B2_ERR_ARG, "%s: set stride must be a multiple of 16 and hold world slots", op);
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = As; a.Bs = Bs; a.bias = bias; a.bias_dtype = bias_dtype;
  a.D = peer_bufs_host[rank]; a.ldd = N; a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K;
  a.rs_peer_bufs = peer_bufs_host;
  a.rs_world = world;
  a.rs_rank = rank;
  a.rs_set_bytes = set_stride_bytes;
  a.rs_counter = local_signal_pad + b2_symm_gather_counter_word();
  char err[256];
  int r = gemm_launch(flavor, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "%s: %s", op, err);
  return B2_OK;
}

int b2_gemm_s8_scatter_symm(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As,
                            const float* Bs, const void* bias, int bias_dtype, void* const* peer_bufs_host,
                            uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world, int out_dtype,
                            int M, int N, int K, b2_stream_t stream) {
  return gemm_scatter_symm(GEMM_S8, "int8_scaled_mm (scatter)", A, lda, W, ldw, As, Bs, bias, bias_dtype, peer_bufs_host,
                           local_signal_pad, set_stride_bytes, rank, world, out_dtype, M, N, K, stream);
}

int b2_gemm_e4m3_scatter_symm(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* As,
                              const float* Bs, const void* bias, int bias_dtype, void* const* peer_bufs_host,
                              uint32_t* local_signal_pad, int64_t set_stride_bytes, int rank, int world,
                              int out_dtype, int M, int N, int K, b2_stream_t stream) {
  return gemm_scatter_symm(GEMM_E4M3, "fp8_scaled_mm (scatter)", A, lda, W, ldw, As, Bs, bias, bias_dtype,
                           peer_bufs_host, local_signal_pad, set_stride_bytes, rank, world, out_dtype, M, N, K, stream);
}

int b2_gemm_e4m3_block128(const uint8_t* A, int64_t lda, const uint8_t* W, int64_t ldw, const float* sfa,
                          int64_t sfa_m_stride, int64_t sfa_k_stride, const float* sfb, int64_t sfb_n_stride,
                          int64_t sfb_k_stride, const void* bias, int bias_dtype, void* D, int64_t ldd,
                          int out_dtype, int M, int N, int K, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16)  // fp8_blockwise_gemm_kernel.cu:405-406
    return fail(B2_ERR_ARG, "fp8_blockwise_scaled_mm: out_dtype must be Half or BFloat16");
  const int ob = 2;
  if (int r = check_gemm_common(A, lda, W, ldw, D, ldd, ob, M, N, K)) return r;
  if (K % 128 != 0) return fail(B2_ERR_ARG, "fp8_blockwise_scaled_mm: K=%d must be a multiple of 128", K);
  if (N % 128 != 0) return fail(B2_ERR_ARG, "fp8_blockwise_scaled_mm: N=%d must be a multiple of 128", N);
  if (!sfa || !sfb) return fail(B2_ERR_ARG, "fp8_blockwise_scaled_mm: scales must not be null");
  if (bias && !aligned16(bias)) return fail(B2_ERR_ARG, "fp8_blockwise_scaled_mm: bias must be 16-byte aligned");
  if (M == 0) return B2_OK;
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = sfa; a.Bs = sfb; a.bias = bias; a.bias_dtype = bias_dtype;
  a.D = D; a.ldd = ldd; a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K;
  a.sfa_m_stride = sfa_m_stride; a.sfa_k_stride = sfa_k_stride;
  a.sfb_n_stride = sfb_n_stride; a.sfb_k_stride = sfb_k_stride;
  char err[256];
  int r = gemm_launch(GEMM_E4M3_BLOCK128, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err,
                      sizeof(err));
  if (r) return fail(r, "fp8_blockwise_scaled_mm: %s", err);
  return B2_OK;
}

int64_t b2_moe_max_padded(int64_t n, int num_experts) {
  // every expert wastes at most 127 rows; rounded up to whole blocks
  const int64_t worst = n + static_cast<int64_t>(num_experts) * 127;
  return (worst + 127) / 128 * 128;
}

int b2_moe_align_block128(const int32_t* topk_ids, int n, int num_experts, int32_t* sorted_ids, int32_t* tile_expert,
                          int32_t* inv_pos, int32_t* num_post_pad, int64_t max_padded, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (n < 0 || num_experts <= 0 || num_experts > 4096 || !topk_ids || !sorted_ids || !tile_expert || !inv_pos ||
      !num_post_pad)
    return fail(B2_ERR_ARG, "moe_align_block_size: bad arguments (n=%d experts=%d)", n, num_experts);
  if (max_padded % 128 != 0 || max_padded < b2_moe_max_padded(n, num_experts) || max_padded > 0x7fffffffll)
    return fail(B2_ERR_ARG, "moe_align_block_size: max_padded must be a multiple of 128 and >= %lld",
                (long long)b2_moe_max_padded(n, num_experts));
  if (int r = check_cuda(moe_align128(topk_ids, n, num_experts, sorted_ids, tile_expert, inv_pos, num_post_pad,
                                      static_cast<int>(max_padded), static_cast<cudaStream_t>(stream)),
                         "moe_align_block_size launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_quant_int8_rows_gather(const void* x, int x_dtype, int64_t ldx, const int32_t* row_map, int row_div,
                              int row_limit, int8_t* q, int64_t ldq, float* s, int64_t P, int K, b2_stream_t stream) {
  const char* op = "per_token_quant_int8 (gather)";
  if (int r = check_fused_common(op, x, x_dtype, ldx, K, q, ldq, s, B2_QMODE_INT8_ROWS, P, K)) return r;
  if (!row_map || row_div <= 0) return fail(B2_ERR_ARG, "%s: row_map / row_div missing", op);
  if (P == 0) return B2_OK;
  FusedQuantArgs a{};
  a.x = x; a.ldx = ldx; a.q = reinterpret_cast<uint8_t*>(q); a.ldq = ldq; a.s = s; a.s_row_stride = 1;
  a.tokens_per_expert = 1; a.M = P; a.K = K; a.q_eps = 1e-10f; a.fmin = -448.f; a.fmax = 448.f;
  a.row_map = row_map; a.row_div = row_div; a.row_limit = row_limit;
  if (int r = check_cuda(fused_quant_launch(0, B2_QMODE_INT8_ROWS, x_dtype, a, static_cast<cudaStream_t>(stream)),
                         "per_token_quant_int8 (gather) launch"))
    return r;
  count_launch();
  return B2_OK;
}

int b2_gemm_s8_grouped(const int8_t* A, int64_t lda, const int8_t* W, int64_t ldw, const float* As, const float* Bs,
                       const int32_t* tile_expert, int num_experts, void* D, int64_t ldd, int out_dtype, int64_t P,
                       int N, int K, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16 && out_dtype != B2_DTYPE_S32)
    return fail(B2_ERR_ARG, "grouped int8 gemm: out_dtype must be Half or BFloat16");
  if (P < 0 || P > 0x7fffffffll || P % 128 != 0)
    return fail(B2_ERR_ARG, "grouped int8 gemm: the row count must be a multiple of 128 (expert-padded layout)");
  if (!tile_expert || num_experts <= 0) return fail(B2_ERR_ARG, "grouped int8 gemm: tile_expert / num_experts missing");
  const int ob = out_dtype == B2_DTYPE_S32 ? 4 : 2;
  if (int r = check_gemm_common(A, lda, W, ldw, D, ldd, ob, static_cast<int>(P), N, K)) return r;
  if (out_dtype != B2_DTYPE_S32 && (!As || !Bs || !aligned16(Bs)))
    return fail(B2_ERR_ARG, "grouped int8 gemm: scales must not be null (Bs 16-byte aligned)");
  if (P == 0) return B2_OK;
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = As; a.Bs = Bs; a.bias = nullptr; a.bias_dtype = B2_DTYPE_F32;
  a.D = D; a.ldd = ldd; a.out_dtype = out_dtype; a.M = static_cast<int>(P); a.N = N; a.K = K;
  a.tile_expert = tile_expert; a.num_experts = num_experts;
  char err[256];
  int r = gemm_launch(GEMM_S8, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "grouped int8 gemm: %s", err);
  return B2_OK;
}

int b2_moe_combine(const void* y, int64_t ldy, const int32_t* inv_pos, const float* topk_weights, void* out,
                   int64_t ldo, int tokens, int topk, int H, int dtype, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(dtype, "moe combine: y")) return r;
  if (tokens < 0 || topk <= 0 || H <= 0 || H % 8 != 0 || !y || !inv_pos || !topk_weights || !out || ldy < H || ldo < H ||
      ldy % 8 != 0 || ldo % 8 != 0 || !aligned16(y) || !aligned16(out))
    return fail(B2_ERR_ARG, "moe combine: bad arguments (tokens=%d topk=%d H=%d)", tokens, topk, H);
  if (tokens == 0) return B2_OK;
  if (int r = check_cuda(moe_combine(y, ldy, inv_pos, topk_weights, out, ldo, tokens, topk, H, dtype,
                                     static_cast<cudaStream_t>(stream)),
                         "moe combine launch"))
    return r;
  count_launch();
  return B2_OK;
}

size_t b2_bmm_e4m3_workspace(int, int) { return 0; }

int b2_bmm_e4m3(const uint8_t* A, int64_t lda, int64_t a_batch_stride, const uint8_t* W, int64_t ldw,
                int64_t w_batch_stride, const float* a_scale, const float* b_scale, void* D, int64_t ldd,
                int64_t d_batch_stride, int out_dtype, int batch, int M, int N, int K, void* workspace,
                size_t workspace_bytes, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (batch < 0 || !a_scale || !b_scale) return fail(B2_ERR_ARG, "bmm_fp8: bad arguments");
  if (batch == 0 || M == 0) return B2_OK;
  (void)workspace;
  (void)workspace_bytes;  // kept in the signature (the reference op passes a cuBLASLt workspace); unused
  if (a_batch_stride % 16 != 0 || w_batch_stride % 16 != 0 || (d_batch_stride * 2) % 16 != 0)
    return fail(B2_ERR_ARG, "bmm_fp8: batch strides must keep every matrix 16-byte aligned");
  if (out_dtype != B2_DTYPE_BF16 && out_dtype != B2_DTYPE_F16 && out_dtype != B2_DTYPE_F32)
    return fail(B2_ERR_ARG, "bmm_fp8: out_dtype must be Half or BFloat16");
  const int ob = out_dtype == B2_DTYPE_F32 ? 4 : 2;
  if (int r = check_gemm_common(A, lda, W, ldw, D, ldd, ob, M, N, K)) return r;
  // ONE launch: the persistent GEMM in batched mode (3-D tensor maps, tiles enumerate batch-major)
  GemmArgs a{};
  a.A = A; a.lda = lda; a.W = W; a.ldw = ldw; a.As = a_scale; a.Bs = b_scale; a.bias = nullptr;
  a.bias_dtype = B2_DTYPE_BF16;
  a.D = D; a.ldd = ldd; a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K;
  a.scalar_scales = 1;  // the epilogue reads the two per-tensor scales directly: no broadcast launch, no workspace
  a.batch = batch; a.a_batch_stride = a_batch_stride; a.w_batch_stride = w_batch_stride; a.d_batch_stride = d_batch_stride;
  char err[256];
  int r = gemm_launch(GEMM_E4M3, a, g_forced_config & 0xff, &g_last_config,
                      static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "bmm_fp8: %s", err);
  return B2_OK;
}

static size_t ws_bytes(int M, int K) {
  // [ q : M x K bytes, padded to 256 ][ scales : M floats ]
  size_t q = (static_cast<size_t>(M) * static_cast<size_t>(K) + 255) & ~static_cast<size_t>(255);
  return q + static_cast<size_t>(M) * sizeof(float) + 256;
}

size_t b2_linear_w8a8_int8_workspace(int M, int K) { return ws_bytes(M, K); }
size_t b2_linear_w8a8_fp8_workspace(int M, int K) { return ws_bytes(M, K); }

// M <= 32: ONE kernel -- the split-K decode GEMM quantizes its own K slice of the activations into shared memory
// (gemm_splitk_sm100.cu, FUSEQ); codes and scales are bit-identical to the two-launch form, the workspace is unused.
// OFF by default: measured slower than the two programmatically chained launches on every decode shape
// (profiles/r02k_decode_fused_ab.log: M = 16 gate_up 35.8 vs 24.1 us, qkv 14.2 vs 9.0 us, even M = 1 10.0 vs 8.4 us) --
// the separate quantizer already runs under the GEMM's first weight round trip, while the fused form serialises two
// L2 passes over the activations in front of the first MMA in EVERY CTA and gives up ring depth for the resident A.
// b2_set_fused_decode(1) (or B2_FUSED_DECODE=1 in the environment) turns it on.
static thread_local int g_fused_decode = [] {
  const char* e = getenv("B2_FUSED_DECODE");
  return (e && e[0] == '1') ? 1 : 0;
}();

static bool fused_decode_ok(const void* x, int64_t ldx, int M, int K) {
  const bool enabled = g_fused_decode != 0;
  return enabled && M <= 32 && K % 16 == 0 && ldx % 8 == 0 && aligned16(x) &&
         ((g_forced_config & 0x80000) != 0 || (g_forced_config & 0xff) == 0);
}

static int linear_fused_decode(GemmFlavor flavor, const char* op, const void* x, int x_dtype, int64_t ldx, const void* W,
                               int64_t ldw, const float* Bs, const void* bias, int bias_dtype, void* D, int64_t ldd,
                               int M, int N, int K, b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (M < 0 || N <= 0 || K <= 0 || !x || !W || !D || !Bs) return fail(B2_ERR_ARG, "%s: bad arguments", op);
  if (N % 8 != 0) return fail(B2_ERR_ARG, "%s: out must be multiple of 16 bytes for memory alignment (N=%d must be a multiple of 8)", op, N);
  if (ldx < K || ldw < K || ldd < N || ldw % 16 != 0 || (ldd * 2) % 16 != 0 || !aligned16(W) || !aligned16(D) || !aligned16(Bs))
    return fail(B2_ERR_ARG, "%s: strides / alignment", op);
  if (bias && !aligned16(bias)) return fail(B2_ERR_ARG, "%s: bias must be 16-byte aligned", op);
  GemmArgs a{};
  a.A = nullptr; a.lda = K; a.W = W; a.ldw = ldw; a.As = nullptr; a.Bs = Bs; a.bias = bias; a.bias_dtype = bias_dtype;
  a.D = D; a.ldd = ldd; a.out_dtype = x_dtype; a.M = M; a.N = N; a.K = K;
  a.x = x; a.ldx = ldx; a.x_dtype = x_dtype;
  char err[256];
  int r = gemm_launch(flavor, a, g_forced_config, &g_last_config, static_cast<cudaStream_t>(stream), err, sizeof(err));
  if (r) return fail(r, "%s: %s", op, err);
  return B2_OK;
}

void b2_set_fused_decode(int enabled) { g_fused_decode = enabled; }
int b2_get_fused_decode(void) { return g_fused_decode; }

int b2_linear_w8a8_int8(const void* x, int x_dtype, int64_t ldx, const int8_t* W, int64_t ldw, const float* Bs,
                        const void* bias, int bias_dtype, void* D, int64_t ldd, int M, int N, int K,
                        void* workspace, size_t workspace_bytes, b2_stream_t stream) {
  if (int r = check_act_dtype(x_dtype, "int8_scaled_mm_with_quant: mat1")) return r;
  if (M == 0) return B2_OK;
  if (fused_decode_ok(x, ldx, M, K))
    return linear_fused_decode(GEMM_S8, "int8_scaled_mm_with_quant", x, x_dtype, ldx, W, ldw, Bs, bias, bias_dtype, D, ldd,
                               M, N, K, stream);
  if (!workspace || workspace_bytes < ws_bytes(M, K) || !aligned16(workspace))
    return fail(B2_ERR_ARG, "int8_scaled_mm_with_quant: workspace too small or misaligned (need %zu bytes)", ws_bytes(M, K));
  int8_t* q = static_cast<int8_t*>(workspace);
  float* s = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) +
                                      ((static_cast<size_t>(M) * K + 255) & ~static_cast<size_t>(255)));
  if (int r = b2_quant_int8_rows(x, x_dtype, ldx, q, K, s, M, K, stream)) return r;
  // reference: out_dtype must equal the input dtype (gemm_int8.cpp:509)
  return b2_gemm_s8(q, K, W, ldw, s, Bs, bias, bias_dtype, D, ldd, x_dtype, M, N, K, stream);
}

int b2_linear_w8a8_fp8(const void* x, int x_dtype, int64_t ldx, const uint8_t* W, int64_t ldw, const float* Bs,
                       const void* bias, int bias_dtype, void* D, int64_t ldd, int M, int N, int K,
                       void* workspace, size_t workspace_bytes, b2_stream_t stream) {
  if (int r = check_act_dtype(x_dtype, "w8a8_fp8 linear: input")) return r;
  if (M == 0) return B2_OK;
  if (fused_decode_ok(x, ldx, M, K))
    return linear_fused_decode(GEMM_E4M3, "w8a8_fp8 linear", x, x_dtype, ldx, W, ldw, Bs, bias, bias_dtype, D, ldd, M, N,
                               K, stream);
  if (!workspace || workspace_bytes < ws_bytes(M, K) || !aligned16(workspace))
    return fail(B2_ERR_ARG, "w8a8_fp8 linear: workspace too small or misaligned (need %zu bytes)", ws_bytes(M, K));
  uint8_t* q = static_cast<uint8_t*>(workspace);
  float* s = reinterpret_cast<float*>(q + ((static_cast<size_t>(M) * K + 255) & ~static_cast<size_t>(255)));
  if (int r = b2_quant_fp8_rows(x, x_dtype, ldx, q, K, s, M, K, stream)) return r;
  return b2_gemm_e4m3(q, K, W, ldw, s, Bs, bias, bias_dtype, D, ldd, x_dtype, M, N, K, stream);
}

// ---- b2_linear_w8a8_int8_static: the fused int8 linear without any allocation by anybody ----
// The scratch of b2_linear_w8a8_int8 (int8 codes + row scales between the quantizer and the GEMM) lives in a file-scope
// device array instead of a caller-provided workspace.  Every device that loads the library holds the array until the
// process exits.  `g_static_scratch` is a host-side stand-in for the symbol here in host code, NOT a device address:
// the address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[B2_STATIC_SCRATCH_BYTES];

static constexpr int kMaxStaticDevices = 64;
static void* g_static_scratch_ptr[kMaxStaticDevices];  // per device, filled on first use
static std::mutex g_static_scratch_mutex;

static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(B2_ERR_CUDA, "int8_scaled_mm_with_quant (static): cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxStaticDevices)
    return fail(B2_ERR_ARG, "int8_scaled_mm_with_quant (static): device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return fail(B2_ERR_CUDA, "int8_scaled_mm_with_quant (static): cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    g_static_scratch_ptr[dev] = ptr;
  }
  *out = g_static_scratch_ptr[dev];
  return B2_OK;
}

size_t b2_linear_w8a8_int8_static_max_bytes(void) { return B2_STATIC_SCRATCH_BYTES; }

int b2_linear_w8a8_int8_static(const void* x, int x_dtype, int64_t ldx, const int8_t* W, int64_t ldw, const float* Bs,
                               const void* bias, int bias_dtype, void* D, int64_t ldd, int M, int N, int K,
                               b2_stream_t stream) {
  if (int r = check_arch()) return r;
  if (int r = check_act_dtype(x_dtype, "int8_scaled_mm_with_quant (static): mat1")) return r;
  if (M < 0 || N <= 0 || K <= 0) return fail(B2_ERR_ARG, "int8_scaled_mm_with_quant (static): bad sizes M=%d N=%d K=%d", M, N, K);
  if (M == 0) return B2_OK;
  const size_t need = ws_bytes(M, K);
  if (need > static_cast<size_t>(B2_STATIC_SCRATCH_BYTES))
    return fail(B2_ERR_ARG,
                "int8_scaled_mm_with_quant (static): M=%d x K=%d needs %zu bytes of scratch, the built-in array holds %zu "
                "(B2_STATIC_SCRATCH_BYTES); use b2_linear_w8a8_int8 with a caller workspace",
                M, K, need, static_cast<size_t>(B2_STATIC_SCRATCH_BYTES));
  void* scratch = nullptr;
  if (int r = static_scratch_for_current_device(&scratch)) return r;
  // Nothing in the scratch has to start at zero: the quantizer writes all M x K codes and all M scales before the GEMM
  // (next in the same stream) reads exactly those; rows past M of the GEMM's activation box are zero-filled by TMA
  // from the tensor map's extent, not read from the array.
  int8_t* q = static_cast<int8_t*>(scratch);
  float* s = reinterpret_cast<float*>(static_cast<uint8_t*>(scratch) +
                                      ((static_cast<size_t>(M) * K + 255) & ~static_cast<size_t>(255)));
  if (int r = b2_quant_int8_rows(x, x_dtype, ldx, q, K, s, M, K, stream)) return r;
  return b2_gemm_s8(q, K, W, ldw, s, Bs, bias, bias_dtype, D, ldd, x_dtype, M, N, K, stream);
}


}  // extern "C"
