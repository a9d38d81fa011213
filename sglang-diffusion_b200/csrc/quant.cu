// Dynamic activation quantizers for the W8A8 linear path (sm_100a).
//
//   quant_rows<INT8>  : per-token symmetric int8, CPU-path arithmetic of the reference
//                       (sgl-kernel/csrc/cpu/vec.h:190-228; driver gemm_int8.cpp:382-410):
//                       amax = max(max|x|, 1e-7); s = amax/127; q = RNE(x * (127/amax)).
//                       Emits signed int8 q (= the reference's u8 code - 128).
//   quant_rows<FP8>   : per-token e4m3 (sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:17-121):
//                       s = max|x|/448; q = e4m3_rn_sat(clamp(x * (s==0 ? 0 : 1/s))).
//   quant_group128    : 1xG group e4m3 with fp32 scales, row- or column-major
//                       (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:19-114):
//                       s = max(eps, max|x|)/fp8_max; q = e4m3(clamp(x / s)).
//
// All divisions are IEEE (__fdiv_rn / __frcp_rn): torch-style reciprocal-multiply
// does not reproduce the reference codes.  The kernels are pure
// streaming work (3 bytes moved per element): one row is owned by TPR threads,
// the row is read ONCE with 16-byte loads and kept in registers between the
// absmax pass and the quantize pass, reductions are warp shuffles (+ one smem hop
// when a row spans warps), stores are 8-byte packed codes.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include <cstdlib>
#include <type_traits>

#include "b200_w8a8_internal.h"
#include "ptx.cuh"
#include "quant_common.cuh"

namespace b2 {

// One row is processed by TPR threads; a CTA of BLOCK threads owns BLOCK/TPR rows.
// Each thread caches up to NCACHE vectors of VEC elements; a row longer than
// TPR*VEC*NCACHE re-reads the overflow from L2 in the second pass.
// SUM: also emit the fp32 sum of the (unquantized) row, stored in T -- per_token_quant_int8(cal_sum=True) of
// python/sglang/srt/layers/quantization/int8_kernel.py:46-56 (QoQ); the sum is free once the row is in registers.
template <typename T, QMode MODE, int VEC, int TPR, int BLOCK, int NCACHE, bool SUM = false>
__global__ void __launch_bounds__(BLOCK) quant_rows_kernel(const T* __restrict__ x, int64_t ldx,
                                                           uint8_t* __restrict__ q, int64_t ldq,
                                                           float* __restrict__ s, int M, int K,
                                                           T* __restrict__ row_sum) {
  using V = typename VecT<VEC * sizeof(T)>::type;
  using QV = typename VecT<VEC>::type;
  constexpr int ROWS = BLOCK / TPR;
  constexpr int WARPS_PER_ROW = (TPR + 31) / 32;
  __shared__ float s_red[ROWS][WARPS_PER_ROW > 1 ? WARPS_PER_ROW : 1];
  __shared__ float s_sum[SUM ? ROWS : 1][WARPS_PER_ROW > 1 ? WARPS_PER_ROW : 1];
  float rsum = 0.f;

  // Release programmatically-launched dependents at once (a no-op for ordinary successors): the decode GEMM that
  // follows starts fetching its weights while these rows are quantized, and waits for this grid's completion
  // (griddepcontrol.wait) before it reads the codes or the scales.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // ... and, when this kernel was itself launched programmatically, wait for its own predecessor before reading x or
  // writing q / s (a no-op under an ordinary launch)
  asm volatile("griddepcontrol.wait;" ::: "memory");

  const int r_in_cta = threadIdx.x / TPR;
  const int t = threadIdx.x % TPR;
  const int row = blockIdx.x * ROWS + r_in_cta;
  const bool active = row < M;
  const int nvec = K / VEC;

  const V* xr = reinterpret_cast<const V*>(x + static_cast<int64_t>(active ? row : 0) * ldx);
  V cache[NCACHE];
  float amax = 0.f;
  if (active) {
#pragma unroll
    for (int i = 0; i < NCACHE; ++i) {
      const int v = t + i * TPR;
      if (v < nvec) cache[i] = ld_stream(xr + v);
    }
#pragma unroll
    for (int i = 0; i < NCACHE; ++i) {
      const int v = t + i * TPR;
      if (v < nvec) {
        const T* e = reinterpret_cast<const T*>(&cache[i]);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          amax = fmaxf(amax, fabsf(to_f32(e[j])));
          if constexpr (SUM) rsum += to_f32(e[j]);
        }
      }
    }
    for (int v = t + NCACHE * TPR; v < nvec; v += TPR) {
      V tmp = ld_stream(xr + v);
      const T* e = reinterpret_cast<const T*>(&tmp);
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        amax = fmaxf(amax, fabsf(to_f32(e[j])));
        if constexpr (SUM) rsum += to_f32(e[j]);
      }
    }
  }
  // reduce over the TPR threads of the row
  constexpr int SHFL_W = TPR < 32 ? TPR : 32;
#pragma unroll
  for (int o = SHFL_W / 2; o > 0; o >>= 1) {
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if constexpr (SUM) rsum += __shfl_xor_sync(0xffffffffu, rsum, o);
  }
  if constexpr (WARPS_PER_ROW > 1) {
    const int w = t / 32;
    if ((t & 31) == 0) {
      s_red[r_in_cta][w] = amax;
      if constexpr (SUM) s_sum[r_in_cta][w] = rsum;
    }
    __syncthreads();
    if constexpr (SUM) {
      rsum = 0.f;
#pragma unroll
      for (int i = 0; i < WARPS_PER_ROW; ++i) rsum += s_sum[r_in_cta][i];
    }
    float m = s_red[r_in_cta][0];
#pragma unroll
    for (int i = 1; i < WARPS_PER_ROW; ++i) m = fmaxf(m, s_red[r_in_cta][i]);
    amax = m;
  }
  if (!active) return;

  float scale, inv;
  if constexpr (MODE == QMode::INT8) {
    amax = fmaxf(amax, 1e-7f);
    scale = __fdiv_rn(amax, 127.f);
    inv = __fdiv_rn(127.f, amax);
  } else {
    scale = __fdiv_rn(amax, 448.f);
    inv = (scale == 0.f) ? 0.f : __frcp_rn(scale);
  }
  if (t == 0) {
    s[row] = scale;
    if constexpr (SUM) {
      if constexpr (std::is_same<T, __nv_bfloat16>::value) row_sum[row] = __float2bfloat16_rn(rsum);
      else row_sum[row] = __float2half_rn(rsum);
    }
  }

  QV* qr = reinterpret_cast<QV*>(q + static_cast<int64_t>(row) * ldq);
#pragma unroll
  for (int i = 0; i < NCACHE; ++i) {
    const int v = t + i * TPR;
    if (v < nvec) {
      const T* e = reinterpret_cast<const T*>(&cache[i]);
      QV out;
      uint8_t* o = reinterpret_cast<uint8_t*>(&out);
      if constexpr (MODE == QMode::FP8 && VEC % 2 == 0) {
        // paired conversion: one cvt.rn.satfinite.e4m3x2 per two elements (the clamp keeps the reference's
        // fminf/fmaxf semantics for non-finite inputs)
        uint16_t* o2 = reinterpret_cast<uint16_t*>(&out);
#pragma unroll
        for (int j = 0; j < VEC; j += 2) {
          const float a = fmaxf(fminf(__fmul_rn(to_f32(e[j]), inv), 448.f), -448.f);
          const float b = fmaxf(fminf(__fmul_rn(to_f32(e[j + 1]), inv), 448.f), -448.f);
          o2[j >> 1] = f32x2_to_e4m3x2(a, b);
        }
      } else {
#pragma unroll
        for (int j = 0; j < VEC; ++j) o[j] = quant_one<MODE>(to_f32(e[j]), inv);
      }
      qr[v] = out;
    }
  }
  for (int v = t + NCACHE * TPR; v < nvec; v += TPR) {
    V tmp = __ldg(xr + v);
    const T* e = reinterpret_cast<const T*>(&tmp);
    QV out;
    uint8_t* o = reinterpret_cast<uint8_t*>(&out);
#pragma unroll
    for (int j = 0; j < VEC; ++j) o[j] = quant_one<MODE>(to_f32(e[j]), inv);
    qr[v] = out;
  }
}

template <typename T, QMode MODE, int VEC, int TPR>
static cudaError_t launch_rows(const void* x, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M, int K,
                               cudaStream_t st) {
  constexpr int BLOCK = TPR > 256 ? TPR : 256;
  constexpr int NCACHE = 8;
  constexpr int ROWS = BLOCK / TPR;
  const int grid = (M + ROWS - 1) / ROWS;
  // Decode chains (b2_set_weight_prefetch): the quantizer itself is launched programmatically, so its launch latency
  // overlaps the tail of the kernel before it; it waits for that kernel (griddepcontrol.wait) before it touches memory.
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(BLOCK);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  static const bool chain_quantizer = [] {  // developer A/B switch: B2_QUANT_PDL=0 launches the quantizer plainly
    const char* e = getenv("B2_QUANT_PDL");
    return !(e && e[0] == '0');
  }();
  cfg.numAttrs = (chain_quantizer && weight_prefetch_enabled() && M <= 128) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, quant_rows_kernel<T, MODE, VEC, TPR, BLOCK, NCACHE>, static_cast<const T*>(x), ldx,
                            q, ldq, s, M, K, static_cast<T*>(nullptr));
}

template <typename T, QMode MODE, int VEC>
static cudaError_t dispatch_tpr(const void* x, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M, int K,
                                cudaStream_t st) {
  // smallest TPR that keeps the whole row in registers (8 vectors / thread) ...
  const int nvec = K / VEC;
  int tpr = 32;
  while (tpr < 1024 && nvec > tpr * 8) tpr *= 2;
  // ... widened when there are too few rows to fill the 148 SMs (decode shapes),
  // as long as every thread still has at least one vector.
  while (tpr < 1024 && static_cast<int64_t>(M) * tpr < 148 * 1024 && nvec >= tpr * 2) tpr *= 2;
  switch (tpr) {
    case 32: return launch_rows<T, MODE, VEC, 32>(x, ldx, q, ldq, s, M, K, st);
    case 64: return launch_rows<T, MODE, VEC, 64>(x, ldx, q, ldq, s, M, K, st);
    case 128: return launch_rows<T, MODE, VEC, 128>(x, ldx, q, ldq, s, M, K, st);
    case 256: return launch_rows<T, MODE, VEC, 256>(x, ldx, q, ldq, s, M, K, st);
    case 512: return launch_rows<T, MODE, VEC, 512>(x, ldx, q, ldq, s, M, K, st);
    default: return launch_rows<T, MODE, VEC, 1024>(x, ldx, q, ldq, s, M, K, st);
  }
}

template <typename T, QMode MODE>
static cudaError_t dispatch_vec(const void* x, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M, int K,
                                cudaStream_t st) {
  const uintptr_t xa = reinterpret_cast<uintptr_t>(x);
  const uintptr_t qa = reinterpret_cast<uintptr_t>(q);
  auto ok = [&](int vec) {
    return K % vec == 0 && ldx % vec == 0 && ldq % vec == 0 && xa % (vec * sizeof(T)) == 0 && qa % vec == 0;
  };
  if (ok(8)) return dispatch_tpr<T, MODE, 8>(x, ldx, q, ldq, s, M, K, st);
  if (ok(4)) return dispatch_tpr<T, MODE, 4>(x, ldx, q, ldq, s, M, K, st);
  if (ok(2)) return dispatch_tpr<T, MODE, 2>(x, ldx, q, ldq, s, M, K, st);
  return dispatch_tpr<T, MODE, 1>(x, ldx, q, ldq, s, M, K, st);
}

cudaError_t quant_rows_int8(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s, int M,
                            int K, cudaStream_t st) {
  uint8_t* qq = reinterpret_cast<uint8_t*>(q);
  if (x_dtype == B2_DTYPE_BF16) return dispatch_vec<__nv_bfloat16, QMode::INT8>(x, ldx, qq, ldq, s, M, K, st);
  return dispatch_vec<__half, QMode::INT8>(x, ldx, qq, ldq, s, M, K, st);
}

// per_token_quant_int8(cal_sum=True): same quantizer, plus the row sums in x's dtype.  K % 8 == 0 and 16-byte aligned rows.
template <typename T>
static cudaError_t launch_rows_sum(const void* x, int64_t ldx, uint8_t* q, int64_t ldq, float* s, void* row_sum, int M,
                                   int K, cudaStream_t st) {
  const int nvec = K / 8;
  auto go = [&](auto tpr_c) {
    constexpr int TPR = decltype(tpr_c)::value;
    constexpr int BLOCK = TPR > 256 ? TPR : 256;
    constexpr int ROWS = BLOCK / TPR;
    quant_rows_kernel<T, QMode::INT8, 8, TPR, BLOCK, 8, true><<<(M + ROWS - 1) / ROWS, BLOCK, 0, st>>>(
        static_cast<const T*>(x), ldx, q, ldq, s, M, K, static_cast<T*>(row_sum));
    return cudaGetLastError();
  };
  if (nvec <= 32 * 8) return go(std::integral_constant<int, 32>{});
  if (nvec <= 128 * 8) return go(std::integral_constant<int, 128>{});
  return go(std::integral_constant<int, 1024>{});
}

cudaError_t quant_rows_int8_sum(const void* x, int x_dtype, int64_t ldx, int8_t* q, int64_t ldq, float* s,
                                void* row_sum, int M, int K, cudaStream_t st) {
  uint8_t* qq = reinterpret_cast<uint8_t*>(q);
  if (x_dtype == B2_DTYPE_BF16) return launch_rows_sum<__nv_bfloat16>(x, ldx, qq, ldq, s, row_sum, M, K, st);
  return launch_rows_sum<__half>(x, ldx, qq, ldq, s, row_sum, M, K, st);
}

cudaError_t quant_rows_fp8(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M,
                           int K, cudaStream_t st) {
  if (x_dtype == B2_DTYPE_BF16) return dispatch_vec<__nv_bfloat16, QMode::FP8>(x, ldx, q, ldq, s, M, K, st);
  return dispatch_vec<__half, QMode::FP8>(x, ldx, q, ldq, s, M, K, st);
}

// ------------------------------------------------------------ 1xG groups ----
// G = 128: one group = 16 threads x 8 elements (one 16-byte load each).  A CTA covers a tile of 16 rows x 4
// adjacent groups (four independent 16-byte loads in flight per thread); amax is reduced with 4 xor-shuffles
// inside each half-warp.  The reference divides every element by the group scale
// (per_token_group_quant_8bit.cu:110); eight IEEE divisions per thread made the first version of this kernel
// ALU-bound (44 % of HBM).  Here the correctly rounded reciprocal r = RN(1/ys) is computed once and each
// quotient is q0 = RN(x r), q = RN(q0 + r (x - q0 ys)) -- Markstein's sequence, which returns RN(x / ys) whenever
// the significand of ys is not all ones (that case, group-uniform, takes the division).  tests/test_gpu_parity.py
// checks every (bf16 x significand, bf16 amax significand, exponent gap) combination against the oracle.
// Scales are staged in shared memory so that the M-major (column-major) scale matrix the block-wise GEMM wants is
// written as 64-byte runs of 16 consecutive rows instead of one 4-byte store per sector.
template <typename T>
B2_DEVICE float group128_quantize(const uint4& v, float eps, float fmin, float fmax, uint2& out) {
  const T* e = reinterpret_cast<const T*>(&v);
  float f[8];
  float amax = eps;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    f[j] = to_f32(e[j]);
    amax = fmaxf(amax, fabsf(f[j]));
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  return group128_quantize_v1(f, amax, fmin, fmax, out);
}

// GQ_GROUPS adjacent groups per CTA tile = independent 16-byte loads in flight per thread
template <typename T, int GQ_GROUPS>
__global__ void __launch_bounds__(256) quant_group128_kernel(const T* __restrict__ x, int64_t ldx,
                                                             uint8_t* __restrict__ q, int64_t ldq,
                                                             float* __restrict__ s, int64_t s_row_stride,
                                                             int64_t s_grp_stride, int M, int groups_per_row,
                                                             float eps, float fmin, float fmax) {
  // one CTA = one tile of 16 rows x GQ_GROUPS groups; the grid covers every tile (no grid-stride loop: with eight
  // CTAs resident per SM the block scheduler keeps ~128 KB of loads in flight per SM without software pipelining)
  __shared__ float s_scale[GQ_GROUPS][16];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // see quant_rows_kernel
  const int sub = threadIdx.x & 15;
  const int r_in = threadIdx.x >> 4;
  const int gtiles = (groups_per_row + GQ_GROUPS - 1) / GQ_GROUPS;
  const int64_t t = blockIdx.x;
  const int64_t rb = t / gtiles;
  const int g0 = static_cast<int>(t - rb * gtiles) * GQ_GROUPS;
  const int64_t row = rb * 16 + r_in;
  const T* xr = x + row * ldx + static_cast<int64_t>(g0) * 128;
  uint4 v[GQ_GROUPS];
  bool act[GQ_GROUPS];
#pragma unroll
  for (int i = 0; i < GQ_GROUPS; ++i) {
    act[i] = row < M && (g0 + i) < groups_per_row;
    v[i] = make_uint4(0, 0, 0, 0);
    if (act[i]) v[i] = ld_stream(reinterpret_cast<const uint4*>(xr + i * 128) + sub);
  }
  uint8_t* qr = q + row * ldq + static_cast<int64_t>(g0) * 128;
  // The IEEE division amax / fmax and the reciprocal of the scale are ~25 instructions each; every thread needs them
  // for GQ_GROUPS groups, and the 16 threads of a group would all compute the same pair.  Instead lane `sub` computes
  // the pair of group sub % GQ_GROUPS only and the 16-lane segment exchanges them with shuffles (profiles/
  // r02i_group_quant.md: issue_active 65 %, the kernel is latency / issue bound, not HBM bound).
  float f[GQ_GROUPS][8], amax[GQ_GROUPS];
#pragma unroll
  for (int i = 0; i < GQ_GROUPS; ++i) {
    const T* e = reinterpret_cast<const T*>(&v[i]);
    float m = eps;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      f[i][j] = to_f32(e[j]);
      m = fmaxf(m, fabsf(f[i][j]));
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    amax[i] = m;
  }
  float my_amax = amax[0];
#pragma unroll
  for (int i = 1; i < GQ_GROUPS; ++i)
    if ((sub & (GQ_GROUPS - 1)) == i) my_amax = amax[i];
  const float my_ys = __fdiv_rn(my_amax, fmax);
  const float my_r = __frcp_rn(my_ys);
  const int seg = threadIdx.x & 16;  // first lane of my 16-lane segment inside the warp
#pragma unroll
  for (int i = 0; i < GQ_GROUPS; ++i) {
    const float ys = __shfl_sync(0xffffffffu, my_ys, seg + i);
    const float r = __shfl_sync(0xffffffffu, my_r, seg + i);
    uint2 o;
    group128_apply_v1(f[i], ys, r, fmin, fmax, o);
    if (act[i]) reinterpret_cast<uint2*>(qr + i * 128)[sub] = o;
    if (sub == 0) s_scale[i][r_in] = ys;
  }
  __syncthreads();
  if (threadIdx.x < 16 * GQ_GROUPS) {
    const int gi = threadIdx.x >> 4, rr = threadIdx.x & 15;
    const int64_t row2 = rb * 16 + rr;
    if (row2 < M && g0 + gi < groups_per_row)
      s[row2 * s_row_stride + static_cast<int64_t>(g0 + gi) * s_grp_stride] = s_scale[gi][rr];
  }
}

template <typename T, int GQ>
static cudaError_t launch_group128(const void* x, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int64_t s_row_stride,
                                   int64_t s_grp_stride, int M, int gpr, float eps, float fmin, float fmax,
                                   cudaStream_t st) {
  const int64_t blocks = static_cast<int64_t>((M + 15) / 16) * ((gpr + GQ - 1) / GQ);
  if (blocks > 0x7fffffffll) return cudaErrorInvalidValue;
  quant_group128_kernel<T, GQ><<<static_cast<int>(blocks), 256, 0, st>>>(static_cast<const T*>(x), ldx, q, ldq, s,
                                                                         s_row_stride, s_grp_stride, M, gpr, eps,
                                                                         fmin, fmax);
  return cudaGetLastError();
}

cudaError_t quant_fp8_group128(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                               int64_t s_row_stride, int64_t s_grp_stride, int M, int K, float eps, float fmin,
                               float fmax, cudaStream_t st) {
  const int gpr = K / 128;
  if (static_cast<int64_t>(M) * gpr == 0) return cudaSuccess;
  static const int forced = [] {
    const char* e = getenv("B2_GQ_GROUPS");
    return e ? atoi(e) : 0;
  }();
  int gq = (gpr >= 4 && static_cast<int64_t>((M + 15) / 16) * (gpr / 4) >= 2 * 148) ? 4 : 2;  // measured (profiles/r02i_group_quant.md): 2 and 8 groups per thread are both slower
  if (forced == 2 || forced == 4 || forced == 8) gq = forced;
#define B2_GQ_CASE(T_)                                                                                             \
  switch (gq) {                                                                                                    \
    case 8: return launch_group128<T_, 8>(x, ldx, q, ldq, s, s_row_stride, s_grp_stride, M, gpr, eps, fmin, fmax, st); \
    case 4: return launch_group128<T_, 4>(x, ldx, q, ldq, s, s_row_stride, s_grp_stride, M, gpr, eps, fmin, fmax, st); \
    default: return launch_group128<T_, 2>(x, ldx, q, ldq, s, s_row_stride, s_grp_stride, M, gpr, eps, fmin, fmax, st); \
  }
  if (x_dtype == B2_DTYPE_BF16) {
    B2_GQ_CASE(__nv_bfloat16)
  }
  B2_GQ_CASE(__half)
#undef B2_GQ_CASE
}

}  // namespace b2
