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

#include "b200_w8a8_internal.h"
#include "ptx.cuh"

namespace b2 {

enum class QMode { INT8, FP8 };

template <typename T>
B2_DEVICE float to_f32(T v);
template <>
B2_DEVICE float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}
template <>
B2_DEVICE float to_f32<__half>(__half v) {
  return __half2float(v);
}

template <int BYTES>
struct VecT;
template <>
struct VecT<16> {
  using type = uint4;
};
template <>
struct VecT<8> {
  using type = uint2;
};
template <>
struct VecT<4> {
  using type = uint32_t;
};
template <>
struct VecT<2> {
  using type = uint16_t;
};
template <>
struct VecT<1> {
  using type = uint8_t;
};

// streaming load: read-only path, do not keep the line in L1
template <typename V>
B2_DEVICE V ld_stream(const V* p) {
  return __ldg(p);
}
template <>
B2_DEVICE uint4 ld_stream<uint4>(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

B2_DEVICE uint8_t f32_to_e4m3_sat(float v) {
  return static_cast<uint8_t>(__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3));
}

// pack two floats -> two e4m3 codes (low = a)
B2_DEVICE uint16_t f32x2_to_e4m3x2(float a, float b) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(r) : "f"(a), "f"(b));
  return r;
}

template <QMode MODE>
B2_DEVICE uint8_t quant_one(float x, float inv) {
  if constexpr (MODE == QMode::INT8) {
    return static_cast<uint8_t>(static_cast<int8_t>(__float2int_rn(__fmul_rn(x, inv))));
  } else {
    float v = __fmul_rn(x, inv);
    v = fmaxf(fminf(v, 448.f), -448.f);
    return f32_to_e4m3_sat(v);
  }
}

// One row is processed by TPR threads; a CTA of BLOCK threads owns BLOCK/TPR rows.
// Each thread caches up to NCACHE vectors of VEC elements; a row longer than
// TPR*VEC*NCACHE re-reads the overflow from L2 in the second pass.
template <typename T, QMode MODE, int VEC, int TPR, int BLOCK, int NCACHE>
__global__ void __launch_bounds__(BLOCK) quant_rows_kernel(const T* __restrict__ x, int64_t ldx,
                                                           uint8_t* __restrict__ q, int64_t ldq,
                                                           float* __restrict__ s, int M, int K) {
  using V = typename VecT<VEC * sizeof(T)>::type;
  using QV = typename VecT<VEC>::type;
  constexpr int ROWS = BLOCK / TPR;
  constexpr int WARPS_PER_ROW = (TPR + 31) / 32;
  __shared__ float s_red[ROWS][WARPS_PER_ROW > 1 ? WARPS_PER_ROW : 1];

  // Release programmatically-launched dependents at once (a no-op for ordinary successors): the decode GEMM that
  // follows starts fetching its weights while these rows are quantized, and waits for this grid's completion
  // (griddepcontrol.wait) before it reads the codes or the scales.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

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
        for (int j = 0; j < VEC; ++j) amax = fmaxf(amax, fabsf(to_f32(e[j])));
      }
    }
    for (int v = t + NCACHE * TPR; v < nvec; v += TPR) {
      V tmp = ld_stream(xr + v);
      const T* e = reinterpret_cast<const T*>(&tmp);
#pragma unroll
      for (int j = 0; j < VEC; ++j) amax = fmaxf(amax, fabsf(to_f32(e[j])));
    }
  }
  // reduce over the TPR threads of the row
  constexpr int SHFL_W = TPR < 32 ? TPR : 32;
#pragma unroll
  for (int o = SHFL_W / 2; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  if constexpr (WARPS_PER_ROW > 1) {
    const int w = t / 32;
    if ((t & 31) == 0) s_red[r_in_cta][w] = amax;
    __syncthreads();
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
  if (t == 0) s[row] = scale;

  QV* qr = reinterpret_cast<QV*>(q + static_cast<int64_t>(row) * ldq);
#pragma unroll
  for (int i = 0; i < NCACHE; ++i) {
    const int v = t + i * TPR;
    if (v < nvec) {
      const T* e = reinterpret_cast<const T*>(&cache[i]);
      QV out;
      uint8_t* o = reinterpret_cast<uint8_t*>(&out);
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = quant_one<MODE>(to_f32(e[j]), inv);
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
  quant_rows_kernel<T, MODE, VEC, TPR, BLOCK, NCACHE>
      <<<grid, BLOCK, 0, st>>>(static_cast<const T*>(x), ldx, q, ldq, s, M, K);
  return cudaGetLastError();
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

cudaError_t quant_rows_fp8(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s, int M,
                           int K, cudaStream_t st) {
  if (x_dtype == B2_DTYPE_BF16) return dispatch_vec<__nv_bfloat16, QMode::FP8>(x, ldx, q, ldq, s, M, K, st);
  return dispatch_vec<__half, QMode::FP8>(x, ldx, q, ldq, s, M, K, st);
}

// ------------------------------------------------------------ 1xG groups ----
// G = 128: one group = 16 threads x 8 elements (one 16-byte load each).  A warp
// covers two groups; amax is reduced with 4 xor-shuffles inside each half-warp.
// Grid-stride over (row, group) pairs so the grid is sized to the machine.
template <typename T>
__global__ void __launch_bounds__(256) quant_group128_kernel(const T* __restrict__ x, int64_t ldx,
                                                             uint8_t* __restrict__ q, int64_t ldq,
                                                             float* __restrict__ s, int64_t s_row_stride,
                                                             int64_t s_grp_stride, int M, int groups_per_row,
                                                             float eps, float fmin, float fmax) {
  const int64_t total = static_cast<int64_t>(M) * groups_per_row;
  const int sub = threadIdx.x & 15;
  for (int64_t g = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 4; g < ((total + 1) & ~1ll);
       g += (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 4) {
    const bool active = g < total;
    const int64_t row = active ? g / groups_per_row : 0;
    const int grp = active ? static_cast<int>(g % groups_per_row) : 0;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (active) v = ld_stream(reinterpret_cast<const uint4*>(x + row * ldx + grp * 128) + sub);
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
    const float ys = __fdiv_rn(amax, fmax);
    if (active) {
      uint2 out;
      uint8_t* o8 = reinterpret_cast<uint8_t*>(&out);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float qv = fminf(fmaxf(__fdiv_rn(f[j], ys), fmin), fmax);
        o8[j] = f32_to_e4m3_sat(qv);
      }
      reinterpret_cast<uint2*>(q + row * ldq + grp * 128)[sub] = out;
      if (sub == 0) s[row * s_row_stride + grp * s_grp_stride] = ys;
    }
  }
}

cudaError_t quant_fp8_group128(const void* x, int x_dtype, int64_t ldx, uint8_t* q, int64_t ldq, float* s,
                               int64_t s_row_stride, int64_t s_grp_stride, int M, int K, float eps, float fmin,
                               float fmax, cudaStream_t st) {
  const int gpr = K / 128;
  const int64_t total = static_cast<int64_t>(M) * gpr;
  if (total == 0) return cudaSuccess;
  const int64_t blocks_needed = (total * 16 + 255) / 256;
  const int grid = static_cast<int>(blocks_needed < 148 * 16 ? blocks_needed : 148 * 16);
  if (x_dtype == B2_DTYPE_BF16)
    quant_group128_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), ldx, q, ldq,
                                                               s, s_row_stride, s_grp_stride, M, gpr, eps, fmin,
                                                               fmax);
  else
    quant_group128_kernel<__half><<<grid, 256, 0, st>>>(static_cast<const __half*>(x), ldx, q, ldq, s,
                                                        s_row_stride, s_grp_stride, M, gpr, eps, fmin, fmax);
  return cudaGetLastError();
}

}  // namespace b2
