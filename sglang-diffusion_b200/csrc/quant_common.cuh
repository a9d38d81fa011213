// Device helpers shared by the activation quantizers (quant.cu, quant_fused.cu).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

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


// ---- 1x128 group quantization of the 8 values one thread holds (16 threads = one group) ----
// v1 arithmetic (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:19-114): s = amax / fmax, q = e4m3(clamp(x / s)).
// The reference divides every element by the group scale (:110); here the correctly rounded reciprocal r = RN(1/ys)
// is computed once and each quotient is q0 = RN(x r), q = RN(q0 + r (x - q0 ys)) -- Markstein's sequence, which
// returns RN(x / ys) whenever the significand of ys is not all ones (that case, group-uniform, takes the division).
// `amax` is already reduced over the 16 lanes and already >= eps.  Returns the scale.
// the codes of 8 values given the group scale ys = amax / fmax and its correctly rounded reciprocal r = RN(1 / ys)
B2_DEVICE void group128_apply_v1(const float (&f)[8], float ys, float r, float fmin, float fmax, uint2& out) {
  float qv[8];
  // ONE group-uniform branch (not one per element): the all-ones significand is the only case the reciprocal
  // sequence does not cover
  if ((__float_as_uint(ys) & 0x7fffffu) != 0x7fffffu) {
    // packed fp32x2 arithmetic (two IEEE roundings per issue slot, each lane exactly the scalar mul.rn / fma.rn): the
    // kernel is issue-bound, and these 24 scalar operations per thread were a third of its instruction stream
    uint64_t r2, nys2;
    asm("mov.b64 %0, {%1, %1};" : "=l"(r2) : "f"(r));
    asm("mov.b64 %0, {%1, %1};" : "=l"(nys2) : "f"(-ys));
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      asm("{\n\t.reg .b64 xx, q0, rem;\n\t"
          "mov.b64 xx, {%2, %3};\n\t"
          "mul.rn.f32x2 q0, xx, %4;\n\t"
          "fma.rn.f32x2 rem, q0, %5, xx;\n\t"
          "fma.rn.f32x2 q0, rem, %4, q0;\n\t"
          "mov.b64 {%0, %1}, q0;\n\t}"
          : "=f"(qv[j]), "=f"(qv[j + 1])
          : "f"(f[j]), "f"(f[j + 1]), "l"(r2), "l"(nys2));
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) qv[j] = __fdiv_rn(f[j], ys);
  }
  // reference: fmaxf(fmin, fminf(q, fmax)) then a saturating cast.  The upper fminf also maps NaN to fmax as the
  // reference does; the lower bound is what cvt.satfinite does by itself when fmin is the format's own minimum.
  const bool lower = fmin > -448.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    qv[j] = fminf(qv[j], fmax);
    if (lower) qv[j] = fmaxf(qv[j], fmin);
  }
  out.x = static_cast<uint32_t>(f32x2_to_e4m3x2(qv[0], qv[1])) | (static_cast<uint32_t>(f32x2_to_e4m3x2(qv[2], qv[3])) << 16);
  out.y = static_cast<uint32_t>(f32x2_to_e4m3x2(qv[4], qv[5])) | (static_cast<uint32_t>(f32x2_to_e4m3x2(qv[6], qv[7])) << 16);
}

B2_DEVICE float group128_quantize_v1(const float (&f)[8], float amax, float fmin, float fmax, uint2& out) {
  const float ys = __fdiv_rn(amax, fmax);
  const float r = __frcp_rn(ys);
  group128_apply_v1(f, ys, r, fmin, fmax, out);
  return ys;
}

// v2 arithmetic (sgl-kernel/csrc/gemm/per_token_group_quant_8bit_v2.cu:69-81,330-372, fp32 scales): the stored scale
// is amax * (1/448), the codes are e4m3(clamp(x * (448 / amax))) -- one multiply per element, no division.
B2_DEVICE float group128_quantize_v2(const float (&f)[8], float amax, uint2& out) {
  const float scale_inv = __fmul_rn(amax, 1.0f / 448.f);
  const float y_scale = __fdiv_rn(448.f, amax);
  float qv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) qv[j] = fminf(fmaxf(__fmul_rn(f[j], y_scale), -448.f), 448.f);
  out.x = static_cast<uint32_t>(f32x2_to_e4m3x2(qv[0], qv[1])) | (static_cast<uint32_t>(f32x2_to_e4m3x2(qv[2], qv[3])) << 16);
  out.y = static_cast<uint32_t>(f32x2_to_e4m3x2(qv[4], qv[5])) | (static_cast<uint32_t>(f32x2_to_e4m3x2(qv[6], qv[7])) << 16);
  return scale_inv;
}

}  // namespace b2
