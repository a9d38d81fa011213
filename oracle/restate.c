/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's W8A8
 * linear path.  Nothing under sglang-diffusion_b200/ may link, import or call this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs do, and only as the checker or the timed CPU baseline.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every int8 function here
 * bit-for-bit against the reference's own CPU code compiled from
 * /root/reference (oracle/_ref/sgl_cpu_oracle.so, recipe oracle/build_ref.py) and
 * against tests/golden/ vectors generated from that .so
 * (tests/golden/make_golden.py).  The reference ships no golden vectors for this
 * path.
 *
 * Each function cites the reference lines (relative to /root/reference) it
 * restates.  Plain C, scalar arithmetic; OpenMP over rows only.  Compile with
 * -ffp-contract=off so that the only fused multiply-add is the explicit fmaf().
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- bf16 ---- */

static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = ((uint32_t)h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

/* round-to-nearest-even, as _mm512_cvtne2ps_pbh (gemm_int8.cpp:57) and
 * __float2bfloat16_rn.  (vcvtne2ps2bf16 flushes bf16-denormal results to zero;
 * no test value gets near 1e-38, documented in DESIGN.md.) */
static inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40u); /* NaN */
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

/* ------------------------------------------------------ int8 quantizer ---- */

/* sgl-kernel/csrc/cpu/vec.h:190-228 (AVX512 bf16 specialisation; generic :170-188)
 *   amax = max(max|x|, 1e-7); scale = amax/127; inv = 127/amax  (IEEE division)
 *   q = round-half-even(x*inv)  (_MM_FROUND_TO_NEAREST_INT) ; stored as u8 = q + 128
 * driver loop: gemm_int8.cpp:382-410 (per_token_quant_int8_cpu), row stride = lda. */
void orc_quant_int8_rows(const uint16_t* x, int64_t ldx, uint8_t* q, float* s, int64_t M, int64_t K) {
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    const uint16_t* xr = x + m * ldx;
    float amax = 0.f;
    for (int64_t k = 0; k < K; ++k) {
      float a = fabsf(bf16_to_f32(xr[k]));
      amax = a > amax ? a : amax;
    }
    amax = amax > 1e-7f ? amax : 1e-7f;
    const float scale = amax / 127.f;
    const float inv = 127.f / amax;
    for (int64_t k = 0; k < K; ++k) {
      float v = bf16_to_f32(xr[k]) * inv;
      int32_t r = (int32_t)nearbyintf(v); /* default rounding mode = nearest-even */
      q[m * K + k] = (uint8_t)(r + 128);
    }
    s[m] = scale;
  }
}

/* --------------------------------------------------------- int8 GEMM ------ */

/* int32 accumulator of the reference's u8 x s8 kernel:
 *   acc = sum_k (q+128)*w  -  128*sum_k w          (== exact s8 x s8 dot)
 * sgl-kernel/csrc/cpu/gemm.cpp:11-36 (s8s8_compensation: comp[n] = 128*sum_k W[n,k]),
 * gemm_int8.cpp:112-121 (_mm512_dpbusd_epi32 over K/4) and :40,175-176 (acc - comp).
 * A is the u8 = q+128 array produced above. */
void orc_gemm_u8s8_acc(const uint8_t* A, const int8_t* W, int32_t* acc, int64_t M, int64_t N, int64_t K) {
  int32_t* comp = (int32_t*)malloc(sizeof(int32_t) * (size_t)N);
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; ++n) {
    int32_t c = 0;
    for (int64_t k = 0; k < K; ++k) c += (int32_t)W[n * K + k];
    comp[n] = 128 * c;
  }
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t n = 0; n < N; ++n) {
      const uint8_t* a = A + m * K;
      const int8_t* w = W + n * K;
      int32_t c = 0;
      for (int64_t k = 0; k < K; ++k) c += (int32_t)a[k] * (int32_t)w[k];
      acc[m * N + n] = c - comp[n];
    }
  }
  free(comp);
}

/* sgl-kernel/csrc/cpu/gemm_int8.cpp:20-60 (scale_C) and :175-186 (tinygemm_kernel_nn::storec)
 *   bias:    out = bf16_rne( fma( f32(acc)*As[m], Bs[n], bias[n] ) )     (_mm512_fmadd_ps(_mm512_mul_ps(..)))
 *   no bias: out = bf16_rne( (f32(acc)*As[m]) * Bs[n] )
 * f32(acc) is _mm512_cvtepi32_ps = round-to-nearest-even of the int32. */
void orc_int8_epilogue(const int32_t* acc, const float* As, const float* Bs, const float* bias, uint16_t* out,
                       int64_t M, int64_t N) {
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t n = 0; n < N; ++n) {
      float c = (float)acc[m * N + n];
      float t = c * As[m];
      float r = bias ? fmaf(t, Bs[n], bias[n]) : t * Bs[n];
      out[m * N + n] = f32_to_bf16_rne(r);
    }
  }
}

/* int8_scaled_mm_with_quant, sgl-kernel/csrc/cpu/gemm_int8.cpp:481-547:
 * quantize rows (:513-533) then int8_scaled_mm_kernel_impl (:286-340).
 * Scratch for q/As/acc is caller-provided so that the accumulators stay visible. */
void orc_int8_linear(const uint16_t* x, int64_t ldx, const int8_t* W, const float* Bs, const float* bias,
                     uint16_t* out, uint8_t* q_scratch, float* as_scratch, int32_t* acc_scratch, int64_t M,
                     int64_t N, int64_t K) {
  orc_quant_int8_rows(x, ldx, q_scratch, as_scratch, M, K);
  orc_gemm_u8s8_acc(q_scratch, W, acc_scratch, M, N, K);
  orc_int8_epilogue(acc_scratch, as_scratch, Bs, bias, out, M, N);
}

/* ---------------------------------------------------------------- e4m3 ---- */

/* OCP e4m3fn: 1-4-3, bias 7, max 448 (0x7e), NaN 0x7f, no inf. */
float orc_e4m3_to_f32(uint8_t c) {
  int sign = c >> 7, e = (c >> 3) & 0xf, m = c & 7;
  float v;
  if (e == 0xf && m == 7) return NAN;
  if (e == 0) v = ldexpf((float)m, -9);
  else v = ldexpf((float)(8 + m), e - 10);
  return sign ? -v : v;
}

/* float -> e4m3fn, round-to-nearest-even, saturate-to-finite
 * (what static_cast<__nv_fp8_e4m3>(float) / cvt.rn.satfinite.e4m3x2.f32 compute). */
uint8_t orc_f32_to_e4m3(float x) {
  uint8_t sign = signbit(x) ? 0x80 : 0;
  float a = fabsf(x);
  if (isnan(a)) return (uint8_t)(sign | 0x7f);
  if (a >= 448.f) return (uint8_t)(sign | 0x7e);
  if (a < 0.015625f) { /* below min normal 2^-6: subnormal grid of 2^-9 */
    int q = (int)nearbyintf(a * 512.f);
    return (uint8_t)(sign | q); /* q == 8 is exactly the min normal code */
  }
  int e;
  float fr = frexpf(a, &e); /* a = fr * 2^e, fr in [0.5,1) */
  int m = (int)nearbyintf(fr * 16.f); /* in [8,16] */
  e -= 1;
  if (m == 16) {
    m = 8;
    e += 1;
  }
  int code = ((e + 7) << 3) | (m - 8);
  if (code > 0x7e) code = 0x7e;
  return (uint8_t)(sign | code);
}

/* ------------------------------------------------- fp8 per-token quant ---- */

/* sgl_per_token_quant_fp8, sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:17-121
 * (warp-per-token kernel; the CTA-per-token kernel :126-198 differs only for an
 * all-zero row):
 *   scale = max|x| / 448 ; inv = scale==0 ? 0 : 1/scale ;
 *   q = e4m3_rn_sat( clamp(x*inv, -448, 448) )
 * The reference is built with -use_fast_math (sgl-kernel/CMakeLists.txt:355), so
 * its own division/reciprocal are approximate; this restatement uses IEEE
 * arithmetic and is the definition the CUDA kernel here is held to bit-for-bit. */
void orc_quant_fp8_rows(const uint16_t* x, int64_t ldx, uint8_t* q, float* s, int64_t M, int64_t K) {
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    const uint16_t* xr = x + m * ldx;
    float amax = 0.f;
    for (int64_t k = 0; k < K; ++k) {
      float a = fabsf(bf16_to_f32(xr[k]));
      amax = a > amax ? a : amax;
    }
    const float scale = amax / 448.f;
    const float inv = scale == 0.f ? 0.f : 1.0f / scale;
    for (int64_t k = 0; k < K; ++k) {
      float v = bf16_to_f32(xr[k]) * inv;
      v = fmaxf(fminf(v, 448.f), -448.f);
      q[m * K + k] = orc_f32_to_e4m3(v);
    }
    s[m] = scale;
  }
}

/* sgl_per_token_group_quant_8bit (fp8, fp32 scales),
 * sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:19-114:
 *   amax = max(eps, max|x| over the 1xG group); y_s = amax / fp8_max;
 *   q = e4m3( clamp(x / y_s, fp8_min, fp8_max) )          (true division, :110)
 * scale element (m, g) is stored at s[g*s_stride + m] (column-major / M-major,
 * :57-66 with num_elems_per_pack == 1) or s[m*(K/G) + g] when s_stride == 0. */
void orc_quant_fp8_group(const uint16_t* x, int64_t ldx, uint8_t* q, float* s, int64_t s_stride, int64_t M,
                         int64_t K, int64_t G, float eps, float fp8_min, float fp8_max) {
  const int64_t ng = K / G;
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t g = 0; g < ng; ++g) {
      const uint16_t* xg = x + m * ldx + g * G;
      float amax = eps;
      for (int64_t k = 0; k < G; ++k) amax = fmaxf(amax, fabsf(bf16_to_f32(xg[k])));
      const float ys = amax / fp8_max;
      for (int64_t k = 0; k < G; ++k) {
        float v = bf16_to_f32(xg[k]) / ys;
        v = fminf(fmaxf(v, fp8_min), fp8_max);
        q[m * K + g * G + k] = orc_f32_to_e4m3(v);
      }
      if (s_stride) s[g * s_stride + m] = ys;
      else s[m * ng + g] = ys;
    }
  }
}

/* ------------------------------------------------------------ fp8 GEMM ---- */

/* fp8_scaled_mm, sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1450-1532, sm100 epilogue
 * :830-880 (EVT: As[m] * (Bs[n] * acc) [+ bias]), fp32 accumulation on the tensor
 * core.  Tensor-core summation order is not specified, so the restatement
 * accumulates the exact e4m3 products in double and rounds once: a
 * tolerance-only oracle (tests use rtol 2e-2 / atol scaled, as
 * sgl-kernel/tests/test_fp8_gemm.py:33-35).  out is fp32 so that tests can
 * measure error before and after the bf16 cast. */
void orc_gemm_e4m3(const uint8_t* A, const uint8_t* W, const float* As, const float* Bs, const float* bias,
                   float* out, int64_t M, int64_t N, int64_t K) {
  float lut[256];
  for (int i = 0; i < 256; ++i) lut[i] = orc_e4m3_to_f32((uint8_t)i);
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t n = 0; n < N; ++n) {
      const uint8_t* a = A + m * K;
      const uint8_t* w = W + n * K;
      double acc = 0.0;
      for (int64_t k = 0; k < K; ++k) acc += (double)(lut[a[k]] * lut[w[k]]);
      float r = As[m] * (Bs[n] * (float)acc);
      if (bias) r += bias[n];
      out[m * N + n] = r;
    }
  }
}

/* fp8_blockwise_scaled_mm, sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:368-465:
 *   D[m,n] = sum_kb SFA[m,kb] * SFB[n/128,kb] * (A[m,kb-block] . W[n,kb-block])
 * SFA element (m,kb) at sfa[kb*sfa_ld + m] (M-major, :396-399);
 * SFB element (nb,kb) at sfb[nb*sfb_ld + kb] (K-major as seen from [K/128,N/128]
 * with stride(0)==1, :400-403).  No bias in the reference op. */
void orc_gemm_e4m3_block128(const uint8_t* A, const uint8_t* W, const float* sfa, int64_t sfa_ld,
                            const float* sfb, int64_t sfb_ld, float* out, int64_t M, int64_t N, int64_t K) {
  float lut[256];
  for (int i = 0; i < 256; ++i) lut[i] = orc_e4m3_to_f32((uint8_t)i);
  const int64_t nkb = (K + 127) / 128;
#pragma omp parallel for schedule(static) collapse(2)
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t n = 0; n < N; ++n) {
      const uint8_t* a = A + m * K;
      const uint8_t* w = W + n * K;
      double tot = 0.0;
      for (int64_t kb = 0; kb < nkb; ++kb) {
        double acc = 0.0;
        int64_t k1 = (kb + 1) * 128 < K ? (kb + 1) * 128 : K;
        for (int64_t k = kb * 128; k < k1; ++k) acc += (double)(lut[a[k]] * lut[w[k]]);
        tot += (double)sfa[kb * sfa_ld + m] * (double)sfb[(n / 128) * sfb_ld + kb] * acc;
      }
      out[m * N + n] = (float)tot;
    }
  }
}

int orc_abi_version(void) { return 1; }
