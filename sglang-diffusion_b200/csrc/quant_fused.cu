// Producer-fused activation quantizers and the per-tensor fp8 quantizer (8f rank 4), sm_100a.
//
//   fused_rows_kernel<PROD, MODE>:  one pass  HBM -> registers -> producer -> quantizer -> HBM
//     PROD = NONE      y = x
//     PROD = SILU_MUL  x is [M, 2K]; y = T(silu(f32(x[:, :K]))) * x[:, K:]  (product rounded to T), the fusion of
//                      sgl-kernel/csrc/gemm/per_token_group_quant_8bit_v2.cu:330-341 incl. its sm_100 silu
//                      0.5 x (1 + tanh.approx(0.5 x)) (:32-40); optional masked layout [E, T, 2K] + masked_m[E] (:196-238)
//     PROD = RMSNORM   s = f32(x) + f32(residual) (residual updated in place, as fused_add_rmsnorm does:
//                      sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu -> flashinfer norm.cuh
//                      FusedAddRMSNormKernel), y = T((s * rsqrt(mean(s^2) + eps)) * w); residual may be NULL (plain
//                      rmsnorm)
//     PROD = GATHER_RMSNORM  the same, but x = T(sum over `world` slots of this rank's gather buffer, fp32 in rank
//                      order) -- the row-parallel partial outputs the preceding GEMM's epilogue pushed to every rank
//                      (b2_gemm_*_push_symm).  The kernel IS the second half of that all-reduce: cross-GPU barrier,
//                      local sum, set-counter advance (allreduce_symm.cu, gathered form), followed in registers by
//                      residual add, RMSNorm and the next layer's quantizer: all-reduce + fused_add_rmsnorm + quantize
//                      in one launch (python/sglang/srt/distributed/parallel_state.py:648-698 fused_allreduce_rmsnorm,
//                      python/sglang/srt/layers/layernorm.py forward_with_allreduce_fusion).  One row per CTA.
//     MODE = INT8_ROWS / FP8_ROWS   per-token scales, arithmetic of quant.cu (vec.h:190-228 / per_token_quant_fp8.cu)
//            FP8_G128_V1            1x128 groups, s = amax/448, q = e4m3(x / s)      (per_token_group_quant_8bit.cu)
//            FP8_G128_V2            1x128 groups, s = amax*(1/448), q = e4m3(x*(448/amax))  (..._8bit_v2.cu)
//   The quantizer sees y rounded to T, exactly what the separate activation / norm kernel would have stored, so
//   fused == unfused bit for bit; y itself is only written when the caller asks for it.
//   Bytes per row: SILU_MUL 2*2K in, K + 4K/128 out (the unfused pair moves 2*2K + 2K + 2K + K);
//                  RMSNORM  2K (+2K residual in, +2K residual out), K out (unfused: + 2K write + 2K read of y).
//
//   per-tensor fp8 (sgl-kernel/csrc/gemm/per_tensor_quant_fp8.cu:11-123): dynamic = absmax pass (atomic max on the
//   caller-zeroed scale) + quantize pass; static = quantize pass with the given scale.  q = e4m3(clamp(x * (1/s))).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include "b200_w8a8_internal.h"
#include "ptx.cuh"
#include "quant_common.cuh"
#include "symm_common.cuh"

namespace b2 {

enum class Prod { NONE, SILU_MUL, RMSNORM, GATHER_RMSNORM };
enum class QM { INT8_ROWS, FP8_ROWS, FP8_G128_V1, FP8_G128_V2 };

B2_DEVICE float tanh_approx(float x) {
  float r;
  asm("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// per_token_group_quant_8bit_v2.cu:32-40 (the __CUDA_ARCH__ >= 1000 branch)
B2_DEVICE float silu_sm100(float v) {
  const float half = 0.5f * v;
  return half * (1.0f + tanh_approx(half));
}

template <typename T>
B2_DEVICE T from_f32(float v);
template <>
B2_DEVICE __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
template <>
B2_DEVICE __half from_f32<__half>(float v) {
  return __float2half_rn(v);
}

template <typename T, Prod PROD, QM MODE, int TPR, int BLOCK>
__global__ void __launch_bounds__(BLOCK) fused_rows_kernel(FusedQuantArgs a) {
  // The row lives in registers between the producer and the quantizer.  SILU_MUL / NONE keep the T-rounded values
  // PACKED (8 per uint4, 16 registers instead of 32: ncu on the fp32-cached form shows issue_active 48 % and
  // long_scoreboard 5.5 warps per issue -- latency-bound at 50 % occupancy).  RMSNORM needs the unrounded fp32 sums
  // for the norm and keeps four fp32 vectors.
  constexpr bool RMS = PROD == Prod::RMSNORM || PROD == Prod::GATHER_RMSNORM;
  constexpr bool PACKED = !RMS;
  constexpr int NC = 4;               // 16-byte vectors cached per thread (8 was measured: worse below K = 8192)
  constexpr int VEC = 8;              // elements per vector
  constexpr int ROWS = BLOCK / TPR;
  constexpr int WPR = (TPR + 31) / 32;  // warps per row
  constexpr bool GROUPS = MODE == QM::FP8_G128_V1 || MODE == QM::FP8_G128_V2;
  __shared__ float s_red[2][ROWS][WPR];

  pdl_launch_dependents();
  pdl_wait_prior_grids();

  int64_t gather_base = 0;  // element offset of the set in use
  uint32_t gather_epoch = 0;
  if constexpr (PROD == Prod::GATHER_RMSNORM) {
    // this rank's GEMM has completed (wait above): its pushes are performed.  One barrier -- every rank's GEMM has
    // completed -- then everything below reads local memory only.
    uint32_t* pad = a.gather_pads[a.gather_rank];
    const uint32_t epoch = pad[kEpochWord + blockIdx.x];
    const uint32_t counter = pad[kGatherCounterWord];
    xgpu_barrier<true>(a.gather_pads, a.gather_rank, a.gather_world, epoch + 1, a.gather_timeout_ns);
    gather_base = static_cast<int64_t>(counter & 1u) * a.gather_set_stride;
    gather_epoch = epoch;
    if (threadIdx.x == 0) {
      if (a.out_peers == nullptr) pad[kEpochWord + blockIdx.x] = epoch + 1;  // (with peers: epoch + 2, at the end)
      __threadfence();
      const uint32_t done = atomicAdd(&pad[kGatherDoneWord], 1u);
      if (done == gridDim.x - 1) {  // last CTA of this call: every CTA has read the counter
        pad[kGatherDoneWord] = 0;
        pad[kGatherCounterWord] = counter + 1;
      }
    }
  }

  const int r_in_cta = threadIdx.x / TPR;
  const int t = threadIdx.x % TPR;
  auto do_rows = [&](int64_t row_base) {
  const int64_t row = row_base + r_in_cta;
  bool active = row < a.M;
  int64_t expert = 0, token = row;
  if (a.masked_m != nullptr) {  // [E, T, ...] layout: rows past masked_m[e] are neither read nor written
    expert = row / a.tokens_per_expert;
    token = row - expert * a.tokens_per_expert;
    if (active) active = token < a.masked_m[expert];
  }
  const int nvec = a.K / VEC;
  const int64_t rr = active ? row : 0;
  // gather mode (MoE): output row r reads input row row_map[r] / row_div; ids >= row_limit are padding rows, which
  // are quantized as all-zero rows (codes 0) so that the grouped GEMM's padding outputs are exact zeros
  int64_t src = rr;
  bool src_ok = true;
  if (a.row_map != nullptr) {
    const int32_t id = active ? a.row_map[rr] : 0;
    src_ok = id < a.row_limit;
    src = src_ok ? id / a.row_div : 0;
  }

  float f[PACKED ? 1 : NC][VEC];
  uint4 yc[PACKED ? NC : 1];
  float sum_sq = 0.f;
  const T* xr = static_cast<const T*>(a.x) + src * a.ldx;
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    const int v = t + i * TPR;
    const bool on = active && src_ok && v < nvec;
    if constexpr (PACKED) {
      yc[i] = make_uint4(0, 0, 0, 0);
    } else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) f[i][j] = 0.f;
    }
    if (on) {
      uint4 raw;
      if constexpr (PROD == Prod::GATHER_RMSNORM) {
        // rank-ordered fp32 sum of the slots, one rounding to T: bit-identical to allreduce_gathered_kernel
        const uint4* sp = reinterpret_cast<const uint4*>(xr + gather_base) + v;
        const int64_t slot_vec = a.M * static_cast<int64_t>(a.K) / VEC;
        float acc[VEC];
        {
          const uint4 p0 = ld_stream(sp);
          const T* pe = reinterpret_cast<const T*>(&p0);
#pragma unroll
          for (int j = 0; j < VEC; ++j) acc[j] = to_f32(pe[j]);
        }
        for (int r = 1; r < a.gather_world; ++r) {
          const uint4 pr = ld_stream(sp + r * slot_vec);
          const T* pe = reinterpret_cast<const T*>(&pr);
#pragma unroll
          for (int j = 0; j < VEC; ++j) acc[j] = __fadd_rn(acc[j], to_f32(pe[j]));
        }
        T* ro = reinterpret_cast<T*>(&raw);
#pragma unroll
        for (int j = 0; j < VEC; ++j) ro[j] = from_f32<T>(acc[j]);
      } else {
        raw = ld_stream(reinterpret_cast<const uint4*>(xr) + v);
      }
      const T* e = reinterpret_cast<const T*>(&raw);
      if constexpr (PROD == Prod::NONE) {
        yc[i] = raw;
      } else if constexpr (PROD == Prod::SILU_MUL) {
        const uint4 raw2 = ld_stream(reinterpret_cast<const uint4*>(xr + a.K) + v);
        const T* u = reinterpret_cast<const T*>(&raw2);
        uint4 yv;
        T* yo = reinterpret_cast<T*>(&yv);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          const T act = from_f32<T>(silu_sm100(to_f32(e[j])));
          yo[j] = from_f32<T>(__fmul_rn(to_f32(act), to_f32(u[j])));  // T * T (operator* of the reference's types)
        }
        yc[i] = yv;
        if (a.y != nullptr) reinterpret_cast<uint4*>(static_cast<T*>(a.y) + rr * a.ldy)[v] = yv;
      } else {  // RMSNORM, first pass: s = x + residual, residual <- T(s), sum of squares of the unrounded s
        if (a.residual != nullptr) {
          T* rp = static_cast<T*>(a.residual) + rr * a.ldr;
          const uint4 rraw = *(reinterpret_cast<const uint4*>(rp) + v);
          const T* re = reinterpret_cast<const T*>(&rraw);
          uint4 rout;
          T* ro = reinterpret_cast<T*>(&rout);
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            const float sx = __fadd_rn(to_f32(e[j]), to_f32(re[j]));
            ro[j] = from_f32<T>(sx);
            f[i][j] = sx;
          }
          reinterpret_cast<uint4*>(rp)[v] = rout;
        } else {
#pragma unroll
          for (int j = 0; j < VEC; ++j) f[i][j] = to_f32(e[j]);
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) sum_sq = __fmaf_rn(f[i][j], f[i][j], sum_sq);
      }
    }
  }

  constexpr int SHFL_W = TPR < 32 ? TPR : 32;
  if constexpr (RMS) {
#pragma unroll
    for (int o = SHFL_W / 2; o > 0; o >>= 1) sum_sq += __shfl_xor_sync(0xffffffffu, sum_sq, o);
    if constexpr (WPR > 1) {
      if ((t & 31) == 0) s_red[0][r_in_cta][t / 32] = sum_sq;
      __syncthreads();
      float m = 0.f;
#pragma unroll
      for (int i = 0; i < WPR; ++i) m += s_red[0][r_in_cta][i];
      sum_sq = m;
    }
    const float rms_rcp = rsqrtf(__fadd_rn(__fdiv_rn(sum_sq, static_cast<float>(a.K)), a.eps));
    const T* wr = static_cast<const T*>(a.weight);
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      const int v = t + i * TPR;
      if (active && v < nvec) {
        const uint4 wraw = __ldg(reinterpret_cast<const uint4*>(wr) + v);
        const T* we = reinterpret_cast<const T*>(&wraw);
        uint4 yv;
        T* yo = reinterpret_cast<T*>(&yv);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          yo[j] = from_f32<T>(__fmul_rn(__fmul_rn(f[i][j], rms_rcp), to_f32(we[j])));
          f[i][j] = to_f32(yo[j]);
        }
        if (a.y != nullptr) reinterpret_cast<uint4*>(static_cast<T*>(a.y) + rr * a.ldy)[v] = yv;
      }
    }
  }

  // ---- quantizer over the T-rounded row held in registers ----
  auto load8 = [&](int i, float (&o)[VEC]) {
    if constexpr (PACKED) {
      const T* e = reinterpret_cast<const T*>(&yc[i]);
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = to_f32(e[j]);
    } else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) o[j] = f[i][j];
    }
  };
  uint8_t* qr = a.q + rr * a.ldq;
  if constexpr (GROUPS) {
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      const int v = t + i * TPR;
      float g8[VEC];
      load8(i, g8);
      float amax = a.q_eps;
#pragma unroll
      for (int j = 0; j < VEC; ++j) amax = fmaxf(amax, fabsf(g8[j]));
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
      uint2 out;
      float sc;
      if constexpr (MODE == QM::FP8_G128_V1) {
        sc = group128_quantize_v1(g8, amax, a.fmin, a.fmax, out);
      } else {
        sc = group128_quantize_v2(g8, amax, out);
      }
      if (active && v < nvec) {
        reinterpret_cast<uint2*>(qr)[v] = out;
        if ((t & 15) == 0)
          a.s[expert * a.s_exp_stride + token * a.s_row_stride + static_cast<int64_t>(v >> 4) * a.s_grp_stride] = sc;
      }
    }
  } else {
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      float g8[VEC];
      load8(i, g8);
#pragma unroll
      for (int j = 0; j < VEC; ++j) amax = fmaxf(amax, fabsf(g8[j]));
    }
#pragma unroll
    for (int o = SHFL_W / 2; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if constexpr (WPR > 1) {
      if ((t & 31) == 0) s_red[1][r_in_cta][t / 32] = amax;
      __syncthreads();
      float m = s_red[1][r_in_cta][0];
#pragma unroll
      for (int i = 1; i < WPR; ++i) m = fmaxf(m, s_red[1][r_in_cta][i]);
      amax = m;
    }
    if (!active) return;
    float scale, inv;
    if constexpr (MODE == QM::INT8_ROWS) {
      amax = fmaxf(amax, 1e-7f);
      scale = __fdiv_rn(amax, 127.f);
      inv = __fdiv_rn(127.f, amax);
    } else {
      scale = __fdiv_rn(amax, 448.f);
      inv = (scale == 0.f) ? 0.f : __frcp_rn(scale);
    }
    if (t == 0) {
      bool local_scale = true;
      if constexpr (PROD == Prod::GATHER_RMSNORM) {
        if (a.out_peers != nullptr) {
          local_scale = false;
          for (int pr = 0; pr < a.gather_world; ++pr)
            reinterpret_cast<float*>(a.out_peers[pr] + a.out_scale_offset)[a.out_row0 + rr] = scale;
        }
      }
      if (local_scale) a.s[expert * a.s_exp_stride + token * a.s_row_stride] = scale;
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      const int v = t + i * TPR;
      if (v < nvec) {
        uint2 out;
        float g8[VEC];
        load8(i, g8);
        if constexpr (MODE == QM::INT8_ROWS) {
          uint8_t* o = reinterpret_cast<uint8_t*>(&out);
#pragma unroll
          for (int j = 0; j < VEC; ++j) o[j] = quant_one<QMode::INT8>(g8[j], inv);
        } else {
          uint16_t* o2 = reinterpret_cast<uint16_t*>(&out);
#pragma unroll
          for (int j = 0; j < VEC; j += 2) {
            const float p = fmaxf(fminf(__fmul_rn(g8[j], inv), 448.f), -448.f);
            const float r = fmaxf(fminf(__fmul_rn(g8[j + 1], inv), 448.f), -448.f);
            o2[j >> 1] = f32x2_to_e4m3x2(p, r);
          }
        }
        if constexpr (PROD == Prod::GATHER_RMSNORM) {
          if (a.out_peers != nullptr) {
#pragma unroll 1
            for (int pi = 0; pi < a.gather_world; ++pi) {
              const int pr = (a.gather_rank + pi) % a.gather_world;  // own copy first, then the peers in rotated order
              reinterpret_cast<uint2*>(a.out_peers[pr] + (a.out_row0 + rr) * a.ldq)[v] = out;
            }
          } else {
            reinterpret_cast<uint2*>(qr)[v] = out;
          }
        } else {
          reinterpret_cast<uint2*>(qr)[v] = out;
        }
      }
    }
  }
  };  // do_rows

  if constexpr (PROD == Prod::GATHER_RMSNORM) {
    // persistent over rows: the grid is bounded by the barrier slots of the signal pad (one per CTA)
    for (int64_t rb = static_cast<int64_t>(blockIdx.x) * ROWS; rb < a.M; rb += static_cast<int64_t>(gridDim.x) * ROWS) {
      do_rows(rb);
      __syncthreads();  // s_red is reused by the next row
    }
    if (a.out_peers != nullptr) {
      // all-gather form: the codes and scales of this CTA's rows are on their way to every rank; after this barrier
      // CTA b of every rank has stored its rows, so when this kernel completes on a rank, ALL rows are in its copy
      xgpu_barrier<true>(a.gather_pads, a.gather_rank, a.gather_world, gather_epoch + 2, a.gather_timeout_ns);
      if (threadIdx.x == 0) a.gather_pads[a.gather_rank][kEpochWord + blockIdx.x] = gather_epoch + 2;
    }
  } else {
    do_rows(static_cast<int64_t>(blockIdx.x) * ROWS);
  }
}

template <typename T, Prod PROD, QM MODE, int TPR, int MIN_BLOCK = 128>
static cudaError_t launch_fused(const FusedQuantArgs& a, cudaStream_t st) {
  constexpr int BLOCK = TPR >= MIN_BLOCK ? TPR : MIN_BLOCK;
  constexpr int ROWS = BLOCK / TPR;
  if constexpr (PROD == Prod::GATHER_RMSNORM && MIN_BLOCK == 128) {
    // more rows than barrier slots (prefill, sequence-parallel norm): the CTAs loop over rows, so give every CTA
    // 512 threads = several rows in flight per iteration
    if (a.M > a.gather_max_blocks && TPR < 512) return launch_fused<T, PROD, MODE, TPR, 512>(a, st);
  }
  int64_t grid = (a.M + ROWS - 1) / ROWS;
  if (PROD == Prod::GATHER_RMSNORM && grid > a.gather_max_blocks) grid = a.gather_max_blocks;  // the kernel loops over rows
  if (grid > 0x7fffffffll) return cudaErrorInvalidValue;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(BLOCK);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = ((weight_prefetch_enabled() && a.M <= 128) || PROD == Prod::GATHER_RMSNORM) ? 1 : 0;  // decode chains, see quant.cu launch_rows
  return cudaLaunchKernelEx(&cfg, fused_rows_kernel<T, PROD, MODE, TPR, BLOCK>, a);
}

template <typename T, Prod PROD, QM MODE>
static cudaError_t dispatch_fused_tpr(const FusedQuantArgs& a, cudaStream_t st) {
  const int nvec = a.K / 8;
  constexpr int NCH = 4;  // vectors a thread caches (fused_rows_kernel::NC)
  int tpr = PROD == Prod::GATHER_RMSNORM ? 128 : 32;  // the gathered form takes a cross-GPU barrier per CTA: one row per CTA
  while (tpr < 1024 && nvec > tpr * NCH) tpr *= 2;
  // few rows (decode): spread a row over more threads as long as every thread keeps a vector
  while (tpr < 1024 && a.M * static_cast<int64_t>(tpr) < 148 * 1024 && nvec >= tpr * 2) tpr *= 2;
  switch (tpr) {
    case 32: return launch_fused<T, PROD, MODE, 32>(a, st);
    case 64: return launch_fused<T, PROD, MODE, 64>(a, st);
    case 128: return launch_fused<T, PROD, MODE, 128>(a, st);
    case 256: return launch_fused<T, PROD, MODE, 256>(a, st);
    case 512: return launch_fused<T, PROD, MODE, 512>(a, st);
    default: return launch_fused<T, PROD, MODE, 1024>(a, st);
  }
}

template <typename T>
static cudaError_t dispatch_fused(int prod, int mode, const FusedQuantArgs& a, cudaStream_t st) {
#define B2_FUSED_CASE(P_, M_) \
  if (prod == static_cast<int>(P_) && mode == static_cast<int>(M_)) return dispatch_fused_tpr<T, P_, M_>(a, st);
  B2_FUSED_CASE(Prod::NONE, QM::FP8_G128_V2)
  B2_FUSED_CASE(Prod::NONE, QM::INT8_ROWS)
  B2_FUSED_CASE(Prod::SILU_MUL, QM::INT8_ROWS)
  B2_FUSED_CASE(Prod::SILU_MUL, QM::FP8_ROWS)
  B2_FUSED_CASE(Prod::SILU_MUL, QM::FP8_G128_V2)
  B2_FUSED_CASE(Prod::RMSNORM, QM::INT8_ROWS)
  B2_FUSED_CASE(Prod::RMSNORM, QM::FP8_ROWS)
  B2_FUSED_CASE(Prod::RMSNORM, QM::FP8_G128_V1)
  B2_FUSED_CASE(Prod::GATHER_RMSNORM, QM::INT8_ROWS)
  B2_FUSED_CASE(Prod::GATHER_RMSNORM, QM::FP8_ROWS)
  B2_FUSED_CASE(Prod::GATHER_RMSNORM, QM::FP8_G128_V1)
#undef B2_FUSED_CASE
  return cudaErrorInvalidValue;
}

cudaError_t fused_quant_launch(int prod, int mode, int x_dtype, const FusedQuantArgs& a, cudaStream_t st) {
  if (x_dtype == B2_DTYPE_BF16) return dispatch_fused<__nv_bfloat16>(prod, mode, a, st);
  return dispatch_fused<__half>(prod, mode, a, st);
}

// ------------------------------------------------------------- per-tensor fp8 ----

template <typename T>
__global__ void __launch_bounds__(256) tensor_absmax_kernel(const T* __restrict__ x, float* __restrict__ s, int64_t n) {
  // per_tensor_quant_fp8.cu:11-47: s = max(s, blockmax / 448) with an atomic max (non-negative floats order like
  // their bit patterns); s must be zero-initialised by the caller (fp8_kernel.py:1671)
  __shared__ float s_red[8];
  pdl_launch_dependents();
  pdl_wait_prior_grids();
  const int64_t nvec = n / 8;
  float m = 0.f;
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += static_cast<int64_t>(gridDim.x) * 256) {
    const uint4 raw = ld_stream(reinterpret_cast<const uint4*>(x) + i);
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int j = 0; j < 8; ++j) m = fmaxf(m, fabsf(to_f32(e[j])));
  }
  for (int64_t i = nvec * 8 + blockIdx.x * 256ll + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * 256)
    m = fmaxf(m, fabsf(to_f32(x[i])));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 1; i < 8; ++i) m = fmaxf(m, s_red[i]);
    atomicMax(reinterpret_cast<int*>(s), __float_as_int(__fdiv_rn(m, 448.f)));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) tensor_quant_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                           const float* __restrict__ s, float* __restrict__ s_repeat,
                                                           int64_t n, int64_t rows) {
  // per_tensor_quant_fp8.cu:49-96: q = e4m3(clamp(x * (1/s), +-448)).  s_repeat (optional, [rows]) receives the
  // scale once per token row: static_quant_fp8(repeat_scale=True) of fp8_kernel.py:649-700.
  pdl_launch_dependents();
  pdl_wait_prior_grids();
  const float sv = *s;
  const float inv = __frcp_rn(sv);
  const int64_t nvec = n / 16;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * 256;
  const int64_t gid = blockIdx.x * 256ll + threadIdx.x;
  for (int64_t i = gid; i < nvec; i += stride) {
    const uint4 r0 = ld_stream(reinterpret_cast<const uint4*>(x) + 2 * i);
    const uint4 r1 = ld_stream(reinterpret_cast<const uint4*>(x) + 2 * i + 1);
    const T* e0 = reinterpret_cast<const T*>(&r0);
    const T* e1 = reinterpret_cast<const T*>(&r1);
    uint4 out;
    uint16_t* o2 = reinterpret_cast<uint16_t*>(&out);
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      o2[j >> 1] = f32x2_to_e4m3x2(fmaxf(fminf(__fmul_rn(to_f32(e0[j]), inv), 448.f), -448.f),
                                   fmaxf(fminf(__fmul_rn(to_f32(e0[j + 1]), inv), 448.f), -448.f));
      o2[4 + (j >> 1)] = f32x2_to_e4m3x2(fmaxf(fminf(__fmul_rn(to_f32(e1[j]), inv), 448.f), -448.f),
                                         fmaxf(fminf(__fmul_rn(to_f32(e1[j + 1]), inv), 448.f), -448.f));
    }
    reinterpret_cast<uint4*>(q)[i] = out;
  }
  for (int64_t i = nvec * 16 + gid; i < n; i += stride)
    q[i] = f32_to_e4m3_sat(fmaxf(fminf(__fmul_rn(to_f32(x[i]), inv), 448.f), -448.f));
  if (s_repeat != nullptr)
    for (int64_t i = gid; i < rows; i += stride) s_repeat[i] = sv;
}

template <typename T>
static cudaError_t tensor_quant_t(const void* x, uint8_t* q, float* s, float* s_repeat, int64_t n, int64_t rows,
                                  int is_static, cudaStream_t st) {
  const int sms = device_sm_count();
  // memory-bound streaming passes: 8 CTAs of 256 threads per SM keep ~64 KB of loads in flight per SM
  const int64_t want = (n / 16 + 255) / 256;
  const int grid = static_cast<int>(want < 1 ? 1 : (want > sms * 8ll ? sms * 8ll : want));
  if (!is_static) {
    tensor_absmax_kernel<T><<<grid, 256, 0, st>>>(static_cast<const T*>(x), s, n);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    count_launch();
  }
  tensor_quant_kernel<T><<<grid, 256, 0, st>>>(static_cast<const T*>(x), q, s, s_repeat, n, rows);
  return cudaGetLastError();
}

cudaError_t quant_fp8_tensor(const void* x, int x_dtype, uint8_t* q, float* s, float* s_repeat, int64_t n,
                             int64_t rows, int is_static, cudaStream_t st) {
  if (x_dtype == B2_DTYPE_BF16) return tensor_quant_t<__nv_bfloat16>(x, q, s, s_repeat, n, rows, is_static, st);
  return tensor_quant_t<__half>(x, q, s, s_repeat, n, rows, is_static, st);
}

}  // namespace b2
