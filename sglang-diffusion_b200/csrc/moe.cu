// W8A8-int8 mixture-of-experts plumbing around the grouped tcgen05 GEMM.
//
// Reference: fused_experts_int8_kernel_impl (sgl-kernel/csrc/cpu/moe_int8.cpp:585-847) as called by
// W8A8Int8MoEMethod.apply (python/sglang/srt/layers/quantization/w8a8_int8.py:334-378):
//   stage 0  per-token int8 quantization of the hidden states            (quantize_row_int8, vec.h:190-228)
//   stage 1  ic1 = silu(x W13_gate^T) * (x W13_up^T) over expert-sorted 'BLOCK_M' row blocks (moe_align_block_size)
//   stage 2  per-token int8 quantization of ic1, ic2 = ic1 W2^T
//   stage 3  out[t] = sum_k topk_weight[t,k] * ic2[row of (t,k)]
// On the GPU the row blocks are 128 tall (one UMMA tile), every block belongs to one expert, and the two GEMMs are
// the library's persistent int8 kernel in grouped mode (gemm_sm100.cu: tile_expert[]).  This file holds the three
// small kernels around them: block alignment, and the weighted combine; the gather + quantize stage is the row
// quantizer of quant_fused.cu with a row map.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <type_traits>

#include "b200_w8a8_internal.h"
#include "ptx.cuh"
#include "quant_common.cuh"

namespace b2 {

// One CTA.  topk_ids[n] in [0, E); position p of the sorted order holds flat index i = token * topk + k.
//   sorted_ids[max_padded]   flat indices grouped by expert, each expert's run padded to a multiple of 128 with `n`
//   tile_expert[max_padded/128]  expert of every 128-row block, -1 past the last used block
//   inv_pos[n]               position of flat index i in sorted_ids
//   num_post_pad[0]          rows in use (multiple of 128)
// (moe_align_block_size of the reference, sgl-kernel/csrc/moe/moe_align_kernel.cu, with block_size = 128; the order
// of the tokens inside one expert's run is not specified there either -- nothing downstream depends on it.)
__global__ void __launch_bounds__(1024) moe_align128_kernel(const int32_t* __restrict__ topk_ids, int n, int E,
                                                            int32_t* __restrict__ sorted_ids,
                                                            int32_t* __restrict__ tile_expert,
                                                            int32_t* __restrict__ inv_pos,
                                                            int32_t* __restrict__ num_post_pad, int max_padded) {
  extern __shared__ int32_t sm[];
  int32_t* counts = sm;          // [E]
  int32_t* offsets = sm + E;     // [E + 1] padded exclusive prefix
  int32_t* cursor = offsets + E + 1;  // [E]
  pdl_launch_dependents();
  pdl_wait_prior_grids();
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    counts[e] = 0;
    cursor[e] = 0;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int e = topk_ids[i];
    if (e >= 0 && e < E) atomicAdd(&counts[e], 1);
  }
  for (int i = threadIdx.x; i < max_padded; i += blockDim.x) sorted_ids[i] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int e = 0; e < E; ++e) {
      offsets[e] = acc;
      acc += (counts[e] + 127) / 128 * 128;
    }
    offsets[E] = acc;
    num_post_pad[0] = acc;
  }
  __syncthreads();
  const int used_tiles = offsets[E] / 128;
  for (int t = threadIdx.x; t < max_padded / 128; t += blockDim.x) {
    int e = -1;
    if (t < used_tiles) {
      // expert whose padded range holds row 128 t: binary search over the prefix
      int lo = 0, hi = E;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (offsets[mid] <= t * 128) lo = mid;
        else hi = mid;
      }
      // skip experts without tokens (empty ranges share an offset with their successor)
      while (lo + 1 <= E - 1 && offsets[lo + 1] <= t * 128) ++lo;
      e = lo;
    }
    tile_expert[t] = e;
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int e = topk_ids[i];
    if (e >= 0 && e < E) {
      const int pos = offsets[e] + atomicAdd(&cursor[e], 1);
      sorted_ids[pos] = i;
      inv_pos[i] = pos;
    } else {
      inv_pos[i] = -1;
    }
  }
}

cudaError_t moe_align128(const int32_t* topk_ids, int n, int E, int32_t* sorted_ids, int32_t* tile_expert,
                         int32_t* inv_pos, int32_t* num_post_pad, int max_padded, cudaStream_t st) {
  const size_t smem = static_cast<size_t>(3 * E + 1) * sizeof(int32_t);
  moe_align128_kernel<<<1, 1024, smem, st>>>(topk_ids, n, E, sorted_ids, tile_expert, inv_pos, num_post_pad,
                                             max_padded);
  return cudaGetLastError();
}

// out[t, :] = T( sum_k w[t,k] * f32(y[inv_pos[t*topk + k], :]) ), k ascending, fp32 accumulation
// (moe_int8.cpp stage 3 / sum_stub: fp32 sum of the weighted rows, one rounding).
template <typename T>
__global__ void __launch_bounds__(256) moe_combine_kernel(const T* __restrict__ y, int64_t ldy,
                                                          const int32_t* __restrict__ inv_pos,
                                                          const float* __restrict__ w, T* __restrict__ out,
                                                          int64_t ldo, int topk, int H) {
  pdl_launch_dependents();
  pdl_wait_prior_grids();
  const int64_t t = blockIdx.x;
  for (int v = threadIdx.x; v < H / 8; v += blockDim.x) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int k = 0; k < topk; ++k) {
      const int pos = inv_pos[t * topk + k];
      if (pos < 0) continue;
      const float wk = w[t * topk + k];
      const uint4 raw = *(reinterpret_cast<const uint4*>(y + static_cast<int64_t>(pos) * ldy) + v);
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = __fmaf_rn(wk, to_f32(e[j]), acc[j]);
    }
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if constexpr (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) oe[j] = __float2bfloat16_rn(acc[j]);
      else oe[j] = __float2half_rn(acc[j]);
    }
    *(reinterpret_cast<uint4*>(out + t * ldo) + v) = o;
  }
}

cudaError_t moe_combine(const void* y, int64_t ldy, const int32_t* inv_pos, const float* w, void* out, int64_t ldo,
                        int tokens, int topk, int H, int dtype, cudaStream_t st) {
  if (tokens == 0) return cudaSuccess;
  if (dtype == B2_DTYPE_BF16)
    moe_combine_kernel<__nv_bfloat16><<<tokens, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(y), ldy, inv_pos, w,
                                                             static_cast<__nv_bfloat16*>(out), ldo, topk, H);
  else
    moe_combine_kernel<__half><<<tokens, 256, 0, st>>>(static_cast<const __half*>(y), ldy, inv_pos, w,
                                                       static_cast<__half*>(out), ldo, topk, H);
  return cudaGetLastError();
}

}  // namespace b2
