// Device + host helpers shared by the tcgen05 GEMM translation units.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cudaTypedefs.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "b200_w8a8_internal.h"
#include "ptx.cuh"

namespace b2 {

constexpr int BM = 128;          // rows per CTA tile (= UMMA M, TMEM lanes)
constexpr int BK = 128;          // bytes of K per pipeline stage (= one 128 B swizzle row)
constexpr int UMMA_K = 32;       // bytes of K per tcgen05.mma for 8-bit kinds
constexpr int NUM_THREADS = 192;
constexpr int EPI_WARP0 = 2;
constexpr int EPI_BUF_BYTES = 32 * 128;  // one warp's staging box: 32 rows x 128 B
constexpr int SMEM_LIMIT = 232448;       // 227 KB opt-in limit per CTA

struct KernelParams {
  const float* As;
  const float* Bs;
  const void* bias;
  int bias_dtype;
  int out_dtype;
  const int32_t* tile_expert;  // grouped (MoE) mode: expert of every 128-row block of A (-1 = skip), else nullptr
  int batch;                   // batched mode (bmm_fp8): > 1 = the tensor maps are 3-D, tiles enumerate batch-major
  int scalar_scales;           // As / Bs point at ONE fp32 value each (per-tensor scales: bmm_fp8)
  // scatter epilogue (row-parallel prefill, RS kernels): rows [r * rs_rows, (r + 1) * rs_rows) of the output belong to
  // rank r and are stored into slot `rs_rank` of set (*rs_counter & 1) of rank r's gather buffer
  const uint32_t* rs_counter;
  int rs_rows, rs_rank, rs_rot;
  int tail_tiles;              // BN = 256: tiles of the ragged last wave that are issued as two half-width work items
  int M, N, K;
  int num_m_blocks, num_n_blocks, num_kb;
  int stages;
  int group_m;     // tile rows (or columns, when band_n) per raster band
  int band_n;      // 0: bands of `group_m` row-blocks sweeping N;  1: bands of `group_m` column-blocks sweeping M
  uint64_t policy_a, policy_b;  // L2 eviction policy of the operand loads (the band-resident operand is kept)
  int a_box_rows;  // rows of A actually fetched per stage (<= BM; small-M shapes fetch fewer)
  // optional per-CTA wait-cycle counters (b2_set_profile_buffer): [cta][0]=MMA waits for operands,
  // [1]=MMA waits for a free accumulator, [2]=producer waits for a free stage, [3]=epilogue waits
  // for an accumulator (warp 2), [4]=CTA lifetime, [5]=epilogue waits for a staging buffer (TMA store read),
  // [6]=epilogue waits for tcgen05.ld, [7]=epilogue tile bodies in total; all in SM clocks
  unsigned long long* prof;
  int debug_flags;  // developer experiments (env B2_GEMM_DEBUG): 1 = skip the MMAs, 2 = skip the epilogue body, 4 = skip the loads, 8 = no tail split
};

B2_DEVICE void tile_coords(int tile, const KernelParams& p, int& m_blk, int& n_blk) {
  // Raster in bands.  band_n == 0: a band is `group_m` row-blocks wide and sweeps all of N, row-block fastest:
  // the band's slice of A stays in L2 while W streams through once per band.  band_n == 1 is the transpose
  // (the band's slice of W is kept, A streams).  The host picks orientation and width from the operand sizes
  // so that the DRAM traffic is as close to "every operand byte once" as the 126 MB L2 allows.
  const int fast_blocks = p.band_n ? p.num_n_blocks : p.num_m_blocks;
  const int slow_blocks = p.band_n ? p.num_m_blocks : p.num_n_blocks;
  const int band_tiles = p.group_m * slow_blocks;
  const int band = tile / band_tiles;
  const int in_band = tile - band * band_tiles;
  const int first = band * p.group_m;
  const int in_band_width = min(p.group_m, fast_blocks - first);
  const int f = first + in_band % in_band_width;
  const int s = in_band / in_band_width;
  m_blk = p.band_n ? s : f;
  n_blk = p.band_n ? f : s;
}

// Host: choose the raster (see tile_coords).  tile_m / tile_n are the rows of A / W one tile covers.
static inline void choose_raster(long long M, long long N, long long K, int tile_m, int tile_n, int num_m_blocks,
                                 int num_n_blocks, int* group, int* band_n, uint64_t* policy_a, uint64_t* policy_b) {
  const double l2_keep = 40e6;  // bytes of one operand we try to keep resident (the output stream shares the L2)
  const double a_bytes = static_cast<double>(M) * K, w_bytes = static_cast<double>(N) * K;
  int gm = static_cast<int>(l2_keep / (static_cast<double>(tile_m) * K));
  int gn = static_cast<int>(l2_keep / (static_cast<double>(tile_n) * K));
  gm = gm < 1 ? 1 : (gm > num_m_blocks ? num_m_blocks : gm);
  gn = gn < 1 ? 1 : (gn > num_n_blocks ? num_n_blocks : gn);
  const double cost_m = a_bytes + w_bytes * ((num_m_blocks + gm - 1) / gm);  // bands over M: W re-streamed per band
  const double cost_n = w_bytes + a_bytes * ((num_n_blocks + gn - 1) / gn);
  bool along_m = cost_m <= cost_n;
  {  // developer override for raster experiments: B2_GEMM_RASTER="<band_n>,<group>"
    static const char* e = getenv("B2_GEMM_RASTER");
    int bn_ = 0, g_ = 0;
    if (e && sscanf(e, "%d,%d", &bn_, &g_) == 2 && g_ >= 1) {
      along_m = bn_ == 0;
      gm = g_ > num_m_blocks ? num_m_blocks : g_;
      gn = g_ > num_n_blocks ? num_n_blocks : g_;
    }
  }
  if (along_m) {
    *band_n = 0;
    *group = gm;
    *policy_a = kEvictLast;
    *policy_b = kEvictNormal;
  } else {
    *band_n = 1;
    *group = gn;
    *policy_a = kEvictNormal;
    *policy_b = kEvictLast;
  }
}

B2_DEVICE float load_bias(const void* bias, int dtype, int col) {
  if (dtype == B2_DTYPE_BF16) return __bfloat162float(__ldg(static_cast<const __nv_bfloat16*>(bias) + col));
  if (dtype == B2_DTYPE_F16) return __half2float(__ldg(static_cast<const __half*>(bias) + col));
  return __ldg(static_cast<const float*>(bias) + col);
}

// 8 consecutive per-column values starting at `col` (col % 8 == 0, col + 8 <= N guaranteed by N % 8 == 0)
B2_DEVICE void load_cols8_f32(const float* p, int col, float (&v)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p + col));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p + col) + 1);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

B2_DEVICE void load_bias8(const void* bias, int dtype, int col, float (&v)[8]) {
  if (dtype == B2_DTYPE_F32) {
    load_cols8_f32(static_cast<const float*>(bias), col, v);
  } else {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(bias) + col));
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (dtype == B2_DTYPE_BF16) {
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
      } else {
        const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
      }
    }
  }
}

// One output element.  `bias` is the bias value or -0.0f when the op has no bias: fma(t, bs, -0) and t + (-0)
// round exactly like the bias-free forms (t * bs, t) including the sign of zero, so one code path serves both
// and stays bit-identical to the reference's two.
template <MmaKind KIND>
B2_DEVICE float epilogue_value(uint32_t raw, float as, float bs, float bias) {
  if constexpr (KIND == MmaKind::I8) {
    // gemm_int8.cpp:40-46: fmadd(mul(cvtepi32_ps(acc), As), Bs, bias)  |  mul(mul(.., As), Bs)
    const float c = __int2float_rn(static_cast<int32_t>(raw));
    return __fmaf_rn(__fmul_rn(c, as), bs, bias);
  } else {
    // fp8_gemm_kernel.cu sm100 EVT: As * (Bs * acc) [+ bias]
    const float c = __uint_as_float(raw);
    return __fadd_rn(__fmul_rn(as, __fmul_rn(bs, c)), bias);
  }
}

// OUT_KIND 0 = bf16, 1 = f16 (compile-time: a run-time dtype doubles every pack in the epilogue's hot loop)
template <int OUT_KIND>
B2_DEVICE uint32_t pack16(float lo, float hi) {
  if constexpr (OUT_KIND == 0) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
  } else {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
  }
}

B2_DEVICE uint32_t pack16(float lo, float hi, int out_dtype) {
  if (out_dtype == B2_DTYPE_BF16) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
  }
  const __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// 2-D row-major tensor [rows, cols] of `esize`-byte elements, row stride `ld_bytes`;
// box = {box_cols elements (128 bytes), box_rows}, 128-byte swizzle.
static inline bool make_tmap_2d(CUtensorMap* m, CUtensorMapDataType dt, int esize, const void* base, uint64_t rows,
                         uint64_t cols, uint64_t ld_bytes, uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  const uint32_t box_bytes = box_cols * static_cast<uint32_t>(esize);
  const CUtensorMapSwizzle swz = box_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                                 : box_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                   : CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swz == CU_TENSOR_MAP_SWIZZLE_NONE) return false;
  CUresult r = enc(m, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// 3-D tensor [batch, rows, cols] (batch stride `batch_bytes`); box = {box_cols, box_rows, 1}: the same shared-memory
// image as the 2-D box, with per-batch clipping of rows (a tile never reads or writes into the next batch entry).
static inline bool make_tmap_3d(CUtensorMap* m, CUtensorMapDataType dt, int esize, const void* base, uint64_t batch,
                                uint64_t rows, uint64_t cols, uint64_t ld_bytes, uint64_t batch_bytes, uint32_t box_rows,
                                uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return false;
  cuuint64_t dims[3] = {cols, rows, batch};
  cuuint64_t strides[2] = {ld_bytes, batch > 1 ? batch_bytes : ld_bytes * rows};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  const uint32_t box_bytes = box_cols * static_cast<uint32_t>(esize);
  const CUtensorMapSwizzle swz = box_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                                 : box_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                   : CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swz == CU_TENSOR_MAP_SWIZZLE_NONE) return false;
  CUresult r = enc(m, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// [2 sets][world slots][rows][cols] gather buffer of one rank (16-bit elements): dims (cols, rows, slots, 2)
static inline bool make_tmap_gather4d(CUtensorMap* m, CUtensorMapDataType dt, const void* base, uint64_t slots,
                                      uint64_t rows, uint64_t cols, uint64_t set_bytes, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return false;
  cuuint64_t dims[4] = {cols, rows, slots, 2};
  cuuint64_t strides[3] = {cols * 2, cols * 2 * rows, set_bytes};
  cuuint32_t box[4] = {64, box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, dt, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// per-peer output maps of the scatter epilogue; the plain kernels carry an empty stand-in
template <bool RS>
struct PeerMaps {
  CUtensorMap m[8];
};
template <>
struct PeerMaps<false> {
  int unused;
};

}  // namespace b2
