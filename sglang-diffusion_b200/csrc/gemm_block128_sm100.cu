// Block-wise (1x128 activation scales, 128x128 weight scales) e4m3 GEMM for sm_100a.
//
//   D[m,n] = out_t( sum_kb SFA[m,kb] * SFB[n/128,kb] * f32(A[m,kb*128:+128] . W[n,kb*128:+128]) [+ bias] )
//
// Reference op: fp8_blockwise_scaled_mm, sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:368-465
// (sm100 CUTLASS kernel :39-196; no bias there -- fp8_utils.py:546-547 adds it in a second kernel).
//
// Same warp roles and TMA ring as gemm_sm100.cu, but the 128-deep partial product of every
// K block lands in one of FOUR 128-column TMEM slots (non-accumulating MMA), and the four
// epilogue warps promote it into fp32 registers with the block's scale while the MMA warp
// is already filling the next slots: the tensor pipe never waits for the rescale as long
// as the epilogue keeps up (128 FFMA per thread per K block).
#include "gemm_common.cuh"

namespace b2 {

constexpr int BW_BN = 128;
constexpr int BW_SLOTS = 4;

struct BlockwiseParams {
  const float* sfa;
  const float* sfb;
  int64_t sfa_m_stride, sfa_k_stride, sfb_n_stride, sfb_k_stride;
  const void* bias;
  int bias_dtype;
  int out_dtype;
  int M, N, K;
  int num_m_blocks, num_n_blocks, num_kb;
  int stages;
  int group_m;
  int a_box_rows;
};

B2_DEVICE void bw_tile_coords(int tile, const BlockwiseParams& p, int& m_blk, int& n_blk) {
  const int band_tiles = p.group_m * p.num_n_blocks;
  const int band = tile / band_tiles;
  const int in_band = tile - band * band_tiles;
  const int m_first = band * p.group_m;
  const int rows_in_band = min(p.group_m, p.num_m_blocks - m_first);
  m_blk = m_first + in_band % rows_in_band;
  n_blk = in_band / rows_in_band;
}

template <int OUT_BYTES>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    gemm_block128_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const __grid_constant__ CUtensorMap tmap_d, const BlockwiseParams p) {
  constexpr int BN = BW_BN;
  constexpr uint32_t A_STAGE_BYTES = BM * BK;
  constexpr uint32_t B_STAGE_BYTES = BN * BK;
  constexpr uint32_t TMEM_COLS = BW_SLOTS * BN;  // 512
  constexpr int COLS_PER_BOX = 128 / OUT_BYTES;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stages = p.stages;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + static_cast<size_t>(stages) * A_STAGE_BYTES;
  uint8_t* smem_epi = smem_b + static_cast<size_t>(stages) * B_STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_epi + 4 * 2 * EPI_BUF_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* part_full = empty_bar + stages;
  uint64_t* part_empty = part_full + BW_SLOTS;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(part_empty + BW_SLOTS);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);  // warp-uniform (see gemm_sm100.cu)
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_d);
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < BW_SLOTS; ++i) {
      mbar_init(&part_full[i], 1);
      mbar_init(&part_empty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<1>(tmem_ptr, TMEM_COLS);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // TMA producer: whole warp walks the ring, one elected lane issues
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t tx_bytes = static_cast<uint32_t>(p.a_box_rows) * BK + B_STAGE_BYTES;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      bw_tile_coords(tile, p, m_blk, n_blk);
      for (int kb = 0; kb < p.num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
          tma_load_2d(smem_a + static_cast<size_t>(stage) * A_STAGE_BYTES, &tmap_a, &full_bar[stage], kb * BK,
                      m_blk * BM, kEvictNormal);
          tma_load_2d(smem_b + static_cast<size_t>(stage) * B_STAGE_BYTES, &tmap_b, &full_bar[stage], kb * BK,
                      n_blk * BN, kEvictNormal);
        }
        __syncwarp();
        if (++stage == stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // MMA issuer: whole warp in the loop, the same elected lane issues every MMA and commit
    constexpr uint32_t idesc = make_idesc<MmaKind::F8>(BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int slot = 0;
    uint32_t slot_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < p.num_kb; ++kb) {
        mbar_wait(&part_empty[slot], slot_phase ^ 1);
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(slot * BN);
        const uint64_t da = make_smem_desc_sw128(smem_u32(smem_a + static_cast<size_t>(stage) * A_STAGE_BYTES));
        const uint64_t db = make_smem_desc_sw128(smem_u32(smem_b + static_cast<size_t>(stage) * B_STAGE_BYTES));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma<MmaKind::F8, 1>(d_tmem, da + static_cast<uint64_t>(k * (UMMA_K >> 4)),
                                 db + static_cast<uint64_t>(k * (UMMA_K >> 4)), idesc, k != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          umma_commit(&part_full[slot]);
        }
        __syncwarp();
        if (++stage == stages) {
          stage = 0;
          phase ^= 1;
        }
        if (++slot == BW_SLOTS) {
          slot = 0;
          slot_phase ^= 1;
        }
      }
    }
  } else {
    const int quarter = warp & 3;
    uint8_t* my_epi = smem_epi + static_cast<size_t>(warp - EPI_WARP0) * 2 * EPI_BUF_BYTES;
    const bool has_bias = p.bias != nullptr;
    int slot = 0;
    uint32_t slot_phase = 0;
    int buf = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      bw_tile_coords(tile, p, m_blk, n_blk);
      const int row0 = m_blk * BM + quarter * 32;
      const int row = row0 + lane;
      const int col_tile = n_blk * BN;
      const bool row_ok = row < p.M;
      const float* sfa_row = p.sfa + static_cast<int64_t>(row_ok ? row : 0) * p.sfa_m_stride;
      const float* sfb_row = p.sfb + static_cast<int64_t>(n_blk) * p.sfb_n_stride;

      float acc[BN];
#pragma unroll
      for (int i = 0; i < BN; ++i) acc[i] = 0.f;

      float sa_next = row_ok ? __ldg(sfa_row) : 0.f;
      float sb_next = __ldg(sfb_row);
      for (int kb = 0; kb < p.num_kb; ++kb) {
        const float scale = __fmul_rn(sa_next, sb_next);
        if (kb + 1 < p.num_kb) {
          sa_next = row_ok ? __ldg(sfa_row + static_cast<int64_t>(kb + 1) * p.sfa_k_stride) : 0.f;
          sb_next = __ldg(sfb_row + static_cast<int64_t>(kb + 1) * p.sfb_k_stride);
        }
        mbar_wait(&part_full[slot], slot_phase);
        tc_fence_after_sync();
        const uint32_t t_row =
            tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(slot * BN);
        // 4 x 32 columns, software-pipelined: the load of chunk c+1 is in flight while chunk c is promoted
        uint32_t r0[32], r1[32];
        tmem_ld_32x32b_x32(t_row, r0);
        tmem_ld_wait();
        tmem_ld_32x32b_x32(t_row + 32, r1);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = __fmaf_rn(__uint_as_float(r0[j]), scale, acc[j]);
        tmem_ld_wait();
        tmem_ld_32x32b_x32(t_row + 64, r0);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[32 + j] = __fmaf_rn(__uint_as_float(r1[j]), scale, acc[32 + j]);
        tmem_ld_wait();
        tmem_ld_32x32b_x32(t_row + 96, r1);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[64 + j] = __fmaf_rn(__uint_as_float(r0[j]), scale, acc[64 + j]);
        tmem_ld_wait();
        // slot fully read -> give it back before the last 32 FMAs
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&part_empty[slot]);
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[96 + j] = __fmaf_rn(__uint_as_float(r1[j]), scale, acc[96 + j]);
        if (++slot == BW_SLOTS) {
          slot = 0;
          slot_phase ^= 1;
        }
      }

      // ---- write the tile: bias, cast, swizzled staging, per-warp TMA store ----
      constexpr int NBOX = BN / COLS_PER_BOX;
#pragma unroll
      for (int bx = 0; bx < NBOX; ++bx) {
        const int col_box = col_tile + bx * COLS_PER_BOX;
        uint8_t* stage_buf = my_epi + buf * EPI_BUF_BYTES;
        tma_store_wait_read<1>();
        __syncwarp();
#pragma unroll
        for (int g = 0; g < COLS_PER_BOX / 8; ++g) {
          const int col = col_box + g * 8;
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = acc[bx * COLS_PER_BOX + g * 8 + j];
          if (has_bias && col < p.N) {
            float bi[8];
            load_bias8(p.bias, p.bias_dtype, col, bi);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __fadd_rn(v[j], bi[j]);
          }
          if constexpr (OUT_BYTES == 2) {
            uint4 o;
            o.x = pack16(v[0], v[1], p.out_dtype);
            o.y = pack16(v[2], v[3], p.out_dtype);
            o.z = pack16(v[4], v[5], p.out_dtype);
            o.w = pack16(v[6], v[7], p.out_dtype);
            const int chunk = g ^ (lane & 7);
            *reinterpret_cast<uint4*>(stage_buf + lane * 128 + chunk * 16) = o;
          } else {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              uint4 o;
              o.x = __float_as_uint(v[4 * c + 0]);
              o.y = __float_as_uint(v[4 * c + 1]);
              o.z = __float_as_uint(v[4 * c + 2]);
              o.w = __float_as_uint(v[4 * c + 3]);
              const int chunk = (g * 2 + c) ^ (lane & 7);
              *reinterpret_cast<uint4*>(stage_buf + lane * 128 + chunk * 16) = o;
            }
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (row0 < p.M && col_box < p.N && elect_one()) {
          tma_store_2d(&tmap_d, stage_buf, col_box, row0);
          tma_store_commit();
        }
        buf ^= 1;
      }
    }
    tma_store_wait_all<0>();
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<1>(tmem_base, TMEM_COLS);
  }
}

template <int OUT_BYTES>
static int launch_bw(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const BlockwiseParams& p,
                     int grid, size_t smem, cudaStream_t st, char* err, size_t err_len) {
  auto kern = gemm_block128_kernel<OUT_BYTES>;
  static int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
    if (e != cudaSuccess) {
      snprintf(err, err_len, "cudaFuncSetAttribute(max dynamic smem): %s", cudaGetErrorString(e));
      return B2_ERR_CUDA;
    }
    attr_dev = dev;
  }
  kern<<<grid, NUM_THREADS, smem, st>>>(ta, tb, td, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(err, err_len, "block-wise gemm kernel launch: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  count_launch();
  return B2_OK;
}

int gemm_launch_blockwise(const GemmArgs& a, int forced_config, const char** used_config, cudaStream_t st,
                          char* err, size_t err_len) {
  (void)forced_config;
  if (used_config) *used_config = "1cta_128x128_block128";
  const int sms = device_sm_count();
  const int out_bytes = (a.out_dtype == B2_DTYPE_BF16 || a.out_dtype == B2_DTYPE_F16) ? 2 : 4;
  CUtensorMapDataType d_dt = a.out_dtype == B2_DTYPE_BF16  ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                             : a.out_dtype == B2_DTYPE_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                           : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  BlockwiseParams p;
  p.sfa = a.As;
  p.sfb = a.Bs;
  p.sfa_m_stride = a.sfa_m_stride;
  p.sfa_k_stride = a.sfa_k_stride;
  p.sfb_n_stride = a.sfb_n_stride;
  p.sfb_k_stride = a.sfb_k_stride;
  p.bias = a.bias;
  p.bias_dtype = a.bias_dtype;
  p.out_dtype = a.out_dtype;
  p.M = a.M;
  p.N = a.N;
  p.K = a.K;
  p.num_m_blocks = (a.M + BM - 1) / BM;
  p.num_n_blocks = (a.N + BW_BN - 1) / BW_BN;
  p.num_kb = (a.K + BK - 1) / BK;
  const int fixed = 4 * 2 * EPI_BUF_BYTES + 1024 + 512;
  int stages = (SMEM_LIMIT - fixed) / (BM * BK + BW_BN * BK);
  if (stages > 8) stages = 8;
  p.stages = stages;
  p.group_m = 16;
  p.a_box_rows = a.M >= BM ? BM : ((a.M + 7) / 8) * 8;

  CUtensorMap ta, tb, td;
  if (!make_tmap_2d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.A, a.M, a.K, a.lda, p.a_box_rows, BK) ||
      !make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.W, a.N, a.K, a.ldw, BW_BN, BK) ||
      !make_tmap_2d(&td, d_dt, out_bytes, a.D, a.M, a.N, a.ldd * out_bytes, 32, 128 / out_bytes)) {
    snprintf(err, err_len, "cuTensorMapEncodeTiled failed (block-wise M=%d N=%d K=%d)", a.M, a.N, a.K);
    return B2_ERR_CUDA;
  }
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  const int grid = tiles < sms ? tiles : sms;
  const size_t smem = static_cast<size_t>(stages) * (BM * BK + BW_BN * BK) + fixed;
  if (out_bytes == 2) return launch_bw<2>(ta, tb, td, p, grid, smem, st, err, err_len);
  return launch_bw<4>(ta, tb, td, p, grid, smem, st, err, err_len);
}

}  // namespace b2
