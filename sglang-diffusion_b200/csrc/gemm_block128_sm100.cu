// Block-wise (1x128 activation scales, 128x128 weight scales) e4m3 GEMM for sm_100a.
//
//   D[m,n] = out_t( sum_kb SFA[m,kb] * SFB[n/128,kb] * f32(A[m,kb*128:+128] . W[n,kb*128:+128]) [+ bias] )
//
// Reference op: fp8_blockwise_scaled_mm, sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:368-465
// (sm100 CUTLASS kernel :39-196; no bias there -- fp8_utils.py:546-547 adds it in a second kernel).
//
// The per-block scales are fp32 (not powers of two), so the hardware block-scaled MMA cannot reproduce the
// reference arithmetic; the rescale is done on the CUDA cores instead, overlapped with the tensor pipe:
//   warp 0     TMA producer  (same ring as gemm_sm100.cu, whole warp walks it, one elected lane issues)
//   warp 1     MMA issuer    -- the 128-deep partial product of every K block lands in one of FOUR 128-column
//                               TMEM slots with a non-accumulating MMA chain; tcgen05.commit -> part_full[slot]
//   warps 2-9  promotion     -- EIGHT warps (two per TMEM lane quarter, 64 columns each): tcgen05.ld the slot,
//                               acc += (SFA[m,kb]*SFB[nb,kb]) * partial in fp32 registers, hand the slot back
//                               (part_empty), and at the end of the tile add bias, cast, stage and TMA-store
// CG = 2: the two CTAs of a cluster share one 256 x 128 tile (cta_group::2): each CTA stages its own 128 rows of
// A and half of the W tile, and promotes its own 128 x 128 half of the partial product.
//
// The barrier poll for the next K block and the loads of the scales four K blocks ahead are issued before the
// FMAs of the current block, so their latencies hide behind the arithmetic.
#include <cstdlib>

#include "gemm_common.cuh"

namespace b2 {

constexpr int BW_BN = 128;
constexpr int BW_SLOTS = 4;
constexpr int BW_EPI_WARPS = 8;
constexpr int BW_SC = 16;  // K blocks per shared-memory scale chunk
constexpr int BW_THREADS = 64 + 32 * BW_EPI_WARPS;  // 320

struct BlockwiseParams {
  const float* sfa;
  const float* sfb;
  int64_t sfa_m_stride, sfa_k_stride, sfb_n_stride, sfb_k_stride;
  const void* bias;
  int bias_dtype;
  int M, N, K;
  int num_m_blocks, num_n_blocks, num_kb;
  int stages;
  int group_m;
  int band_n;
  uint64_t policy_a, policy_b;
  int a_box_rows;
  int ksplit;   // CG == 1 only: K split over the CTAs of a {ksplit,1,1} cluster (decode shapes), 1 = off
  void* D;      // narrow tiles (BN < 128) store straight from registers
  int64_t ldd;  // elements
  unsigned long long* prof;  // see KernelParams::prof ([0] MMA/full [1] MMA/slot [2] producer [3] promotion/partial [4] life)
  int debug_flags;           // 1 = no MMAs, 2 = no promotion arithmetic (handshakes only)
};

// acc[0..n) += x[0..n) * scale with the packed fp32x2 FMA of sm_100 (two IEEE fma.rn per issue slot; the result of
// each lane is exactly that of a scalar fma.rn, so the promotion arithmetic is unchanged).  The promotion warps are
// issue-bound -- 64 FMAs per thread per K block against a 256-clock MMA -- and this halves their FMA slots.
template <int N>
B2_DEVICE void promote_fma2(float* acc, int base, const uint32_t* x, float scale) {
  uint64_t s2;
  asm("mov.b64 %0, {%1, %1};" : "=l"(s2) : "f"(scale));
#pragma unroll
  for (int j = 0; j < N; j += 2) {
    asm("{\n\t.reg .b64 xx, cc;\n\t"
        "mov.b64 xx, {%2, %3};\n\t"
        "mov.b64 cc, {%0, %1};\n\t"
        "fma.rn.f32x2 cc, xx, %4, cc;\n\t"
        "mov.b64 {%0, %1}, cc;\n\t}"
        : "+f"(acc[base + j]), "+f"(acc[base + j + 1])
        : "r"(x[j]), "r"(x[j + 1]), "l"(s2));
  }
}

B2_DEVICE void st_dsmem_v4(uint32_t local_addr, uint32_t cta, float a, float b, float c, float d) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "st.shared::cluster.v4.f32 [ra], {%2, %3, %4, %5};\n\t}" ::"r"(local_addr),
      "r"(cta), "f"(a), "f"(b), "f"(c), "f"(d)
      : "memory");
}

B2_DEVICE void bw_tile_coords(int tile, const BlockwiseParams& p, int& m_blk, int& n_blk) {
  const int fast_blocks = p.band_n ? p.num_n_blocks : p.num_m_blocks;
  const int slow_blocks = p.band_n ? p.num_m_blocks : p.num_n_blocks;
  const int band_tiles = p.group_m * slow_blocks;
  const int band = tile / band_tiles;
  const int in_band = tile - band * band_tiles;
  const int first = band * p.group_m;
  const int width = min(p.group_m, fast_blocks - first);
  const int f = first + in_band % width;
  const int s = in_band / width;
  m_blk = p.band_n ? s : f;
  n_blk = p.band_n ? f : s;
}

// BN = 128: the prefill tile.  BN = 64 / 32 (CG = 1 only): decode shapes -- a 128-wide tile leaves N/128 CTAs for the
// whole weight matrix (32 of 148 SMs for N = 4096); narrow tiles inside ONE 128-column scale block put a CTA on every
// SM.  Each promotion warp then owns BN/2 columns and stores them straight from registers.
template <int OUT_KIND, int CG, int BN>
__global__ void __launch_bounds__(BW_THREADS, 1)
    gemm_block128_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const __grid_constant__ CUtensorMap tmap_d, const BlockwiseParams p) {
  constexpr int HC = BN / 2;  // columns per promotion warp
  static_assert(BN == 128 || (CG == 1 && (BN == 64 || BN == 32)), "narrow tiles are 1-CTA");
  constexpr int B_ROWS = BN / CG;
  constexpr uint32_t A_STAGE_BYTES = BM * BK;
  constexpr uint32_t B_STAGE_BYTES = B_ROWS * BK;
  constexpr uint32_t TMEM_COLS = BW_SLOTS * BN;  // 512

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int stages = p.stages;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + static_cast<size_t>(stages) * A_STAGE_BYTES;
  uint8_t* smem_epi = smem_b + static_cast<size_t>(stages) * B_STAGE_BYTES;  // 8 warps x 4 KB
  // K split: the partial sums of ranks 1..S-1, pushed into rank 0's shared memory ([S-1][256 threads][HC] fp32)
  float* smem_x = reinterpret_cast<float*>(smem_epi + BW_EPI_WARPS * EPI_BUF_BYTES);
  const int S = CG == 1 ? p.ksplit : 1;
  float* smem_sfa = smem_x + static_cast<size_t>(S - 1) * 256 * HC;  // [2][BW_SC][128]
  float* smem_sfb = smem_sfa + 2 * BW_SC * BM;                                           // [2][BW_SC]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_sfb + 2 * BW_SC);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* part_full = empty_bar + stages;
  uint64_t* part_empty = part_full + BW_SLOTS;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(part_empty + BW_SLOTS);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);  // warp-uniform (see gemm_sm100.cu)
  const int lane = threadIdx.x & 31;
  const long long t_start = p.prof ? clock64() : 0;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;  // tiles of (BM*CG) x BN
  const int cta_rank = CG == 1 ? 0 : static_cast<int>(cluster_ctarank());
  const int unit = CG == 1 ? static_cast<int>(blockIdx.x) / S : static_cast<int>(blockIdx.x >> 1);
  const int num_units = CG == 1 ? static_cast<int>(gridDim.x) / S : static_cast<int>(gridDim.x >> 1);
  const int krank = (CG == 1 && S > 1) ? static_cast<int>(cluster_ctarank()) : 0;
  const int kb0 = static_cast<int>(static_cast<long long>(krank) * p.num_kb / S);      // my K blocks: [kb0, kb1)
  const int kb1 = static_cast<int>(static_cast<long long>(krank + 1) * p.num_kb / S);

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
      mbar_init(&part_empty[i], BW_EPI_WARPS * CG);  // one arrival per promotion warp of every CTA of the cluster
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<CG>(tmem_ptr, TMEM_COLS);
  tc_fence_before_sync();
  if constexpr (CG == 1) {
    __syncthreads();
  } else {
    __syncwarp();
    cluster_sync_all();
  }
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================================================== TMA producer ====
    // two K blocks per step: lanes 0 / 1 wait for the two ring slots with ONE try_wait round
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t tx_bytes = (static_cast<uint32_t>(p.a_box_rows) * BK + B_STAGE_BYTES) * CG;
    long long prof_wait = 0;
    for (int tile = unit; tile < num_tiles; tile += num_units) {
      int m_blk, n_blk;
      bw_tile_coords(tile, p, m_blk, n_blk);
      const int m_row = (m_blk * CG + cta_rank) * BM;
      const int n_row = n_blk * BN + cta_rank * B_ROWS;
      for (int kb = kb0; kb < kb1; kb += 2) {
        const int n2 = kb + 1 < kb1 ? 2 : 1;
        const int stage1 = stage + 1 == stages ? 0 : stage + 1;
        const uint32_t phase1 = stage + 1 == stages ? phase ^ 1 : phase;
        const long long w0 = p.prof ? clock64() : 0;
        mbar_wait_lanes(&empty_bar[lane == 1 ? stage1 : stage], (lane == 1 ? phase1 : phase) ^ 1, lane < n2);
        if (p.prof) prof_wait += clock64() - w0;
        if (elect_one()) {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            if (j < n2) {
              const int st_j = j ? stage1 : stage;
              uint8_t* a_dst = smem_a + static_cast<size_t>(st_j) * A_STAGE_BYTES;
              uint8_t* b_dst = smem_b + static_cast<size_t>(st_j) * B_STAGE_BYTES;
              if constexpr (CG == 1) {
                mbar_arrive_expect_tx(&full_bar[st_j], tx_bytes);
                tma_load_2d(a_dst, &tmap_a, &full_bar[st_j], (kb + j) * BK, m_row, p.policy_a);
                tma_load_2d(b_dst, &tmap_b, &full_bar[st_j], (kb + j) * BK, n_row, p.policy_b);
              } else {
                if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[st_j], tx_bytes);
                const uint32_t bar0 = smem_u32(&full_bar[st_j]) & kPeerBitMask;
                tma_load_2d_cg2(smem_u32(a_dst), &tmap_a, bar0, (kb + j) * BK, m_row, p.policy_a);
                tma_load_2d_cg2(smem_u32(b_dst), &tmap_b, bar0, (kb + j) * BK, n_row, p.policy_b);
              }
            }
          }
        }
        __syncwarp();
        for (int j = 0; j < n2; ++j) {
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      if (S > 1) {  // the two cluster barriers of the tile's K-split reduction (see the promotion warps)
        __syncwarp();
        cluster_sync_all();
        cluster_sync_all();
      }
    }
    if (p.prof && lane == 0) p.prof[blockIdx.x * 8 + 2] = static_cast<unsigned long long>(prof_wait);
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    // two K blocks per step: lanes 0-1 wait for the two operand stages, lanes 2-3 for the two TMEM slots, in
    // ONE try_wait round (measured: 1966 vs 1812 TOPS against one block per step on 4096 x 36864 x 7168)
    if (cta_rank == 0) {
      constexpr uint32_t idesc = make_idesc<MmaKind::F8>(BM * CG, BN);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;  // running K-block counter: slot = it & 3, slot phase = (it >> 2) & 1
      long long prof_full = 0, prof_issue = 0;
      const int mma_step = (p.debug_flags & 4) ? 1 : 2;  // experiment: one K block per MMA hand-shake
      for (int tile = unit; tile < num_tiles; tile += num_units) {
        for (int kb = kb0; kb < kb1; kb += mma_step) {
          const int n2 = (mma_step == 2 && kb + 1 < kb1) ? 2 : 1;
          const int stage1 = stage + 1 == stages ? 0 : stage + 1;
          const uint32_t phase1 = stage + 1 == stages ? phase ^ 1 : phase;
          {
            const int j = lane & 1;
            const uint32_t itj = it + j;
            uint64_t* bar = lane < 2 ? &full_bar[j ? stage1 : stage] : &part_empty[itj & (BW_SLOTS - 1)];
            const uint32_t par = lane < 2 ? (j ? phase1 : phase) : (((itj >> 2) & 1) ^ 1);
            const long long w0 = p.prof ? clock64() : 0;
            mbar_wait_lanes(bar, par, lane < 4 && j < n2);
            if (p.prof) prof_full += clock64() - w0;
          }
          tc_fence_after_sync();
          const long long wi0 = p.prof ? clock64() : 0;
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              if (j < n2) {
                const int st_j = j ? stage1 : stage;
                const int slot = static_cast<int>((it + j) & (BW_SLOTS - 1));
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(slot * BN);
                const uint64_t da = make_smem_desc_sw128(smem_u32(smem_a + static_cast<size_t>(st_j) * A_STAGE_BYTES));
                const uint64_t db = make_smem_desc_sw128(smem_u32(smem_b + static_cast<size_t>(st_j) * B_STAGE_BYTES));
                if (!(p.debug_flags & 1)) {
#pragma unroll
                  for (int k = 0; k < BK / UMMA_K; ++k) {
                    umma<MmaKind::F8, CG>(d_tmem, da + static_cast<uint64_t>(k * (UMMA_K >> 4)),
                                          db + static_cast<uint64_t>(k * (UMMA_K >> 4)), idesc, k != 0 ? 1u : 0u);
                  }
                }
                if constexpr (CG == 1) {
                  umma_commit(&empty_bar[st_j]);
                  umma_commit(&part_full[slot]);
                } else {
                  umma_commit_cg2(&empty_bar[st_j], 0x3);
                  umma_commit_cg2(&part_full[slot], 0x3);
                }
              }
            }
          }
          __syncwarp();
          if (p.prof) prof_issue += clock64() - wi0;
          it += n2;
          for (int j = 0; j < n2; ++j) {
            if (++stage == stages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        if (S > 1) {
          __syncwarp();
          cluster_sync_all();
          cluster_sync_all();
        }
      }
      if (p.prof && lane == 0) {
        p.prof[blockIdx.x * 8 + 0] = static_cast<unsigned long long>(prof_full);
        p.prof[blockIdx.x * 8 + 1] = static_cast<unsigned long long>(prof_issue);
      }
    }
  } else {
    // ============================================== promotion + epilogue ====
    const int ew = warp - EPI_WARP0;  // 0..7
    const int quarter = warp & 3;     // TMEM lane quarter this warp may read
    const int half = ew >> 2;         // which 64 of the slot's 128 columns
    const int ptid = ew * 32 + lane;  // 0..255
    uint8_t* stage_buf = smem_epi + static_cast<size_t>(ew) * EPI_BUF_BYTES;
    const bool has_bias = p.bias != nullptr;
    uint32_t it = 0;
    long long prof_part = 0, prof_sc = 0, prof_ld = 0, prof_body = 0;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(half * HC);

    // Scales travel through shared memory in chunks of BW_SC K blocks (SFA: 128 rows x BW_SC, SFB: BW_SC),
    // double-buffered and fetched with cp.async one chunk ahead -- across tile boundaries too -- so no
    // register waits on a global load inside the promotion loop.
    auto issue_scales = [&](int m_blk_, int n_blk_, int chunk, int b) {
      const int m_base = (m_blk_ * CG + cta_rank) * BM;
      float* dst_a = smem_sfa + b * (BW_SC * BM);
      float* dst_b = smem_sfb + b * BW_SC;
      const int r = ptid & (BM - 1);
      const bool r_ok = m_base + r < p.M;
      const float* src_row = p.sfa + static_cast<int64_t>(r_ok ? m_base + r : 0) * p.sfa_m_stride;
#pragma unroll
      for (int i = 0; i < BW_SC / 2; ++i) {
        const int kl = (ptid >> 7) + 2 * i;
        const int kb = kb0 + chunk * BW_SC + kl;
        if (r_ok && kb < kb1) cp_async_4(dst_a + kl * BM + r, src_row + static_cast<int64_t>(kb) * p.sfa_k_stride);
      }
      if (ptid < BW_SC) {
        const int kb = kb0 + chunk * BW_SC + ptid;
        if (kb < kb1)
          cp_async_4(dst_b + ptid, p.sfb + static_cast<int64_t>(n_blk_ * BN / 128) * p.sfb_n_stride +
                                       static_cast<int64_t>(kb) * p.sfb_k_stride);
      }
      cp_async_commit();
    };
    const int nkb = kb1 - kb0;
    const int nchunks = (nkb + BW_SC - 1) / BW_SC;
    int sbuf = 0;
    if (unit < num_tiles) {
      int m0, n0;
      bw_tile_coords(unit, p, m0, n0);
      issue_scales(m0, n0, 0, 0);
    }
    for (int tile = unit; tile < num_tiles; tile += num_units) {
      int m_blk, n_blk;
      bw_tile_coords(tile, p, m_blk, n_blk);
      const int row0 = (m_blk * CG + cta_rank) * BM + quarter * 32;
      const int col0 = n_blk * BN + half * HC;

      float acc[HC];
#pragma unroll
      for (int i = 0; i < HC; ++i) acc[i] = 0.f;
      uint32_t r0[32], r1[32];

      for (int chunk = 0; chunk < nchunks; ++chunk) {
        {
          const long long w0 = p.prof ? clock64() : 0;
          cp_async_wait<0>();  // my copies of this chunk have landed ...
          promo_bar_sync();    // ... so have everybody's, and everybody is done with the other buffer
          if (p.prof) prof_sc += clock64() - w0;
        }
        if (chunk + 1 < nchunks) {
          issue_scales(m_blk, n_blk, chunk + 1, sbuf ^ 1);
        } else if (tile + num_units < num_tiles) {
          int m1, n1;
          bw_tile_coords(tile + num_units, p, m1, n1);
          issue_scales(m1, n1, 0, sbuf ^ 1);
        }
        const float* sa_s = smem_sfa + sbuf * (BW_SC * BM) + quarter * 32 + lane;
        const float* sb_s = smem_sfb + sbuf * BW_SC;
        const int kb_begin = chunk * BW_SC;
        const int kb_end = min(nkb, kb_begin + BW_SC);
        for (int kb = kb_begin; kb < kb_end; kb += 2) {
          const bool two = kb + 1 < kb_end;
          const uint32_t ita = it, itb = it + 1;
          const uint32_t col_a = static_cast<uint32_t>((ita & (BW_SLOTS - 1)) * BN);
          const uint32_t col_b = static_cast<uint32_t>((itb & (BW_SLOTS - 1)) * BN);
          const float scale_a = __fmul_rn(sa_s[(kb - kb_begin) * BM], sb_s[kb - kb_begin]);
          const float scale_b = two ? __fmul_rn(sa_s[(kb + 1 - kb_begin) * BM], sb_s[kb + 1 - kb_begin]) : 0.f;
          {
            // one try_wait round for both partial products (lane 0: block kb, lane 1: block kb+1)
            const uint32_t itj = lane == 1 ? itb : ita;
            const long long w0 = p.prof ? clock64() : 0;
            mbar_wait_lanes(&part_full[itj & (BW_SLOTS - 1)], (itj >> 2) & 1, lane == 0 || (lane == 1 && two));
            if (p.prof) prof_part += clock64() - w0;
          }
          tc_fence_after_sync();
          const long long wb0 = p.prof ? clock64() : 0;
          if constexpr (HC == 64) {
            tmem_ld_32x32b_x32(t_lane + col_a, r0);
            tmem_ld_32x32b_x32(t_lane + col_a + 32, r1);
            tmem_ld_wait();
            if (p.prof) prof_ld += clock64() - wb0;
            // slot a is in registers: hand it back
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) {
              if constexpr (CG == 1) mbar_arrive(&part_empty[ita & (BW_SLOTS - 1)]);
              else mbar_arrive_cluster(&part_empty[ita & (BW_SLOTS - 1)], 0);
            }
            if (!(p.debug_flags & 2)) {
              promote_fma2<32>(acc, 0, r0, scale_a);
            }
            if (two) tmem_ld_32x32b_x32(t_lane + col_b, r0);
            if (!(p.debug_flags & 2)) {
              promote_fma2<32>(acc, 32, r1, scale_a);
            }
            if (two) {
              tmem_ld_32x32b_x32(t_lane + col_b + 32, r1);
              tmem_ld_wait();
              tc_fence_before_sync();
              __syncwarp();
              if (lane == 0) {
                if constexpr (CG == 1) mbar_arrive(&part_empty[itb & (BW_SLOTS - 1)]);
                else mbar_arrive_cluster(&part_empty[itb & (BW_SLOTS - 1)], 0);
              }
              if (!(p.debug_flags & 2)) {
                promote_fma2<32>(acc, 0, r0, scale_b);
                promote_fma2<32>(acc, 32, r1, scale_b);
              }
            }
          } else {
            // narrow tile: both partial products of the step in one TMEM round trip (HC = 32 or 16 columns each)
            uint32_t(&r0lo)[16] = *reinterpret_cast<uint32_t(*)[16]>(&r0[0]);
            uint32_t(&r1lo)[16] = *reinterpret_cast<uint32_t(*)[16]>(&r1[0]);
            if constexpr (HC == 32) {
              tmem_ld_32x32b_x32(t_lane + col_a, r0);
              if (two) tmem_ld_32x32b_x32(t_lane + col_b, r1);
            } else {
              tmem_ld_32x32b_x16(t_lane + col_a, r0lo);
              if (two) tmem_ld_32x32b_x16(t_lane + col_b, r1lo);
            }
            tmem_ld_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&part_empty[ita & (BW_SLOTS - 1)]);
            else if (lane == 1 && two) mbar_arrive(&part_empty[itb & (BW_SLOTS - 1)]);
            if (!(p.debug_flags & 2)) {
              promote_fma2<HC>(acc, 0, r0, scale_a);
              if (two) promote_fma2<HC>(acc, 0, r1, scale_b);
            }
          }
          it += two ? 2 : 1;
          if (p.prof) prof_body += clock64() - wb0;
        }
        sbuf ^= 1;
      }

      if (S > 1) {
        // K-split reduction: ranks 1..S-1 push their partial sums into rank 0's shared memory (DSMEM stores), one
        // cluster barrier, rank 0 adds them in rank order (deterministic) and owns the store; a second barrier keeps
        // the next tile's pushes behind rank 0's reads.  The producer and MMA warps take part in both barriers.
        if (krank != 0) {
          const uint32_t dst = smem_u32(smem_x + (static_cast<size_t>(krank - 1) * 256 + ptid) * HC);
#pragma unroll
          for (int j = 0; j < HC; j += 4) st_dsmem_v4(dst + j * 4, 0, acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
        }
        __syncwarp();
        cluster_sync_all();
        if (krank == 0) {
          for (int r = 1; r < S; ++r) {
            const float4* src = reinterpret_cast<const float4*>(smem_x + (static_cast<size_t>(r - 1) * 256 + ptid) * HC);
#pragma unroll
            for (int j = 0; j < HC / 4; ++j) {
              const float4 v4 = src[j];
              acc[4 * j] = __fadd_rn(acc[4 * j], v4.x);
              acc[4 * j + 1] = __fadd_rn(acc[4 * j + 1], v4.y);
              acc[4 * j + 2] = __fadd_rn(acc[4 * j + 2], v4.z);
              acc[4 * j + 3] = __fadd_rn(acc[4 * j + 3], v4.w);
            }
          }
        }
        __syncwarp();
        cluster_sync_all();
      }
      if (S > 1 && krank != 0) continue;  // rank 0 owns the tile's store
      if constexpr (HC == 64) {
        // ---- write the warp's 32 x 64 block: bias, cast, swizzled staging, one TMA store ----
        tma_store_wait_read<0>();  // the previous tile's store has long drained the staging buffer
        __syncwarp();
  #pragma unroll
        for (int g = 0; g < 8; ++g) {
          const int col = col0 + g * 8;
          float v[8];
  #pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = acc[g * 8 + j];
          if (has_bias && col < p.N) {
            float bi[8];
            load_bias8(p.bias, p.bias_dtype, col, bi);
  #pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __fadd_rn(v[j], bi[j]);
          }
          uint4 o;
          o.x = pack16<OUT_KIND>(v[0], v[1]);
          o.y = pack16<OUT_KIND>(v[2], v[3]);
          o.z = pack16<OUT_KIND>(v[4], v[5]);
          o.w = pack16<OUT_KIND>(v[6], v[7]);
          const int chunk16 = g ^ (lane & 7);
          *reinterpret_cast<uint4*>(stage_buf + lane * 128 + chunk16 * 16) = o;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (row0 < p.M && col0 < p.N && elect_one()) {
          tma_store_2d(&tmap_d, stage_buf, col0, row0);
          tma_store_commit();
        }

      } else {
        // narrow tile: each thread owns HC consecutive columns (64 or 32 bytes) of one row -> 16-byte global stores
        const int row = row0 + lane;
        if (row < p.M && col0 < p.N) {
          uint8_t* drow = static_cast<uint8_t*>(p.D) + (static_cast<int64_t>(row) * p.ldd + col0) * 2;
#pragma unroll
          for (int g = 0; g < HC / 8; ++g) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = acc[g * 8 + j];
            if (has_bias) {
              float bi[8];
              load_bias8(p.bias, p.bias_dtype, col0 + g * 8, bi);
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = __fadd_rn(v[j], bi[j]);
            }
            uint4 o;
            o.x = pack16<OUT_KIND>(v[0], v[1]);
            o.y = pack16<OUT_KIND>(v[2], v[3]);
            o.z = pack16<OUT_KIND>(v[4], v[5]);
            o.w = pack16<OUT_KIND>(v[6], v[7]);
            *reinterpret_cast<uint4*>(drow + g * 16) = o;
          }
        }
      }
    }
    tma_store_wait_all<0>();
    if (p.prof && warp == EPI_WARP0 && lane == 0) {
      p.prof[blockIdx.x * 8 + 3] = static_cast<unsigned long long>(prof_part);
      p.prof[blockIdx.x * 8 + 5] = static_cast<unsigned long long>(prof_sc);
      p.prof[blockIdx.x * 8 + 6] = static_cast<unsigned long long>(prof_ld);
      p.prof[blockIdx.x * 8 + 7] = static_cast<unsigned long long>(prof_body);
    }
  }
  if (p.prof && threadIdx.x == 0) p.prof[blockIdx.x * 8 + 4] = static_cast<unsigned long long>(clock64() - t_start);

  tc_fence_before_sync();
  if constexpr (CG == 1) {
    __syncthreads();
  } else {
    __syncwarp();
    cluster_sync_all();
  }
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<CG>(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host ----

static int bw_fixed_smem(int xbytes = 0) {
  return BW_EPI_WARPS * EPI_BUF_BYTES + xbytes + 2 * BW_SC * BM * 4 + 2 * BW_SC * 4 + 256;
}

template <int OUT_KIND, int CG, int BN>
static int launch_bw(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const BlockwiseParams& p,
                     int grid, size_t smem, cudaStream_t st, char* err, size_t err_len) {
  auto kern = gemm_block128_kernel<OUT_KIND, CG, BN>;
  static thread_local int attr_dev = -1;
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
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(BW_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = CG == 2 ? 2 : p.ksplit;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, td, p);
  if (e != cudaSuccess) {
    snprintf(err, err_len, "block-wise gemm kernel launch: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  count_launch();
  return B2_OK;
}

// forced_config (low byte): 0 = automatic, 1 = 1-CTA 128x128 tiles, 2 = 2-CTA 256x128 tiles, 7 / 8 = 1-CTA 128x64 /
// 128x32 tiles (decode shapes)
int gemm_launch_blockwise(const GemmArgs& a, int forced_config, const char** used_config, cudaStream_t st,
                          char* err, size_t err_len) {
  const int forced_debug = forced_config >> 8;
  forced_config &= 0xff;
  const int sms = device_sm_count();
  int cg = a.M > BM ? 2 : 1;  // a pair needs more than one 128-row block to fill
  int bn = BW_BN, ks = 1;
  const int num_kb_h = (a.K + BK - 1) / BK;
  auto stages_of = [&](int bn_, int ks_) {
    const int x = (ks_ - 1) * 256 * (bn_ / 2) * 4;
    return (SMEM_LIMIT - bw_fixed_smem(x)) / (BM * BK + bn_ * BK);
  };
  if (cg == 1) {
    // M <= 128 is a stream of the weight bytes through a per-K-block hand-shake loop (~600 clocks per K block, or the
    // HBM time of the block's bytes on one SM if that is longer).  Pick (tile width, K split) minimising
    //   waves x K blocks per CTA x time per K block (+ ~1500 clocks for the split's two cluster barriers):
    // narrow tiles put a CTA on every SM, the split shortens the serial loop.  N = 4096, K = 7168: 128x64 tiles, split 2.
    long best = -1;
    for (int cand : {128, 64, 32}) {
      for (int s_ : {1, 2}) {  // measured (profiles/r02m_block_decode_plans.log): splits of 4 / 8 lose to their barriers
        if (s_ > 1 && (num_kb_h / s_ < 4 || stages_of(cand, s_) < 4)) continue;
        if (s_ > 1 && ((a.N + cand - 1) / cand) * s_ > sms) continue;  // the split only pays inside ONE wave
        const long tiles = (a.N + cand - 1) / cand;
        const long waves = (tiles * s_ + sms - 1) / sms;
        const long per_kb = cand * 56 / 10 > 600 ? cand * 56 / 10 : 600;
        const long cost = waves * ((num_kb_h + s_ - 1) / s_) * per_kb + (s_ > 1 ? 1500 : 0);
        if (best < 0 || cost < best) {
          best = cost;
          bn = cand;
          ks = s_;
        }
      }
    }
  }
  // forced ids: 1 / 7 / 8 = 1-CTA 128 / 64 / 32 wide, unsplit; 2 = 2-CTA; 9 + 16*b + s: b = 0/1/2 -> 128/64/32 wide,
  // split 2^(s+1) (s = 0..2)  -- e.g. 9 = 128 wide split 2, 26 = 64 wide split 4
  if (forced_config == 1) { cg = 1; bn = 128; ks = 1; }
  if (forced_config == 2) { cg = 2; bn = 128; ks = 1; }
  if (forced_config == 7) { cg = 1; bn = 64; ks = 1; }
  if (forced_config == 8) { cg = 1; bn = 32; ks = 1; }
  if (forced_config >= 9 && forced_config < 9 + 48) {
    const int b = (forced_config - 9) / 16, sidx = (forced_config - 9) % 16;
    cg = 1;
    bn = 128 >> b;
    ks = 2 << sidx;
    if (b > 2 || sidx > 2 || num_kb_h / ks < 1 || stages_of(bn, ks) < 2) {
      snprintf(err, err_len, "block-wise forced config %d does not fit (bn=%d split=%d)", forced_config, bn, ks);
      return B2_ERR_ARG;
    }
  }
  static const int dbg_env = [] {
    const char* e = getenv("B2_GEMM_DEBUG");
    return e ? atoi(e) : 0;
  }();
  static thread_local char cfg_name[64];
  if (cg == 2) snprintf(cfg_name, sizeof(cfg_name), "2cta_256x128_block128");
  else if (ks == 1) snprintf(cfg_name, sizeof(cfg_name), "1cta_128x%d_block128", bn);
  else snprintf(cfg_name, sizeof(cfg_name), "1cta_128x%d_block128_ks%d", bn, ks);
  if (used_config) *used_config = cfg_name;
  const int b_rows = bn / cg;

  BlockwiseParams p;
  p.sfa = a.As;
  p.sfb = a.Bs;
  p.sfa_m_stride = a.sfa_m_stride;
  p.sfa_k_stride = a.sfa_k_stride;
  p.sfb_n_stride = a.sfb_n_stride;
  p.sfb_k_stride = a.sfb_k_stride;
  p.bias = a.bias;
  p.bias_dtype = a.bias_dtype;
  p.M = a.M;
  p.N = a.N;
  p.K = a.K;
  p.D = a.D;
  p.ldd = a.ldd;
  p.ksplit = ks;
  p.num_m_blocks = (a.M + BM * cg - 1) / (BM * cg);
  p.num_n_blocks = (a.N + bn - 1) / bn;
  p.num_kb = (a.K + BK - 1) / BK;
  const int xbytes = (ks - 1) * 256 * (bn / 2) * 4;
  int stages = (SMEM_LIMIT - bw_fixed_smem(xbytes)) / (BM * BK + b_rows * BK);
  if (stages > 10) stages = 10;
  p.stages = stages;
  choose_raster(a.M, a.N, a.K, BM * cg, bn, p.num_m_blocks, p.num_n_blocks, &p.group_m, &p.band_n, &p.policy_a,
                &p.policy_b);
  p.a_box_rows = (a.M >= BM || cg == 2) ? BM : ((a.M + 7) / 8) * 8;
  p.prof = static_cast<unsigned long long*>(profile_buffer());
  p.debug_flags = dbg_env | forced_debug;

  const CUtensorMapDataType d_dt =
      a.out_dtype == B2_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUtensorMap ta, tb, td;
  if (!make_tmap_2d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.A, a.M, a.K, a.lda, p.a_box_rows, BK) ||
      !make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.W, a.N, a.K, a.ldw, b_rows, BK) ||
      !make_tmap_2d(&td, d_dt, 2, a.D, a.M, a.N, a.ldd * 2, 32, 64)) {
    snprintf(err, err_len, "cuTensorMapEncodeTiled failed (block-wise M=%d N=%d K=%d)", a.M, a.N, a.K);
    return B2_ERR_CUDA;
  }
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  const int units = sms / (cg * ks);
  const int grid = (tiles < units ? tiles : units) * cg * ks;
  const size_t smem = static_cast<size_t>(stages) * (BM * BK + b_rows * BK) + bw_fixed_smem(xbytes);
  const bool f16 = a.out_dtype == B2_DTYPE_F16;
  if (cg == 2) {
    if (f16) return launch_bw<1, 2, 128>(ta, tb, td, p, grid, smem, st, err, err_len);
    return launch_bw<0, 2, 128>(ta, tb, td, p, grid, smem, st, err, err_len);
  }
  if (bn == 64) {
    if (f16) return launch_bw<1, 1, 64>(ta, tb, td, p, grid, smem, st, err, err_len);
    return launch_bw<0, 1, 64>(ta, tb, td, p, grid, smem, st, err, err_len);
  }
  if (bn == 32) {
    if (f16) return launch_bw<1, 1, 32>(ta, tb, td, p, grid, smem, st, err, err_len);
    return launch_bw<0, 1, 32>(ta, tb, td, p, grid, smem, st, err, err_len);
  }
  if (f16) return launch_bw<1, 1, 128>(ta, tb, td, p, grid, smem, st, err, err_len);
  return launch_bw<0, 1, 128>(ta, tb, td, p, grid, smem, st, err, err_len);
}

}  // namespace b2
