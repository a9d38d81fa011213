// Persistent, warp-specialised tcgen05 GEMM for the W8A8 linear path (sm_100a).
//
//   D[M,N] = epilogue( A[M,K] (8-bit, K-major)  x  W[N,K]^T (8-bit, K-major) )
//
// kind::i8     (s8 x s8 -> s32 in TMEM): int8_scaled_mm
//              (reference: sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747, no sm_100 path;
//               bit-exact epilogue of sgl-kernel/csrc/cpu/gemm_int8.cpp:20-60)
// kind::f8f6f4 (e4m3 x e4m3 -> f32 in TMEM): fp8_scaled_mm
//              (reference: sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:796-1170, 1450-1532)
//
// Structure (one CTA per SM, 192 threads, static round-robin tile schedule):
//   warp 0   : TMA producer  -- cp.async.bulk.tensor 2-D boxes {128 B of K, rows} with the
//              128-byte swizzle into a STAGES-deep ring, completion on `full[s]` mbarriers
//   warp 1   : MMA issuer    -- one lane issues 4 x tcgen05.mma (K = 32 B each) per stage,
//              tcgen05.commit -> `empty[s]` releases the stage, -> `acc_full[a]` hands the
//              accumulator to the epilogue; owns the TMEM allocation
//   warps 2-5: epilogue      -- tcgen05.ld 32 lanes x 32 columns, dequant/scale/bias in fp32,
//              cast, swizzled st.shared, per-warp TMA store of 32 x 128 B boxes
// Two TMEM accumulator buffers (2 x BN columns) let the epilogue of tile i overlap the
// main loop of tile i+1.  M/N/K tails are handled by TMA zero-fill on loads and clipping
// on stores; no padding copies are ever made.
//
// CG = 2 (large M): the two CTAs of a cluster (one TPC) work on ONE 256 x BN tile with
// tcgen05.mma.cta_group::2.  Each CTA loads its own 128 rows of A and only HALF of the W tile
// (BN/2 rows), the tensor cores of both SMs read both halves, so the L2->SM operand traffic per
// MAC drops by a third against two independent 128 x BN tiles.  CTA 0 issues every MMA; the
// TMA loads of both CTAs complete on CTA 0's `full` barriers; tcgen05.commit multicasts the
// `empty` / `acc_full` arrivals to both CTAs; the epilogue warps of both CTAs hand their TMEM
// buffer back on CTA 0's `acc_empty` barrier.
#include <cstdlib>

#include "gemm_common.cuh"

namespace b2 {

// OUT_KIND: 0 = bf16, 1 = f16, 2 = 32-bit (f32, or the raw s32 accumulators)
// RS: scatter epilogue -- every 32-row store box goes to the rank that owns its rows (peers.m[owner], a map over that
// rank's gather buffer, possibly across NVLink) instead of tmap_d; 16-bit outputs only
template <MmaKind KIND, int OUT_KIND, int BN, int CG, bool RS = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    gemm_w8a8_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                     const __grid_constant__ CUtensorMap tmap_d, const KernelParams p,
                     const __grid_constant__ PeerMaps<RS> peers) {
  static_assert(!RS || OUT_KIND != 2, "the scatter epilogue stores 16-bit outputs");
  constexpr int OUT_BYTES = OUT_KIND == 2 ? 4 : 2;
  constexpr int B_ROWS = BN / CG;  // rows of the W tile this CTA stages
  constexpr uint32_t A_STAGE_BYTES = BM * BK;
  constexpr uint32_t B_STAGE_BYTES = B_ROWS * BK;
  constexpr uint32_t TMEM_COLS = (2 * BN) < 32 ? 32 : (2 * BN);
  static_assert((TMEM_COLS & (TMEM_COLS - 1)) == 0 && TMEM_COLS <= 512, "TMEM columns must be a power of two");
  constexpr int COLS_PER_BOX = 128 / OUT_BYTES;  // output columns in one 128-byte staging row
  static_assert(BN % COLS_PER_BOX == 0 || BN < COLS_PER_BOX, "tile width vs store box");

  // The dynamic window is the only shared memory of the kernel, so it starts on a 1024-byte boundary (what
  // the 128-byte swizzle needs); no slack is budgeted for rounding up, a misplaced base traps instead.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int stages = p.stages;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + static_cast<size_t>(stages) * A_STAGE_BYTES;
  uint8_t* smem_epi = smem_b + static_cast<size_t>(stages) * B_STAGE_BYTES;  // 4 warps x 2 bufs x 4 KB
  float* smem_bs = reinterpret_cast<float*>(smem_epi + 4 * 2 * EPI_BUF_BYTES);  // this tile's Bs[BN], bias[BN] (fp32)
  float* smem_bias = smem_bs + BN;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_bias + BN);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* acc_full = empty_bar + stages;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(acc_empty + 2);

  // warp-uniform role index (the shuffle tells the compiler so): everything the producer and MMA warps
  // compute then lives in uniform registers, and UTMALDG / UTCIMMA / UTCBAR issue straight from them
  // (a divergent `if (lane == 0)` region costs a 5 x R2UR + ELECT waterfall loop per instruction instead)
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const long long t_start = p.prof ? clock64() : 0;
  const int tiles_pb = p.num_m_blocks * p.num_n_blocks;  // tiles of (BM*CG) x BN per batch entry
  const int num_tiles_full = tiles_pb * (p.batch > 1 ? p.batch : 1);
  // Tail split (BN = 256 only): the last `tail_tiles` tiles of the raster order -- the ragged last wave -- are issued
  // as two half-width (BN/2) work items each, so a last wave that would occupy less than half of the CTAs for a whole
  // tile time takes half a tile time instead.  Work item i >= full_items is half (i - full_items) & 1 of tile
  // full_items + (i - full_items) / 2.  The B box is still loaded whole (row offset shifted), the MMA runs with N/2.
  const int full_items = num_tiles_full - p.tail_tiles;
  const int num_tiles = num_tiles_full + p.tail_tiles;  // work items
  const int cta_rank = CG == 1 ? 0 : static_cast<int>(cluster_ctarank());
  const int unit = CG == 1 ? static_cast<int>(blockIdx.x) : static_cast<int>(blockIdx.x >> 1);
  const int num_units = CG == 1 ? static_cast<int>(gridDim.x) : static_cast<int>(gridDim.x >> 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_d);
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4 * CG);  // one arrival per epilogue warp of every CTA of the cluster
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<CG>(tmem_ptr, TMEM_COLS);
  tc_fence_before_sync();
  if constexpr (CG == 1) {
    __syncthreads();
  } else {
    __syncwarp();
    cluster_sync_all();  // barrier inits of both CTAs visible before any remote arrival / TMA completion
  }
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================================================== TMA producer ====
    // the whole warp walks the ring (all lanes poll the barrier); one elected lane issues
    {
      // static-weight launches (b2_set_weight_prefetch) start this kernel under the tail of the quantizer: barrier
      // init, TMEM allocation and the cluster handshake above are already done when its codes become readable
      pdl_wait_prior_grids();
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = (static_cast<uint32_t>(p.a_box_rows) * BK + B_STAGE_BYTES) * CG;
      long long prof_wait = 0;
      for (int tile = unit; tile < num_tiles; tile += num_units) {
        int m_blk, n_blk;
        const int half = tile >= full_items ? ((tile - full_items) & 1) : -1;
        const int tidx = tile >= full_items ? full_items + ((tile - full_items) >> 1) : tile;
        const int bidx = p.batch > 1 ? tidx / tiles_pb : 0;  // batched mode: tiles enumerate batch-major
        tile_coords(tidx - bidx * tiles_pb, p, m_blk, n_blk);
        if constexpr (RS) {  // rank-rotated row order, see the epilogue
          m_blk += p.rs_rot;
          if (m_blk >= p.num_m_blocks) m_blk -= p.num_m_blocks;
        }
        const int expert = p.tile_expert ? __ldg(p.tile_expert + m_blk) : 0;  // grouped (MoE) mode: CG == 1
        if (expert < 0) continue;  // a padding block past the last routed token
        const int m_row = (m_blk * CG + cta_rank) * BM;
        const int n_row = half < 0 ? expert * p.N + n_blk * BN + cta_rank * B_ROWS
                                   : n_blk * BN + half * (BN / 2) + cta_rank * (B_ROWS / 2);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          const long long w0 = p.prof ? clock64() : 0;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (p.prof) prof_wait += clock64() - w0;
          uint8_t* a_dst = smem_a + static_cast<size_t>(stage) * A_STAGE_BYTES;
          uint8_t* b_dst = smem_b + static_cast<size_t>(stage) * B_STAGE_BYTES;
          if (p.debug_flags & 4) {  // experiment: no loads, the MMAs run over whatever the ring holds
            if (cta_rank == 0 && elect_one()) mbar_arrive(&full_bar[stage]);
          } else if constexpr (CG == 1) {
            if (elect_one()) {
              mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
              if (p.batch > 1) {
                tma_load_3d(a_dst, &tmap_a, &full_bar[stage], kb * BK, m_row, bidx, p.policy_a);
                tma_load_3d(b_dst, &tmap_b, &full_bar[stage], kb * BK, n_row, bidx, p.policy_b);
              } else {
                tma_load_2d(a_dst, &tmap_a, &full_bar[stage], kb * BK, m_row, p.policy_a);
                tma_load_2d(b_dst, &tmap_b, &full_bar[stage], kb * BK, n_row, p.policy_b);
              }
            }
          } else {
            // both CTAs' bytes are expected on (and complete on) the leader CTA's barrier
            if (elect_one()) {
              if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
              const uint32_t bar0 = smem_u32(&full_bar[stage]) & kPeerBitMask;
              if (p.batch > 1) {
                tma_load_3d_cg2(smem_u32(a_dst), &tmap_a, bar0, kb * BK, m_row, bidx, p.policy_a);
                tma_load_3d_cg2(smem_u32(b_dst), &tmap_b, bar0, kb * BK, n_row, bidx, p.policy_b);
              } else {
                tma_load_2d_cg2(smem_u32(a_dst), &tmap_a, bar0, kb * BK, m_row, p.policy_a);
                tma_load_2d_cg2(smem_u32(b_dst), &tmap_b, bar0, kb * BK, n_row, p.policy_b);
              }
            }
          }
          __syncwarp();
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      if (p.prof && lane == 0) p.prof[blockIdx.x * 8 + 2] = static_cast<unsigned long long>(prof_wait);
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    // whole warp in the loop, one elected lane (always the same one) issues every MMA and commit
    if (cta_rank == 0) {
      constexpr uint32_t idesc_full = make_idesc<KIND>(BM * CG, BN);
      constexpr uint32_t idesc_half = make_idesc<KIND>(BM * CG, BN >= 64 ? BN / 2 : BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      long long prof_full = 0, prof_acc = 0;
      for (int tile = unit; tile < num_tiles; tile += num_units) {
        const uint32_t idesc = tile >= full_items ? idesc_half : idesc_full;
        if (p.tile_expert) {  // (never together with the batched mode or the tail split)
          int m_blk, n_blk;
          tile_coords(tile, p, m_blk, n_blk);
          if (__ldg(p.tile_expert + m_blk) < 0) continue;
        }
        long long w0 = p.prof ? clock64() : 0;
        mbar_wait(&acc_empty[acc], acc_phase ^ 1);
        if (p.prof) prof_acc += clock64() - w0;
        tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          w0 = p.prof ? clock64() : 0;
          mbar_wait(&full_bar[stage], phase);
          if (p.prof) prof_full += clock64() - w0;
          tc_fence_after_sync();
          const uint64_t da = make_smem_desc_sw128(smem_u32(smem_a + static_cast<size_t>(stage) * A_STAGE_BYTES));
          const uint64_t db = make_smem_desc_sw128(smem_u32(smem_b + static_cast<size_t>(stage) * B_STAGE_BYTES));
          if (elect_one()) {
            if (!(p.debug_flags & 1)) {
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k) {
                // advance 32 B along K inside the 128 B swizzle row: +2 in the (addr >> 4) field
                umma<KIND, CG>(d_tmem, da + static_cast<uint64_t>(k * (UMMA_K >> 4)),
                               db + static_cast<uint64_t>(k * (UMMA_K >> 4)), idesc, (kb | k) != 0 ? 1u : 0u);
              }
            }
            if constexpr (CG == 1) umma_commit(&empty_bar[stage]);
            else umma_commit_cg2(&empty_bar[stage], 0x3);
          }
          __syncwarp();
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (elect_one()) {
          if constexpr (CG == 1) umma_commit(&acc_full[acc]);
          else umma_commit_cg2(&acc_full[acc], 0x3);
        }
        __syncwarp();
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      if (p.prof && lane == 0) {
        p.prof[blockIdx.x * 8 + 0] = static_cast<unsigned long long>(prof_full);
        p.prof[blockIdx.x * 8 + 1] = static_cast<unsigned long long>(prof_acc);
      }
    }
  } else {
    // ========================================================= epilogue ====
    pdl_wait_prior_grids();        // scales come from, and the output may still be read by, the preceding kernels
    const int quarter = warp & 3;  // TMEM lane quarter this warp may read
    const int epi_tid = (warp - EPI_WARP0) * 32 + lane;
    uint8_t* my_epi = smem_epi + static_cast<size_t>(warp - EPI_WARP0) * 2 * EPI_BUF_BYTES;
    const bool has_bias = p.bias != nullptr;
    const bool raw_out = (KIND == MmaKind::I8) && p.out_dtype == B2_DTYPE_S32;
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    long long prof_epi = 0, prof_wr = 0, prof_ld = 0, prof_body = 0;
    constexpr int BOX_COLS = COLS_PER_BOX < BN ? COLS_PER_BOX : BN;
    constexpr int BOX_BYTES = BOX_COLS * OUT_BYTES;  // 128 (SW128 box) or 64 (SW64 box, BN = 32 @ 16 bit)
    static_assert(BOX_BYTES == 128 || BOX_BYTES == 64, "store box is 128 or 64 bytes wide");
    constexpr int CH_PER_BOX = BOX_COLS / 32;  // 32-column TMEM loads per store box
    constexpr int NCH = BN / 32;
    // TMA swizzle of the staging box: 16-byte chunk index ^= (byte offset >> 7) & (BOX_BYTES/16 - 1)
    const int swz = BOX_BYTES == 128 ? (lane & 7) : ((lane >> 1) & 3);
    int rs_set = 0;
    if constexpr (RS) rs_set = static_cast<int>(*reinterpret_cast<const volatile uint32_t*>(p.rs_counter) & 1u);  // after the wait above: the previous sum has advanced it
    for (int tile = unit; tile < num_tiles; tile += num_units) {
      int m_blk, n_blk;
      const int half = tile >= full_items ? ((tile - full_items) & 1) : -1;
      const int tidx = tile >= full_items ? full_items + ((tile - full_items) >> 1) : tile;
      const int bidx = p.batch > 1 ? tidx / tiles_pb : 0;
      tile_coords(tidx - bidx * tiles_pb, p, m_blk, n_blk);
      if constexpr (RS) {
        // every rank walks the row blocks starting behind its own share: at any time rank r stores to owner
        // (r + 1 + t) % world, a permutation -- no owner's NVLink ingress takes all the ranks at once -- and ends on
        // its own rows, so nothing remote is outstanding at the end of the grid
        m_blk += p.rs_rot;
        if (m_blk >= p.num_m_blocks) m_blk -= p.num_m_blocks;
      }
      const int expert = p.tile_expert ? __ldg(p.tile_expert + m_blk) : 0;
      if (expert < 0) continue;
      const float* Bs_e = p.Bs + static_cast<size_t>(expert) * p.N;  // per-expert channel scales [E, N]
      const int row0 = (m_blk * CG + cta_rank) * BM + quarter * 32;
      const int row = row0 + lane;
      const int col_tile = n_blk * BN + (half > 0 ? BN / 2 : 0);
      const int nch = half < 0 ? NCH : NCH / 2;  // 32-column chunks of this work item

      // Per-column scale and bias of this tile -> shared memory, once, while the main loop of the tile is
      // still running (the four warps read them back as conflict-free broadcasts).  Columns beyond N get 0;
      // the TMA store clips them.
      epi_bar_sync();  // every warp is done with the previous tile's values
      if (!raw_out) {
        for (int c = epi_tid; c < nch * 32; c += 128) {
          const int col = col_tile + c;
          smem_bs[c] = col < p.N ? __ldg(Bs_e + (p.scalar_scales ? 0 : col)) : 0.f;
          smem_bias[c] = (has_bias && col < p.N) ? load_bias(p.bias, p.bias_dtype, col) : -0.f;  // x + (-0) == x bit for bit
        }
      }
      epi_bar_sync();
      float as = 0.f;
      if (!raw_out && row < p.M) as = __ldg(p.As + (p.scalar_scales ? 0 : row));

      const long long w0 = p.prof ? clock64() : 0;
      mbar_wait(&acc_full[acc], acc_phase);
      if (p.prof) prof_epi += clock64() - w0;
      tc_fence_after_sync();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(acc * BN);

      const long long body0 = p.prof ? clock64() : 0;
      if (!(p.debug_flags & 2)) {
        // 32-column TMEM loads, software-pipelined over two register sets: the load of chunk c+1 is in flight
        // while chunk c is scaled, cast and staged.  The loop over PAIRS of chunks stays rolled -- fully unrolled
        // the body is ~40 KB of SASS and the four epilogue warps stall on instruction fetch.
        uint32_t r[2][32];
        constexpr int CH_PER_PAIR = NCH >= 2 ? 2 : 1;
        const int npair = nch >= 2 ? nch / 2 : 1;
        tmem_ld_32x32b_x32(t_row, r[0]);
#pragma unroll 1
        for (int pr = 0; pr < npair; ++pr) {
#pragma unroll
          for (int hh = 0; hh < CH_PER_PAIR; ++hh) {
            const int c = pr * CH_PER_PAIR + hh;      // chunk index inside the tile
            const int h = CH_PER_BOX == 2 ? hh : 0;   // chunk index inside its store box
            uint8_t* stage_buf = my_epi + buf * EPI_BUF_BYTES;
            {
              const long long l0 = p.prof ? clock64() : 0;
              tmem_ld_wait();
              if (p.prof) prof_ld += clock64() - l0;
            }
            if (c + 1 < nch) {
              tmem_ld_32x32b_x32(t_row + static_cast<uint32_t>((c + 1) * 32), r[(hh + 1) & 1]);
            } else {
              // accumulator fully read: hand the TMEM buffer back to the MMA warp before the last chunk's math
              tc_fence_before_sync();
              __syncwarp();
              if (lane == 0) {
                if constexpr (CG == 1) mbar_arrive(&acc_empty[acc]);
                else mbar_arrive_cluster(&acc_empty[acc], 0);
              }
            }
            if (h == 0) {
              // the TMA store that last read this staging buffer must have drained it
              const long long r0 = p.prof ? clock64() : 0;
              tma_store_wait_read<1>();  // a no-op for the lanes that never committed a store group
              __syncwarp();
              if (p.prof) prof_wr += clock64() - r0;
            }
            const uint32_t(&rc)[32] = r[hh & 1];
            const float4* bs4 = reinterpret_cast<const float4*>(smem_bs + c * 32);
            const float4* bi4 = reinterpret_cast<const float4*>(smem_bias + c * 32);
            // per-column scale / bias of the next 8 columns are fetched before the current 8 are stored, so the
            // shared-memory load latency hides behind the math instead of following every STS
            float4 nb[4];
            if (!raw_out) {
              nb[0] = bs4[0]; nb[1] = bs4[1]; nb[2] = bi4[0]; nb[3] = bi4[1];
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) {  // 8 columns per group
              float v[8];
              if (raw_out) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(rc[g * 8 + j]);
              } else {
                const float bs[8] = {nb[0].x, nb[0].y, nb[0].z, nb[0].w, nb[1].x, nb[1].y, nb[1].z, nb[1].w};
                const float bi[8] = {nb[2].x, nb[2].y, nb[2].z, nb[2].w, nb[3].x, nb[3].y, nb[3].z, nb[3].w};
                if (g < 3) {
                  nb[0] = bs4[2 * g + 2]; nb[1] = bs4[2 * g + 3]; nb[2] = bi4[2 * g + 2]; nb[3] = bi4[2 * g + 3];
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = epilogue_value<KIND>(rc[g * 8 + j], as, bs[j], bi[j]);
              }
              if constexpr (OUT_BYTES == 2) {
                // 8 columns -> one 16-byte chunk; chunk index inside the box row = h*4 + g
                uint4 o;
                o.x = pack16<OUT_KIND>(v[0], v[1]);
                o.y = pack16<OUT_KIND>(v[2], v[3]);
                o.z = pack16<OUT_KIND>(v[4], v[5]);
                o.w = pack16<OUT_KIND>(v[6], v[7]);
                const int chunk = (h * 4 + g) ^ swz;
                *reinterpret_cast<uint4*>(stage_buf + lane * BOX_BYTES + chunk * 16) = o;
              } else {
                // 8 columns -> two 16-byte chunks 2g, 2g+1
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                  uint4 o;
                  o.x = __float_as_uint(v[4 * cc + 0]);
                  o.y = __float_as_uint(v[4 * cc + 1]);
                  o.z = __float_as_uint(v[4 * cc + 2]);
                  o.w = __float_as_uint(v[4 * cc + 3]);
                  const int chunk = (g * 2 + cc) ^ swz;
                  *reinterpret_cast<uint4*>(stage_buf + lane * BOX_BYTES + chunk * 16) = o;
                }
              }
            }
            if (h == CH_PER_BOX - 1) {
              const int col_box = col_tile + (c / CH_PER_BOX) * BOX_COLS;
              fence_proxy_async_smem();
              __syncwarp();
              // A box that lies outside the output is not stored and does NOT take a staging buffer: the wait above
              // only guarantees "at most one store still reading shared memory", which protects the buffer about to
              // be written only while committed stores strictly alternate between the two buffers.
              if (row0 < p.M && col_box < p.N) {  // warp-uniform
                if (elect_one()) {  // the same lane every time
                  if constexpr (RS) {
                    const int owner = row0 / p.rs_rows;
                    tma_store_4d(&peers.m[owner], stage_buf, col_box, row0 - owner * p.rs_rows, p.rs_rank, rs_set);
                  } else
                  if (p.batch > 1) tma_store_3d(&tmap_d, stage_buf, col_box, row0, bidx);
                  else tma_store_2d(&tmap_d, stage_buf, col_box, row0);
                  tma_store_commit();
                }
                buf ^= 1;
              }
            }
          }
        }
      } else {
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CG == 1) mbar_arrive(&acc_empty[acc]);
          else mbar_arrive_cluster(&acc_empty[acc], 0);
        }
      }
      if (p.prof) prof_body += clock64() - body0;
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    tma_store_wait_all<0>();
    if constexpr (RS) __threadfence_system();  // the boxes written to the peers are performed before this grid completes
    if (p.prof && warp == EPI_WARP0 && lane == 0) {
      p.prof[blockIdx.x * 8 + 3] = static_cast<unsigned long long>(prof_epi);
      p.prof[blockIdx.x * 8 + 5] = static_cast<unsigned long long>(prof_wr);
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
    cluster_sync_all();  // the peer's smem, TMEM and barriers stay alive until both CTAs are done
  }
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<CG>(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host ----

struct TileConfig {
  int id;
  int bn;
  int cg;
  double rate;  // relative tensor-side efficiency of the tile shape (measured, tools/gpu_probe.py)
  const char* name;
};
static const TileConfig kConfigs[] = {
    {1, 32, 1, 0.35, "1cta_128x32"},   {2, 64, 1, 0.55, "1cta_128x64"},   {3, 128, 1, 0.66, "1cta_128x128"},
    {4, 256, 1, 1.0, "1cta_128x256"},  {5, 256, 2, 1.08, "2cta_256x256"}, {6, 128, 2, 0.8, "2cta_256x128"}};
constexpr int kNumConfigs = sizeof(kConfigs) / sizeof(kConfigs[0]);

// fixed part of the shared-memory window: store staging, the tile's Bs / bias (fp32), barriers + TMEM pointer
static int fixed_smem(int bn) { return 4 * 2 * EPI_BUF_BYTES + 2 * bn * 4 + 256; }

static int stages_for(int bn, int b_rows) {
  int s = (SMEM_LIMIT - fixed_smem(bn)) / (BM * BK + b_rows * BK);
  return s > 12 ? 12 : s;
}

static size_t smem_bytes_for(int bn, int b_rows, int stages) {
  return static_cast<size_t>(stages) * (BM * BK + b_rows * BK) + fixed_smem(bn);
}

static const TileConfig& pick_config(int M, int N, int sms, int batch = 1) {
  int best = 3;
  double best_score = -1;
  for (int i = 0; i < kNumConfigs; ++i) {
    const TileConfig& c = kConfigs[i];
    if (c.cg == 2 && M <= BM) continue;  // the second CTA of the pair would only see padding rows
    const int units = sms / c.cg;
    if (c.bn > 32 && c.bn / 2 >= N) continue;  // a tile more than twice as wide as the problem is mostly padding
    const long tiles = static_cast<long>((M + BM * c.cg - 1) / (BM * c.cg)) * ((N + c.bn - 1) / c.bn) * batch;
    const long waves = (tiles + units - 1) / units;
    // fraction of useful rows in the M tiles (a 256-row tile on M = 300 is 41 % padding)
    const double m_fill = static_cast<double>(M) / (static_cast<double>((M + BM * c.cg - 1) / (BM * c.cg)) * BM * c.cg);
    const double score = static_cast<double>(tiles) / static_cast<double>(waves * units) * c.rate * m_fill;
    if (score > best_score + 1e-9) {
      best_score = score;
      best = i;
    }
  }
  return kConfigs[best];
}

template <MmaKind KIND, int OUT_KIND, int BN, int CG, bool RS = false>
static int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const KernelParams& p,
                      int grid, size_t smem, cudaStream_t st, char* err, size_t err_len,
                      const PeerMaps<RS>& peers = PeerMaps<RS>{}) {
  auto kern = gemm_w8a8_kernel<KIND, OUT_KIND, BN, CG, RS>;
  static thread_local int attr_dev = -1;  // per instantiation; the attribute is per-function, per-device context
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
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = CG;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = weight_prefetch_enabled() ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, td, p, peers);
  if (e != cudaSuccess) {
    snprintf(err, err_len, "gemm kernel launch: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  count_launch();
  return B2_OK;
}

template <MmaKind KIND, int OUT_KIND>
static int launch_cfg(const TileConfig& c, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td,
                      const KernelParams& p, int grid, size_t smem, cudaStream_t st, char* err, size_t err_len) {
  switch (c.id) {
    case 1: return launch_one<KIND, OUT_KIND, 32, 1>(ta, tb, td, p, grid, smem, st, err, err_len);
    case 2: return launch_one<KIND, OUT_KIND, 64, 1>(ta, tb, td, p, grid, smem, st, err, err_len);
    case 3: return launch_one<KIND, OUT_KIND, 128, 1>(ta, tb, td, p, grid, smem, st, err, err_len);
    case 4: return launch_one<KIND, OUT_KIND, 256, 1>(ta, tb, td, p, grid, smem, st, err, err_len);
    case 5: return launch_one<KIND, OUT_KIND, 256, 2>(ta, tb, td, p, grid, smem, st, err, err_len);
    default: return launch_one<KIND, OUT_KIND, 128, 2>(ta, tb, td, p, grid, smem, st, err, err_len);
  }
}

int gemm_launch_blockwise(const GemmArgs& a, int forced_config, const char** used_config, cudaStream_t st,
                          char* err, size_t err_len);
int gemm_launch_splitk(GemmFlavor flavor, const GemmArgs& a, int forced, const char** used_config, cudaStream_t st,
                       char* err, size_t err_len);

int gemm_launch(GemmFlavor flavor, const GemmArgs& a, int forced_config, const char** used_config,
                cudaStream_t st, char* err, size_t err_len) {
  if (flavor == GEMM_E4M3_BLOCK128) return gemm_launch_blockwise(a, forced_config, used_config, st, err, err_len);
  const int sms = device_sm_count();
  // decode shapes (HBM-bound on the weight bytes): cluster split-K kernel of gemm_splitk_sm100.cu
  // (measured, profiles/r01c_sweep_splitk_planned.log: it wins whenever the activation box is small, and for any M <= 128
  // when N is too narrow to give the persistent kernel one tile per SM)
  // M in (32, 64]: still ahead on layers up to ~64 rows of W per SM (profiles/r01e_m64_probe.log: 9.4 vs 12.0 us on
  // 6144 x 4096, 15.7 vs 19.9 us on 7168 x 8192), behind on the widest ones where 128 x 256 persistent tiles stream well
  // (short K -- 8 K-blocks -- leaves the decode kernel nothing to pipeline at M > 32: 8.2 vs 6.8 us on 8192 x 1024)
  const bool decode_shape = a.tile_expert == nullptr && a.batch <= 1 && !a.scalar_scales && !a.rs_peer_bufs && (
      a.M <= 32 || (a.M <= 64 && a.N <= 64 * sms && a.K >= 2048) || (a.M <= BM && a.N < 32 * sms));
  if (a.mc_slot != nullptr && !(decode_shape || a.M <= BM)) {
    snprintf(err, err_len, "multicast push epilogue is built for decode shapes (M <= 128) only");
    return B2_ERR_ARG;
  }
  if (a.x != nullptr) return gemm_launch_splitk(flavor, a, forced_config & 0x80000 ? forced_config : 0, used_config, st, err, err_len);
  if (a.mc_slot != nullptr) return gemm_launch_splitk(flavor, a, forced_config & 0x80000 ? forced_config : 0, used_config, st, err, err_len);
  if ((forced_config & 0x80000) ? (a.M <= BM && !a.tile_expert && a.batch <= 1 && !a.scalar_scales && !a.rs_peer_bufs) : ((forced_config & 0xff) == 0 && decode_shape))
    return gemm_launch_splitk(flavor, a, forced_config, used_config, st, err, err_len);
  const TileConfig* cfg = &pick_config(a.M, a.N, sms, a.batch > 1 ? a.batch : 1);
  const int forced_debug = forced_config >> 8;  // developer experiments ride in the upper bits of the forced id
  forced_config &= 0xff;
  if (forced_config >= 1 && forced_config <= kNumConfigs) cfg = &kConfigs[forced_config - 1];
  if (a.tile_expert) {
    // grouped (MoE) mode: every 128-row block of A belongs to ONE expert (rows sorted and padded per expert), so the
    // tile must be one block tall: 1-CTA configurations only
    if (cfg->cg != 1) cfg = &kConfigs[a.N >= 256 ? 3 : (a.N >= 128 ? 2 : 1)];
    if (a.bias) {
      snprintf(err, err_len, "grouped GEMM takes no bias");
      return B2_ERR_ARG;
    }
  }
  const bool scatter = a.rs_peer_bufs != nullptr;
  if (scatter) {
    // scatter epilogue (row-parallel prefill): 256-wide tiles only, 16-bit output, every rank's rows a whole number of
    // 32-row store boxes
    if (a.tile_expert || a.batch > 1 || a.rs_world < 2 || a.rs_world > 8 || a.M % (32 * a.rs_world) != 0 || a.N % 8 != 0 ||
        (a.out_dtype != B2_DTYPE_BF16 && a.out_dtype != B2_DTYPE_F16)) {
      snprintf(err, err_len, "scatter epilogue: needs M %% (32 * world) == 0, N %% 8 == 0, 16-bit output, 2..8 ranks");
      return B2_ERR_ARG;
    }
    if (cfg->bn != 256) cfg = &kConfigs[a.M > BM ? 4 : 3];
  }
  if (used_config) *used_config = cfg->name;
  const int bn = cfg->bn;
  const int cg = cfg->cg;
  const int b_rows = bn / cg;

  const int out_bytes = (a.out_dtype == B2_DTYPE_BF16 || a.out_dtype == B2_DTYPE_F16) ? 2 : 4;
  CUtensorMapDataType d_dt = a.out_dtype == B2_DTYPE_BF16  ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                             : a.out_dtype == B2_DTYPE_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                             : a.out_dtype == B2_DTYPE_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                                           : CU_TENSOR_MAP_DATA_TYPE_INT32;
  KernelParams p;
  p.As = a.As;
  p.Bs = a.Bs;
  p.bias = a.bias;
  p.bias_dtype = a.bias_dtype;
  p.out_dtype = a.out_dtype;
  p.tile_expert = a.tile_expert;
  p.batch = a.batch > 1 ? a.batch : 1;
  p.scalar_scales = a.scalar_scales;
  p.tail_tiles = 0;
  p.rs_counter = a.rs_counter;
  p.rs_rows = scatter ? a.M / a.rs_world : 0;
  p.rs_rank = a.rs_rank;
  p.M = a.M;
  p.N = a.N;
  p.K = a.K;
  p.num_m_blocks = (a.M + BM * cg - 1) / (BM * cg);
  p.num_n_blocks = (a.N + bn - 1) / bn;
  p.num_kb = (a.K + BK - 1) / BK;
  p.stages = stages_for(bn, b_rows);
  {  // developer override for pipeline-depth experiments (tools/gemm_prof.py); never deepens beyond what fits
    static const int forced_stages = [] {
      const char* e = getenv("B2_GEMM_STAGES");
      return e ? atoi(e) : 0;
    }();
    if (forced_stages >= 2 && forced_stages < p.stages) p.stages = forced_stages;
    static const int dbg = [] {
      const char* e = getenv("B2_GEMM_DEBUG");
      return e ? atoi(e) : 0;
    }();
    p.debug_flags = dbg | forced_debug;
  }
  p.rs_rot = 0;
  if (scatter) {
    const int owner0 = (a.rs_rank + 1) % a.rs_world;
    p.rs_rot = static_cast<int>((static_cast<long long>(owner0) * p.rs_rows) / (BM * cg));  // row block holding that owner's first row
  }
  choose_raster(a.M, a.N, a.K, BM * cg, bn, p.num_m_blocks, p.num_n_blocks, &p.group_m, &p.band_n, &p.policy_a,
                &p.policy_b);
  p.prof = static_cast<unsigned long long*>(profile_buffer());
  // decode shapes: fetch only the rows that exist (rounded to the 8-row swizzle atom);
  // the untouched tail of the smem tile feeds MMA rows whose outputs are clipped by the store.
  p.a_box_rows = (a.M >= BM || cg == 2) ? BM : ((a.M + 7) / 8) * 8;

  CUtensorMap ta, tb, td;
  const uint32_t d_box_cols = (128 / out_bytes) < bn ? (128 / out_bytes) : bn;
  if (a.batch > 1) {
    // batched (bmm_fp8): 3-D maps, one depth-1 box per tile; rows are clipped per batch entry
    if (a.tile_expert || !make_tmap_3d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.A, a.batch, a.M, a.K, a.lda,
                                       a.a_batch_stride, p.a_box_rows, BK) ||
        !make_tmap_3d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.W, a.batch, a.N, a.K, a.ldw, a.w_batch_stride, b_rows, BK) ||
        !make_tmap_3d(&td, d_dt, out_bytes, a.D, a.batch, a.M, a.N, a.ldd * out_bytes, a.d_batch_stride * out_bytes, 32,
                      d_box_cols)) {
      snprintf(err, err_len, "cuTensorMapEncodeTiled failed (batched, batch=%d M=%d N=%d K=%d)", a.batch, a.M, a.N, a.K);
      return B2_ERR_CUDA;
    }
  } else
  if (!make_tmap_2d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.A, a.M, a.K, a.lda, p.a_box_rows, BK) ||
      !make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.W,
                    static_cast<uint64_t>(a.N) * (a.tile_expert ? a.num_experts : 1), a.K, a.ldw, b_rows, BK) ||
      !make_tmap_2d(&td, d_dt, out_bytes, a.D, a.M, a.N, a.ldd * out_bytes, 32,
                    (128 / out_bytes) < bn ? (128 / out_bytes) : bn)) {  // 64-byte boxes use SWIZZLE_64B
    snprintf(err, err_len, "cuTensorMapEncodeTiled failed (M=%d N=%d K=%d lda=%lld ldw=%lld ldd=%lld)", a.M, a.N,
             a.K, (long long)a.lda, (long long)a.ldw, (long long)a.ldd);
    return B2_ERR_CUDA;
  }
  const int tiles = p.num_m_blocks * p.num_n_blocks * p.batch;
  const int units = sms / cg;
  {
    // tail split: when the ragged last wave has at most half a wave of 256-wide tiles, issue them as twice as many
    // 128-wide work items (see the kernel).  N = 5120, K = 13824, M = 4096: 320 tiles on 74 CTA pairs = 4 waves + 24 ->
    // 48 half tiles: 4.5 tile times instead of 5.
    static const bool tail_on = [] {
      const char* e = getenv("B2_GEMM_TAIL_SPLIT");
      return !(e && e[0] == '0');
    }();
    const int rem = tiles % units;
    if (tail_on && !(p.debug_flags & 8) && bn == 256 && !a.tile_expert && p.batch == 1 && rem > 0 && rem * 2 <= units && a.N % 128 == 0)
      p.tail_tiles = rem;
  }
  const int items = tiles + p.tail_tiles;
  const int grid = (items < units ? items : units) * cg;
  const size_t smem = smem_bytes_for(bn, b_rows, p.stages);
  const int out_kind = a.out_dtype == B2_DTYPE_BF16 ? 0 : (a.out_dtype == B2_DTYPE_F16 ? 1 : 2);
  if (scatter) {
    PeerMaps<true> peers;
    for (int r = 0; r < 8; ++r) {
      const int src = r < a.rs_world ? r : 0;  // unused entries repeat a valid map
      if (!make_tmap_gather4d(&peers.m[r], d_dt, a.rs_peer_bufs[src], a.rs_world, p.rs_rows, a.N, a.rs_set_bytes, 32)) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled failed (scatter map of rank %d: rows=%d N=%d set=%lld)", src,
                 p.rs_rows, a.N, a.rs_set_bytes);
        return B2_ERR_CUDA;
      }
    }
#define B2_RS_LAUNCH(KIND, OK)                                                                                        \
  (cg == 2 ? launch_one<KIND, OK, 256, 2, true>(ta, tb, td, p, grid, smem, st, err, err_len, peers)                   \
           : launch_one<KIND, OK, 256, 1, true>(ta, tb, td, p, grid, smem, st, err, err_len, peers))
    if (flavor == GEMM_S8) return out_kind == 0 ? B2_RS_LAUNCH(MmaKind::I8, 0) : B2_RS_LAUNCH(MmaKind::I8, 1);
    return out_kind == 0 ? B2_RS_LAUNCH(MmaKind::F8, 0) : B2_RS_LAUNCH(MmaKind::F8, 1);
#undef B2_RS_LAUNCH
  }
  if (flavor == GEMM_S8) {
    if (out_kind == 0) return launch_cfg<MmaKind::I8, 0>(*cfg, ta, tb, td, p, grid, smem, st, err, err_len);
    if (out_kind == 1) return launch_cfg<MmaKind::I8, 1>(*cfg, ta, tb, td, p, grid, smem, st, err, err_len);
    return launch_cfg<MmaKind::I8, 2>(*cfg, ta, tb, td, p, grid, smem, st, err, err_len);
  }
  if (out_kind == 0) return launch_cfg<MmaKind::F8, 0>(*cfg, ta, tb, td, p, grid, smem, st, err, err_len);
  if (out_kind == 1) return launch_cfg<MmaKind::F8, 1>(*cfg, ta, tb, td, p, grid, smem, st, err, err_len);
  return launch_cfg<MmaKind::F8, 2>(*cfg, ta, tb, td, p, grid, smem, st, err, err_len);
}

}  // namespace b2
