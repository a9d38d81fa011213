// Decode-shape (M <= 128) W8A8 GEMM for sm_100a: the weight matrix is streamed from HBM exactly once by
// MANY small CTAs, K split across the CTAs of a thread-block cluster and reduced through distributed
// shared memory.
//
//   D[M,N] = epilogue( A[M,K] x W[N,K]^T ),  M <= 128  -- HBM-bound on the N*K weight bytes
//
// Same arithmetic contract as gemm_sm100.cu (int8_scaled_mm / fp8_scaled_mm; reference
// sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747, fp8_gemm_kernel.cu:1450-1532, epilogue order of
// sgl-kernel/csrc/cpu/gemm_int8.cpp:20-60).  What differs is the schedule.  The persistent kernel gives one
// 128 x BN tile with the whole K extent to one SM: at M <= 128 there are fewer tiles than SMs (or a ragged
// second wave) and each SM has at most a dozen 4-32 KB boxes in flight, a third of what HBM3e needs.  Here
//
//   * a work unit is BN weight rows x (K / S) bytes; S CTAs of one cluster (cluster dims {S,1,1}, S <= 8) share a
//     tile, each running the usual TMA -> tcgen05.mma pipeline over its own K range into its own TMEM;
//   * only the `a_box_rows` (M rounded up to 8) activation rows are fetched per stage -- the MMA still reads 128
//     rows of shared memory, the rows past the box are whatever the ring holds and land in TMEM lanes nobody reads;
//   * units are small (tens of KB) so several CTAs are resident per SM and the block scheduler balances them;
//   * reduction: every CTA spills its raw 32-bit partial accumulators (rows < M only) to its own shared memory,
//     one cluster barrier, then CTA r sums columns [r*BN/S, (r+1)*BN/S) over all S peers IN RANK ORDER through
//     ld.shared::cluster (DSMEM), applies scale / bias / cast and stores to global.  No workspace, no atomics, no
//     second launch; int32 partial sums are exact so the int8 result is bit-identical to the unsplit kernel, and
//     the fp32 sums of the fp8 path are deterministic.
#include <cstdlib>

#include <cuda_fp8.h>

#include "gemm_common.cuh"

namespace b2 {

struct SplitParams {
  const float* As;
  const float* Bs;
  const void* bias;
  int bias_dtype;
  int out_dtype;
  void* D;
  long long ldd;  // elements
  int M, N, K;
  int num_kb;      // 128-byte K blocks in total
  int split;       // S = CTAs sharing one tile (K split)
  int amc;         // C = adjacent tiles sharing their activation loads: cluster = C x S CTAs, cluster rank = c * S + s;
                   // the C CTAs with the same s load a_box_rows / C rows of the A stage each and multicast them
  int stages;
  int a_box_rows;  // activation rows fetched per stage (multiple of 8, >= M)
  int bn;          // weight rows per tile = UMMA N (multiple of 16, 16..256): chosen at run time so that the grid is
                   // ONE balanced wave (e.g. N = 28672 -> 138 tiles of 208 rows on 148 SMs)
  uint32_t tmem_cols;   // power of two >= max(32, bn)
  uint32_t part_pitch;  // bytes per row of the partial buffer: bn rounded up to 32 words (the XOR swizzle's span)
  uint32_t bar_offset;  // byte offset of the barrier block inside the dynamic shared-memory window
  // Row-parallel tensor-parallel layers: instead of D, the epilogue stores this rank's partial output through the
  // NVSwitch multicast mapping into slot `rank` of EVERY rank's gather buffer (multimem.st), set (counter & 1) of two
  // -- the reduction that follows (allreduce_symm.cu, gathered form) then only reads local memory.  nullptr = off.
  // Fused activation quantizer (M <= 32; b2_linear_w8a8_int8 / _fp8): x is the bf16 / f16 activation matrix; the CTA
  // quantizes the rows of its own K slice into a RESIDENT shared-memory copy of A (no A loads, no codes round trip
  // through L2 / HBM, no separate launch) and keeps the row scales in shared memory.  nullptr = A comes by TMA.
  const void* x;
  long long ldx;       // elements
  int x_is_bf16;
  int a_res_blocks;    // K blocks of resident A per CTA (>= the longest K slice)
  uint8_t* mc_slot;             // multicast address of this rank's slot in set 0
  long long mc_set_bytes;       // distance between the two sets
  const uint32_t* mc_counter;   // device-resident call counter (local signal pad); read after griddepcontrol.wait
};

B2_DEVICE void multimem_st_v2(void* mc, uint32_t a, uint32_t b) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc), "r"(a), "r"(b) : "memory");
}

B2_DEVICE uint32_t ld_dsmem_u32(uint32_t local_addr, uint32_t cta) {
  uint32_t v;
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %1, %2;\n\t"
      "ld.shared::cluster.u32 %0, [ra];\n\t}"
      : "=r"(v)
      : "r"(local_addr), "r"(cta)
      : "memory");
  return v;
}

B2_DEVICE uint4 ld_dsmem_v4(uint32_t local_addr, uint32_t cta) {
  uint4 v;
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %4, %5;\n\t"
      "ld.shared::cluster.v4.u32 {%0, %1, %2, %3}, [ra];\n\t}"
      : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
      : "r"(local_addr), "r"(cta)
      : "memory");
  return v;
}

// partial accumulators in shared memory: [row][BN] words, 16-byte chunk index XORed with (row & 7) -- the
// row-per-lane 16-byte stores of the spill and the chunk-per-lane 16-byte loads of the reduction are both
// bank-conflict free
B2_DEVICE uint32_t part_offset(int row, int chunk, uint32_t pitch_bytes) {
  return static_cast<uint32_t>(row) * pitch_bytes + static_cast<uint32_t>((chunk ^ (row & 7)) << 4);
}

B2_DEVICE float x_to_f32(uint16_t bits, bool bf16) {
  return bf16 ? __uint_as_float(static_cast<uint32_t>(bits) << 16) : __half2float(__ushort_as_half(bits));
}

// The activation quantizer of quant.cu inside the GEMM (arithmetic of sgl-kernel/csrc/cpu/vec.h:190-228 for int8,
// sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:17-121 for e4m3 -- the same IEEE divisions and roundings, so codes and
// scales are bit-identical to the separate launch).  Called by the 128 epilogue threads.
//   pass 1  absmax of rows [0, M) over this CTA's K slice, reduced over the S CTAs of the cluster through DSMEM
//           (every CTA of a cluster owns a different slice of the same rows)
//   pass 2  re-read the slice (L1 / L2 hits), quantize, store 16-byte chunks into the resident A tiles in the
//           128-byte-swizzled K-major layout the UMMA descriptor expects (what TMA would have written)
// `amax_part` = 32 floats of this CTA visible cluster-wide; `smem_as` receives the row scales.
template <MmaKind KIND>
B2_DEVICE void fused_quant_phase(const SplitParams& p, uint8_t* smem_a, uint32_t a_stage_bytes, float* amax_part,
                                 float* smem_as, int kb0, int kb1, int S, int epi_warp, int lane) {
  const bool bf16 = p.x_is_bf16 != 0;
  const int L = (kb1 - kb0) * BK;  // elements of K in my slice (K % 16 == 0; the last block may be partial)
  const int k_begin = kb0 * BK;
  const int k_len = min(L, p.K - k_begin);
  const uint16_t* xb = static_cast<const uint16_t*>(p.x);
  for (int r = epi_warp; r < 32; r += 4) {
    float m = 0.f;
    if (r < p.M) {
      const uint4* row = reinterpret_cast<const uint4*>(xb + static_cast<long long>(r) * p.ldx + k_begin);
      const int nv = k_len / 8;
      for (int v0 = lane; v0 < nv; v0 += 32 * 8) {  // eight independent 16-byte loads in flight per thread
        uint4 raw[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) raw[u] = v0 + 32 * u < nv ? __ldg(row + v0 + 32 * u) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const uint16_t* e = reinterpret_cast<const uint16_t*>(&raw[u]);
#pragma unroll
          for (int j = 0; j < 8; ++j) m = fmaxf(m, fabsf(x_to_f32(e[j], bf16)));
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) amax_part[r] = m;
  }
  // every CTA's partial maxima are visible cluster-wide after this barrier (all 192 threads take part, see caller)
  __syncwarp();
}

template <MmaKind KIND>
B2_DEVICE void fused_quant_phase2(const SplitParams& p, uint8_t* smem_a, uint32_t a_stage_bytes, float* amax_part,
                                  float* smem_as, int kb0, int kb1, int S, int epi_warp, int lane) {
  const bool bf16 = p.x_is_bf16 != 0;
  const int k_begin = kb0 * BK;
  const int k_len = min((kb1 - kb0) * BK, p.K - k_begin);
  const uint16_t* xb = static_cast<const uint16_t*>(p.x);
  const uint32_t amax_addr = smem_u32(amax_part);
  for (int r = epi_warp; r < p.a_box_rows; r += 4) {
    float amax = 0.f;
    if (r < p.M) {
      if (S > 1) {
        for (int c = 0; c < S; ++c) amax = fmaxf(amax, __uint_as_float(ld_dsmem_u32(amax_addr + r * 4, c)));
      } else {
        amax = amax_part[r];
      }
    }
    float scale, inv;
    if constexpr (KIND == MmaKind::I8) {
      amax = fmaxf(amax, 1e-7f);
      scale = __fdiv_rn(amax, 127.f);
      inv = __fdiv_rn(127.f, amax);
    } else {
      scale = __fdiv_rn(amax, 448.f);
      inv = (scale == 0.f) ? 0.f : __frcp_rn(scale);
    }
    if (lane == 0 && r < p.M) smem_as[r] = scale;
    const uint4* row = reinterpret_cast<const uint4*>(xb + static_cast<long long>(r < p.M ? r : 0) * p.ldx + k_begin);
    uint8_t* a_row = smem_a + (r >> 3) * 1024 + (r & 7) * 128;
    const int nc = k_len / 16;
    for (int c0 = lane; c0 < nc; c0 += 32 * 4) {  // four chunks (eight 16-byte loads) in flight per thread
      uint4 ra[4], rb[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool on = r < p.M && c0 + 32 * u < nc;
        ra[u] = on ? __ldg(row + 2 * (c0 + 32 * u)) : make_uint4(0, 0, 0, 0);
        rb[u] = on ? __ldg(row + 2 * (c0 + 32 * u) + 1) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
      const int c = c0 + 32 * u;
      if (c >= nc) break;
      uint4 out = make_uint4(0, 0, 0, 0);
      if (r < p.M) {
        const uint4 r0 = ra[u], r1 = rb[u];
        const uint16_t* e0 = reinterpret_cast<const uint16_t*>(&r0);
        const uint16_t* e1 = reinterpret_cast<const uint16_t*>(&r1);
        uint8_t* o = reinterpret_cast<uint8_t*>(&out);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float a = __fmul_rn(x_to_f32(e0[j], bf16), inv), b = __fmul_rn(x_to_f32(e1[j], bf16), inv);
          if constexpr (KIND == MmaKind::I8) {
            o[j] = static_cast<uint8_t>(static_cast<int8_t>(__float2int_rn(a)));
            o[8 + j] = static_cast<uint8_t>(static_cast<int8_t>(__float2int_rn(b)));
          } else {
            o[j] = static_cast<uint8_t>(__nv_cvt_float_to_fp8(fmaxf(fminf(a, 448.f), -448.f), __NV_SATFINITE, __NV_E4M3));
            o[8 + j] = static_cast<uint8_t>(__nv_cvt_float_to_fp8(fmaxf(fminf(b, 448.f), -448.f), __NV_SATFINITE, __NV_E4M3));
          }
        }
      }
      const int kbl = c >> 3, c16 = c & 7;  // K block inside the slice, chunk inside its 128-byte row
      *reinterpret_cast<uint4*>(a_row + static_cast<size_t>(kbl) * a_stage_bytes + ((c16 ^ (r & 7)) << 4)) = out;
      }
    }
  }
}

template <MmaKind KIND, bool FUSEQ>
__global__ void __launch_bounds__(NUM_THREADS)
    gemm_splitk_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                       const SplitParams p) {
  const int BN = p.bn;
  const uint32_t B_STAGE_BYTES = static_cast<uint32_t>(BN) * BK;
  const uint32_t TMEM_COLS = p.tmem_cols;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int stages = p.stages;
  const uint32_t a_stage_bytes = static_cast<uint32_t>(p.a_box_rows) * BK;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + static_cast<size_t>(FUSEQ ? p.a_res_blocks : stages) * a_stage_bytes;
  uint32_t* part = reinterpret_cast<uint32_t*>(smem);  // aliases the operand ring once every MMA has retired

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int S = p.split;
  const int C = p.amc;
  const int crank = S * C > 1 ? static_cast<int>(cluster_ctarank()) : 0;
  const int rank = crank % S;   // K-split rank inside the tile
  const int cidx = crank / S;   // position inside the activation-multicast group
  const int tile = (static_cast<int>(blockIdx.x) / (S * C)) * C + cidx;
  uint16_t amc_mask = 0;
  for (int c = 0; c < C; ++c) amc_mask |= static_cast<uint16_t>(1u << (c * S + rank));
  const int a_piece_rows = p.a_box_rows / C;
  const int kb0 = static_cast<int>(static_cast<long long>(rank) * p.num_kb / S);
  const int kb1 = static_cast<int>(static_cast<long long>(rank + 1) * p.num_kb / S);

  // dynamic window layout: [ data region (host computes its size) | full[stages] empty[stages] acc_full tmem_ptr ]
  // -- the barriers live above everything the ring, the MMA's 128-row reads and the partials can touch
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.bar_offset);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* acc_full = empty_bar + stages;
  uint64_t* a_ready = acc_full + 1;  // FUSEQ: the 128 epilogue threads have written the resident A tiles
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(a_ready + 1);
  float* smem_as = reinterpret_cast<float*>(smem + p.bar_offset + 512);  // As[0..M) staged by the epilogue warps
  float* amax_part = smem_as + 64;  // FUSEQ (M <= 32): this CTA's partial row maxima, read by the cluster

  // Warp 0 initialises the barriers and puts the first ring of loads in flight BEFORE the block-wide sync: the
  // TMEM allocation and the sync are off the critical path of the first HBM round trip.
  const uint32_t tx_bytes = (FUSEQ ? 0u : a_stage_bytes) + B_STAGE_BYTES;
  const int first_pass = min(stages, kb1 - kb0);
  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmap_a);
      tma_prefetch_desc(&tmap_b);
      for (int i = 0; i < stages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], static_cast<uint32_t>(C));  // every CTA the stage's A tile went to has consumed it
      }
      mbar_init(acc_full, 1);
      mbar_init(a_ready, 128);
      fence_mbar_init();
      // The weights are touched exactly once (evict-first) and -- when the caller has declared them static
      // (b2_set_weight_prefetch) -- are requested before the preceding kernel of the stream has finished; the
      // activations come from that kernel, so their loads sit behind griddepcontrol.wait.  They are re-read by
      // every tile: keep them in L2.
      for (int i = 0; i < first_pass; ++i) {
        mbar_arrive_expect_tx(&full_bar[i], tx_bytes);
        tma_load_2d(smem_b + static_cast<size_t>(i) * B_STAGE_BYTES, &tmap_b, &full_bar[i], (kb0 + i) * BK, tile * BN,
                    kEvictFirst);
      }
      if (C == 1) {
        if constexpr (!FUSEQ) {
          pdl_wait_prior_grids();
          for (int i = 0; i < first_pass; ++i)
            tma_load_2d(smem_a + static_cast<size_t>(i) * a_stage_bytes, &tmap_a, &full_bar[i], (kb0 + i) * BK, 0,
                        kEvictLast);
        }
        pdl_launch_dependents();
      }
    }
    __syncwarp();
  }
  if (warp == 1) tmem_alloc<1>(tmem_ptr, TMEM_COLS);
  tc_fence_before_sync();
  if (C > 1) {
    // multicast activation loads land in the PEERS' shared memory and complete on their barriers: those must be
    // initialised first (the weight loads above are local and already in flight)
    __syncwarp();
    cluster_sync_all();
    if (warp == 0 && lane == 0) {
      pdl_wait_prior_grids();
      for (int i = 0; i < first_pass; ++i)
        tma_load_2d_mc(smem_a + static_cast<size_t>(i) * a_stage_bytes + static_cast<size_t>(cidx * a_piece_rows) * BK,
                       &tmap_a, &full_bar[i], (kb0 + i) * BK, cidx * a_piece_rows, amc_mask, kEvictLast);
      pdl_launch_dependents();
    }
    __syncwarp();
  } else {
    __syncthreads();
  }
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;

  if constexpr (FUSEQ) {
    // quantize my K slice of the activations into the resident A tiles while the first ring of weight tiles is in
    // flight (see fused_quant_phase); the cluster barrier in the middle carries the partial row maxima
    if (warp >= EPI_WARP0) {
      pdl_wait_prior_grids();  // x is the output of the preceding kernel
      fused_quant_phase<KIND>(p, smem_a, a_stage_bytes, amax_part, smem_as, kb0, kb1, S, warp - EPI_WARP0, lane);
    }
    if (S > 1) {
      __syncwarp();
      cluster_sync_all();
    } else {
      __syncthreads();
    }
    if (warp >= EPI_WARP0) {
      fused_quant_phase2<KIND>(p, smem_a, a_stage_bytes, amax_part, smem_as, kb0, kb1, S, warp - EPI_WARP0, lane);
      fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
      mbar_arrive(a_ready);
    }
  }

  // column slice of the tile this CTA reduces and writes (work item = row x group of 4 columns)
  const int epi_tid = (warp - EPI_WARP0) * 32 + lane;
  const int cols_per = (((BN + S - 1) / S) + 3) & ~3;
  const int c0 = rank * cols_per;
  const int width = max(0, min(cols_per, BN - c0));
  const int gpr = width >> 2;  // groups per row
  const bool has_bias = p.bias != nullptr;
  const bool raw_out = (KIND == MmaKind::I8) && p.out_dtype == B2_DTYPE_S32;
  const bool fixed_group = gpr > 0 && (128 % gpr) == 0;  // then a thread keeps its column group for every row
  float bs[4] = {0.f, 0.f, 0.f, 0.f}, bi[4] = {-0.f, -0.f, -0.f, -0.f};
  int cached_col = -1;

  if (warp == 0) {
    // ========================================== TMA producer (ring reuse) ====
    int stage = 0;
    uint32_t phase = 1;  // the first pass over the ring is already in flight
    for (int kb = kb0 + first_pass; kb < kb1; ++kb) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
        tma_load_2d(smem_b + static_cast<size_t>(stage) * B_STAGE_BYTES, &tmap_b, &full_bar[stage], kb * BK,
                    tile * BN, kEvictFirst);
        if constexpr (!FUSEQ) {
          if (C == 1)
            tma_load_2d(smem_a + static_cast<size_t>(stage) * a_stage_bytes, &tmap_a, &full_bar[stage], kb * BK, 0,
                        kEvictLast);
          else
            tma_load_2d_mc(smem_a + static_cast<size_t>(stage) * a_stage_bytes + static_cast<size_t>(cidx * a_piece_rows) * BK,
                           &tmap_a, &full_bar[stage], kb * BK, cidx * a_piece_rows, amc_mask, kEvictLast);
        }
      }
      __syncwarp();
      if (++stage == stages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    const uint32_t idesc = make_idesc<KIND>(BM, static_cast<uint32_t>(BN));
    int stage = 0;
    uint32_t phase = 0;
    if constexpr (FUSEQ) mbar_wait(a_ready, 0);
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after_sync();
      const uint64_t da = make_smem_desc_sw128(
          smem_u32(smem_a + static_cast<size_t>(FUSEQ ? kb - kb0 : stage) * a_stage_bytes));
      const uint64_t db = make_smem_desc_sw128(smem_u32(smem_b + static_cast<size_t>(stage) * B_STAGE_BYTES));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k)
          umma<KIND, 1>(tmem_base, da + static_cast<uint64_t>(k * (UMMA_K >> 4)),
                        db + static_cast<uint64_t>(k * (UMMA_K >> 4)), idesc, (kb > kb0 || k != 0) ? 1u : 0u);
        if (C == 1) umma_commit(&empty_bar[stage]);
        else umma_commit_mc(&empty_bar[stage], amc_mask);
      }
      __syncwarp();
      if (++stage == stages) {
        stage = 0;
        phase ^= 1;
      }
    }
    if (elect_one()) umma_commit(acc_full);
    __syncwarp();
  } else {
    // ============================== accumulators: TMEM -> own shared memory ====
    const int quarter = warp & 3;
    pdl_wait_prior_grids();  // As below and D at the end belong to / may still be read by the preceding kernels
    {
      const int t = (warp - EPI_WARP0) * 32 + lane;
      if (!FUSEQ && t < p.M && p.As != nullptr) smem_as[t] = __ldg(p.As + t);
    }
    // per-column scale / bias of this thread's column group: fetched while the main loop runs
    if (fixed_group && !raw_out && epi_tid < p.M * gpr) {
      const int col = tile * BN + c0 + (epi_tid % gpr) * 4;
      if (col < p.N) {
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.Bs + col));
        bs[0] = b4.x; bs[1] = b4.y; bs[2] = b4.z; bs[3] = b4.w;
        if (has_bias) {
#pragma unroll
          for (int j = 0; j < 4; ++j) bi[j] = load_bias(p.bias, p.bias_dtype, col + j);
        }
        cached_col = col;
      }
    }
    if (quarter * 32 < p.a_box_rows) {
      mbar_wait(acc_full, 0);
      tc_fence_after_sync();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
      const int row = quarter * 32 + lane;
      uint8_t* part_bytes = reinterpret_cast<uint8_t*>(part);
#pragma unroll 1
      for (int c = 0; c < BN / 16; ++c) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(t_row + static_cast<uint32_t>(c * 16), r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(part_bytes + part_offset(row, c * 4 + j, p.part_pitch)) =
              make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
      }
    }
  }

  // every CTA's partials are in its shared memory and visible cluster-wide after this barrier
  tc_fence_before_sync();
  __syncwarp();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<1>(tmem_base, TMEM_COLS);
  }

  if (warp >= EPI_WARP0) {
    // ===================== reduce my column slice over the cluster, epilogue ====
    // work item = (row, group of 4 columns); consecutive threads take consecutive groups of one row, so the DSMEM
    // loads are conflict free and the global stores of a warp cover whole contiguous row segments
    const uint32_t part_base = smem_u32(part);
    const int items = p.M * gpr;
    uint8_t* mc_out = nullptr;
    if (p.mc_slot != nullptr) mc_out = p.mc_slot + static_cast<long long>(__ldg(p.mc_counter) & 1u) * p.mc_set_bytes;
    for (int idx = epi_tid; idx < items; idx += 128) {
      const int row = idx / gpr;
      const int g = idx - row * gpr;
      const int c = c0 + g * 4;
      const int col = tile * BN + c;
      if (col >= p.N) continue;  // N % 8 == 0: a group of 4 columns is inside or outside as a whole
      const uint32_t addr = part_base + part_offset(row, c >> 2, p.part_pitch);
      uint4 acc = ld_dsmem_v4(addr, cidx * S);
      for (int pr = 1; pr < S; ++pr) {
        const uint4 o = ld_dsmem_v4(addr, cidx * S + pr);
        if constexpr (KIND == MmaKind::I8) {
          acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;  // two's complement: exact int32 sums
        } else {
          acc.x = __float_as_uint(__fadd_rn(__uint_as_float(acc.x), __uint_as_float(o.x)));
          acc.y = __float_as_uint(__fadd_rn(__uint_as_float(acc.y), __uint_as_float(o.y)));
          acc.z = __float_as_uint(__fadd_rn(__uint_as_float(acc.z), __uint_as_float(o.z)));
          acc.w = __float_as_uint(__fadd_rn(__uint_as_float(acc.w), __uint_as_float(o.w)));
        }
      }
      const size_t o = static_cast<size_t>(row) * static_cast<size_t>(p.ldd) + static_cast<size_t>(col);
      if (raw_out) {
        *reinterpret_cast<uint4*>(static_cast<uint32_t*>(p.D) + o) = acc;
        continue;
      }
      if (!fixed_group || cached_col != col) {
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.Bs + col));
        bs[0] = b4.x; bs[1] = b4.y; bs[2] = b4.z; bs[3] = b4.w;
        if (has_bias) {
#pragma unroll
          for (int j = 0; j < 4; ++j) bi[j] = load_bias(p.bias, p.bias_dtype, col + j);
        }
        cached_col = col;
      }
      const float as = smem_as[row];
      const uint32_t a4[4] = {acc.x, acc.y, acc.z, acc.w};
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = epilogue_value<KIND>(a4[j], as, bs[j], bi[j]);
      if (p.out_dtype == B2_DTYPE_F32) {
        *reinterpret_cast<float4*>(static_cast<float*>(p.D) + o) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
        uint2 w;
        w.x = pack16(v[0], v[1], p.out_dtype);
        w.y = pack16(v[2], v[3], p.out_dtype);
        if (mc_out != nullptr) multimem_st_v2(mc_out + o * 2, w.x, w.y);
        else *reinterpret_cast<uint2*>(static_cast<uint16_t*>(p.D) + o) = w;
      }
    }
  }
  // nobody leaves while a peer may still be reading its partials
  __syncwarp();
  cluster_sync_all();
}

// ------------------------------------------------------------------ host ----

struct SplitPlan {
  int bn, split, stages;
  size_t smem;
  int amc = 1;
};

// a_res > 0: the fused-quantizer layout -- `a_res` resident A tiles in front of a ring of weight stages only
static size_t splitk_smem(int bn, int a_rows, int stages, int a_res = 0) {
  const size_t a_stage = static_cast<size_t>(a_rows) * BK, b_stage = static_cast<size_t>(bn) * BK;
  size_t data = a_res > 0 ? a_res * a_stage + stages * b_stage : stages * (a_stage + b_stage);
  // 128-row reads of the last A tile
  const size_t mma_reach = (a_res > 0 ? (a_res - 1) * a_stage : stages * a_stage) + static_cast<size_t>(BM) * BK;
  if (mma_reach > data) data = mma_reach;
  const size_t rows_p = static_cast<size_t>((a_rows + 31) / 32) * 32;
  const size_t partial = rows_p * static_cast<size_t>((bn + 31) / 32 * 32) * 4;
  if (partial > data) data = partial;
  data = (data + 1023) / 1024 * 1024;
  return data + 1024;  // 512 B of barriers + TMEM pointer (2 * 24 + 2 slots of 8 bytes max), 512 B of staged As
}

static int deepest_ring(int bn, int a_rows, int kb_per, int a_res = 0) {
  int stages = kb_per < 24 ? kb_per : 24;
  while (stages > 1 && splitk_smem(bn, a_rows, stages, a_res) > static_cast<size_t>(SMEM_LIMIT)) --stages;
  return stages;
}

template <MmaKind KIND, bool FUSEQ>
static void configure_kernel(char* err, size_t err_len, int* status) {
  static thread_local int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev == dev) return;
  cudaError_t e =
      cudaFuncSetAttribute(gemm_splitk_kernel<KIND, FUSEQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) {
    snprintf(err, err_len, "cudaFuncSetAttribute(max dynamic smem): %s", cudaGetErrorString(e));
    *status = B2_ERR_CUDA;
    return;
  }
  attr_dev = dev;
}

// Schedule heuristic, fitted to tools/perf_sweep.py splitk --all (profiles/r01c_sweep_splitk_all.log):
//   * tile width 128 while the activation box is small (<= 32 rows), 64 above (the 16 KB activation stage then
//     dominates the ring);
//   * the smallest power-of-two split that puts >= ~0.85 x #SMs CTAs in the grid with >= 4 K-blocks each;
//   * the deepest ring (<= 12 stages) that still lets EVERY CTA of the grid be resident at once, and 4 stages when
//     the grid needs more than 3 CTAs per SM.
// Measured and rejected (profiles/r01d_sweep_splitk_balanced.log): one balanced wave of <= #SM CTAs with odd tile
// widths (208 rows for N = 28672, 176 x split 4 for N = 6144) and the deepest ring that fits.  It is no faster on
// the wide layer and 1.5x slower on the narrow ones: a CTA that owns a whole SM's shared memory leaves no room for
// the programmatically launched successor, so nothing hides the ~4 us launch / first-round-trip / drain cost.  The
// kernel keeps the run-time tile width (any multiple of 16) for shapes whose N is not a multiple of 128.
static SplitPlan plan_splitk(int N, int K, int a_rows, bool fuseq = false) {
  const int sms = device_sm_count();
  const int num_kb = (K + BK - 1) / BK;
  int bn = a_rows <= 32 ? 128 : 64;
  if (N < bn) bn = (N + 15) / 16 * 16;
  const int tiles = (N + bn - 1) / bn;
  // grid target: ~0.85 x #SMs CTAs for the narrow activation boxes; 1.5 x #SMs once the box is 64+ rows (two
  // resident CTAs per SM then cover each other's 8-16 KB activation loads and the DSMEM reduction tail --
  // profiles/r02g_splitk_plans_{int8,fp8}.log: C1 M=128 4096x4096 10.1 -> 9.2 us, M=64 8192x3584 11.6 -> 10.6 us)
  const int target = a_rows > 32 ? (sms * 3) / 2 : (sms * 85) / 100;
  int s = 1;
  while (s < 8 && tiles * s < target && num_kb / (2 * s) >= 4) s *= 2;
  const int kb_per = (num_kb + s - 1) / s;
  const int a_res = fuseq ? kb_per : 0;
  const long ctas = static_cast<long>(tiles) * s;
  const int need_occ = static_cast<int>((ctas + sms - 1) / sms);
  int stages = 4;
  if (need_occ <= 3) {
    const size_t budget = (228u * 1024u) / need_occ - 1024u;
    stages = 12;
    while (stages > 2 && splitk_smem(bn, a_rows, stages, a_res) > budget) --stages;
  }
  if (stages > kb_per) stages = kb_per;
  while (stages > 1 && splitk_smem(bn, a_rows, stages, a_res) > static_cast<size_t>(SMEM_LIMIT)) --stages;
  return SplitPlan{bn, s, stages, splitk_smem(bn, a_rows, stages, a_res)};
}

template <MmaKind KIND, bool FUSEQ>
static int launch_splitk(const CUtensorMap& ta, const CUtensorMap& tb, const SplitParams& p, int grid, size_t smem,
                         bool pdl, cudaStream_t st, char* err, size_t err_len) {
  int status = B2_OK;
  configure_kernel<KIND, FUSEQ>(err, err_len, &status);
  if (status != B2_OK) return status;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = static_cast<unsigned>(p.split * p.amc);
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = pdl ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_splitk_kernel<KIND, FUSEQ>, ta, tb, p);
  if (e != cudaSuccess) {
    snprintf(err, err_len, "split-K gemm kernel launch (cluster %d, smem %zu): %s", p.split, smem,
             cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  count_launch();
  return B2_OK;
}

// forced: 0 = plan; otherwise 0x80000 | (bn / 16 - 1) | (S - 1) << 4, optionally | ring depth cap << 8
int gemm_launch_splitk(GemmFlavor flavor, const GemmArgs& a, int forced, const char** used_config, cudaStream_t st,
                       char* err, size_t err_len) {
  const int a_rows = ((a.M + 7) / 8) * 8;
  const int num_kb = (a.K + BK - 1) / BK;
  const bool fuseq = a.x != nullptr;  // activations arrive as bf16 / f16 and are quantized inside the kernel
  SplitPlan plan;
  if (forced & 0x80000) {
    plan.bn = ((forced & 15) + 1) * 16;
    plan.split = ((forced >> 4) & 7) + 1;
    if (plan.split > num_kb) plan.split = num_kb;
    const int kb_per_f = (num_kb + plan.split - 1) / plan.split;
    plan.stages = deepest_ring(plan.bn, a_rows, kb_per_f, fuseq ? kb_per_f : 0);
    const int cap = (forced >> 8) & 31;  // developer sweeps: ring depth cap in bits 8-12
    if (cap >= 1 && cap < plan.stages) plan.stages = cap;
    plan.smem = splitk_smem(plan.bn, a_rows, plan.stages, fuseq ? kb_per_f : 0);
    plan.amc = 1 << ((forced >> 13) & 3);  // activation multicast group in bits 13-14 (log2)
  } else {
    plan = plan_splitk(a.N, a.K, a_rows, fuseq);
  }
  const int bn = plan.bn;
  {
    // the multicast group needs whole groups of tiles, 8-row pieces of the activation box, a cluster of <= 8 CTAs
    const int tiles_n = (a.N + bn - 1) / bn;
    while (plan.amc > 1 && (fuseq || plan.split * plan.amc > 8 || tiles_n % plan.amc != 0 || a_rows % (8 * plan.amc) != 0))
      plan.amc >>= 1;
  }
  if (plan.smem > static_cast<size_t>(SMEM_LIMIT)) {
    snprintf(err, err_len, "split-K plan does not fit shared memory (bn=%d rows=%d)", bn, a_rows);
    return B2_ERR_ARG;
  }
  static thread_local char name[64];
  snprintf(name, sizeof(name), "splitk_%dx%d_s%d_st%d%s%s", a_rows, bn, plan.split, plan.stages, fuseq ? "_fusedq" : "",
           plan.amc == 4 ? "_a4" : (plan.amc == 2 ? "_a2" : ""));
  if (used_config) *used_config = name;

  SplitParams p;
  p.As = a.As;
  p.Bs = a.Bs;
  p.bias = a.bias;
  p.bias_dtype = a.bias_dtype;
  p.out_dtype = a.out_dtype;
  p.D = a.D;
  p.ldd = a.ldd;
  p.M = a.M;
  p.N = a.N;
  p.K = a.K;
  p.num_kb = num_kb;
  p.split = plan.split;
  p.amc = plan.amc;
  p.stages = plan.stages;
  p.a_box_rows = a_rows;
  p.bn = bn;
  p.tmem_cols = 32;
  while (p.tmem_cols < static_cast<uint32_t>(bn)) p.tmem_cols *= 2;
  p.part_pitch = static_cast<uint32_t>((bn + 31) / 32 * 32) * 4u;
  p.bar_offset = static_cast<uint32_t>(plan.smem - 1024);
  p.x = a.x;
  p.ldx = a.ldx;
  p.x_is_bf16 = a.x_dtype == B2_DTYPE_BF16;
  p.a_res_blocks = fuseq ? (num_kb + plan.split - 1) / plan.split : 0;
  p.mc_slot = static_cast<uint8_t*>(a.mc_slot);
  p.mc_set_bytes = a.mc_set_bytes;
  p.mc_counter = a.mc_counter;

  CUtensorMap ta, tb;
  if (!make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.W, a.N, a.K, a.ldw, bn, BK) ||
      (fuseq ? (ta = tb, false)  // no A loads in the fused form: the parameter slot carries a valid (unused) map
             : !make_tmap_2d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a.A, a.M, a.K, a.lda, a_rows / plan.amc, BK))) {
    snprintf(err, err_len, "cuTensorMapEncodeTiled failed (M=%d N=%d K=%d lda=%lld ldw=%lld)", a.M, a.N, a.K,
             (long long)a.lda, (long long)a.ldw);
    return B2_ERR_CUDA;
  }
  const int tiles = (a.N + bn - 1) / bn;
  const int grid = tiles * plan.split;
  const bool pdl = weight_prefetch_enabled();
  if (fuseq) {
    if (flavor == GEMM_S8) return launch_splitk<MmaKind::I8, true>(ta, tb, p, grid, plan.smem, pdl, st, err, err_len);
    return launch_splitk<MmaKind::F8, true>(ta, tb, p, grid, plan.smem, pdl, st, err, err_len);
  }
  if (flavor == GEMM_S8) return launch_splitk<MmaKind::I8, false>(ta, tb, p, grid, plan.smem, pdl, st, err, err_len);
  return launch_splitk<MmaKind::F8, false>(ta, tb, p, grid, plan.smem, pdl, st, err, err_len);
}

}  // namespace b2
