// TMA feed-rate microbenchmark (developer tool): how many bytes per clock can one SM pull through
// cp.async.bulk.tensor into a ring of shared-memory slots, as a function of box shape, slots in flight,
// cluster multicast -- with nothing consuming the data.  Explains the operand-feed ceiling of the GEMM.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sglang-diffusion_b200/csrc -I include \
//        tools/tma_peak.cu -o tools/bin/tma_peak -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "gemm_common.cuh"

using namespace b2;

// 3-D view of a row-major [rows, K] byte matrix: dims {128 B, rows, K/128}; box {128, R, C}
static bool make_tmap_3d(CUtensorMap* m, const void* base, uint64_t rows, uint64_t K, uint32_t R, uint32_t C) {
  EncodeTiledFn enc = get_encode_fn();
  cuuint64_t dims[3] = {128, rows, K / 128};
  cuuint64_t strides[2] = {K, 128};
  cuuint32_t box[3] = {128, R, C};
  cuuint32_t estr[3] = {1, 1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

B2_DEVICE void tma_load_3d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

B2_DEVICE void tma_load_3d_mc(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "h"(mask)
      : "memory");
}

// Each CTA streams `iters` boxes from its own row block (K-chunks in order, wrapping), SLOTS in flight.
// MC = 1: plain.  MC = 2: cluster of 2, each CTA loads HALF the rows of every box and multicasts it to both.
template <int MC>
__global__ void __launch_bounds__(128, 1)
    tma_stream(const __grid_constant__ CUtensorMap tmap, int R, int C, int slots, int iters, int kchunks, int row_blocks,
               unsigned long long* out, bool two_d) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t box_bytes = 128u * R * C;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(slots) * box_bytes);
  const int rank = MC == 1 ? 0 : static_cast<int>(cluster_ctarank());
  if (threadIdx.x == 0) {
    for (int i = 0; i < slots; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  if (MC == 1) __syncthreads(); else cluster_sync_all();
  if (threadIdx.x == 0) {
    const int unit = MC == 1 ? blockIdx.x : blockIdx.x / 2;
    const long long t0 = clock64();
    int issued = 0, done = 0;
    uint32_t phase = 0;
    int slot_i = 0, slot_d = 0;
    while (done < iters) {
      while (issued < iters && issued - done < slots) {
        const int kc = (issued * C) % kchunks;
        const int rb = (unit + (issued * C) / kchunks) % row_blocks;
        mbar_arrive_expect_tx(&bars[slot_i], box_bytes);
        uint8_t* dst = smem + static_cast<size_t>(slot_i) * box_bytes;
        if (MC == 1 && two_d) {
          tma_load_2d(dst, &tmap, &bars[slot_i], kc * 128, rb * R, kEvictNormal);
        } else if (MC == 1) {
          tma_load_3d(dst, &tmap, &bars[slot_i], 0, rb * R, kc);
        } else {
          // half the rows each, multicast to both CTAs: the tensor map's box is R/2 rows
          tma_load_3d_mc(dst + rank * (box_bytes / 2), &tmap, &bars[slot_i], 0, rb * R + rank * (R / 2), kc, 0x3);
        }
        ++issued;
        if (++slot_i == slots) slot_i = 0;
      }
      mbar_wait(&bars[slot_d], phase);
      ++done;
      if (++slot_d == slots) { slot_d = 0; phase ^= 1; }
    }
    out[blockIdx.x] = static_cast<unsigned long long>(clock64() - t0);
  }
  if (MC == 1) __syncthreads(); else cluster_sync_all();
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const uint64_t rows = 28672, K = 4096;  // the gate_up weight: 117 MB, larger than half of L2
  uint8_t* w;
  cudaMalloc(&w, rows * K);
  cudaMemset(w, 1, rows * K);
  unsigned long long* out;
  cudaMalloc(&out, sizeof(unsigned long long) * sms);
  std::vector<unsigned long long> h(sms);
  printf("%-10s %5s %3s %6s %12s %12s %10s\n", "mode", "rows", "C", "slots", "clk/box", "B/clk/SM", "TB/s");
  struct Cfg { int mc, R, C, slots; bool two_d = false; };
  const Cfg cfgs[] = {
      {1, 128, 1, 4, true}, {1, 128, 1, 8, true}, {1, 256, 1, 4, true}, {1, 256, 1, 6, true}, {1, 64, 1, 8, true},
      {1, 128, 1, 4}, {1, 128, 1, 8}, {1, 128, 1, 12}, {1, 256, 1, 4}, {1, 256, 1, 6}, {1, 64, 1, 8}, {1, 64, 1, 16},
      {1, 128, 2, 4}, {1, 128, 2, 6}, {1, 256, 2, 3}, {1, 128, 4, 3}, {1, 32, 1, 16}, {1, 32, 1, 32},
      {2, 128, 1, 8}, {2, 128, 1, 12}, {2, 256, 1, 6}, {2, 128, 2, 6}, {2, 256, 2, 3},
  };
  for (const Cfg& c : cfgs) {
    CUtensorMap tm;
    const int box_rows = c.mc == 1 ? c.R : c.R / 2;
    if (c.two_d) {
      if (!make_tmap_2d(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, w, rows, K, K, box_rows, 128)) { printf("tensor map failed\n"); continue; }
    } else if (!make_tmap_3d(&tm, w, rows, K, box_rows, c.C)) { printf("tensor map failed\n"); continue; }
    const size_t smem = static_cast<size_t>(c.slots) * 128 * c.R * c.C + 1024 + 8 * c.slots + 64;
    if (smem > 232448) { printf("skip (smem)\n"); continue; }
    const int iters = 4096 / c.C;
    const int row_blocks = rows / c.R, kchunks = K / 128;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (c.mc == 1) {
        cudaFuncSetAttribute(tma_stream<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        tma_stream<1><<<sms, 128, smem>>>(tm, c.R, c.C, c.slots, iters, kchunks, row_blocks, out, c.two_d);
      } else {
        cudaFuncSetAttribute(tma_stream<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(sms); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, tma_stream<2>, tm, c.R, c.C, c.slots, iters, kchunks, row_blocks, out, false);
      }
      cudaEventRecord(e1);
      cudaError_t err = cudaDeviceSynchronize();
      if (err != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(err)); return 1; }
      cudaEventElapsedTime(&ms, e0, e1);
    }
    cudaMemcpy(h.data(), out, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost);
    double clk = 0;
    for (auto v : h) clk += (double)v;
    clk /= sms;
    const double box_bytes = 128.0 * c.R * c.C;          // bytes landing in each SM's smem per box
    printf("%-10s %5d %3d %6d %12.1f %12.1f %10.2f\n", c.mc == 2 ? "mcast2" : (c.two_d ? "unicast2d" : "unicast"), c.R, c.C, c.slots,
           clk / iters, box_bytes * iters / clk, box_bytes * iters * sms / (ms * 1e-3) / 1e12);
    fflush(stdout);
  }
  return 0;
}
