// MMA-only ceiling of the B200 tensor pipe for the 8-bit kinds (developer tool; SURVEY section 6 asks for it).
// Every CTA (or CTA pair) loops tcgen05.mma on operand tiles that are already resident in shared memory --
// no TMA, no epilogue -- so the number is what the tensor cores sustain under the board's power cap with
// nothing else in the way.  Operand bytes are random (zeros would draw unrealistically little power).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sglang-diffusion_b200/csrc -I include \
//        tools/mma_peak.cu -o tools/bin/mma_peak && tools/bin/mma_peak
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

#include "ptx.cuh"

using namespace b2;

template <MmaKind KIND, int CG, int BN>
__global__ void __launch_bounds__(128, 1) mma_loop(const uint8_t* __restrict__ seed, int batches) {
  constexpr int B_ROWS = BN / CG;
  constexpr int K_STEPS = 4;   // 4 x 32 B along the 128 B swizzle row
  constexpr int NBUF = 4;      // operand buffers cycled through (as a pipeline ring would)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = 128 * 128, B_BYTES = B_ROWS * 128;
  uint8_t* sa = smem;
  uint8_t* sb = smem + NBUF * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sb + NBUF * B_BYTES);
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2);
  // A tiles come from the first half of the 1 MiB seed buffer, W tiles from the second half
  for (uint32_t i = threadIdx.x; i < NBUF * (A_BYTES + B_BYTES) / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(
        seed)[((i + blockIdx.x * 977u) & 0x7fffu) + (i < NBUF * A_BYTES / 16 ? 0u : 0x8000u)];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = CG == 1 ? 0 : static_cast<int>(cluster_ctarank());
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc<CG>(tmem_ptr, 512);
  tc_fence_before_sync();
  if (CG == 1) __syncthreads(); else { __syncwarp(); cluster_sync_all(); }
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  if (warp == 0 && lane == 0 && rank == 0) {
    constexpr uint32_t idesc = make_idesc<KIND>(128 * CG, BN);
    uint32_t phase[2] = {0, 0};
    for (int b = 0; b < batches; ++b) {
      const int which = b & 1;
      if (b >= 2) {  // keep at most two batches (2 x 32 MMAs) in flight
        mbar_wait(&bars[which], phase[which]);
        phase[which] ^= 1;
      }
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const int buf = s & (NBUF - 1);
        const uint64_t da = make_smem_desc_sw128(smem_u32(sa + buf * A_BYTES));
        const uint64_t db = make_smem_desc_sw128(smem_u32(sb + buf * B_BYTES));
        const uint32_t d = tmem_base + ((s & 1) ? BN : 0);
#pragma unroll
        for (int k = 0; k < K_STEPS; ++k) umma<KIND, CG>(d, da + 2 * k, db + 2 * k, idesc, (b | s | k) != 0);
      }
      if (CG == 1) umma_commit(&bars[which]); else umma_commit_cg2(&bars[which], 0x1);
    }
    const int last = (batches - 1) & 1;
    if (batches >= 2) mbar_wait(&bars[last ^ 1], phase[last ^ 1]);
    mbar_wait(&bars[last], phase[last]);
  }
  tc_fence_before_sync();
  if (CG == 1) __syncthreads(); else { __syncwarp(); cluster_sync_all(); }
  if (warp == 1) { tc_fence_after_sync(); tmem_dealloc<CG>(tmem_base, 512); }
}

template <MmaKind KIND, int CG, int BN>
static void run(const char* name, const uint8_t* seed, int sms, int batches) {
  auto kern = mma_loop<KIND, CG, BN>;
  const size_t smem = 4 * (128 * 128 + (BN / CG) * 128) + 1024 + 64;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(sms);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaLaunchKernelEx(&cfg, kern, seed, 64);  // warm-up
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  cudaLaunchKernelEx(&cfg, kern, seed, batches);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double macs = (double)(sms / CG) * batches * 32.0 * (128.0 * CG) * BN * 32.0;
  printf("%-28s batches=%-7d %9.3f ms  %8.1f TOPS  (%s)\n", name, batches, ms, 2.0 * macs / (ms * 1e-3) / 1e12,
         cudaGetErrorString(err));
  fflush(stdout);
}


// Ring variant: the issue structure of the real GEMM without its producer -- groups of G MMAs, one
// tcgen05.commit per group onto a ring of S barriers, and a group may only be issued once the group that
// used its ring slot S groups earlier has completed (+ `hop` extra SM clocks, standing in for the
// empty -> producer -> full round trip).  Shows how deep the ring must be to keep the tensor pipe busy.
template <MmaKind KIND, int CG, int BN>
__global__ void __launch_bounds__(128, 1) mma_ring(const uint8_t* __restrict__ seed, int groups, int G, int S, int hop,
                                                   int same_acc, int variant, long long* tout) {
  // variant bits: 1 = no tcgen05.fence after the wait, 2 = no wait at all, 4 = commit only every other group
  long long t_wait = 0, t_fence = 0, t_mma = 0, t_commit = 0;
  constexpr int B_ROWS = BN / CG;
  constexpr int NBUF = 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t A_BYTES = 128 * 128, B_BYTES = B_ROWS * 128;
  uint8_t* sa = smem;
  uint8_t* sb = smem + NBUF * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sb + NBUF * B_BYTES);
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 32);
  for (uint32_t i = threadIdx.x; i < NBUF * (A_BYTES + B_BYTES) / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(seed)[(i + blockIdx.x * 977u) & 0xffffu];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = CG == 1 ? 0 : static_cast<int>(cluster_ctarank());
  if (threadIdx.x == 0) {
    for (int i = 0; i < 32; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc<CG>(tmem_ptr, 512);
  tc_fence_before_sync();
  if (CG == 1) __syncthreads(); else { __syncwarp(); cluster_sync_all(); }
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  if (warp == 0 && lane == 0 && rank == 0) {
    constexpr uint32_t idesc = make_idesc<KIND>(128 * CG, BN);
    int slot = 0;
    uint32_t phase = 0;
    for (int g = 0; g < groups; ++g) {
      long long c0 = clock64();
      if (g >= S && !(variant & 2)) {
        if (variant & 16) { while (!mbar_test_wait(&bars[slot], phase ^ 1)) {} }
        else mbar_wait(&bars[slot], phase ^ 1);  // the commit of the previous trip round this ring slot
        if (hop) { const long long t0 = clock64(); while (clock64() - t0 < hop) {} }
      }
      long long c1 = clock64();
      if (!(variant & 1)) tc_fence_after_sync();
      long long c2 = clock64();
      const int buf = g & (NBUF - 1);
      const uint64_t da = make_smem_desc_sw128(smem_u32(sa + buf * A_BYTES));
      const uint64_t db = make_smem_desc_sw128(smem_u32(sb + buf * B_BYTES));
      uint32_t d = tmem_base + ((same_acc || !((g >> 5) & 1)) ? 0 : BN);  // switch accumulator every 32 groups
      if (variant & 8) d = tmem_base + (g & (512 / BN - 1)) * BN;  // block-wise pattern: a fresh slot per group
      for (int k = 0; k < G; ++k)
        umma<KIND, CG>(d, da + 2 * (k & 3), db + 2 * (k & 3), idesc, (variant & 8) ? k != 0 : (g | k) != 0);
      long long c3 = clock64();
      if (CG == 1) umma_commit(&bars[slot]); else umma_commit_cg2(&bars[slot], 0x1);
      long long c4 = clock64();
      t_wait += c1 - c0; t_fence += c2 - c1; t_mma += c3 - c2; t_commit += c4 - c3;
      if (++slot == S) { slot = 0; phase ^= 1; }
    }
    if (blockIdx.x == 0 && tout) { tout[0] = t_wait; tout[1] = t_fence; tout[2] = t_mma; tout[3] = t_commit; }
    // drain: wait for the last S groups
    for (int i = 0; i < (groups < S ? groups : S); ++i) {
      mbar_wait(&bars[slot], phase ^ 1);
      if (++slot == S) { slot = 0; phase ^= 1; }
    }
  }
  tc_fence_before_sync();
  if (CG == 1) __syncthreads(); else { __syncwarp(); cluster_sync_all(); }
  if (warp == 1) { tc_fence_after_sync(); tmem_dealloc<CG>(tmem_base, 512); }
}

template <MmaKind KIND, int CG, int BN>
static void run_ring(const char* name, const uint8_t* seed, int sms, int G, int S, int hop, int same_acc, int variant = 0) {
  static long long* tout = nullptr;
  if (!tout) cudaMalloc(&tout, 64);
  auto kern = mma_ring<KIND, CG, BN>;
  const size_t smem = 4 * (128 * 128 + (BN / CG) * 128) + 1024 + 512;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(sms);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  const int groups = 64000 / G;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaLaunchKernelEx(&cfg, kern, seed, 256, G, S, hop, same_acc, variant, (long long*)nullptr);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  cudaLaunchKernelEx(&cfg, kern, seed, groups, G, S, hop, same_acc, variant, tout);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double macs = (double)(sms / CG) * groups * G * (128.0 * CG) * BN * 32.0;
  long long ht[4] = {0, 0, 0, 0};
  cudaMemcpy(ht, tout, 32, cudaMemcpyDeviceToHost);
  printf("ring %-14s G=%-2d S=%-2d hop=%-4d same_acc=%d var=%d %8.3f ms  %8.1f TOPS | clk/group: wait %5.0f fence %5.0f mma-issue %5.0f commit %5.0f (%s)\n",
         name, G, S, hop, same_acc, variant, ms, 2.0 * macs / (ms * 1e-3) / 1e12, (double)ht[0] / groups, (double)ht[1] / groups,
         (double)ht[2] / groups, (double)ht[3] / groups, cudaGetErrorString(err));
  fflush(stdout);
}

int main(int argc, char** argv) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  std::vector<uint8_t> h(1 << 20);
  srand(1);
  for (auto& b : h) b = (uint8_t)(rand() & 0x77);  // e4m3-safe (no NaN codes), mixed signs via int8 view irrelevant
  if (argc > 1 && !strcmp(argv[1], "bench")) {
    // bench.py's operand distribution: activations ~ round(N(0, 32)) (first half), weights uniform in [-127, 127]
    for (size_t i = 0; i < h.size() / 2; ++i) {
      double u1 = (rand() + 1.0) / (RAND_MAX + 2.0), u2 = (rand() + 1.0) / (RAND_MAX + 2.0);
      double g = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2) * 32.0;
      int v = (int)lrint(g);
      v = v > 127 ? 127 : (v < -127 ? -127 : v);
      h[i] = (uint8_t)(int8_t)v;
    }
    for (size_t i = h.size() / 2; i < h.size(); ++i) h[i] = (uint8_t)(int8_t)(rand() % 255 - 127);
    printf("operands: bench.py distribution (A ~ N(0,32) int8, W uniform +-127)\n");
  }
  uint8_t* d;
  cudaMalloc(&d, h.size());
  cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  if (argc > 1 && !strcmp(argv[1], "ring128")) {
    // the block-wise fp8 GEMM's issue pattern: groups of 4 MMAs (one 128-deep K block), a commit per group
    for (int variant : {8, 8 | 16, 8 | 2})
      for (int G : {4})
        for (int S : {4, 8}) {
          run_ring<MmaKind::F8, 1, 128>("f8 1cta128x128", d, sms, G, S, 0, 1, variant);
          run_ring<MmaKind::F8, 2, 128>("f8 2cta256x128", d, sms, G, S, 0, 1, variant);
          run_ring<MmaKind::F8, 1, 256>("f8 1cta128x256", d, sms, G, S, 0, 1, variant);
          run_ring<MmaKind::F8, 2, 256>("f8 2cta256x256", d, sms, G, S, 0, 1, variant);
        }
    return 0;
  }
  if (argc > 1 && !strcmp(argv[1], "ring")) {
    const bool zero = argc > 2 && !strcmp(argv[2], "zeros");
    if (zero) cudaMemset(d, 0, h.size());
    for (int variant : {0, 1, 2, 3})
      for (int G : {2, 4, 6, 8, 16})
        for (int S : {2, 6}) {
          run_ring<MmaKind::I8, 1, 256>("i8 1cta128x256", d, sms, G, S, 0, 1, variant);
          run_ring<MmaKind::I8, 2, 256>("i8 2cta256x256", d, sms, G, S, 0, 1, variant);
        }
    if (argc > 3)
    for (int same = 0; same < 2; ++same)
      for (int hop : {0, 400, 1000})
        for (int G : {4, 8})
          for (int S : {2, 3, 4, 6, 8, 12}) {
            run_ring<MmaKind::I8, 1, 256>("i8 1cta128x256", d, sms, G, S, hop, same);
            if (same == 1) run_ring<MmaKind::I8, 2, 256>("i8 2cta256x256", d, sms, G, S, hop, same);
          }
    return 0;
  }
  for (int rep = 0; rep < 2; ++rep) {
    const int burst = 2000, sustained = 400000;  // ~1.5 ms and ~0.3 s per launch at full rate
    run<MmaKind::I8, 1, 256>("i8  1cta 128x256 burst", d, sms, burst);
    run<MmaKind::I8, 2, 256>("i8  2cta 256x256 burst", d, sms, burst);
    if (argc > 1 && !strcmp(argv[1], "bench")) {  // int8 only: the e4m3 view of these bytes contains NaN codes
      run<MmaKind::I8, 1, 256>("i8  1cta 128x256 sustained", d, sms, sustained);
      run<MmaKind::I8, 2, 256>("i8  2cta 256x256 sustained", d, sms, sustained);
      run<MmaKind::I8, 2, 256>("i8  2cta 256x256 sustained x4", d, sms, sustained * 4);
      continue;
    }
    run<MmaKind::F8, 1, 256>("f8  1cta 128x256 burst", d, sms, burst);
    run<MmaKind::F8, 2, 256>("f8  2cta 256x256 burst", d, sms, burst);
    run<MmaKind::I8, 1, 128>("i8  1cta 128x128 burst", d, sms, burst);
    run<MmaKind::I8, 1, 64>("i8  1cta 128x64  burst", d, sms, burst);
    run<MmaKind::I8, 1, 32>("i8  1cta 128x32  burst", d, sms, burst);
    run<MmaKind::I8, 1, 256>("i8  1cta 128x256 sustained", d, sms, sustained);
    run<MmaKind::I8, 2, 256>("i8  2cta 256x256 sustained", d, sms, sustained);
    run<MmaKind::F8, 1, 256>("f8  1cta 128x256 sustained", d, sms, sustained);
    run<MmaKind::F8, 2, 256>("f8  2cta 256x256 sustained", d, sms, sustained);
  }
  return 0;
}
