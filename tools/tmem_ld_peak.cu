// TMEM -> register read bandwidth per SM (developer tool): W warps (4 or 8, two per lane quarter) loop over
// tcgen05.ld.32x32b of width x16 / x32 / x64 on their own lane quarter.  Decides how the block-wise GEMM may
// promote partial sums (every promoted element is one TMEM read).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sglang-diffusion_b200/csrc tools/tmem_ld_peak.cu -o tools/bin/tmem_ld_peak
#include <cstdio>
#include "ptx.cuh"
using namespace b2;

template <int X>
__device__ __forceinline__ void ld(uint32_t taddr, uint32_t* r);
template <>
__device__ __forceinline__ void ld<16>(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr) : "memory");
}
template <>
__device__ __forceinline__ void ld<32>(uint32_t taddr, uint32_t* r) {
  uint32_t (&q)[32] = *reinterpret_cast<uint32_t (*)[32]>(r);
  tmem_ld_32x32b_x32(taddr, q);
}
template <>
__device__ __forceinline__ void ld<64>(uint32_t taddr, uint32_t* r) {
  uint32_t (&a)[32] = *reinterpret_cast<uint32_t (*)[32]>(r);
  uint32_t (&b)[32] = *reinterpret_cast<uint32_t (*)[32]>(r + 32);
  tmem_ld_32x32b_x32(taddr, a);       // two back-to-back x32 (one wait) stand in for x64
  tmem_ld_32x32b_x32(taddr + 32, b);
}

template <int X>
__global__ void __launch_bounds__(320, 1) k(int iters, int nwarps, unsigned long long* out, uint32_t* sink) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc<1>(&tptr, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t base = tptr;
  uint32_t acc = 0;
  long long t0 = 0, t1 = 0;
  if (warp >= 2 && warp < 2 + nwarps) {
    const uint32_t t = base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + ((warp - 2) >> 2) * 64;
    uint32_t r[64];
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      ld<X>(t + ((i & 3) * 128), r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < X; ++j) acc ^= r[j];
    }
    t1 = clock64();
    if (lane == 0) out[blockIdx.x * 8 + (warp - 2)] = t1 - t0;
  }
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) { tc_fence_after_sync(); tmem_dealloc<1>(base, 512); }
}

template <int X>
void run(int nwarps) {
  unsigned long long* d; uint32_t* sink;
  cudaMalloc(&d, 148 * 8 * 8); cudaMalloc(&sink, 4);
  const int iters = 20000;
  k<X><<<148, 320>>>(100, nwarps, d, sink);
  k<X><<<148, 320>>>(iters, nwarps, d, sink);
  cudaError_t e = cudaDeviceSynchronize();
  unsigned long long h[8];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const double clk = (double)h[0] / iters;
  const double bytes = (double)nwarps * 32 * X * 4;
  printf("32x32b.x%-3d warps=%d : %7.1f clk per load round, %7.1f B/clk/SM  (%s)\n", X, nwarps, clk, bytes / clk,
         cudaGetErrorString(e));
}
int main() {
  for (int w : {1, 4, 8}) { run<16>(w); run<32>(w); run<64>(w); }
  return 0;
}
