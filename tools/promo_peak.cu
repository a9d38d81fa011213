// Promotion ceiling of the block-wise fp8 GEMM (developer tool): how fast can W warps drain 128 x 128 fp32 partial
// products out of TMEM (tcgen05.ld) and fold them into register accumulators (fma.rn.f32x2) -- alone, and while the
// tensor pipe is running kind::f8f6f4 MMAs into the same TMEM at full rate?  No TMA, no barriers between the two
// sides: this is the throughput bound the real kernel's handshakes can only approach.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sglang-diffusion_b200/csrc -I include \
//        tools/promo_peak.cu -o tools/bin/promo_peak && tools/bin/promo_peak
#include <cstdio>
#include <cstdlib>

#include "ptx.cuh"

using namespace b2;

__device__ __forceinline__ void ld16(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr) : "memory");
}

template <int N>
__device__ __forceinline__ void fma2(float* acc, const uint32_t* x, float scale) {
  uint64_t s2;
  asm("mov.b64 %0, {%1, %1};" : "=l"(s2) : "f"(scale));
#pragma unroll
  for (int j = 0; j < N; j += 2) {
    asm("{\n\t.reg .b64 xx, cc;\n\t"
        "mov.b64 xx, {%2, %3};\n\t"
        "mov.b64 cc, {%0, %1};\n\t"
        "fma.rn.f32x2 cc, xx, %4, cc;\n\t"
        "mov.b64 {%0, %1}, cc;\n\t}"
        : "+f"(acc[j]), "+f"(acc[j + 1])
        : "r"(x[j]), "r"(x[j + 1]), "l"(s2));
  }
}

// PW promotion warps (multiple of 4), each owning CW = 128 / (PW / 4) columns of its lane quarter; loads are X
// columns wide (16 or 32), double-buffered: the load of chunk c+1 is in flight during the FMAs of chunk c.
template <int PW, int X>
__global__ void __launch_bounds__(64 + 32 * PW, 1)
    promo(const uint8_t* __restrict__ seed, int iters, int do_mma, int do_fma, unsigned long long* out, float* sink) {
  constexpr int CW = 128 / (PW / 4);
  constexpr int NCH = CW / X;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 4 * 32768);
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 4);
  for (uint32_t i = threadIdx.x; i < 4 * 32768 / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(seed)[(i + blockIdx.x * 977u) & 0xffffu];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc<1>(tmem_ptr, 512);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  long long t0 = clock64();
  if (warp == 0) {
    if (do_mma && lane == 0) {
      constexpr uint32_t idesc128 = make_idesc<MmaKind::F8>(128, 128);
      constexpr uint32_t idesc256 = make_idesc<MmaKind::F8>(128, 256);
      const uint32_t idesc = do_mma == 3 ? idesc256 : idesc128;
      uint32_t phase[4] = {0, 0, 0, 0};
      if (do_mma == 3) iters /= 2;  // N = 256: each K block carries twice the flops
      for (int kb = 0; kb < iters; ++kb) {
        const int slot = do_mma == 2 ? 0 : (do_mma == 3 ? (kb & 1) * 2 : (kb & 3));
        const int nk = do_mma == 5 ? 8 : 4;  // 5: chains of 8 MMAs (a 256-deep K block)
        if (do_mma == 5 && kb >= iters / 2) break;
        if ((do_mma == 1 || do_mma >= 4) ? kb >= 4 : (do_mma == 3 ? kb >= 2 : kb >= 1)) {  // at most 4 K blocks (16 MMAs) in flight
          mbar_wait(&bars[slot], phase[slot]);
          phase[slot] ^= 1;
        }
        const uint64_t da = make_smem_desc_sw128(smem_u32(smem + slot * 32768));
        const uint64_t db = make_smem_desc_sw128(smem_u32(smem + slot * 32768 + 16384));
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (k < nk)
            umma<MmaKind::F8, 1>(tmem_base + slot * 128, da + 2 * (k & 3), db + 2 * (k & 3), idesc,
                                 do_mma == 4 ? (kb | k) != 0 : k != 0);
        umma_commit(&bars[slot]);
      }
      if (do_mma == 1 || do_mma >= 4) {
        if (do_mma == 5) iters /= 2;
        for (int s = 0; s < 4 && s < iters; ++s) {
          const int slot = (iters - 1 - s) & 3;
          mbar_wait(&bars[slot], phase[slot]);
          phase[slot] ^= 1;
        }
      } else if (do_mma == 2) {
        mbar_wait(&bars[0], phase[0]);
      } else {
        mbar_wait(&bars[0], phase[0]);
        mbar_wait(&bars[2], phase[2]);
      }
      out[blockIdx.x * 2 + 0] = clock64() - t0;
    }
  } else if (warp >= 2 && do_fma != 2) {
    const int ew = warp - 2;
    const uint32_t t = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (ew >> 2) * CW;
    float acc[CW];
#pragma unroll
    for (int i = 0; i < CW; ++i) acc[i] = 0.f;
    uint32_t r[2][X];
    float scale = 1.0f + lane * 1e-3f;
    for (int kb = 0; kb < iters; ++kb) {
      const uint32_t tb = t + (kb & 3) * 128;
      if constexpr (X == 32) tmem_ld_32x32b_x32(tb, *reinterpret_cast<uint32_t(*)[32]>(r[0]));
      else ld16(tb, r[0]);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        tmem_ld_wait();
        if (c + 1 < NCH) {
          if constexpr (X == 32) tmem_ld_32x32b_x32(tb + (c + 1) * X, *reinterpret_cast<uint32_t(*)[32]>(r[(c + 1) & 1]));
          else ld16(tb + (c + 1) * X, r[(c + 1) & 1]);
        }
        if (do_fma) fma2<X>(acc + c * X, r[c & 1], scale);
        else {
#pragma unroll
          for (int j = 0; j < X; ++j) acc[c * X + j] = __uint_as_float(__float_as_uint(acc[c * X + j]) ^ r[c & 1][j]);
        }
      }
      scale += 1e-6f;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < CW; ++i) s += acc[i];
    if (s == 123.456f) sink[0] = s;
    if (ew == 0 && lane == 0) out[blockIdx.x * 2 + 1] = clock64() - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

template <int PW, int X>
static void run(const uint8_t* seed, int do_mma, int do_fma) {
  auto kern = promo<PW, X>;
  const size_t smem = 4 * 32768 + 1024 + 256;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  unsigned long long* d;
  float* sink;
  cudaMalloc(&d, 148 * 2 * 8);
  cudaMalloc(&sink, 4);
  cudaMemset(d, 0, 148 * 2 * 8);
  const int iters = 20000;
  kern<<<148, 64 + 32 * PW, smem>>>(seed, 200, do_mma, do_fma, d, sink);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  kern<<<148, 64 + 32 * PW, smem>>>(seed, iters, do_mma, do_fma, d, sink);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  unsigned long long h[2];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const double flops = 148.0 * iters * 2.0 * 128 * 128 * 128;
  printf("warps=%-2d x%-2d mma=%d fma=%d : promotion %7.1f clk/Kblock (%6.1f B/clk/SM)  mma %7.1f clk/Kblock  | %8.3f ms -> %7.1f TFLOPS-equivalent (%s)\n",
         PW, X, do_mma, do_fma, (double)h[1] / iters, 65536.0 * iters / (double)(h[1] ? h[1] : 1), (double)h[0] / iters, ms,
         flops / (ms * 1e-3) / 1e12, cudaGetErrorString(e));
  fflush(stdout);
  cudaFree(d);
  cudaFree(sink);
}

int main() {
  uint8_t* seed;
  cudaMalloc(&seed, 1 << 20);
  {
    uint8_t* h = (uint8_t*)malloc(1 << 20);
    srand(1);
    // e4m3 codes with small exponents: finite products, realistic switching activity
    for (int i = 0; i < (1 << 20); ++i) h[i] = (uint8_t)((rand() & 0x3f) | (rand() & 0x80));
    cudaMemcpy(seed, h, 1 << 20, cudaMemcpyHostToDevice);
    free(h);
  }
  // MMA alone in the promotion pattern (no TMEM reader): 4 rotating 128-col slots / one slot / two 256-col slots
  run<8, 32>(seed, 1, 2);
  run<8, 32>(seed, 2, 2);
  run<8, 32>(seed, 3, 2);
  run<8, 32>(seed, 4, 2);  // always accumulate
  run<8, 32>(seed, 5, 2);  // chains of 8
  for (int mma = 0; mma < 2; ++mma)
    for (int fma = 0; fma < 2; ++fma) {
      run<4, 32>(seed, mma, fma);
      run<8, 32>(seed, mma, fma);
      run<8, 16>(seed, mma, fma);
      run<16, 32>(seed, mma, fma);
      run<16, 16>(seed, mma, fma);
    }
  return 0;
}
