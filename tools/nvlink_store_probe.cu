// How fast can SMs store staged tiles into a PEER GPU's memory over NVLink, and does the shape of the store matter?
// (developer tool behind the scatter epilogue of gemm_sm100.cu: that kernel reaches 420-550 GB/s per rank out of 900.)
// GPU 0 runs the kernel, the destination lives on GPU 1 (cudaDeviceEnablePeerAccess); `local` = same test into GPU 0.
//   mode 0  TMA tensor store, 32 rows x 128 bytes per box into a row-major [R][pitch] destination (the epilogue's form)
//   mode 1  cp.async.bulk 1-D store of one contiguous 4 KB block per box
//   mode 2  32 cp.async.bulk 1-D stores of 128 bytes per box (the same segments as mode 0, issued one by one)
//   mode 3  st.global.v4 by the whole warp (512 contiguous bytes per instruction), 4 KB per "box"
// Every warp keeps DEPTH boxes in flight (wait_group.read DEPTH-1), 4 warps per CTA, one or two CTAs per SM.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sglang-diffusion_b200/csrc -I include \
//        tools/nvlink_store_probe.cu -o tools/bin/nvlink_store_probe && tools/bin/nvlink_store_probe
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#include "ptx.cuh"

using namespace b2;

#define CK(x)                                                                                \
  do {                                                                                       \
    cudaError_t e_ = (x);                                                                    \
    if (e_ != cudaSuccess) {                                                                 \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);        \
      exit(1);                                                                               \
    }                                                                                        \
  } while (0)

__device__ __forceinline__ void bulk_store_1d(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}

template <int N>
__device__ __forceinline__ void wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

constexpr int kPitch = 8192;  // bytes per destination row (N = 4096 bf16)

template <int MODE, int DEPTH>
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tmap, uint8_t* dst, long long boxes_total,
                                             int rows_total) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* my = smem + warp * DEPTH * 4096;
  for (int i = lane; i < DEPTH * 4096 / 16; i += 32) reinterpret_cast<uint4*>(my)[i] = make_uint4(i, warp, blockIdx.x, 7);
  fence_proxy_async_smem();
  __syncwarp();
  const long long stride = static_cast<long long>(gridDim.x) * 4;
  int buf = 0;
  const int boxes_per_row = kPitch / 128;
  for (long long b = static_cast<long long>(blockIdx.x) * 4 + warp; b < boxes_total; b += stride) {
    uint8_t* src = my + buf * 4096;
    if (MODE != 3) {
      wait_read<DEPTH - 1>();
      __syncwarp();
    }
    // (a real epilogue rewrites the staging buffer here)
    if constexpr (MODE == 0) {
      const int rb = static_cast<int>(b / boxes_per_row), cb = static_cast<int>(b % boxes_per_row);
      if (lane == 0) {
        tma_store_2d(&tmap, src, cb * 64, (rb * 32) % rows_total);
        tma_store_commit();
      }
    } else if constexpr (MODE == 1) {
      if (lane == 0) {
        bulk_store_1d(dst + b * 4096, src, 4096);
        tma_store_commit();
      }
    } else if constexpr (MODE == 2) {
      const long long rb = b / boxes_per_row, cb = b % boxes_per_row;
      bulk_store_1d(dst + (rb * 32 + lane) * kPitch + cb * 128, src + lane * 128, 128);
      tma_store_commit();
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        reinterpret_cast<uint4*>(dst + b * 4096)[i * 32 + lane] = reinterpret_cast<const uint4*>(src)[i * 32 + lane];
    }
    buf = (buf + 1) % DEPTH;
  }
  if (MODE != 3) tma_store_wait_all<0>();
  __threadfence_system();
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int MODE, int DEPTH>
static float run(const CUtensorMap& tm, uint8_t* dst, long long bytes, int ctas_per_sm, int rows_total) {
  auto k = probe<MODE, DEPTH>;
  const int smem = 4 * DEPTH * 4096;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int grid = 148 * ctas_per_sm;
  const long long boxes = bytes / 4096;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e9f;
  for (int it = 0; it < 4; ++it) {
    CK(cudaEventRecord(e0));
    k<<<grid, 128, smem>>>(tm, dst, boxes, rows_total);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (it > 0 && ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return static_cast<float>(bytes / (best * 1e-3) / 1e9);
}

int main() {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const long long bytes = 256ll << 20;
  const int rows_total = static_cast<int>(bytes / kPitch);
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaSetDevice(0));
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", reinterpret_cast<void**>(&enc), cudaEnableDefault, &qres));
  for (int target = 0; target < (ndev > 1 ? 2 : 1); ++target) {
    uint8_t* dst = nullptr;
    if (target == 1) {
      int can = 0;
      CK(cudaDeviceCanAccessPeer(&can, 0, 1));
      if (!can) {
        printf("no peer access 0 -> 1\n");
        break;
      }
      CK(cudaSetDevice(1));
      CK(cudaMalloc(&dst, bytes));
      CK(cudaSetDevice(0));
      cudaError_t pe = cudaDeviceEnablePeerAccess(1, 0);
      if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) CK(pe);
      (void)cudaGetLastError();
    } else {
      CK(cudaMalloc(&dst, bytes));
    }
    CUtensorMap tm;
    cuuint64_t dims[2] = {kPitch / 2, static_cast<cuuint64_t>(rows_total)};
    cuuint64_t strides[1] = {kPitch};
    cuuint32_t box[2] = {64, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dst, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      printf("encode failed %d\n", (int)r);
      return 1;
    }
    const char* where = target ? "peer (NVLink)" : "local HBM";
    for (int cps = 1; cps <= 2; ++cps) {
      printf("%s, %d CTA/SM x 4 warps:\n", where, cps);
      printf("  TMA tensor 32x128B   depth 2: %7.1f GB/s   depth 4: %7.1f GB/s\n", run<0, 2>(tm, dst, bytes, cps, rows_total),
             run<0, 4>(tm, dst, bytes, cps, rows_total));
      printf("  bulk 1-D 4 KB        depth 2: %7.1f GB/s   depth 4: %7.1f GB/s\n", run<1, 2>(tm, dst, bytes, cps, rows_total),
             run<1, 4>(tm, dst, bytes, cps, rows_total));
      printf("  bulk 1-D 32 x 128 B  depth 2: %7.1f GB/s   depth 4: %7.1f GB/s\n", run<2, 2>(tm, dst, bytes, cps, rows_total),
             run<2, 4>(tm, dst, bytes, cps, rows_total));
      printf("  st.global.v4 warp             %7.1f GB/s\n", run<3, 2>(tm, dst, bytes, cps, rows_total));
    }
  }
  return 0;
}
