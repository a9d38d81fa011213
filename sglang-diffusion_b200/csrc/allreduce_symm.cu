// Row-parallel partial-output all-reduce over NVLink 5 / NVSwitch peer memory (sm_100a), for the decode-shape
// W8A8 linear layers whose [M,N] bf16 partial products are summed across tensor-parallel ranks
// (reference: python/sglang/srt/layers/linear.py:1486-1490 -> tensor_model_parallel_all_reduce; the reference's own
// small-message path is sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:204-260).
//
// The GEMM of every rank writes its partial output straight into a SYMMETRIC buffer (same size on every rank, mapped
// into every peer's address space and into one NVSwitch multicast address; the mapping is set up by the host --
// torch.distributed._symmetric_memory -- and handed over as plain device pointers).  One kernel then does
//
//   two-shot / NVLS  (multicast pointer available):
//     barrier | rank r:  multimem.ld_reduce (the SWITCH adds the 8 partials in fp32) over slice r
//             |          multimem.st of the bf16 sums to slice r of every rank's buffer
//     barrier | every rank copies the reduced buffer to its private output tensor
//   one-shot / P2P   (no multicast, or tiny messages):
//     barrier | every rank loads all peers' partials (16-byte peer loads), adds them in fp32 IN RANK ORDER and
//             | writes its private output  (bit-identical on every rank)
//     barrier | (buffers may be overwritten by the next layer)
//
// Cross-GPU barriers are per CTA (CTA b of every rank handles the same elements): thread t < world posts the
// barrier number to peer t's signal pad with st.release.sys and polls its own pad with ld.acquire.sys; barrier
// numbers only grow (the per-CTA counter lives in the pad, so graph replays advance it), nothing is reset.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdio>

#include "b200_w8a8_internal.h"
#include "symm_common.cuh"

namespace b2 {

template <bool BF16>
AR_DEVICE uint4 multimem_ld_reduce_add(const void* mc) {
  uint4 v;
  if constexpr (BF16) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
  }
  return v;
}

AR_DEVICE void multimem_st(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

template <bool BF16>
AR_DEVICE void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if constexpr (BF16) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    } else {
      const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
      f[2 * i] = __low2float(h);
      f[2 * i + 1] = __high2float(h);
    }
  }
}

template <bool BF16>
AR_DEVICE uint4 pack8(const float (&f)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if constexpr (BF16) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    } else {
      const __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// nvec = number of 16-byte vectors of the message.  Vector v belongs to CTA (v / blockDim.x) % gridDim.x on every
// rank -- the per-CTA barriers rely on that -- and, in the two-shot path, to the slice of rank
// (v / (blockDim.x * gridDim.x)) % world.
template <bool BF16>
__global__ void __launch_bounds__(512) allreduce_two_shot_kernel(uint4* const* bufs, uint32_t* const* pads,
                                                                  uint4* mc, uint4* out, long long nvec, int rank,
                                                                  int world, unsigned long long timeout_ns) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // the next quantizer may come up; it waits for us
  asm volatile("griddepcontrol.wait;" ::: "memory");               // this rank's GEMM (and everything before it) is done
  const uint32_t epoch = pads[rank][kEpochWord + blockIdx.x];
  xgpu_barrier<false>(pads, rank, world, epoch + 1, timeout_ns);  // every rank's GEMM has written its partials (stream order)
  const long long stride = static_cast<long long>(blockDim.x) * gridDim.x;
  const long long lane0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  // four independent switch reductions in flight per thread: large (prefill) messages are NVLink-latency x
  // bytes-in-flight bound, not issue bound
  const long long step = stride * world;
  for (long long v = lane0 + stride * rank; v < nvec; v += 4 * step) {
    uint4 sum[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v + u * step < nvec) sum[u] = multimem_ld_reduce_add<BF16>(mc + v + u * step);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v + u * step < nvec) multimem_st(mc + v + u * step, sum[u]);
  }
  xgpu_barrier<true>(pads, rank, world, epoch + 2, timeout_ns);  // the sums of every slice have landed in this rank's buffer
  if (threadIdx.x == 0) pads[rank][kEpochWord + blockIdx.x] = epoch + 2;
  if (out == nullptr) return;  // in place: the caller consumes the sums from its symmetric buffer
  const uint4* mine = bufs[rank];
  for (long long v = lane0; v < nvec; v += 4 * stride) {
    uint4 t[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v + u * stride < nvec) t[u] = mine[v + u * stride];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (v + u * stride < nvec) out[v + u * stride] = t[u];
  }
  // no trailing barrier: a peer can only overwrite this rank's buffer through a later multimem.st, which sits
  // behind the first barrier of the NEXT call, and that barrier needs this CTA to have left this kernel
}

template <bool BF16>
__global__ void __launch_bounds__(512) allreduce_one_shot_kernel(uint4* const* bufs, uint32_t* const* pads, uint4* out,
                                                                  long long nvec, int rank, int world,
                                                                  unsigned long long timeout_ns) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const uint32_t epoch = pads[rank][kEpochWord + blockIdx.x];
  xgpu_barrier<false>(pads, rank, world, epoch + 1, timeout_ns);
  const long long stride = static_cast<long long>(blockDim.x) * gridDim.x;
  for (long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float acc[8];
    unpack8<BF16>(bufs[0][v], acc);
    for (int r = 1; r < world; ++r) {  // rank order: the same fp32 sum, bit for bit, on every rank
      float f[8];
      unpack8<BF16>(bufs[r][v], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
    }
    out[v] = pack8<BF16>(acc);
  }
  xgpu_barrier<false>(pads, rank, world, epoch + 2, timeout_ns);  // every peer has read this rank's partials: the buffer may be reused
  if (threadIdx.x == 0) pads[rank][kEpochWord + blockIdx.x] = epoch + 2;
}

// ---- gathered form: the GEMM epilogue has already PUSHED every rank's partials into every rank's buffer ----
// (gemm_splitk_sm100.cu, SplitParams::mc_*: multimem.st through the NVSwitch multicast mapping, slot `rank` of set
// counter & 1).  What is left of the all-reduce is one barrier -- "every rank's GEMM has completed" -- and a sum over
// `world` slots of LOCAL memory in rank order (fp32, one rounding: bit-identical on every rank and equal to the
// one-shot kernel above).  Two sets alternate so that a rank may start pushing call n+2 while a slow peer still sums
// call n+1; call n's readers are done by then: the pusher has passed barrier n+1, which every peer only enters after
// its call-n kernel completed.  The counter lives in the signal pad so that CUDA-graph replays advance it.
// The two sets sit a FIXED stride apart (set_vec, half the buffer on the Python side): with a size-dependent stride a
// call of a different size could land its set (n+1)&1 inside the region a slow peer is still summing for call n.
// The same kernel finishes the scatter epilogue of the prefill GEMM (gemm_sm100.cu, RS): there every rank's slots hold
// only the rows it owns, and `out` is this rank's [M / world, N] share of the sum (a reduce-scatter).

template <bool BF16>
__global__ void __launch_bounds__(512) allreduce_gathered_kernel(const uint4* __restrict__ local_buf,
                                                                  uint32_t* const* pads, uint4* __restrict__ out,
                                                                  long long nvec, long long set_vec, int rank,
                                                                  int world, unsigned long long timeout_ns) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");  // this rank's GEMM has completed: its pushes are performed
  uint32_t* pad = pads[rank];
  const uint32_t epoch = pad[kEpochWord + blockIdx.x];
  const uint32_t counter = pad[kGatherCounterWord];
  xgpu_barrier<true>(pads, rank, world, epoch + 1, timeout_ns);
  const uint4* set = local_buf + static_cast<long long>(counter & 1u) * set_vec;
  const long long stride = static_cast<long long>(blockDim.x) * gridDim.x;
  long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  // prefill-sized sums (the scatter form): four vectors per thread, 4 x world independent 16-byte loads in flight --
  // with one vector per thread 128 CTAs keep ~2 MB in flight, a third of what HBM needs
  for (; v + 3 * stride < nvec; v += 4 * stride) {
    float acc[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) unpack8<BF16>(__ldcs(set + v + u * stride), acc[u]);
    for (int r = 1; r < world; ++r) {
      uint4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) t[u] = __ldcs(set + static_cast<long long>(r) * nvec + v + u * stride);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8<BF16>(t[u], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[u][i] = __fadd_rn(acc[u][i], f[i]);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) out[v + u * stride] = pack8<BF16>(acc[u]);
  }
  for (; v < nvec; v += stride) {
    float acc[8];
    unpack8<BF16>(set[v], acc);
    for (int r = 1; r < world; ++r) {  // rank order: the same fp32 sum, bit for bit, on every rank
      float f[8];
      unpack8<BF16>(set[static_cast<long long>(r) * nvec + v], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
    }
    out[v] = pack8<BF16>(acc);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    pad[kEpochWord + blockIdx.x] = epoch + 1;
    __threadfence();
    const uint32_t t = atomicAdd(&pad[kGatherDoneWord], 1u);
    if (t == gridDim.x - 1) {  // last CTA of this call: every CTA has read the counter
      pad[kGatherDoneWord] = 0;
      pad[kGatherCounterWord] = counter + 1;
    }
  }
}


// ---- scatter GEMM + sum + all-gather = an all-reduce whose reduce-scatter half ran inside the GEMM ----
// After the scatter epilogue (gemm_sm100.cu, RS) this rank holds `world` partial copies of ITS rows.  One barrier, the
// rank-ordered fp32 sum of those slots (local memory), and the result is stored into rows [rank * rows, ...) of EVERY
// rank's copy of a second symmetric buffer -- plain peer stores, one per peer: the NVSwitch multicast would also loop
// the rows back to the sender, and it is the ingress side that is busy here ((world-1)/world * M*N*2 bytes per rank
// either way).  A second barrier, then the full [M, N] sum is in this rank's copy (full[rank]); the other ranks' rows
// are copied from there to `out` (this rank's rows went to `out` directly) unless the caller consumes it in place.
// Vector v of a slice belongs to CTA (v / blockDim.x) % gridDim.x on every rank, which is what lets the per-CTA
// barriers order the copy-out behind the peers' stores.
template <bool BF16>
__global__ void __launch_bounds__(512) allreduce_gathered_bcast_kernel(const uint4* __restrict__ local_buf,
                                                                        uint32_t* const* pads, uint4* const* full,
                                                                        uint4* __restrict__ out, long long nvec,
                                                                        long long set_vec, int rank, int world,
                                                                        unsigned long long timeout_ns) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");  // this rank's GEMM has completed: its stores are performed
  uint32_t* pad = pads[rank];
  const uint32_t epoch = pad[kEpochWord + blockIdx.x];
  const uint32_t counter = pad[kGatherCounterWord];
  xgpu_barrier<true>(pads, rank, world, epoch + 1, timeout_ns);
  const uint4* set = local_buf + static_cast<long long>(counter & 1u) * set_vec;
  const long long mine = static_cast<long long>(rank) * nvec;
  uint4* dst[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) dst[r] = r < world ? full[r] + mine : nullptr;
  if (out != nullptr) dst[rank] = out + mine;  // this rank's own rows go straight to the private output
  const long long stride = static_cast<long long>(blockDim.x) * gridDim.x;
  const long long v0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  long long v = v0;
  for (; v + 3 * stride < nvec; v += 4 * stride) {
    float acc[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) unpack8<BF16>(__ldcs(set + v + u * stride), acc[u]);
    for (int r = 1; r < world; ++r) {
      uint4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) t[u] = __ldcs(set + static_cast<long long>(r) * nvec + v + u * stride);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8<BF16>(t[u], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[u][i] = __fadd_rn(acc[u][i], f[i]);
      }
    }
    uint4 o[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) o[u] = pack8<BF16>(acc[u]);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      if (r < world) {
        const int p = (rank + r) % world;  // start with this rank, then the peers in rotated order
#pragma unroll
        for (int u = 0; u < 4; ++u) dst[p][v + u * stride] = o[u];
      }
    }
  }
  for (; v < nvec; v += stride) {
    float acc[8];
    unpack8<BF16>(set[v], acc);
    for (int r = 1; r < world; ++r) {
      float f[8];
      unpack8<BF16>(set[static_cast<long long>(r) * nvec + v], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
    }
    const uint4 o = pack8<BF16>(acc);
    for (int r = 0; r < world; ++r) dst[r][v] = o;
  }
  xgpu_barrier<true>(pads, rank, world, epoch + 2, timeout_ns);  // every rank's rows have landed in this rank's copy
  if (out != nullptr) {
    const uint4* local_full = full[rank];
    for (int r = 0; r < world; ++r) {
      if (r == rank) continue;
      const uint4* src = local_full + static_cast<long long>(r) * nvec;
      uint4* d = out + static_cast<long long>(r) * nvec;
      long long w = v0;
      for (; w + 3 * stride < nvec; w += 4 * stride) {
        uint4 t[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) t[u] = __ldcg(src + w + u * stride);
#pragma unroll
        for (int u = 0; u < 4; ++u) d[w + u * stride] = t[u];
      }
      for (; w < nvec; w += stride) d[w] = __ldcg(src + w);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    pad[kEpochWord + blockIdx.x] = epoch + 2;
    __threadfence();
    const uint32_t t = atomicAdd(&pad[kGatherDoneWord], 1u);
    if (t == gridDim.x - 1) {
      pad[kGatherDoneWord] = 0;
      pad[kGatherCounterWord] = counter + 1;
    }
  }
}

}  // namespace b2

using namespace b2;

extern "C" int b2_allreduce_symm_gathered(const void* local_buf, uint32_t* const* signal_pads_dev, void* out,
                                          int64_t numel, int64_t set_stride_bytes, int dtype, int rank, int world,
                                          int max_blocks, b2_stream_t stream);

#include <atomic>
static std::atomic<long long> g_ar_timeout_ms{0};

extern "C" void b2_set_allreduce_timeout_ms(long long ms) { g_ar_timeout_ms.store(ms < 0 ? 0 : ms); }
extern "C" long long b2_get_allreduce_timeout_ms(void) { return g_ar_timeout_ms.load(); }

extern "C" int b2_allreduce_symm(void* const* peer_bufs_dev, uint32_t* const* signal_pads_dev, void* multicast_ptr,
                                 void* out, int64_t numel, int dtype, int rank, int world, int algo,
                                 int max_blocks, b2_stream_t stream) {
  if (!peer_bufs_dev || !signal_pads_dev) return fail_msg(B2_ERR_ARG, "allreduce_symm: null pointer");
  if (dtype != B2_DTYPE_BF16 && dtype != B2_DTYPE_F16)
    return fail_msg(B2_ERR_ARG, "allreduce_symm: dtype must be bfloat16 or float16");
  if (world < 2 || world > 8 || rank < 0 || rank >= world)
    return fail_msg(B2_ERR_ARG, "allreduce_symm: world size must be 2..8 (one NVSwitch domain), rank inside it");
  if (numel <= 0 || numel % 8 != 0) return fail_msg(B2_ERR_ARG, "allreduce_symm: numel must be a positive multiple of 8");
  if (reinterpret_cast<uintptr_t>(out) % 16 != 0) return fail_msg(B2_ERR_ARG, "allreduce_symm: out must be 16-byte aligned");
  if (!out && (algo == 1 || !multicast_ptr))
    return fail_msg(B2_ERR_ARG, "allreduce_symm: in-place (out == NULL) needs the two-shot multicast path");
  if (algo == 2 && !multicast_ptr) return fail_msg(B2_ERR_ARG, "allreduce_symm: two-shot needs a multicast pointer");
  const long long nvec = numel / 8;
  // auto (measured, profiles/r01d_symm_allreduce_{2,8}gpu.log): the switch reduction wins at every size on 8 ranks
  // (1 MB: 15 us vs 33 us one-shot vs 34 us NCCL); between two ranks the ordered peer loads are ahead up to 4 MB+
  const bool two_shot =
      algo == 2 || !out || (algo == 0 && multicast_ptr != nullptr && (world > 2 || numel * 2 > (8ll << 20)));
  const int threads = 512;
  // per-CTA flags: world u32 per block in every rank's signal pad; the host says how many blocks the pad can take
  long long blocks = (nvec + threads - 1) / threads;
  if (two_shot) blocks = (blocks + world - 1) / world;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks > kEpochWord / world) blocks = kEpochWord / world;  // the flag words must stay below the epoch words
  if (blocks < 1) blocks = 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint4* const* bufs = reinterpret_cast<uint4* const*>(peer_bufs_dev);
  // Launched programmatically: the CTAs come up under the tail of the GEMM that produces the partials and wait for it
  // (griddepcontrol.wait, first instruction) -- the launch latency is off the critical path of the decode step.
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(blocks));
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  uint4* mc = static_cast<uint4*>(multicast_ptr);
  uint4* o = static_cast<uint4*>(out);
  const unsigned long long tmo = static_cast<unsigned long long>(g_ar_timeout_ms.load()) * 1000000ull;
  cudaError_t le;
  if (two_shot) {
    le = dtype == B2_DTYPE_BF16
             ? cudaLaunchKernelEx(&cfg, allreduce_two_shot_kernel<true>, bufs, signal_pads_dev, mc, o, nvec, rank, world, tmo)
             : cudaLaunchKernelEx(&cfg, allreduce_two_shot_kernel<false>, bufs, signal_pads_dev, mc, o, nvec, rank, world, tmo);
  } else {
    le = dtype == B2_DTYPE_BF16
             ? cudaLaunchKernelEx(&cfg, allreduce_one_shot_kernel<true>, bufs, signal_pads_dev, o, nvec, rank, world, tmo)
             : cudaLaunchKernelEx(&cfg, allreduce_one_shot_kernel<false>, bufs, signal_pads_dev, o, nvec, rank, world, tmo);
  }
  if (le != cudaSuccess) return fail_msg(B2_ERR_CUDA, "allreduce_symm launch: %s", cudaGetErrorString(le));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail_msg(B2_ERR_CUDA, "allreduce_symm launch: %s", cudaGetErrorString(e));
  count_launch();
  return B2_OK;
}

extern "C" int b2_allreduce_symm_gathered(const void* local_buf, uint32_t* const* signal_pads_dev, void* out,
                                          int64_t numel, int64_t set_stride_bytes, int dtype, int rank, int world,
                                          int max_blocks, b2_stream_t stream) {
  if (!local_buf || !signal_pads_dev || !out) return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered: null pointer");
  if (dtype != B2_DTYPE_BF16 && dtype != B2_DTYPE_F16)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered: dtype must be bfloat16 or float16");
  if (world < 2 || world > 8 || rank < 0 || rank >= world)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered: world size must be 2..8, rank inside it");
  if (numel <= 0 || numel % 8 != 0) return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered: numel must be a positive multiple of 8");
  if (reinterpret_cast<uintptr_t>(out) % 16 != 0 || reinterpret_cast<uintptr_t>(local_buf) % 16 != 0)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered: buffers must be 16-byte aligned");
  const long long nvec = numel / 8;
  if (set_stride_bytes == 0) set_stride_bytes = numel * 2 * world;
  if (set_stride_bytes % 16 != 0 || set_stride_bytes < numel * 2 * world)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered: set stride must be a multiple of 16 and hold world slots");
  const long long set_vec = set_stride_bytes / 16;
  const int threads = 512;
  long long blocks = (nvec + threads - 1) / threads;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks > kEpochWord / world) blocks = kEpochWord / world;
  if (blocks < 1) blocks = 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(blocks));
  cfg.blockDim = dim3(threads);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const unsigned long long tmo = static_cast<unsigned long long>(g_ar_timeout_ms.load()) * 1000000ull;
  const uint4* lb = static_cast<const uint4*>(local_buf);
  uint4* o = static_cast<uint4*>(out);
  cudaError_t le = dtype == B2_DTYPE_BF16
                       ? cudaLaunchKernelEx(&cfg, allreduce_gathered_kernel<true>, lb, signal_pads_dev, o, nvec, set_vec, rank, world, tmo)
                       : cudaLaunchKernelEx(&cfg, allreduce_gathered_kernel<false>, lb, signal_pads_dev, o, nvec, set_vec, rank, world, tmo);
  if (le != cudaSuccess) return fail_msg(B2_ERR_CUDA, "allreduce_symm_gathered launch: %s", cudaGetErrorString(le));
  count_launch();
  return B2_OK;
}

extern "C" int b2_allreduce_symm_gathered_bcast(const void* local_buf, uint32_t* const* signal_pads_dev,
                                                void* const* full_bufs_dev, void* out, int64_t numel_slice,
                                                int64_t set_stride_bytes, int dtype, int rank, int world,
                                                int max_blocks, b2_stream_t stream) {
  if (!local_buf || !signal_pads_dev || !full_bufs_dev)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered_bcast: null pointer");
  if (dtype != B2_DTYPE_BF16 && dtype != B2_DTYPE_F16)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered_bcast: dtype must be bfloat16 or float16");
  if (world < 2 || world > 8 || rank < 0 || rank >= world)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered_bcast: world size must be 2..8, rank inside it");
  if (numel_slice <= 0 || numel_slice % 8 != 0)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered_bcast: numel_slice must be a positive multiple of 8");
  if (reinterpret_cast<uintptr_t>(out) % 16 != 0 || reinterpret_cast<uintptr_t>(local_buf) % 16 != 0)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered_bcast: buffers must be 16-byte aligned");
  const long long nvec = numel_slice / 8;
  if (set_stride_bytes == 0) set_stride_bytes = numel_slice * 2 * world;
  if (set_stride_bytes % 16 != 0 || set_stride_bytes < numel_slice * 2 * world)
    return fail_msg(B2_ERR_ARG, "allreduce_symm_gathered_bcast: set stride must be a multiple of 16 and hold world slots");
  const long long set_vec = set_stride_bytes / 16;
  const int threads = 512;
  long long blocks = (nvec + threads - 1) / threads;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks > kEpochWord / world) blocks = kEpochWord / world;
  if (blocks < 1) blocks = 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>(blocks));
  cfg.blockDim = dim3(threads);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const unsigned long long tmo = static_cast<unsigned long long>(g_ar_timeout_ms.load()) * 1000000ull;
  const uint4* lb = static_cast<const uint4*>(local_buf);
  uint4* const* fb = reinterpret_cast<uint4* const*>(full_bufs_dev);
  uint4* o = static_cast<uint4*>(out);
  cudaError_t le =
      dtype == B2_DTYPE_BF16
          ? cudaLaunchKernelEx(&cfg, allreduce_gathered_bcast_kernel<true>, lb, signal_pads_dev, fb, o, nvec, set_vec, rank, world, tmo)
          : cudaLaunchKernelEx(&cfg, allreduce_gathered_bcast_kernel<false>, lb, signal_pads_dev, fb, o, nvec, set_vec, rank, world, tmo);
  if (le != cudaSuccess) return fail_msg(B2_ERR_CUDA, "allreduce_symm_gathered_bcast launch: %s", cudaGetErrorString(le));
  count_launch();
  return B2_OK;
}

extern "C" int b2_allreduce_symm_gathered_rmsnorm_quant(const void* local_buf, uint32_t* const* signal_pads_dev,
                                                        void* residual, int64_t ldr, const void* weight, float eps,
                                                        int qmode, uint8_t* q, int64_t ldq, float* s,
                                                        int64_t s_row_stride, int64_t s_grp_stride, void* y_out,
                                                        int64_t ldy, int M, int N, int64_t set_stride_bytes, int dtype,
                                                        int rank, int world, int max_blocks, b2_stream_t stream) {
  const char* op = "allreduce_symm_gathered_rmsnorm_quant";
  if (!local_buf || !signal_pads_dev || !weight || !q || !s) return fail_msg(B2_ERR_ARG, "%s: null pointer", op);
  if (dtype != B2_DTYPE_BF16 && dtype != B2_DTYPE_F16) return fail_msg(B2_ERR_ARG, "%s: dtype must be bfloat16 or float16", op);
  if (world < 2 || world > 8 || rank < 0 || rank >= world) return fail_msg(B2_ERR_ARG, "%s: world size must be 2..8, rank inside it", op);
  if (qmode != B2_QMODE_INT8_ROWS && qmode != B2_QMODE_FP8_ROWS && qmode != B2_QMODE_FP8_GROUP128)
    return fail_msg(B2_ERR_ARG, "%s: qmode must be INT8_ROWS, FP8_ROWS or FP8_GROUP128", op);
  if (M <= 0 || N <= 0 || N % 8 != 0 || (qmode == B2_QMODE_FP8_GROUP128 && N % 128 != 0))
    return fail_msg(B2_ERR_ARG, "%s: M=%d N=%d (N must be a multiple of 8; of 128 for the group mode)", op, M, N);
  if (N > 1024 * 4 * 8) return fail_msg(B2_ERR_ARG, "%s: N=%d exceeds the 32768 columns a CTA can hold in registers", op, N);
  int blocks_max = max_blocks < kEpochWord / world ? max_blocks : kEpochWord / world;
  if (blocks_max < 1) blocks_max = 1;
  const uintptr_t al = reinterpret_cast<uintptr_t>(local_buf) | reinterpret_cast<uintptr_t>(weight) |
                       reinterpret_cast<uintptr_t>(residual) | reinterpret_cast<uintptr_t>(y_out);
  if (al % 16 != 0 || reinterpret_cast<uintptr_t>(q) % 8 != 0) return fail_msg(B2_ERR_ARG, "%s: buffers must be 16-byte aligned", op);
  if ((residual && (ldr < N || ldr % 8 != 0)) || (y_out && (ldy < N || ldy % 8 != 0)) || ldq < N || ldq % 8 != 0)
    return fail_msg(B2_ERR_ARG, "%s: row strides must cover N and keep rows 16-byte (codes: 8-byte) aligned", op);
  const long long slot_bytes = static_cast<long long>(M) * N * 2;
  if (set_stride_bytes == 0) set_stride_bytes = slot_bytes * world;
  if (set_stride_bytes % 16 != 0 || set_stride_bytes < slot_bytes * world)
    return fail_msg(B2_ERR_ARG, "%s: set stride must be a multiple of 16 and hold world slots", op);
  FusedQuantArgs a{};
  a.x = local_buf; a.ldx = N; a.residual = residual; a.ldr = ldr; a.weight = weight; a.eps = eps; a.y = y_out; a.ldy = ldy;
  a.q = q; a.ldq = ldq; a.s = s; a.s_row_stride = s_row_stride; a.s_grp_stride = s_grp_stride; a.s_exp_stride = 0;
  a.masked_m = nullptr; a.tokens_per_expert = 1; a.M = M; a.K = N; a.q_eps = 1e-10f; a.fmin = -448.f; a.fmax = 448.f;
  a.gather_world = world; a.gather_rank = rank; a.gather_set_stride = set_stride_bytes / 2;
  a.gather_pads = signal_pads_dev;
  a.gather_timeout_ns = static_cast<unsigned long long>(g_ar_timeout_ms.load()) * 1000000ull;
  a.gather_max_blocks = blocks_max;
  cudaError_t le = fused_quant_launch(3, qmode, dtype, a, static_cast<cudaStream_t>(stream));
  if (le != cudaSuccess) return fail_msg(B2_ERR_CUDA, "%s launch: %s", op, cudaGetErrorString(le));
  count_launch();
  return B2_OK;
}

extern "C" int b2_reduce_scatter_symm_rmsnorm_quant_allgather(
    const void* local_buf, uint32_t* const* signal_pads_dev, void* residual_local, int64_t ldr, const void* weight,
    float eps, int qmode, void* const* act_bufs_dev, int64_t scale_offset_bytes, void* y_local, int64_t ldy, int M, int N,
    int64_t set_stride_bytes, int dtype, int rank, int world, int max_blocks, b2_stream_t stream) {
  const char* op = "reduce_scatter_symm_rmsnorm_quant_allgather";
  if (!local_buf || !signal_pads_dev || !weight || !act_bufs_dev) return fail_msg(B2_ERR_ARG, "%s: null pointer", op);
  if (dtype != B2_DTYPE_BF16 && dtype != B2_DTYPE_F16) return fail_msg(B2_ERR_ARG, "%s: dtype must be bfloat16 or float16", op);
  if (world < 2 || world > 8 || rank < 0 || rank >= world) return fail_msg(B2_ERR_ARG, "%s: world size must be 2..8, rank inside it", op);
  if (qmode != B2_QMODE_INT8_ROWS && qmode != B2_QMODE_FP8_ROWS)
    return fail_msg(B2_ERR_ARG, "%s: qmode must be INT8_ROWS or FP8_ROWS", op);
  if (M <= 0 || M % world != 0 || N <= 0 || N % 16 != 0 || N > 32768)
    return fail_msg(B2_ERR_ARG, "%s: M=%d must be a positive multiple of world, N=%d a multiple of 16 up to 32768", op, M, N);
  const int64_t codes = static_cast<int64_t>(M) * N;
  if (scale_offset_bytes < codes || scale_offset_bytes % 16 != 0)
    return fail_msg(B2_ERR_ARG, "%s: the scales must start at a 16-byte aligned offset behind the M*N codes", op);
  const uintptr_t al = reinterpret_cast<uintptr_t>(local_buf) | reinterpret_cast<uintptr_t>(weight) |
                       reinterpret_cast<uintptr_t>(residual_local) | reinterpret_cast<uintptr_t>(y_local);
  if (al % 16 != 0) return fail_msg(B2_ERR_ARG, "%s: buffers must be 16-byte aligned", op);
  const int rows = M / world;
  if ((residual_local && (ldr < N || ldr % 8 != 0)) || (y_local && (ldy < N || ldy % 8 != 0)))
    return fail_msg(B2_ERR_ARG, "%s: row strides must cover N and keep rows 16-byte aligned", op);
  const long long slot_bytes = static_cast<long long>(rows) * N * 2;
  if (set_stride_bytes == 0) set_stride_bytes = slot_bytes * world;
  if (set_stride_bytes % 16 != 0 || set_stride_bytes < slot_bytes * world)
    return fail_msg(B2_ERR_ARG, "%s: set stride must be a multiple of 16 and hold world slots", op);
  int blocks_max = max_blocks < kEpochWord / world ? max_blocks : kEpochWord / world;
  if (blocks_max < 1) blocks_max = 1;
  FusedQuantArgs a{};
  a.x = local_buf; a.ldx = N; a.residual = residual_local; a.ldr = ldr; a.weight = weight; a.eps = eps;
  a.y = y_local; a.ldy = ldy;
  a.q = nullptr; a.ldq = N; a.s = nullptr; a.s_row_stride = 1; a.s_grp_stride = 0; a.s_exp_stride = 0;
  a.masked_m = nullptr; a.tokens_per_expert = 1; a.M = rows; a.K = N; a.q_eps = 1e-10f; a.fmin = -448.f; a.fmax = 448.f;
  a.gather_world = world; a.gather_rank = rank; a.gather_set_stride = set_stride_bytes / 2;
  a.gather_pads = signal_pads_dev;
  a.gather_timeout_ns = static_cast<unsigned long long>(g_ar_timeout_ms.load()) * 1000000ull;
  a.gather_max_blocks = blocks_max;
  a.out_peers = reinterpret_cast<uint8_t* const*>(act_bufs_dev);
  a.out_row0 = static_cast<int64_t>(rank) * rows;
  a.out_scale_offset = scale_offset_bytes;
  cudaError_t le = fused_quant_launch(3, qmode, dtype, a, static_cast<cudaStream_t>(stream));
  if (le != cudaSuccess) return fail_msg(B2_ERR_CUDA, "%s launch: %s", op, cudaGetErrorString(le));
  count_launch();
  return B2_OK;
}

extern "C" int b2_symm_gather_counter_word(void) { return kGatherCounterWord; }
