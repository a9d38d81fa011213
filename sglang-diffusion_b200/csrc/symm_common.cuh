// Cross-GPU barrier and signal-pad layout shared by the peer-memory reduction kernels (allreduce_symm.cu) and the fused
// gathered-sum + RMSNorm + quantizer (quant_fused.cu).  See allreduce_symm.cu for the protocol.
#pragma once
#include <cstdint>

namespace b2 {

#define AR_DEVICE __device__ __forceinline__

AR_DEVICE void st_release_sys(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
AR_DEVICE void st_relaxed_sys(uint32_t* addr, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
AR_DEVICE uint32_t ld_acquire_sys(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

// Signal pad of a rank (u32 words): [0, world * blocks) flags, flag (b, r) = last barrier number rank r announced to
// this rank for CTA b;  [kEpochWord + b] = barrier number CTA b of THIS rank has completed (kept on the device so
// that a CUDA-graph replay, whose kernel arguments are frozen, still advances).
constexpr int kEpochWord = 1024;
// [kTimeoutWord] = barrier number at which a wait of THIS rank gave up (0 = never); see b2_set_allreduce_timeout_ms
constexpr int kTimeoutWord = 1024 + 512;

// Barrier number `seq` between CTA `blockIdx.x` of all ranks.  A posted release-store to every peer (no NVLink round
// trip), then an acquire-poll of the local pad.  Barrier numbers only grow and a peer can be at most one barrier
// ahead, so "flag - seq >= 0" is the arrival test and nothing is ever reset.  Everything this CTA wrote before the
// barrier (bar.sync makes the release cumulative over the whole CTA) is visible to the peers after it.
// RELEASE = false when this CTA has written nothing a peer will read since the last barrier (the partials of the
// preceding GEMM kernel are already visible at its completion; "I have finished READING" needs no release either:
// the loads have returned before the dependent stores that precede the bar.sync were issued).
// A peer that is only late (first-iteration skew, a weight load or a compile on one rank, a debugger) must not poison
// this rank's CUDA context, so the wait never traps.  timeout_ns == 0 (default) waits for ever, like the reference's
// custom all-reduce (custom_all_reduce.cuh:204-260) and leaves hangs to the host watchdog; otherwise the wait gives up
// after timeout_ns of %globaltimer, records the barrier number in word kTimeoutWord of the local pad for the host to
// read (the result of that call is then undefined) and lets the kernel finish.
template <bool RELEASE>
AR_DEVICE void xgpu_barrier(uint32_t* const* pads, int rank, int world, uint32_t seq, unsigned long long timeout_ns) {
  __syncthreads();
  if (static_cast<int>(threadIdx.x) < world) {
    const int peer = static_cast<int>(threadIdx.x);
    uint32_t* send = pads[peer] + static_cast<size_t>(blockIdx.x) * world + rank;
    if constexpr (RELEASE) st_release_sys(send, seq);
    else st_relaxed_sys(send, seq);
    const uint32_t* recv = pads[rank] + static_cast<size_t>(blockIdx.x) * world + peer;
    unsigned int spins = 0;
    unsigned long long t0 = 0;
    while (static_cast<int32_t>(ld_acquire_sys(recv) - seq) < 0) {
      if (timeout_ns != 0 && (++spins & 1023u) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > timeout_ns) {
          pads[rank][kTimeoutWord] = seq;
          break;
        }
      }
    }
  }
  __syncthreads();
}

// gathered form (GEMM-epilogue push / scatter): call counter selecting the set in use, and the CTA-completion count
// that advances it -- both in the signal pad so that CUDA-graph replays stay consistent
constexpr int kGatherCounterWord = 1024 + 520;
constexpr int kGatherDoneWord = 1024 + 521;

}  // namespace b2
