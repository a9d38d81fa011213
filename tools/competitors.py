"""Library yardsticks for the W8A8 GEMMs on the same box, same shapes, interleaved with ours (developer tool; feeds
profiles/).  The reference's own sm_100 CUDA kernels (sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:796-1170,
fp8_blockwise_gemm_kernel.cu:39-196) cannot be built here (CMake FetchContent, no network); SURVEY 8(c) names the
stand-ins that ARE in the image.  None of them is the reference -- they are labelled as such.

    python tools/competitors.py [int8] [fp8] [block] [--libs torch,vllm,flashinfer] [--quick]

  int8   ours int8_scaled_mm            vs torch._int_mm (cuBLASLt s8->s32, no epilogue), vLLM cutlass_scaled_mm (s8)
  fp8    ours fp8_scaled_mm             vs torch._scaled_mm row-wise, vLLM cutlass_scaled_mm (e4m3 per-token/channel)
  block  ours fp8_blockwise_scaled_mm   vs vLLM cutlass_scaled_mm with 1x128 / 128x128 group scales,
                                           flashinfer gemm_fp8_nt_groupwise (call shape of
                                           python/sglang/srt/layers/quantization/fp8_utils.py:187-243)

Method: every candidate is captured in a CUDA graph over enough rotating operand sets to defeat the L2
(bytes between two uses of a buffer > 160 MB); rounds are INTERLEAVED (ours, lib A, lib B, ours, ...), 5 rounds, min
and median per candidate -- single-shot tables cannot rank kernels at the +-10 % board/power noise of this pool.
"""
from __future__ import annotations

import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sglang_diffusion_b200 as b2  # noqa: E402

dev = torch.device("cuda:0")
QUICK = "--quick" in sys.argv
LIBS = "torch,vllm,flashinfer"
for i, a in enumerate(sys.argv):
    if a == "--libs":
        LIBS = sys.argv[i + 1]
LIBS = set(LIBS.split(","))
want = [a for a in sys.argv[1:] if a in ("int8", "fp8", "block")] or ["int8", "fp8", "block"]
L2_BYTES = 160e6
ROUNDS = 3 if QUICK else 5


def nsets(bytes_per_set):
    return max(1, min(16, int(L2_BYTES // max(bytes_per_set, 1)) + 1))


class Cand:
    def __init__(self, name, fns):
        self.name, self.fns, self.ts, self.err = name, fns, [], None
        self.graph = None
        try:
            for f in fns[:2]:
                f()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            s_ = torch.cuda.Stream()
            s_.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s_):
                for f in fns:
                    f()
                with torch.cuda.graph(g, stream=s_):
                    for f in fns:
                        f()
            torch.cuda.current_stream().wait_stream(s_)
            torch.cuda.synchronize()
            self.graph = g
        except Exception as ex:  # noqa: BLE001
            try:
                torch.cuda.synchronize()
            except Exception:  # noqa: BLE001
                pass
            # eager fallback if the library call cannot be captured
            try:
                for f in fns[:2]:
                    f()
                torch.cuda.synchronize()
                self.err = f"graph capture failed ({type(ex).__name__}), eager timing"
            except Exception as ex2:  # noqa: BLE001
                self.err = f"unavailable: {type(ex2).__name__}: {str(ex2)[:160]}"
                self.fns = None

    def round(self, min_ms=25.0):
        if self.fns is None:
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 1
        while True:
            e0.record()
            for _ in range(reps):
                if self.graph is not None:
                    self.graph.replay()
                else:
                    for f in self.fns:
                        f()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            if ms >= min_ms or reps >= 2048:
                self.ts.append(ms * 1e-3 / (reps * len(self.fns)))
                return
            reps = max(reps * 2, int(reps * min_ms / max(ms, 1e-3)) + 1)


def report(tag, M, N, K, cands, ops):
    for _ in range(ROUNDS):
        for c in cands:
            c.round()
    base = None
    for c in cands:
        if c.fns is None:
            print(f"{tag:6s} M={M:<6d} N={N:<6d} K={K:<6d} {c.name:44s} {c.err}", flush=True)
            continue
        tmin, tmed = min(c.ts), statistics.median(c.ts)
        if base is None:
            base = tmed
        print(f"{tag:6s} M={M:<6d} N={N:<6d} K={K:<6d} {c.name:44s} min {tmin*1e6:9.1f} us  med {tmed*1e6:9.1f} us  "
              f"{ops/tmed/1e12:8.1f} T(FL)OPS med  ours/this = {tmed/base:5.2f}x speed-up"
              f"{'  [' + c.err + ']' if c.err else ''}", flush=True)


def vllm_ops():
    if "vllm" not in LIBS:
        return None
    try:
        import vllm._custom_ops as vo  # noqa: F401

        return vo
    except Exception as ex:  # noqa: BLE001
        print(f"# vllm unavailable: {type(ex).__name__}: {ex}")
        return None


def run_int8(M, N, K, vo):
    n = nsets(N * K + M * K + 2 * M * N)
    sets = [(torch.randint(-127, 128, (M, K), dtype=torch.int8, device=dev),
             torch.randint(-127, 128, (N, K), dtype=torch.int8, device=dev)) for _ in range(n)]
    As = torch.rand(M, 1, device=dev)
    Bs = torch.rand(N, 1, device=dev) * 1e-2
    cands = [Cand("ours int8_scaled_mm [" + "cfg" + "]", [(lambda A=A, W=W: b2.int8_scaled_mm(A, W.t(), As, Bs, torch.bfloat16, None)) for A, W in sets])]
    cands[0].name = f"ours int8_scaled_mm [{b2.capi.last_gemm_config()}]"
    if "torch" in LIBS and M > 16:
        cands.append(Cand("torch._int_mm (cuBLASLt, raw s32 out, no epilogue)", [(lambda A=A, W=W: torch._int_mm(A, W.t())) for A, W in sets]))
    if vo is not None:
        cands.append(Cand("vLLM cutlass_scaled_mm s8 (per-token x per-channel)",
                          [(lambda A=A, W=W: vo.cutlass_scaled_mm(A, W.t(), As, Bs, torch.bfloat16)) for A, W in sets]))
    report("int8", M, N, K, cands, 2.0 * M * N * K)


def run_fp8(M, N, K, vo):
    n = nsets(N * K + M * K + 2 * M * N)
    sets = [(torch.randn(M, K, device=dev).to(torch.float8_e4m3fn), torch.randn(N, K, device=dev).to(torch.float8_e4m3fn))
            for _ in range(n)]
    As = torch.rand(M, 1, device=dev)
    Bs = torch.rand(N, 1, device=dev) * 1e-2
    BsT = Bs.t().contiguous()
    cands = [Cand("ours", [(lambda A=A, W=W: b2.fp8_scaled_mm(A, W.t(), As, Bs, torch.bfloat16, None)) for A, W in sets])]
    cands[0].name = f"ours fp8_scaled_mm [{b2.capi.last_gemm_config()}]"
    if "torch" in LIBS:
        cands.append(Cand("torch._scaled_mm row-wise (cuBLASLt/CUTLASS)",
                          [(lambda A=A, W=W: torch._scaled_mm(A, W.t(), scale_a=As, scale_b=BsT, out_dtype=torch.bfloat16))
                           for A, W in sets]))
    if vo is not None:
        cands.append(Cand("vLLM cutlass_scaled_mm e4m3 (per-token x per-channel)",
                          [(lambda A=A, W=W: vo.cutlass_scaled_mm(A, W.t(), As, Bs, torch.bfloat16)) for A, W in sets]))
    report("fp8", M, N, K, cands, 2.0 * M * N * K)


def run_block(M, N, K, vo):
    n = nsets(N * K + M * K + 2 * M * N)
    sets = [(torch.randn(M, K, device=dev).to(torch.float8_e4m3fn), torch.randn(N, K, device=dev).to(torch.float8_e4m3fn))
            for _ in range(n)]
    sfa = torch.rand(K // 128, M, device=dev).t()           # [M, K/128] M-major
    sfb = torch.rand(N // 128, K // 128, device=dev).t()    # [K/128, N/128] K-major
    cands = [Cand("ours", [(lambda A=A, W=W: b2.fp8_blockwise_scaled_mm(A, W.t(), sfa, sfb, torch.bfloat16)) for A, W in sets])]
    cands[0].name = f"ours fp8_blockwise_scaled_mm [{b2.capi.last_gemm_config()}]"
    if vo is not None:
        # vLLM's group-broadcast rule: scale_a [M, K/128], scale_b [K/128, N/128]
        cands.append(Cand("vLLM cutlass_scaled_mm block 1x128 / 128x128",
                          [(lambda A=A, W=W: vo.cutlass_scaled_mm(A, W.t(), sfa, sfb, torch.bfloat16)) for A, W in sets]))
    if "flashinfer" in LIBS:
        try:
            from flashinfer.gemm import gemm_fp8_nt_groupwise

            sfa_mn = sfa.t().contiguous()                    # [K/128, M]
            sfb_mn = sfb.contiguous()                        # [K/128, N/128]
            for sm in (1, 2):
                cands.append(Cand(f"flashinfer gemm_fp8_nt_groupwise cutlass MN mma_sm={sm}",
                                  [(lambda A=A, W=W, sm=sm: gemm_fp8_nt_groupwise(A, W, sfa_mn, sfb_mn, scale_major_mode="MN",
                                                                                  mma_sm=sm, out_dtype=torch.bfloat16,
                                                                                  backend="cutlass")) for A, W in sets]))
        except Exception as ex:  # noqa: BLE001
            print(f"# flashinfer unavailable: {type(ex).__name__}: {ex}")
    report("block", M, N, K, cands, 2.0 * M * N * K)


def main():
    vo = vllm_ops()
    print(f"# {torch.cuda.get_device_name(0)}; rounds={ROUNDS} interleaved, graph-replayed; libs={sorted(LIBS)}", flush=True)
    if "int8" in want:
        shapes = [(8192, 6144, 4096), (8192, 28672, 4096), (8192, 4096, 14336), (4096, 5120, 5120), (4096, 13824, 5120),
                  (4096, 5120, 13824), (128, 4096, 4096), (16, 28672, 4096)]
        if QUICK:
            shapes = shapes[1:2] + shapes[5:6]
        for M, N, K in shapes:
            run_int8(M, N, K, vo)
    if "fp8" in want:
        shapes = [(8192, 1280, 8192), (8192, 7168, 8192), (8192, 8192, 1024), (8192, 8192, 3584), (64, 7168, 8192), (64, 8192, 3584)]
        if QUICK:
            shapes = shapes[1:2]
        for M, N, K in shapes:
            run_fp8(M, N, K, vo)
    if "block" in want:
        shapes = [(4096, 24576, 1536), (4096, 32768, 512), (4096, 7168, 16384), (4096, 36864, 7168), (4096, 7168, 18432),
                  (4096, 4096, 7168), (4096, 7168, 2048), (16, 4096, 7168), (128, 7168, 2048)]
        if "--decode" in sys.argv:  # decode-shape block-wise only (the narrow-tile schedule)
            shapes = [(1, 4096, 7168), (16, 4096, 7168), (16, 7168, 2048), (16, 24576, 1536), (64, 7168, 16384),
                      (128, 7168, 2048), (128, 4096, 7168)]
        if QUICK:
            shapes = [shapes[3], shapes[0]]
        for M, N, K in shapes:
            run_block(M, N, K, vo)


if __name__ == "__main__":
    main()
