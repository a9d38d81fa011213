"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/sgl_cpu_oracle.so, the
reference's own CPU code compiled in place from /root/reference by oracle/build_ref.py).

Run in the build container (it needs /root/reference and an AVX512-BF16 + AMX host):
    python tests/golden/make_golden.py
The reference ships no golden vectors for this path; these files pin the
oracle restatement (oracle/restate.c) and the CUDA kernels to the reference's own outputs
on machines where the reference cannot run (e.g. the GPU box has no /root/reference).

Cases follow the reference's own test grid and input distributions
(test/srt/cpu/test_gemm.py:17,131-174: seed 1234, A = randn/10 bf16, W = ((rand-0.5)*2*127) -> int8,
Bs = rand*1e-2; M in {2,128} (VNNI and AMX paths), N=384, K=544; test_gemm.py:176-233 for fp8)
plus edge rows: an all-zero row (eps path), a row with a +-max outlier, codes at +-127, weights at -128.
Files hold inputs AND outputs (bf16 stored as uint16 bit patterns) so nothing depends on RNG parity.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import build_ref  # noqa: E402


def bits(t: torch.Tensor) -> np.ndarray:
    if t.dtype == torch.bfloat16:
        return t.contiguous().view(torch.int16).numpy().view(np.uint16)
    if t.dtype == torch.float8_e4m3fn:
        return t.contiguous().view(torch.uint8).numpy()
    return t.contiguous().numpy()


def int8_case(ref, M, N, K, has_bias, seed, edge=False):
    torch.manual_seed(seed)
    A = torch.randn((M, K), dtype=torch.bfloat16) / 10
    if edge:
        A[0].zero_()                      # all-zero row: amax -> eps 1e-7
        if M > 1:
            A[1, 7] = 30.0                # outlier: every other code in the row is tiny
            A[1, 8] = -30.0
        if M > 2:
            A[2] = 1e-8                   # below eps; bf16 keeps it non-zero
        if M > 3:
            A[3, ::2] = 0.5               # exact ties: 0.5/1.0*127 = 63.5 -> RNE 64
            A[3, 1::2] = -1.0
    B = (torch.rand((N, K), dtype=torch.float32) - 0.5) * 2
    Bq = (B * 127).clamp(min=-128, max=127).to(torch.int8)
    if edge:
        Bq[0].fill_(-128)
        Bq[1].fill_(127)
    Bs = torch.rand(N) * 1e-2
    bias = torch.randn(N).bfloat16().float() if has_bias else None  # bf16-representable
    Aq, As = ref.per_token_quant_int8_cpu(A)
    out = ref.int8_scaled_mm_cpu(Aq, Bq, As, Bs, bias, torch.bfloat16, False)
    Bp = ref.convert_weight_packed(Bq)
    fused = ref.int8_scaled_mm_with_quant(A, Bp, Bs, bias, torch.bfloat16, True)
    assert torch.equal(out.view(torch.int16), fused.view(torch.int16)), "reference fused != unfused"
    d = dict(A=bits(A), W=bits(Bq), Bs=bits(Bs), Aq_u8=bits(Aq), As=bits(As), out=bits(out))
    if has_bias:
        d["bias"] = bits(bias)
    return d


def fp8_case(ref, M, N, K, has_bias, seed):
    """fp8_scaled_mm_cpu is W8A16: a TOLERANCE oracle for the W8A8 GPU ops.
    128x128 blocks; weight codes are kept normal and non-zero (the reference mis-decodes +-0 and subnormals)."""
    torch.manual_seed(seed)
    A = torch.randn(M, K, dtype=torch.bfloat16)
    w = torch.randn(N, K) * 0.05
    wb = w.view(N // 128, 128, K // 128, 128)
    sfb = (wb.abs().amax(dim=(1, 3)) / 448.0).float()
    wq = (wb / sfb[:, None, :, None]).reshape(N, K)
    wq = torch.where(wq.abs() < 2.0 ** -6, torch.full_like(wq, 2.0 ** -6) * torch.sign(wq + 1e-30), wq)
    Wq = wq.to(torch.float8_e4m3fn)
    bias = torch.randn(N).bfloat16().float() if has_bias else None
    Wp = ref.convert_weight_packed(Wq)
    out = ref.fp8_scaled_mm_cpu(A, Wp, sfb, [128, 128], bias, torch.bfloat16, True)
    d = dict(A=bits(A), W=bits(Wq), sfb=bits(sfb), out=bits(out))
    if has_bias:
        d["bias"] = bits(bias)
    return d


def main():
    so = build_ref.build()
    assert so and build_ref.load(), "the reference CPU oracle could not be built/loaded on this host"
    ref = torch.ops.sgl_kernel_ref
    cases = {}
    seed = 1234
    for M in (2, 128):
        for hb in (False, True):
            cases[f"int8_M{M}_N384_K544_bias{int(hb)}"] = int8_case(ref, M, 384, 544, hb, seed)
            seed += 1
    cases["int8_edge_M8_N64_K512_bias1"] = int8_case(ref, 8, 64, 512, True, 99, edge=True)
    cases["int8_M1_N128_K4096_bias0"] = int8_case(ref, 1, 128, 4096, False, 100)
    cases["int8_M33_N160_K1024_bias1"] = int8_case(ref, 33, 160, 1024, True, 101)
    for (M, N, K, hb) in ((1, 128, 512, False), (11, 256, 1024, True)):
        cases[f"fp8w8a16_M{M}_N{N}_K{K}_bias{int(hb)}"] = fp8_case(ref, M, N, K, hb, seed)
        seed += 1
    total = 0
    for name, d in cases.items():
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **d)
        total += os.path.getsize(path)
        print(f"{name}: {os.path.getsize(path) / 1024:.0f} KiB")
    print(f"total {total / 1024:.0f} KiB in {len(cases)} files")


if __name__ == "__main__":
    main()
