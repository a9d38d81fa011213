"""CI-sized pass over every kernel family of the library, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck  python tools/sanitize_cases.py
    compute-sanitizer --tool racecheck python tools/sanitize_cases.py
    compute-sanitizer --tool synccheck python tools/sanitize_cases.py

Each case is also checked against the oracle so that a sanitizer-clean run is a run that computed the right thing.
(tools/sanitize.sh drives the three tools and writes profiles/<round>_sanitizer.log.)
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import sglang_diffusion_b200 as b2  # noqa: E402
from sglang_diffusion_b200 import capi  # noqa: E402
from conftest import mk_fp8_block_weight, mk_fp8_weight, mk_int8_problem  # noqa: E402
from oracle import restate  # noqa: E402  (checker)

DEV = torch.device("cuda:0")
only = set(sys.argv[1:])


def case(name):
    def deco(fn):
        if only and name not in only:
            return fn
        fn()
        torch.cuda.synchronize()
        print(f"[sanitize] {name}: ok", flush=True)
        return fn
    return deco


@case("quant_rows")
def _():
    for M, K in ((3, 544), (17, 4096), (2, 16384)):
        x = (torch.randn(M, K) / 10).bfloat16()
        q, s = b2.per_token_quant_int8(x.to(DEV))
        qr, sr = restate.per_token_quant_int8_cpu(x)
        assert torch.equal((q.cpu().to(torch.int16) + 128).to(torch.uint8), qr) and torch.equal(s.cpu().view(-1), sr)
        q8, s8 = b2.sglang_per_token_quant_fp8(x.to(DEV))
        q8r, s8r = restate.per_token_quant_fp8(x)
        assert torch.equal(q8.cpu().view(torch.uint8), q8r)


@case("quant_group128")
def _():
    for M, K in ((5, 512), (33, 1536)):
        x = (torch.randn(M, K) / 10).bfloat16()
        for cm in (True, False):
            q, s = b2.per_token_group_quant_fp8(x.to(DEV), 128, column_major_scales=cm)
            qr, sr = restate.per_token_group_quant_fp8(x, 128, 1e-10, cm)
            assert torch.equal(q.cpu().view(torch.uint8), qr) and torch.equal(s.cpu(), sr)


@case("gemm_persistent")
def _():
    x, W, Bs, b = mk_int8_problem(300, 520, 544, seed=3)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float())
    for cfg in (1, 2, 3, 4, 5, 6):
        capi.set_gemm_config(cfg)
        try:
            out = b2.int8_scaled_mm_with_quant(x.to(DEV), W.to(DEV), Bs.to(DEV), b.to(DEV), torch.bfloat16, False)
        finally:
            capi.set_gemm_config(0)
        assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16)), cfg


@case("gemm_splitk")
def _():
    x, W, Bs, b = mk_int8_problem(16, 1000, 4096, seed=4)
    ref = restate.int8_scaled_mm_with_quant(x, W, Bs, b.float())
    with capi.static_weights():
        for _ in range(2):
            out = b2.int8_scaled_mm_with_quant(x.to(DEV), W.to(DEV), Bs.to(DEV), b.to(DEV), torch.bfloat16, False)
    assert capi.last_gemm_config().startswith("splitk_"), capi.last_gemm_config()
    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))


@case("gemm_fp8")
def _():
    x = (torch.randn(130, 1024) / 10).bfloat16()
    Wq, ws = mk_fp8_weight(520, 1024)
    out = b2.apply_fp8_linear(x.to(DEV), Wq.to(DEV).t(), ws.to(DEV))
    q, s = restate.per_token_quant_fp8(x)
    ref = restate.fp8_scaled_mm(q, Wq.view(torch.uint8), s, ws, None)
    assert ((out.float().cpu() - ref).abs().max() / ref.abs().max()).item() < 1e-2


@case("gemm_block128")
def _():
    for M, N, K in ((5, 256, 640), (300, 512, 1024), (600, 384, 2176)):
        x = (torch.randn(M, K) / 10).bfloat16()
        Wq, sfb = mk_fp8_block_weight(N, K)
        out = b2.cutlass_w8a8_block_fp8_linear_with_fallback(x.to(DEV), Wq.to(DEV), [128, 128], sfb.to(DEV))
        q, sfa = restate.per_token_group_quant_fp8(x, 128, 1e-10, True)
        ref = restate.fp8_blockwise_scaled_mm(q, Wq.view(torch.uint8), sfa, sfb)
        assert ((out.float().cpu() - ref).abs().max() / ref.abs().max()).item() < 1e-2, (M, N, K)


print("[sanitize] all cases done", flush=True)
