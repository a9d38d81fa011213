"""pytest configuration: `gpu` marker, repo root on sys.path, shared helpers.

`-m "not gpu"`: oracle vs golden vectors / the compiled reference, host logic, C-ABI symbol
export -- runs on a CPU-only box.  `-m gpu`: the parity tests proper (CUDA path vs oracle),
all through the C ABI or the torch ops registered over it.
"""
from __future__ import annotations

import glob
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Binaries are built in-tree by __graft_entry__.build(); make sure they exist (no JIT)."""
    import __graft_entry__ as ge

    ge.build(oracle_ref=False)


def golden_names(prefix: str):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, prefix + "*.npz")))


def load_golden(name: str) -> dict:
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    out = {}
    for k in z.files:
        a = z[k]
        t = torch.from_numpy(a.copy())
        if k in ("A", "out") and a.dtype == np.uint16:
            t = t.view(torch.int16).view(torch.bfloat16)
        out[k] = t
    return out


def bf16_bits(t: torch.Tensor) -> torch.Tensor:
    return t.contiguous().view(torch.int16)


def mk_int8_problem(M, N, K, seed=1234, bias=True):
    """Input distributions of the reference's own test (test/srt/cpu/test_gemm.py:131-144)."""
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, K, generator=g) / 10).bfloat16()
    W = ((torch.rand(N, K, generator=g) - 0.5) * 2 * 127).clamp(-128, 127).to(torch.int8)
    Bs = torch.rand(N, generator=g) * 1e-2
    b = torch.randn(N, generator=g).bfloat16() if bias else None
    return x, W, Bs, b


def mk_fp8_weight(N, K, seed=7):
    """randn*0.05 quantized per output channel, s = amax/448."""
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(N, K, generator=g) * 0.05
    ws = (w.abs().amax(dim=1, keepdim=True) / 448.0).float()
    Wq = (w / ws).clamp(-448, 448).to(torch.float8_e4m3fn)
    return Wq, ws


def mk_fp8_block_weight(N, K, seed=11):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(N, K, generator=g) * 0.05
    wb = w.view(N // 128, 128, K // 128, 128)
    sfb = (wb.abs().amax(dim=(1, 3)) / 448.0).float()
    Wq = (wb / sfb[:, None, :, None]).clamp(-448, 448).reshape(N, K).to(torch.float8_e4m3fn)
    return Wq, sfb
