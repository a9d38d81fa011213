"""Multi-GPU parity of the peer-memory all-reduce (b2_allreduce_symm) and of the TP linear layers that use it.
Needs >= 2 GPUs on the box: skipped on the single-GPU boxes the default `-m gpu` run uses; `tools/scale_round.sh`
and `gpurun --gpus 2 -- python -m pytest tests/test_multi_gpu.py -m gpu` run it."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _torchrun(n, script, *args, port=29541):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, script), *args]
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_symmetric_all_reduce_matches_ordered_fp32_sum():
    n = min(torch.cuda.device_count(), 8)
    r = _torchrun(n, "tools/symm_allreduce_check.py")
    assert r.returncode == 0 and "SYMM_ALLREDUCE_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_gemm_epilogue_fused_all_reduce_equals_one_shot_bitwise():
    n = min(torch.cuda.device_count(), 8)
    r = _torchrun(n, "tools/symm_push_check.py", port=29543)
    assert r.returncode == 0 and ("SYMM_PUSH_CHECK PASS" in r.stdout or "SYMM_PUSH_CHECK SKIP" in r.stdout), \
        r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_gemm_scatter_epilogue_reduce_scatter_equals_one_shot_rows_bitwise():
    n = min(torch.cuda.device_count(), 8)
    r = _torchrun(n, "tools/symm_scatter_check.py", port=29544)
    assert r.returncode == 0 and "SYMM_SCATTER_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("flags", [("--check",), ("--check", "--symm"), ("--check", "--symm", "--int8"),
                                   ("--check", "--symm", "--prefill-symm"), ("--check", "--symm", "--prefill-symm", "--int8")])
def test_tp_linear_layers_against_fp64_shard_products(flags):
    n = min(torch.cuda.device_count(), 8)
    r = _torchrun(n, "tools/tp_bench.py", *flags, port=29542)
    assert r.returncode == 0 and r.stdout.count('"layer"') == 8, r.stdout[-3000:] + r.stderr[-3000:]
