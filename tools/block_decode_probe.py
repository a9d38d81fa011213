"""Block-wise fp8 GEMM at decode M (informational): python tools/block_decode_probe.py"""
import os, sys
sys.argv = [sys.argv[0], "none", "--graph"]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import perf_sweep as ps  # noqa: E402

for N, K in ((24576, 1536), (7168, 16384), (4096, 7168), (7168, 2048)):
    for M in (16, 128, 1024):
        ps.gemm_block(M, N, K)
