"""Decode kernel vs persistent kernel at M = 32..128 on wide layers (dispatch-rule experiment)."""
import os, sys
sys.argv = [sys.argv[0], "none", "--graph", "--pdl"]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import perf_sweep as ps  # noqa: E402  (runs nothing for want = ["none"])

for N, K in ((6144, 4096), (28672, 4096), (7168, 8192), (8192, 3584), (8192, 1024)):
    for M in (48, 64, 128):
        print(f"# M={M} N={N} K={K}")
        ps.gemm_int8(M, N, K)
        for bn in (64, 128):
            for S in (1, 2, 4):
                for cap in (4, 8):
                    ps.gemm_int8(M, N, K, cfg=0x80000 | (bn // 16 - 1) | ((S - 1) << 4) | (cap << 8))
