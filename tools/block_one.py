"""ncu target: one block-wise fp8 GEMM of the DeepSeek-V3 C3 list (M=4096, N=7168, K=16384)."""
import os, sys
sys.argv = [sys.argv[0], "none"]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import perf_sweep as ps  # noqa: E402
ps.gemm_block(4096, 7168, 16384)
