#!/bin/bash
# Wait-cycle sweep of the per-token GEMM: tile configs x shapes x pipeline depth x debug ablations.
mkdir -p gpurun_out
{
for shape in "8192 28672 4096" "8192 6144 4096" "8192 4096 14336"; do
  python tools/gemm_prof.py s8 $shape 4 5
done
python tools/gemm_prof.py f8 8192 28672 4096 4 5
echo "== stages sweep (cfg 5)"
for s in 3 4 5; do echo "stages=$s"; B2_GEMM_STAGES=$s python tools/gemm_prof.py s8 8192 28672 4096 5; done
echo "== skip epilogue body"; B2_GEMM_DEBUG=2 python tools/gemm_prof.py s8 8192 28672 4096 4 5
echo "== skip MMAs";          B2_GEMM_DEBUG=1 python tools/gemm_prof.py s8 8192 28672 4096 4 5
echo "== skip both";          B2_GEMM_DEBUG=3 python tools/gemm_prof.py s8 8192 28672 4096 4 5
} 2>&1 | grep -v Warning | tee gpurun_out/prof_sweep.log
