#!/bin/bash
# Structural ablations of the per-token GEMM on ZERO operands (no power cap in the way):
# debug flags 1 = no MMAs, 2 = no epilogue body, 4 = no loads.
mkdir -p gpurun_out
{
for d in ${ABLATE_FLAGS:-0 2 4 6 1 3 5}; do
  echo "== zeros, B2_GEMM_DEBUG=$d"
  B2_PROF_DATA=zeros B2_GEMM_DEBUG=$d python tools/gemm_prof.py s8 8192 28672 4096 4 5
done
echo "== random, B2_GEMM_DEBUG=0"
python tools/gemm_prof.py s8 8192 28672 4096 4 5
} 2>&1 | grep -v Warning | tee gpurun_out/ablate.log
