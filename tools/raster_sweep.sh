#!/bin/bash
# sustained TOPS of one shape under different raster overrides
mkdir -p gpurun_out
M=$1; N=$2; K=$3; shift 3
for r in "$@"; do
  echo "== $M x $N x $K raster=$r"
  if [ "$r" = auto ]; then unset B2_GEMM_RASTER; else export B2_GEMM_RASTER=$r; fi
  B2_PROBE_MODES=bench python tools/power_probe.py s8 $M $N $K 5 2>&1 | grep -E "^s8"
done
