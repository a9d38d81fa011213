#!/bin/bash
# Under the power cap, does throughput depend on how many SMs the persistent GEMM occupies?
mkdir -p gpurun_out
for n in 148 132 120 104; do
  echo "== B2_GEMM_MAX_SMS=$n"
  B2_GEMM_MAX_SMS=$n B2_PROBE_MODES=bench python tools/power_probe.py s8 8192 28672 4096 5 2>&1 | grep -E "^s8"
done | tee gpurun_out/sm_sweep.log
