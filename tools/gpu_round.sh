#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, ncu launch list + one full capture of the GEMM kernel.
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
SHORT="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-decode"
echo "== ncu launch list"
$SHORT > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_list.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_list.log
echo "== ncu full"
$SHORT > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemm_w8a8 -s 9 -c 3 -o gpurun_out/prof_gemm -f $SHORT > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_full.log
