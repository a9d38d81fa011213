#!/bin/bash
# One multi-GPU box visit (N = $1): M-sharded default bench, the Wan2.1 DiT workload, TP = N fp8 shards with NCCL.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
echo "== bench --gpus $N (llama3-8b int8, weak)"
timeout 300 $TR bench.py --gpus $N --steps 50 --warmup 5 --no-cpu-baseline 2> gpurun_out/scale_bench_$N.err | tee gpurun_out/scale_bench_$N.json
echo "== bench --gpus $N --workload wan2.1-14b-dit-int8 (M=32768 sharded)"
timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --workload wan2.1-14b-dit-int8 2> gpurun_out/scale_wan_$N.err | tee gpurun_out/scale_wan_$N.json
echo "== tp_bench TP=$N fp8"
timeout 300 $TR tools/tp_bench.py --check > gpurun_out/tp${N}_fp8_full.log 2>&1; grep '^{' gpurun_out/tp${N}_fp8_full.log | tee gpurun_out/tp${N}_fp8.log; grep -iE 'error|assert|Traceback' gpurun_out/tp${N}_fp8_full.log | head -5
