#!/bin/bash
# usage: gpu_multi_caps.sh N — DDP check, then bench.py at N GPUs with the default bucket caps and with smaller ones
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29542 scripts/ddp_check.py resnet18 > gpurun_out/ddp_check_$N.log 2>&1; echo "ddp_check rc=$?"
grep -E "DDP_CHECK|FAIL|algorithms" gpurun_out/ddp_check_$N.log
for cap in 25 4 8; do
  B200DDP_BUCKET_CAP_MB=$cap timeout 300 $TR --master-port 29551 bench.py --gpus $N --steps 60 --warmup 5 > gpurun_out/bench_${N}_cap$cap.log 2>&1; echo "bench cap=$cap rc=$?"
  tail -n 1 gpurun_out/bench_${N}_cap$cap.log | cut -c1-330
done
timeout 300 $TR --master-port 29544 bench.py --gpus $N --steps 60 --warmup 5 --model custom > gpurun_out/bench_custom_$N.log 2>&1; echo "bench custom rc=$?"
tail -n 1 gpurun_out/bench_custom_$N.log | cut -c1-330
