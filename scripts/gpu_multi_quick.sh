#!/bin/bash
# usage: gpu_multi_quick.sh N — DDP semantics + bench (resnet18 and the workshop Net) on N GPUs, no sweeps
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 scripts/comm_check.py --out gpurun_out/comm_check_$N.json > gpurun_out/comm_check_$N.log 2>&1; echo "comm_check rc=$?"
timeout 300 $TR --master-port 29542 scripts/ddp_check.py resnet18 > gpurun_out/ddp_check_$N.log 2>&1; echo "ddp_check rc=$?"
timeout 300 $TR --master-port 29546 scripts/ddp_check.py custom > gpurun_out/ddp_check_custom_$N.log 2>&1; echo "ddp_check custom rc=$?"
timeout 300 $TR --master-port 29543 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_$N.log 2>&1; echo "bench rc=$?"
timeout 300 $TR --master-port 29544 bench.py --gpus $N --steps 50 --warmup 5 --model custom > gpurun_out/bench_custom_$N.log 2>&1; echo "bench custom rc=$?"
grep -E "COMM_CHECK|FAIL" gpurun_out/comm_check_$N.log | head -5; grep -E "DDP_CHECK|FAIL|algorithms" gpurun_out/ddp_check_$N.log gpurun_out/ddp_check_custom_$N.log
tail -n 1 gpurun_out/bench_$N.log | cut -c1-420; tail -n 1 gpurun_out/bench_custom_$N.log | cut -c1-420
