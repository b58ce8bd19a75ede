#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29542 scripts/ddp_check.py resnet18 > gpurun_out/ddp_check_$N.log 2>&1; echo "ddp_check rc=$?"
grep -E "DDP_CHECK|FAIL|bucket|algorithms" gpurun_out/ddp_check_$N.log
timeout 300 $TR --master-port 29543 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_$N.log 2>&1; echo "bench rc=$?"
tail -n 1 gpurun_out/bench_$N.log | cut -c1-330
timeout 500 $TR --master-port 29544 scripts/comm_check.py --sweep --graph --max-mb 64 --out gpurun_out/sweep_graph_$N.json > gpurun_out/sweep_graph_$N.log 2>&1; echo "sweep rc=$?"
grep -E "COMM_CHECK|FAIL" gpurun_out/sweep_graph_$N.log | head -5
