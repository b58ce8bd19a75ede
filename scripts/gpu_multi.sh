#!/bin/bash
# usage: gpu_multi.sh N  — correctness, DDP semantics, all-reduce sweep vs NCCL, bench (own + reference) on N GPUs
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 scripts/comm_check.py --out gpurun_out/comm_check_$N.json > gpurun_out/comm_check_$N.log 2>&1; echo "comm_check rc=$?" | tee gpurun_out/rc_multi_$N.txt
timeout 300 $TR --master-port 29542 scripts/ddp_check.py resnet18 > gpurun_out/ddp_check_$N.log 2>&1; echo "ddp_check rc=$?" | tee -a gpurun_out/rc_multi_$N.txt
timeout 300 $TR --master-port 29543 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_$N.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/rc_multi_$N.txt
timeout 500 $TR --master-port 29544 scripts/comm_check.py --sweep --nccl --max-mb 256 --out gpurun_out/sweep_$N.json > gpurun_out/sweep_$N.log 2>&1; echo "sweep rc=$?" | tee -a gpurun_out/rc_multi_$N.txt
NCCL_DEBUG=INFO timeout 400 $TR --master-port 29545 bench.py --impl reference --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_ref_$N.log 2>&1; echo "ref rc=$?" | tee -a gpurun_out/rc_multi_$N.txt
grep -E "COMM_CHECK|FAIL" gpurun_out/comm_check_$N.log | head; grep -E "DDP_CHECK|FAIL|algorithms" gpurun_out/ddp_check_$N.log
tail -n 1 gpurun_out/bench_$N.log | cut -c1-400; grep '"impl"' gpurun_out/bench_ref_$N.log | cut -c1-300; grep -c NVLS gpurun_out/bench_ref_$N.log
