#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29541 scripts/comm_check.py > gpurun_out/cc.log 2>&1; grep -E "COMM_CHECK|FAIL" gpurun_out/cc.log | head
B200DDP_AUX_STREAM=1 timeout 300 $TR --master-port 29542 scripts/ddp_check.py resnet18 2>&1 | grep -E "bucket|DDP_CHECK|FAIL|rank 0"
echo "--- aux off"
B200DDP_AUX_STREAM=0 timeout 300 $TR --master-port 29543 scripts/ddp_check.py resnet18 2>&1 | grep -E "bucket|DDP_CHECK|FAIL|rank 0"
