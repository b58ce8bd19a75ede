#!/bin/bash
# both arms of bench.py at N GPUs (reference first, as the driver does)
mkdir -p gpurun_out
N=${1:-4}
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --impl reference --gpus 1 --steps 30 --warmup 5 2>/dev/null | grep '"impl"' > gpurun_out/scale_ref_$N.json
  timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 2>/dev/null | grep '"impl"' > gpurun_out/scale_own_$N.json
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
  timeout 600 $TR --master-port 29551 bench.py --impl reference --gpus $N --steps 30 --warmup 5 2>/dev/null | grep '"impl"' > gpurun_out/scale_ref_$N.json
  timeout 300 $TR --master-port 29552 bench.py --gpus $N --steps 50 --warmup 5 2>/dev/null | grep '"impl"' > gpurun_out/scale_own_$N.json
  timeout 300 $TR --master-port 29553 scripts/comm_check.py 2>&1 | grep -E "COMM_CHECK|FAIL" | head -3
fi
cut -c1-260 gpurun_out/scale_ref_$N.json; cut -c1-330 gpurun_out/scale_own_$N.json
