#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_graph.log 2>&1; echo "graph1 rc=$?" | tee gpurun_out/rc3.txt
timeout 300 $TR --master-port 29541 scripts/comm_check.py --out gpurun_out/comm_check_$N.json > gpurun_out/comm_check_$N.log 2>&1; echo "comm_check rc=$?" | tee -a gpurun_out/rc3.txt
timeout 300 $TR --master-port 29542 scripts/ddp_check.py resnet18 > gpurun_out/ddp_check_$N.log 2>&1; echo "ddp_check rc=$?" | tee -a gpurun_out/rc3.txt
timeout 300 $TR --master-port 29543 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_$N.log 2>&1; echo "bench$N rc=$?" | tee -a gpurun_out/rc3.txt
timeout 400 $TR --master-port 29544 scripts/comm_check.py --sweep --nccl --max-mb 256 --out gpurun_out/sweep_$N.json > gpurun_out/sweep_$N.log 2>&1; echo "sweep rc=$?" | tee -a gpurun_out/rc3.txt
for f in bench_graph comm_check_$N ddp_check_$N bench_$N sweep_$N; do echo "== $f"; grep -v "^$" gpurun_out/$f.log | tail -n 12; done
