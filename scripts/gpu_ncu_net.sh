#!/bin/bash
# launch list (per-kernel durations) of the workshop Net step, eager so that ncu sees every kernel
mkdir -p gpurun_out
CMD="python bench.py --model custom --steps 4 --warmup 3 --no-graph"
$CMD > gpurun_out/plain_net.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 60 --csv --log-file gpurun_out/launches_net.csv $CMD > gpurun_out/ncu_net.log 2>&1
echo "rc=$?"; tail -n 2 gpurun_out/ncu_net.log | cut -c1-300; wc -l gpurun_out/launches_net.csv
