#!/bin/bash
# usage: gpu_ncu_full.sh <kernel-regex> <skip> <count> <outname>
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $3 -f -o gpurun_out/$4 $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/ncu_full.log
