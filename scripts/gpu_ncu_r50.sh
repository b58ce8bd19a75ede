#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --model resnet50 --global-batch 128 --steps 1 --warmup 3 --no-graph"
$CMD > gpurun_out/plain50.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 1900 -c 480 --csv --log-file gpurun_out/launches50.csv $CMD > gpurun_out/ncu50.log 2>&1
echo "rc=$?"; tail -n 1 gpurun_out/plain50.log | cut -c1-300
