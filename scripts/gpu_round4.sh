#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "conv or linear" -p no:cacheprovider > gpurun_out/t_conv.log 2>&1; echo "conv rc=$?" | tee gpurun_out/rc4.txt
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_graph.log 2>&1; echo "graph1 rc=$?" | tee -a gpurun_out/rc4.txt
CMD="python bench.py --steps 2 --warmup 3 --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 320 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?" | tee -a gpurun_out/rc4.txt
tail -n 4 gpurun_out/t_conv.log; tail -n 1 gpurun_out/bench_graph.log | cut -c1-400
