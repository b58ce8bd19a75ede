#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "batchnorm" -p no:cacheprovider > gpurun_out/t_bn.log 2>&1; echo "bn rc=$?" | tee gpurun_out/rc6.txt
timeout 900 python -m pytest tests/test_gpu_models.py -m gpu -q -p no:cacheprovider > gpurun_out/t_models.log 2>&1; echo "models rc=$?" | tee -a gpurun_out/rc6.txt
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_graph.log 2>&1; echo "graph1 rc=$?" | tee -a gpurun_out/rc6.txt
tail -n 3 gpurun_out/t_bn.log; grep -v "^$" gpurun_out/t_models.log | tail -n 25; tail -n 1 gpurun_out/bench_graph.log | cut -c1-330
