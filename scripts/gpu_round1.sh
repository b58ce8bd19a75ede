#!/bin/bash
# first GPU contact: memory-bound kernels, then tensor-core kernels, each in its own process
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "not conv and not linear" -p no:cacheprovider > gpurun_out/t_mem.log 2>&1; echo "mem rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "linear" -p no:cacheprovider > gpurun_out/t_linear.log 2>&1; echo "linear rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "conv" -p no:cacheprovider > gpurun_out/t_conv.log 2>&1; echo "conv rc=$?" | tee -a gpurun_out/rc.txt
tail -5 gpurun_out/t_mem.log gpurun_out/t_linear.log gpurun_out/t_conv.log
