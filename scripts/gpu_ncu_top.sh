#!/bin/bash
# ncu --set full on the top kernels of the resnet18 train step (one GPU, eager step so kernel names map 1:1)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"conv_tc_kernel|bn_finalize_kernel|bn_apply_kernel|bn_bwd_cluster_kernel|bn_bwd_reduce_kernel|sgd_kernel" -s 300 -c 36 -f -o gpurun_out/prof_top $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/ncu_full.log
