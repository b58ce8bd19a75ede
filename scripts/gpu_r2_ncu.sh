#!/bin/bash
# ncu --set full on the top kernels of the fp32/TF32 resnet18 step (one GPU, eager step so kernel names map 1:1)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-like-for-like --no-secondary"
$CMD > gpurun_out/r2n_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"conv_tc_kernel|sgd_kernel|bn_bwd_reduce_kernel" -s 330 -c 30 -f -o gpurun_out/r2n_prof_top $CMD > gpurun_out/r2n_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/r2n_ncu_full.log; ls -la gpurun_out/r2n_prof_top.ncu-rep
