#!/bin/bash
# ncu --set full on the workshop Net megakernel (one GPU, eager step), then the launch list of a resnet18 step
mkdir -p gpurun_out
CMD="python bench.py --model custom --steps 2 --warmup 3 --no-graph"
$CMD > gpurun_out/plain_net.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"net_fwd_bwd_kernel|net_fc_wgrad_kernel" -s 6 -c 2 -f -o gpurun_out/prof_net $CMD > gpurun_out/ncu_net_full.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/ncu_net_full.log | cut -c1-200
python scripts/net_breakdown.py 256 > gpurun_out/net_breakdown.json 2> gpurun_out/net_breakdown.err; tail -n 24 gpurun_out/net_breakdown.json
python scripts/step_breakdown.py 256 > gpurun_out/step_breakdown.json 2> gpurun_out/step_breakdown.err; tail -n 40 gpurun_out/step_breakdown.json
