#!/bin/bash
# end-of-round verification on one GPU: full GPU test tier, smoke, both bench arms, both models
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider 2>&1 | tail -n 3
timeout 300 python __graft_entry__.py smoke 2>&1 | grep smoke
timeout 600 python bench.py --steps 100 --warmup 5 > gpurun_out/final_bench_resnet18.log 2>&1; tail -n 1 gpurun_out/final_bench_resnet18.log | cut -c1-520
timeout 600 python bench.py --model custom --steps 200 --warmup 5 > gpurun_out/final_bench_custom.log 2>&1; tail -n 1 gpurun_out/final_bench_custom.log | cut -c1-520
timeout 600 python bench.py --model resnet50 --global-batch 128 --steps 10 --warmup 3 > gpurun_out/final_bench_resnet50.log 2>&1; tail -n 1 gpurun_out/final_bench_resnet50.log | cut -c1-330
timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/final_ref_resnet18.log 2>&1; tail -n 1 gpurun_out/final_ref_resnet18.log | cut -c1-330
