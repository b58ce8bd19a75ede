#!/bin/bash
# round 2, call A: fp32/TF32 kernel tests, regression of the bf16 tier, first like-for-like bench numbers (1 GPU)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fp32.py -q -p no:cacheprovider > gpurun_out/r2a_fp32.log 2>&1; echo "fp32 rc=$?" | tee gpurun_out/r2a_rc.txt
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider --deselect tests/test_gpu_fp32.py > gpurun_out/r2a_all.log 2>&1; echo "all rc=$?" | tee -a gpurun_out/r2a_rc.txt
timeout 600 python bench.py --dtype bf16 --steps 50 --warmup 5 > gpurun_out/r2a_bench_bf16.log 2>&1; echo "bench bf16 rc=$?" | tee -a gpurun_out/r2a_rc.txt
timeout 600 python bench.py --dtype fp32 --steps 50 --warmup 5 --no-like-for-like > gpurun_out/r2a_bench_fp32.log 2>&1; echo "bench fp32 rc=$?" | tee -a gpurun_out/r2a_rc.txt
grep -E "passed|failed|error" gpurun_out/r2a_fp32.log | tail -n 3
grep -E "^FAILED|^ERROR" gpurun_out/r2a_fp32.log | head -n 30
grep -E "passed|failed|error" gpurun_out/r2a_all.log | tail -n 3
grep -E "^FAILED|^ERROR" gpurun_out/r2a_all.log | head -n 10
tail -n 1 gpurun_out/r2a_bench_bf16.log | cut -c1-1500
tail -n 1 gpurun_out/r2a_bench_fp32.log | cut -c1-700
