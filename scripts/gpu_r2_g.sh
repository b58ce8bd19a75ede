#!/bin/bash
# round 2, call G (1 GPU): pooled-BN test + model tests, bench fp32 with the pooled BN statistics, MNTD meta-training epoch time
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_models.py tests/test_gpu_chain.py tests/test_gpu_fp32.py -q -p no:cacheprovider > gpurun_out/r2g_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2g_rc.txt
grep -E "passed|failed" gpurun_out/r2g_tests.log | tail -n 1; grep -E "^FAILED|^E  " gpurun_out/r2g_tests.log | head -n 8 | cut -c1-400
timeout 600 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary > gpurun_out/r2g_bench_fp32.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/r2g_rc.txt
B200DDP_POOL_BN_SUMS=0 timeout 600 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary > gpurun_out/r2g_bench_fp32_nopool.log 2>&1
timeout 600 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary --global-batch 32 > gpurun_out/r2g_bench_fp32_b32.log 2>&1
B200DDP_POOL_BN_SUMS=0 timeout 600 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary --global-batch 32 > gpurun_out/r2g_bench_fp32_b32_nopool.log 2>&1
timeout 600 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary --dtype bf16 > gpurun_out/r2g_bench_bf16.log 2>&1
for f in bench_fp32 bench_fp32_nopool bench_fp32_b32 bench_fp32_b32_nopool bench_bf16; do grep '^{' gpurun_out/r2g_$f.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f: %.0f img/s  %.3f ms/step  e2e %.0f launches %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['launches_per_step']))" 2>&1 | tail -n 1; done
timeout 600 python scripts/meta_bench.py > gpurun_out/r2g_meta_bench.log 2>&1; echo "meta rc=$?" | tee -a gpurun_out/r2g_rc.txt; grep META_BENCH gpurun_out/r2g_meta_bench.log || tail -n 5 gpurun_out/r2g_meta_bench.log
