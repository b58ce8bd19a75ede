#!/bin/bash
# round 2 (1 GPU): 16-byte split-K join through distributed shared memory, 64-wide tiles for small grids, 8-way clusters for a
# handful of tiles: kernel tests in the default and in the forced configuration, phase stamps, step time at batch 256 and 32
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_fp32.py tests/test_gpu_kernels.py -q -p no:cacheprovider > gpurun_out/r2q_tests_default.log 2>&1; echo "tests default rc=$?" | tee gpurun_out/r2q_rc.txt
B200DDP_NARROW_TILES=100000 B200DDP_CLUSTER8_TILES=8 timeout 300 python -m pytest tests/test_gpu_fp32.py tests/test_gpu_kernels.py -q -p no:cacheprovider > gpurun_out/r2q_tests_forced.log 2>&1; echo "tests forced rc=$?" | tee -a gpurun_out/r2q_rc.txt
for f in default forced; do grep -E "passed|failed" gpurun_out/r2q_tests_$f.log | tail -n 1; grep -E "^FAILED|^E  " gpurun_out/r2q_tests_$f.log | head -n 12 | cut -c1-300; done
python scripts/conv_phase_profile.py fp32 32 > gpurun_out/r2q_conv_phases_b32.txt 2>&1; grep -E "fwd \||dgrad" gpurun_out/r2q_conv_phases_b32.txt | cut -c1-200
for cfg in "0 0" "18 0" "40 0" "40 8"; do
  set -- $cfg
  for gb in 256 32; do
    B200DDP_NARROW_TILES=$1 B200DDP_CLUSTER8_TILES=$2 timeout 200 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary --global-batch $gb > gpurun_out/r2q_bench_n$1_c$2_b$gb.json 2> gpurun_out/r2q_bench_n$1_c$2_b$gb.err
    grep '^{' gpurun_out/r2q_bench_n$1_c$2_b$gb.json | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('narrow $1 cluster8 $2 batch $gb: %.4f ms  %.0f img/s' % (d['ms_per_step'], d['value']))" 2>&1 | cut -c1-200
  done
done
