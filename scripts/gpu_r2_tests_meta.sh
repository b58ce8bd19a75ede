#!/bin/bash
# round 2, call H (1 GPU): full GPU test tier once more, MNTD meta-training epoch time with the graphed step
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2h_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2h_rc.txt
grep -E "passed|failed" gpurun_out/r2h_tests.log | tail -n 1; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r2h_tests.log | head -n 12 | cut -c1-400
timeout 600 python scripts/meta_bench.py > gpurun_out/r2h_meta_bench.log 2>&1; echo "meta rc=$?" | tee -a gpurun_out/r2h_rc.txt; grep META_BENCH gpurun_out/r2h_meta_bench.log || tail -n 8 gpurun_out/r2h_meta_bench.log
