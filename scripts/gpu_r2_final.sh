#!/bin/bash
# round 2, last call (1 GPU): what the driver runs at round end -- full GPU test tier, smoke(), the default bench line -- on the final tree
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2z_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2z_rc.txt
grep -E "passed|failed" gpurun_out/r2z_tests.log | tail -n 2; grep -E "^FAILED|^ERROR" gpurun_out/r2z_tests.log | head -n 20; grep -E "^E  " gpurun_out/r2z_tests.log | head -n 20 | cut -c1-400
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r2z_rc.txt; tail -n 3 gpurun_out/r2z_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/r2z_bench_default.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/r2z_rc.txt
grep '^{' gpurun_out/r2z_bench_default.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; l=c.get('like_for_like') or {}
print('value %.0f e2e %.0f ms %.4f dtype %s launches %s clocks %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['dtype'], d.get('gpu_launches'), d.get('clocks')))
print('ratios', l.get('ratio_product_over_torch')); print('secondary', c.get('secondary'))" 2>&1 | cut -c1-900
timeout 200 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary --global-batch 32 > gpurun_out/r2z_bench_b32.log 2>&1; echo "bench b32 rc=$?" | tee -a gpurun_out/r2z_rc.txt
grep '^{' gpurun_out/r2z_bench_b32.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('batch 32 (what each rank of an 8-GPU job runs): %.4f ms  %.0f img/s' % (d['ms_per_step'], d['value']))" 2>&1 | cut -c1-200
