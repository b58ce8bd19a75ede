#!/bin/bash
# round 2, call E (1 GPU): full GPU test tier, smoke, headline bench (fp32 + bf16) with like-for-like, per-kernel launch lists (ncu)
mkdir -p gpurun_out
python baseline/install_reference.py > gpurun_out/r2e_install.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2e_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2e_rc.txt
grep -E "passed|failed" gpurun_out/r2e_tests.log | tail -n 2; grep -E "^FAILED|^ERROR" gpurun_out/r2e_tests.log | head -n 20; grep -E "^E  " gpurun_out/r2e_tests.log | head -n 10 | cut -c1-400
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2e_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r2e_rc.txt; grep smoke gpurun_out/r2e_smoke.log
timeout 600 python bench.py --steps 100 --warmup 5 > gpurun_out/r2e_bench_fp32.log 2>&1; echo "bench fp32 rc=$?" | tee -a gpurun_out/r2e_rc.txt
timeout 600 python bench.py --steps 100 --warmup 5 --dtype bf16 --no-secondary > gpurun_out/r2e_bench_bf16.log 2>&1; echo "bench bf16 rc=$?" | tee -a gpurun_out/r2e_rc.txt
for f in fp32 bf16; do grep '^{' gpurun_out/r2e_bench_$f.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; l=c.get('like_for_like')
print('$f', 'value %.0f e2e %.0f ms %.3f launches %d phases %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], c['launches_per_step'], c.get('phase_ms_eager')))
print('   like_for_like', {k:(round(v.get('images_per_sec',0)), round(v.get('e2e_images_per_sec',0))) if isinstance(v,dict) and 'images_per_sec' in v else v for k,v in (l or {}).items()})
" 2>&1 | cut -c1-900; done
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-like-for-like --no-secondary"
$CMD > gpurun_out/r2e_plain256.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 500 -c 1200 --csv --log-file gpurun_out/r2e_launches_fp32_b256.csv $CMD > gpurun_out/r2e_ncu256.log 2>&1
echo "ncu b256 rc=$?" | tee -a gpurun_out/r2e_rc.txt
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-like-for-like --no-secondary --global-batch 32"
$CMD > gpurun_out/r2e_plain32.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 500 -c 1200 --csv --log-file gpurun_out/r2e_launches_fp32_b32.csv $CMD > gpurun_out/r2e_ncu32.log 2>&1
echo "ncu b32 rc=$?" | tee -a gpurun_out/r2e_rc.txt
python scripts/agg_launches.py gpurun_out/r2e_launches_fp32_b256.csv 24 2>&1 | head -n 30
python scripts/agg_launches.py gpurun_out/r2e_launches_fp32_b32.csv 24 2>&1 | head -n 30
