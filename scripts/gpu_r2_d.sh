#!/bin/bash
# round 2, call D (1 GPU): chained test re-check, ResNet-50 product vs stock torch (bf16 and fp32), per-GPU-batch-32 step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_chain.py -q -s -p no:cacheprovider > gpurun_out/r2d_chain.log 2>&1; echo "chain rc=$?" | tee gpurun_out/r2d_rc.txt
grep -E "passed|failed" gpurun_out/r2d_chain.log | tail -n 1; grep -E "^E  " gpurun_out/r2d_chain.log | head -n 4 | cut -c1-500
timeout 900 python bench.py --model resnet50 --dtype bf16 --global-batch 128 --steps 10 --warmup 3 > gpurun_out/r2d_r50_bf16.log 2>&1; echo "r50 bf16 rc=$?" | tee -a gpurun_out/r2d_rc.txt
timeout 900 python bench.py --model resnet50 --dtype fp32 --global-batch 128 --steps 10 --warmup 3 --no-like-for-like > gpurun_out/r2d_r50_fp32.log 2>&1; echo "r50 fp32 rc=$?" | tee -a gpurun_out/r2d_rc.txt
timeout 600 python bench.py --global-batch 32 --steps 50 --warmup 5 --no-secondary > gpurun_out/r2d_r18_b32.log 2>&1; echo "r18 b32 rc=$?" | tee -a gpurun_out/r2d_rc.txt
for f in r50_bf16 r50_fp32 r18_b32; do grep '^{' gpurun_out/r2d_$f.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; l=c.get('like_for_like')
print('$f', 'value %.0f e2e %.0f ms %.3f launches %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], c['launches_per_step']))
print('   like_for_like', {k:(round(v.get('images_per_sec',0)), round(v.get('e2e_images_per_sec',0)), v.get('note','')) if isinstance(v,dict) and 'images_per_sec' in v else v for k,v in (l or {}).items()})
" 2>&1 | cut -c1-900; done
