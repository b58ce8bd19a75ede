#!/bin/bash
# round 2, call H (1 GPU): graphed forward/backward behind the caller-owned loop (tests + the reference script on this framework,
# graphed vs launch-by-launch, global batch 256 and 32), and the rows-per-CTA knob of the BN reductions
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_graphed.py tests/test_gpu_smddp.py -q -p no:cacheprovider > gpurun_out/r2h_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2h_rc.txt
grep -E "passed|failed" gpurun_out/r2h_tests.log | tail -n 2; grep -E "^FAILED|^ERROR" gpurun_out/r2h_tests.log | head -n 20; grep -E "^E  " gpurun_out/r2h_tests.log | head -n 30 | cut -c1-600
for gb in 256 32; do
  timeout 300 python bench.py --impl reference --steps 20 --warmup 3 --global-batch $gb > gpurun_out/r2h_ref_b$gb.json 2> gpurun_out/r2h_ref_b$gb.err; echo "ref b$gb rc=$?" | tee -a gpurun_out/r2h_rc.txt
  timeout 300 python bench.py --impl reference_on_b200 --steps 20 --warmup 3 --global-batch $gb > gpurun_out/r2h_onb200_graph_b$gb.json 2> gpurun_out/r2h_onb200_graph_b$gb.err; echo "on_b200 graph b$gb rc=$?" | tee -a gpurun_out/r2h_rc.txt
  B200DDP_SHIM_GRAPH=0 timeout 300 python bench.py --impl reference_on_b200 --steps 20 --warmup 3 --global-batch $gb > gpurun_out/r2h_onb200_eager_b$gb.json 2> gpurun_out/r2h_onb200_eager_b$gb.err; echo "on_b200 eager b$gb rc=$?" | tee -a gpurun_out/r2h_rc.txt
done
for f in gpurun_out/r2h_ref_b*.json gpurun_out/r2h_onb200_*.json; do echo $f; grep '^{' $f | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  value %.0f ms %.3f launches %s %s' % (d.get('value',0), d.get('ms_per_step',0), d.get('gpu_launches'), d.get('config',{}).get('graphed_forward_backward')))" 2>&1 | cut -c1-400; done
for r in 256 128; do
  B200DDP_BN_REDUCE_ROWS=$r timeout 300 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary > gpurun_out/r2h_bench_bnrows$r.json 2> gpurun_out/r2h_bench_bnrows$r.err; echo "bench bnrows $r rc=$?" | tee -a gpurun_out/r2h_rc.txt
  grep '^{' gpurun_out/r2h_bench_bnrows$r.json | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bnrows $r value %.0f ms %.4f e2e %.0f' % (d['value'], d['ms_per_step'], d['e2e']['value']))" 2>&1 | cut -c1-300
done
tail -n 5 gpurun_out/r2h_onb200_graph_b256.err | cut -c1-400
