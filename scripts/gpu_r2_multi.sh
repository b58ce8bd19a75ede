#!/bin/bash
# round 2, multi-GPU call: N = number of GPUs of this lease (2, 4 or 8).  Multi-GPU pytest (collectives, DDP semantics, skipped
# collective), bench at N with ddp_check (fp32, fused bucket step on and off), all-reduce sweep 1 KiB..1 GiB vs NCCL.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_gpu_multi.py -q -p no:cacheprovider > gpurun_out/r2m_tests_$N.log 2>&1; echo "multi tests rc=$?" | tee gpurun_out/r2m_rc_$N.txt
grep -E "passed|failed|skipped" gpurun_out/r2m_tests_$N.log | tail -n 2; grep -E "CHECK.*FAIL|^FAILED|^E  " gpurun_out/r2m_tests_$N.log | head -n 20 | cut -c1-300
timeout 600 $TR --master-port 29701 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r2m_bench_fp32_$N.log 2>&1; echo "bench fused rc=$?" | tee -a gpurun_out/r2m_rc_$N.txt
B200DDP_FUSED_SGD=0 timeout 600 $TR --master-port 29702 bench.py --gpus $N --steps 50 --warmup 5 --no-like-for-like --no-secondary > gpurun_out/r2m_bench_fp32_unfused_$N.log 2>&1; echo "bench unfused rc=$?" | tee -a gpurun_out/r2m_rc_$N.txt
timeout 600 $TR --master-port 29703 bench.py --gpus $N --steps 50 --warmup 5 --dtype bf16 --no-like-for-like --no-secondary > gpurun_out/r2m_bench_bf16_$N.log 2>&1; echo "bench bf16 rc=$?" | tee -a gpurun_out/r2m_rc_$N.txt
for f in fp32 fp32_unfused bf16; do grep '^{' gpurun_out/r2m_bench_${f}_$N.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$f', 'value %.0f e2e %.0f ms %.3f launches %d algos %s ddp_check %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], c['launches_per_step'], c['allreduce_algos'], c['ddp_check']))
print('   secondary', c.get('secondary')); l=c.get('like_for_like')
print('   like_for_like', {k:(round(v.get('images_per_sec',0)), round(v.get('e2e_images_per_sec',0))) if isinstance(v,dict) and 'images_per_sec' in v else v for k,v in (l or {}).items()})
" 2>&1 | cut -c1-900; done
B200DDP_ARENA_MB=2560 timeout 900 $TR --master-port 29704 scripts/comm_check.py --sweep --nccl --graph --max-mb 1024 --iters 10 --ctas 32,96,128,148 --out gpurun_out/r2m_sweep_$N.json > gpurun_out/r2m_sweep_$N.log 2>&1; echo "sweep rc=$?" | tee -a gpurun_out/r2m_rc_$N.txt
grep -E "COMM_CHECK|FAIL" gpurun_out/r2m_sweep_$N.log | head; grep '"bytes": \(4194304\|16777216\|268435456\|1073741824\)' gpurun_out/r2m_sweep_$N.log | cut -c1-900
