#!/bin/bash
# round 2, evidence run at N GPUs (N = 2, 4 or 8): multi-GPU pytest, headline bench (fp32 fused / unfused, bf16) with ddp_check,
# like-for-like and the workshop Net, ResNet-50 (weak scaling, 128 images per GPU) against stock PyTorch, the unmodified
# reference script on NCCL and on this framework, all-reduce sweep 1 KiB..1 GiB against NCCL.  Everything lands in gpurun_out/r2f_*_N.*
N=${1:-8}
mkdir -p gpurun_out
python baseline/install_reference.py > gpurun_out/r2f_install_$N.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
RC=gpurun_out/r2f_rc_$N.txt; : > $RC
if [ "$2" != "notests" ]; then
timeout 900 python -m pytest tests/test_gpu_multi.py -q -p no:cacheprovider > gpurun_out/r2f_tests_$N.log 2>&1; echo "multi tests rc=$?" | tee -a $RC
grep -E "passed|failed|skipped" gpurun_out/r2f_tests_$N.log | tail -n 2; grep -E "CHECK.*FAIL|^FAILED|^E  " gpurun_out/r2f_tests_$N.log | head -n 20 | cut -c1-300
fi
timeout 600 $TR --master-port 29701 bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/r2f_bench_fp32_$N.log 2>&1; echo "bench fp32 rc=$?" | tee -a $RC
B200DDP_FUSED_SGD=0 timeout 600 $TR --master-port 29702 bench.py --gpus $N --steps 100 --warmup 5 --no-like-for-like --no-secondary > gpurun_out/r2f_bench_fp32_unfused_$N.log 2>&1; echo "bench fp32 unfused rc=$?" | tee -a $RC
timeout 600 $TR --master-port 29703 bench.py --gpus $N --steps 100 --warmup 5 --dtype bf16 --no-like-for-like --no-secondary > gpurun_out/r2f_bench_bf16_$N.log 2>&1; echo "bench bf16 rc=$?" | tee -a $RC
timeout 900 $TR --master-port 29704 bench.py --gpus $N --model resnet50 --dtype bf16 --global-batch $((128*N)) --steps 10 --warmup 3 --lfl-modes fp32,bf16,bf16_graph > gpurun_out/r2f_r50_bf16_$N.log 2>&1; echo "r50 rc=$?" | tee -a $RC
timeout 600 $TR --master-port 29705 bench.py --impl reference --gpus $N --steps 20 --warmup 3 > gpurun_out/r2f_ref_$N.log 2>&1; echo "reference rc=$?" | tee -a $RC
timeout 600 $TR --master-port 29706 bench.py --impl reference_on_b200 --gpus $N --steps 20 --warmup 3 > gpurun_out/r2f_ref_on_b200_$N.log 2>&1; echo "reference_on_b200 rc=$?" | tee -a $RC
for f in bench_fp32 bench_fp32_unfused bench_bf16 r50_bf16 ref ref_on_b200; do grep '^{' gpurun_out/r2f_${f}_$N.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d.get('config',{}); l=c.get('like_for_like')
if 'value' not in d: print('$f', d); sys.exit()
print('$f N=$N', 'value %.0f e2e %.0f ms %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step']), 'launches', c.get('launches_per_step'), 'ddp_check', c.get('ddp_check'), 'phases', c.get('phase_ms_eager'))
if c.get('secondary'): print('   secondary', c.get('secondary'))
if l: print('   like_for_like', {k:(round(v.get('images_per_sec',0)), round(v.get('e2e_images_per_sec',0))) if isinstance(v,dict) and 'images_per_sec' in v else v for k,v in l.items()})
" 2>&1 | cut -c1-1100; done
B200DDP_ARENA_MB=2560 timeout 900 $TR --master-port 29707 scripts/comm_check.py --sweep --nccl --graph --max-mb 1024 --iters 10 --ctas 32 --out gpurun_out/r2f_sweep_$N.json > gpurun_out/r2f_sweep_$N.log 2>&1; echo "sweep rc=$?" | tee -a $RC
grep -E "COMM_CHECK|FAIL" gpurun_out/r2f_sweep_$N.log | head -n 5
