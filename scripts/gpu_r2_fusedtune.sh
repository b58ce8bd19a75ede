#!/bin/bash
# fused bucket step: CTA budget experiment at N GPUs (fp32, 30 steps each)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { # label, env...
  label=$1; shift
  env "$@" timeout 300 $TR --master-port 29711 bench.py --gpus $N --steps 30 --warmup 5 --no-like-for-like --no-secondary > gpurun_out/r2t_${label}_$N.log 2>&1
  grep '^{' gpurun_out/r2t_${label}_$N.log | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$label N=$N: %.0f img/s  %.3f ms/step  e2e %.0f' % (d['value'], d['ms_per_step'], d['e2e']['value']))" 2>&1 | tail -n 1
}
run unfused B200DDP_FUSED_SGD=0
run unfused_tail32 B200DDP_FUSED_SGD=0 B200DDP_TAIL_CTAS=32
run unfused_tail64 B200DDP_FUSED_SGD=0 B200DDP_TAIL_CTAS=64
run fused_24_96 B200DDP_FUSED_CTAS=24 B200DDP_FUSED_TAIL_CTAS=96
run fused_16_64 B200DDP_FUSED_CTAS=16 B200DDP_FUSED_TAIL_CTAS=64
run fused_32_128 B200DDP_FUSED_CTAS=32 B200DDP_FUSED_TAIL_CTAS=128
run fused_48_148 B200DDP_FUSED_CTAS=48 B200DDP_FUSED_TAIL_CTAS=148
run fused_8_96 B200DDP_FUSED_CTAS=8 B200DDP_FUSED_TAIL_CTAS=96
