mkdir -p gpurun_out
for cfg in "0 0" "18 0" "40 0" "150 0" "40 8"; do
  set -- $cfg
  for gb in 256 32; do
    B200DDP_NARROW_TILES=$1 B200DDP_CLUSTER8_TILES=$2 timeout 200 python bench.py --steps 100 --warmup 5 --no-like-for-like --no-secondary --global-batch $gb > gpurun_out/r2r_bench_n$1_c$2_b$gb.json 2> gpurun_out/r2r_bench_n$1_c$2_b$gb.err
    grep '^{' gpurun_out/r2r_bench_n$1_c$2_b$gb.json | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('narrow $1 cluster8 $2 batch $gb: %.4f ms  %.0f img/s' % (d['ms_per_step'], d['value']))" 2>&1 | cut -c1-200
  done
done
