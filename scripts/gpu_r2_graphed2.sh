#!/bin/bash
# round 2, call I (2 GPUs): graphed forward/backward with the collectives inside the graphs (tests at world 2), the single-GPU
# graphed test, and the reference script on this framework at 2 ranks (graphed vs the reference itself)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_graphed.py "tests/test_gpu_multi.py::test_graphed_forward_backward_on_nvlink" -q -p no:cacheprovider -s > gpurun_out/r2i_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2i_rc.txt
grep -E "passed|failed" gpurun_out/r2i_tests.log | tail -n 2; grep -E "^FAILED|^ERROR|^CHECK|worst relative" gpurun_out/r2i_tests.log | head -n 40 | cut -c1-300; grep -E "^E  " gpurun_out/r2i_tests.log | head -n 30 | cut -c1-600
for impl in reference reference_on_b200; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --impl $impl --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2i_${impl}_2gpu.json 2> gpurun_out/r2i_${impl}_2gpu.err; echo "$impl 2gpu rc=$?" | tee -a gpurun_out/r2i_rc.txt
  grep '^{' gpurun_out/r2i_${impl}_2gpu.json | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  $impl value %.0f ms %.3f launches %s %s' % (d.get('value',0), d.get('ms_per_step',0), d.get('gpu_launches'), d.get('config',{}).get('graphed_forward_backward')))" 2>&1 | cut -c1-400
done
tail -n 4 gpurun_out/r2i_reference_on_b200_2gpu.err | cut -c1-300
