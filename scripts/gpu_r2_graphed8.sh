#!/bin/bash
# round 2, call M (8 GPUs): the reference script on this framework with the graphed forward/backward at 8 and 4 ranks, and the
# multi-GPU test of the graphed path + SGD.sync_state at world 8
mkdir -p gpurun_out
for n in 8 4; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2970$n bench.py --impl reference_on_b200 --gpus $n --steps 20 --warmup 3 > gpurun_out/r2m_ref_on_b200_graph_${n}gpu.json 2> gpurun_out/r2m_ref_on_b200_graph_${n}gpu.err; echo "reference_on_b200 ${n}gpu rc=$?" | tee -a gpurun_out/r2m_rc.txt
  grep '^{' gpurun_out/r2m_ref_on_b200_graph_${n}gpu.json | tail -n 1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  N=$n value %.0f ms %.3f launches %s %s' % (d.get('value',0), d.get('ms_per_step',0), d.get('gpu_launches'), d.get('config',{}).get('graphed_forward_backward')))" 2>&1 | cut -c1-400
done
timeout 300 python -m pytest "tests/test_gpu_multi.py::test_graphed_forward_backward_on_nvlink[8]" -q -p no:cacheprovider -s > gpurun_out/r2m_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r2m_rc.txt
grep -E "passed|failed" gpurun_out/r2m_tests.log | tail -n 2; grep -E "^CHECK" gpurun_out/r2m_tests.log | cut -c1-200; grep -E "^E  " gpurun_out/r2m_tests.log | head -n 20 | cut -c1-400
