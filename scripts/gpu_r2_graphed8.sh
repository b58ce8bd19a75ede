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
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 tests/mp_gpu_worker.py graphed > gpurun_out/r2m_graphed_world8.log 2>&1; echo "graphed worker world 8 rc=$?" | tee -a gpurun_out/r2m_rc.txt
grep -E "^CHECK|^MP_GPU_RESULT" gpurun_out/r2m_graphed_world8.log | cut -c1-220; grep -E "Error|error" gpurun_out/r2m_graphed_world8.log | head -n 10 | cut -c1-300
