#!/bin/bash
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee gpurun_out/rc2.txt
timeout 300 python bench.py --steps 20 --warmup 3 --no-graph > gpurun_out/bench_eager.log 2>&1; echo "eager rc=$?" | tee -a gpurun_out/rc2.txt
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_graph.log 2>&1; echo "graph rc=$?" | tee -a gpurun_out/rc2.txt
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?" | tee -a gpurun_out/rc2.txt
for f in smoke bench_eager bench_graph bench_ref; do echo "== $f"; tail -n 6 gpurun_out/$f.log; done
