#!/bin/bash
# round 2, call B (1 GPU): new single-GPU tests (per-block chain, convergence, smddp contract, graph-vs-eager), reference_on_b200 arm
mkdir -p gpurun_out
python baseline/install_reference.py > gpurun_out/r2b_install.log 2>&1
timeout 1200 python -m pytest tests/test_gpu_chain.py tests/test_gpu_convergence.py tests/test_gpu_smddp.py tests/test_gpu_models.py tests/test_zz_no_alloc.py -q -s -p no:cacheprovider > gpurun_out/r2b_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2b_rc.txt
timeout 600 python bench.py --impl reference_on_b200 --steps 20 --warmup 3 > gpurun_out/r2b_ref_on_b200.log 2>&1; echo "ref_on_b200 rc=$?" | tee -a gpurun_out/r2b_rc.txt
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2b_ref.log 2>&1; echo "ref rc=$?" | tee -a gpurun_out/r2b_rc.txt
grep -E "passed|failed" gpurun_out/r2b_tests.log | tail -n 3
grep -E "^FAILED|^ERROR|^E  " gpurun_out/r2b_tests.log | head -n 40
grep -E "^fp32 \{|^bf16 \{|windows|logits rel" gpurun_out/r2b_tests.log | cut -c1-1800
tail -n 1 gpurun_out/r2b_ref_on_b200.log | cut -c1-600
tail -n 1 gpurun_out/r2b_ref.log | cut -c1-400
