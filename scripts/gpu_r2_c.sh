#!/bin/bash
# round 2, call C (1 GPU): re-run of the tests fixed after call B + reference_on_b200
mkdir -p gpurun_out
python baseline/install_reference.py > gpurun_out/r2c_install.log 2>&1
timeout 1200 python -m pytest tests/test_gpu_chain.py tests/test_gpu_convergence.py tests/test_gpu_smddp.py -q -s -p no:cacheprovider > gpurun_out/r2c_tests.log 2>&1; echo "tests rc=$?" | tee gpurun_out/r2c_rc.txt
timeout 600 python bench.py --impl reference_on_b200 --steps 20 --warmup 3 > gpurun_out/r2c_ref_on_b200.log 2>&1; echo "ref_on_b200 rc=$?" | tee -a gpurun_out/r2c_rc.txt
grep -E "passed|failed" gpurun_out/r2c_tests.log | tail -n 3
grep -E "^FAILED|^ERROR" gpurun_out/r2c_tests.log | head -n 20
grep -E "^E  " gpurun_out/r2c_tests.log | head -n 12 | cut -c1-600
grep -E "^fp32 \{|^bf16 \{|windows" gpurun_out/r2c_tests.log | cut -c1-2500
tail -n 1 gpurun_out/r2c_ref_on_b200.log | cut -c1-700
