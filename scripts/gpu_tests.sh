#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/t_zz_alone.log 2>&1; echo "zz alone rc=$?" | tee gpurun_out/rc_tests.txt
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/t_all.log 2>&1; echo "all rc=$?" | tee -a gpurun_out/rc_tests.txt
tail -n 3 gpurun_out/t_zz_alone.log; grep -v "^$" gpurun_out/t_all.log | tail -n 15
