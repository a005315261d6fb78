#!/bin/bash
# round 2, A/B: next-step ring prefetch in the multi-stripe kernels too (vpf2) against the in-tree build (R = 24, 25 multi-stripe now at ptxas -O1)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_plans.py -m gpu -x -q > gpurun_out/r2y_pytest_plans.txt 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r2y_pytest_plans.txt
timeout 1800 python scripts/ab_lib.py --libs seq-align-gpu_b200/libswgpu.so,seq-align-gpu_b200/libswgpu_vpf2.so --rounds 1 > gpurun_out/r2y_ab5.log 2>&1; echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2y_ab5.log | cut -c1-220 | tail -12
