#!/bin/bash
# round 2, A/B y: K1S slot pointer in registers (flush without constant-bank loads) and IDP selectors in registers (multi-stripe)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_plans.py -m gpu -x -q > gpurun_out/r2y_pytest_plans.txt 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r2y_pytest_plans.txt
timeout 1800 python scripts/ab_lib.py --libs seq-align-gpu_b200/libswgpu_vold.so,seq-align-gpu_b200/libswgpu_vslot.so,seq-align-gpu_b200/libswgpu.so --rounds 1 > gpurun_out/r2y_ab.log 2>&1; echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2y_ab.log | cut -c1-220
