#!/bin/bash
# parity suite on the current build, then A/B against the previous build (seq-align-gpu_b200/libswgpu_a.so)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 5 gpurun_out/pytest_gpu.log
timeout 900 python scripts/ab_lib.py --a seq-align-gpu_b200/libswgpu_a.so --b seq-align-gpu_b200/libswgpu.so --rounds ${ROUNDS:-2} > gpurun_out/ab_ring.log 2>&1
cat gpurun_out/ab_ring.log
