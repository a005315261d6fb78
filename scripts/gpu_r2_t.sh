#!/bin/bash
# round 2, batch t: the new row-bound test, then the ncu launch list of a C2 bench step (after the plain run of the same command)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_topk.py -m gpu -x -q > gpurun_out/r2t_pytest_topk.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2t_pytest_topk.txt
bash scripts/gpu_r2_launches.sh c2
