#!/bin/bash
# round 2, batch 1: register-bank microbenchmark + A/B of K1S build variants (built here by scripts/build_variant.sh)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2_box.txt 2>&1
timeout 120 bench_micro/ubench_bank > gpurun_out/r2_ubench_bank.jsonl 2>&1; echo "ubench_bank rc=$?"
cat gpurun_out/r2_ubench_bank.jsonl
L=seq-align-gpu_b200
timeout 1500 python scripts/ab_lib.py --libs $L/libswgpu.so,$L/libswgpu_u2.so,$L/libswgpu_imm.so,$L/libswgpu_immu2.so,$L/libswgpu_o1.so --rounds ${ROUNDS:-2} > gpurun_out/r2_ab1.log 2>&1
echo "ab rc=$?"; cat gpurun_out/r2_ab1.log
