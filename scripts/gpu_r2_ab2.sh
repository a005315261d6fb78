#!/bin/bash
# round 2, batch 2: forced-bank microbenchmark + A/B of the best-from-G variants
mkdir -p gpurun_out
timeout 120 bench_micro/ubench_bank > gpurun_out/r2_ubench_bank.jsonl 2>&1; echo "ubench_bank rc=$?"
grep forced gpurun_out/r2_ubench_bank.jsonl
L=seq-align-gpu_b200
timeout 1500 python scripts/ab_lib.py --libs $L/libswgpu.so,$L/libswgpu_imm.so,$L/libswgpu_bg.so,$L/libswgpu_bgimm.so,$L/libswgpu_immu2.so,$L/libswgpu_bgimmu2.so --rounds ${ROUNDS:-2} > gpurun_out/r2_ab2.log 2>&1
echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2_ab2.log
