#!/bin/bash
# round 2, A/B v: multi-stripe K1S column loop as blocks of 32 steps without the per-step block test
mkdir -p gpurun_out
timeout 1500 python scripts/ab_lib.py --a seq-align-gpu_b200/libswgpu.so --b seq-align-gpu_b200/libswgpu_vnested.so --rounds 1 > gpurun_out/r2v_ab_nested.log 2>&1; echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2v_ab_nested.log | cut -c1-200
