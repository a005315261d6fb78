#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/r2i_k3x_trace.txt
SWGPU_LIB=seq-align-gpu_b200/libswgpu_vtrace.so SWGPU_K3X_TRACE=gpurun_out/r2i_k3x_trace.txt timeout 300 python scripts/k3x_steps.py --lqs 4096,35213 --reps 1 > gpurun_out/r2i_steps.jsonl 2> gpurun_out/r2i_steps.err; echo "rc=$?"
grep ms_rescore gpurun_out/r2i_steps.jsonl; wc -l gpurun_out/r2i_k3x_trace.txt
