#!/bin/bash
# round 2, probe E: K3X column-step time / stripe hand-over cost, integer-pipe mixes at 4 and 8 warps per scheduler
mkdir -p gpurun_out
timeout 300 python scripts/k3x_steps.py > gpurun_out/r2e_k3x_steps.jsonl 2> gpurun_out/r2e_k3x_steps.err; echo "k3x_steps rc=$?"
timeout 300 python scripts/k3x_steps.py --targets 64 --lqs 128,1024,4096,35213 --reps 2 > gpurun_out/r2e_k3x_steps_64.jsonl 2>> gpurun_out/r2e_k3x_steps.err; echo "k3x_steps64 rc=$?"
cat gpurun_out/r2e_k3x_steps.jsonl gpurun_out/r2e_k3x_steps_64.jsonl; tail -3 gpurun_out/r2e_k3x_steps.err
for c in 2 3 4; do timeout 200 bench_micro/ubench_int $c > gpurun_out/r2e_ubench_int_ctas$c.jsonl 2>&1; echo "ubench $c rc=$?"; done
grep -h '"op2":"IDP.4A"\|"op2":"none"' gpurun_out/r2e_ubench_int_ctas2.jsonl | cut -c1-60,230-400 | head -20
