#!/bin/bash
# round 2: ncu --set full of the K1S bulk launch in its two modes + every admissible plan per query length
set -x
mkdir -p gpurun_out
for lq in 1000 2005; do
timeout 300 python scripts/prof_one.py --lq $lq --reps 2 > gpurun_out/plain_$lq.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 1 -c 1 -f -o gpurun_out/r2_k1s_lq$lq \
    python scripts/prof_one.py --lq $lq --reps 2 > gpurun_out/ncu_k1s_$lq.log 2>&1
done
timeout 900 python scripts/tune_plan.py --lqs 144,189,222,375,464,567,657,729,850,1000,1500,2005,2504,3005,3564,4061,4548,4743,5147,5478 > gpurun_out/r2_tune_plan.log 2>&1
tail -n 2 gpurun_out/plain_1000.log gpurun_out/plain_2005.log; cut -c1-200 gpurun_out/r2_tune_plan.log
