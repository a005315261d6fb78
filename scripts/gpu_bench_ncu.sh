#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
for lq in 1000 144; do
timeout 300 python scripts/prof_one.py --lq $lq --reps 2 > gpurun_out/plain_$lq.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 2 -c 1 -f -o gpurun_out/k1s_lq$lq \
    python scripts/prof_one.py --lq $lq --reps 2 > gpurun_out/ncu_k1s_$lq.log 2>&1
done
cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
