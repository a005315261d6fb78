#!/bin/bash
# refresh after the K3X rework: plain bench, launch list of the same command, full capture of k3x on C4
mkdir -p gpurun_out
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_bench.log 2>&1
timeout 300 python scripts/prof_c4.py --reps 2 > gpurun_out/plain_c4.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k3x$" -s 1 -c 1 -f -o gpurun_out/k3x_c4 \
    python scripts/prof_c4.py --reps 2 > gpurun_out/ncu_k3x_c4.log 2>&1
cut -c1-300 gpurun_out/bench.json; tail -n 2 gpurun_out/bench.err; tail -n 1 gpurun_out/plain_c4.log | cut -c1-200
