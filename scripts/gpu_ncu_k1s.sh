#!/bin/bash
# full ncu capture of the K1S bulk launch: lq=1000 (g=32, single launch per search) and lq=144 (g=8, after the long-stream launch)
set -x
mkdir -p gpurun_out
timeout 300 python scripts/prof_one.py --lq 1000 --reps 2 > gpurun_out/plain_1000.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 1 -c 1 -f -o gpurun_out/k1s_lq1000 \
    python scripts/prof_one.py --lq 1000 --reps 2 > gpurun_out/ncu_k1s_1000.log 2>&1
timeout 300 python scripts/prof_one.py --lq 144 --reps 2 > gpurun_out/plain_144.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 3 -c 1 -f -o gpurun_out/k1s_lq144 \
    python scripts/prof_one.py --lq 144 --reps 2 > gpurun_out/ncu_k1s_144.log 2>&1
timeout 300 python scripts/prof_one.py --lq 375 --reps 2 --topk 100 > gpurun_out/plain_375.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_lq375.csv \
    python scripts/prof_one.py --lq 375 --reps 2 --topk 100 > gpurun_out/ncu_launches.log 2>&1
ls -la gpurun_out
