#!/bin/bash
# ncu --set full of ONE K1S bulk launch with the final code (LQ=2005: multi-stripe, nested loop), after the same command ran plain
LQ=${1:-2005}
mkdir -p gpurun_out
timeout 300 python scripts/prof_one.py --lq $LQ --reps 2 > gpurun_out/r2z_plain_$LQ.log 2>&1 || { echo "plain failed"; tail -3 gpurun_out/r2z_plain_$LQ.log; exit 1; }
tail -1 gpurun_out/r2z_plain_$LQ.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 1 -c 1 -f -o gpurun_out/r2z_k1s_lq$LQ python scripts/prof_one.py --lq $LQ --reps 2 > gpurun_out/r2z_ncu_k1s_$LQ.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r2z_k1s_lq$LQ.ncu-rep
