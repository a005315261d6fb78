#!/bin/bash
# Round-1 evidence run (one GPU): plain bench, ncu launch list of the same bench command (cold-cache, serialised:
# compare SHARES, not absolute times), and one `--set full` capture of the dominant kernel in its two modes.
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_bench.log 2>&1
for lq in 1000 2005; do
timeout 300 python scripts/prof_one.py --lq $lq --reps 2 > gpurun_out/plain_$lq.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 1 -c 1 -f -o gpurun_out/k1s_lq$lq \
    python scripts/prof_one.py --lq $lq --reps 2 > gpurun_out/ncu_k1s_$lq.log 2>&1
done
timeout 300 python scripts/prof_c4.py --reps 2 > gpurun_out/plain_c4.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^k3x$" -s 1 -c 1 -f -o gpurun_out/k3x_c4 \
    python scripts/prof_c4.py --reps 2 > gpurun_out/ncu_k3x_c4.log 2>&1
cat gpurun_out/bench.json | cut -c1-400; tail -n 3 gpurun_out/bench.err; ls -la gpurun_out | tail -n 12
