#!/bin/bash
# ncu launch list (gpu__time_duration per launch, cold-cache and serialised) of one bench.py step: W=c2|c4 ...
W=${1:-c2}
mkdir -p gpurun_out
timeout 900 python bench.py --workload $W --steps 1 --warmup 3 --no-cpu > gpurun_out/r2q_bench_${W}_plain.json 2> gpurun_out/r2q_bench_${W}_plain.err || { echo "plain failed"; tail -3 gpurun_out/r2q_bench_${W}_plain.err; exit 1; }
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r2q_launches_${W}.csv python bench.py --workload $W --steps 1 --warmup 3 --no-cpu > gpurun_out/r2q_ncu_${W}.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/r2q_launches_${W}.csv
