#!/bin/bash
# GPU call: parity suite, microbenchmark, launch list of a short bench, full ncu capture of the top kernels
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 ./bench_micro/ubench_int 4 > gpurun_out/ubench_int.jsonl 2> gpurun_out/ubench_int.err
for lq in 144 375 1000 5478; do timeout 300 python scripts/prof_one.py --lq $lq --reps 3 >> gpurun_out/prof_one.log 2>&1; done
timeout 300 python scripts/prof_one.py --lq 375 --reps 2 --topk 100 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_lq375.csv \
    python scripts/prof_one.py --lq 375 --reps 2 --topk 100 > gpurun_out/ncu_launches.log 2>&1
timeout 300 python scripts/prof_one.py --lq 375 --reps 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k1_inter -s 1 -c 1 -f -o gpurun_out/k1_lq375 \
    python scripts/prof_one.py --lq 375 --reps 2 > gpurun_out/ncu_k1.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log; cat gpurun_out/prof_one.log
