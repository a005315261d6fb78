#!/bin/bash
# first GPU call of a session: smoke, parity suite, microbenchmark, a short bench, launch list
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 120 ./bench_micro/ubench_int 4 > gpurun_out/ubench_int.jsonl 2> gpurun_out/ubench_int.err
timeout 600 python bench.py --steps 1 --warmup 1 > gpurun_out/bench_first.json 2> gpurun_out/bench_first.err; echo "bench rc=$?" >> gpurun_out/bench_first.err
tail -3 gpurun_out/smoke.log gpurun_out/pytest_gpu.log gpurun_out/bench_first.err; cat gpurun_out/bench_first.json
