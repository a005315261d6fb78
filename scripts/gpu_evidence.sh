#!/bin/bash
# Evidence run (one GPU): GPU parity suite, plain bench, ncu launch list of the same bench command (cold-cache,
# serialised: compare SHARES, not absolute times), one `--set full` capture of the dominant kernel in its two modes
# and of the 32-bit wavefront kernel, and the per-length serial sweep.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
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
timeout 600 python scripts/tune_plan.py --lqs 144,189,222,375,464,567,657,729,850,1000,1500,2005,2504,3005,3564,4061,4548,4743,5147,5478 > gpurun_out/tune_plan3.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json | cut -c1-400; tail -n 3 gpurun_out/bench.err; tail -n 2 gpurun_out/plain_c4.log; ls -la gpurun_out | tail -n 14
