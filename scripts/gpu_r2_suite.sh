#!/bin/bash
# round 2: the no-allocation entry point alone in a fresh process, then the whole GPU suite (it runs last there), then smoke + a short bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > gpurun_out/r2_box.txt 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_noalloc_alone.log
tail -n 4 gpurun_out/r2_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2_pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_gpu.log
tail -n 16 gpurun_out/r2_pytest_gpu.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 6 gpurun_out/r2_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/r2_bench_n1.json; tail -n 3 gpurun_out/r2_bench_n1.err
