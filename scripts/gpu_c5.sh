#!/bin/bash
# C5 evidence run (one GPU): the GPU parity suite, the C5 bench line at 1/10 size, then the named size (10 G residues
# generated on the device, 64 queries) through tests/test_gpu_c5_device.py (sampled oracle parity + planted best hits).
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/box.txt; nproc >> gpurun_out/box.txt; nvidia-smi --query-gpu=name,memory.total --format=csv >> gpurun_out/box.txt
timeout 1200 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 16 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload c5 --scale ${C5_BENCH_SCALE:-0.1} --steps 1 --warmup 3 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?" >> gpurun_out/bench_c5.err
cut -c1-600 gpurun_out/bench_c5.json; tail -n 3 gpurun_out/bench_c5.err
if [ "${C5_FULL:-1}" = "1" ]; then
  SWGPU_C5_SCALE=1.0 timeout 1500 python -m pytest tests/test_gpu_c5_device.py -x -q -s -k device_generated > gpurun_out/c5_full.log 2>&1; echo "c5 full rc=$?" >> gpurun_out/c5_full.log
  tail -n 8 gpurun_out/c5_full.log
fi
