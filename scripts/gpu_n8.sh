#!/bin/bash
# eight GPUs of one box: the C2 bench line (weak scaling) and C5 at the NAMED size (10 G residues dealt over 8 ranks),
# NCCL top-k merge in both
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512"
timeout 400 $T bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "rc=$?" >> gpurun_out/bench_n8.err
timeout 600 $T bench.py --gpus 8 --workload c5 --scale 1.0 --steps 1 --warmup 3 > gpurun_out/bench_c5_n8.json 2> gpurun_out/bench_c5_n8.err; echo "rc=$?" >> gpurun_out/bench_c5_n8.err
cut -c1-300 gpurun_out/bench_n8.json; tail -n 2 gpurun_out/bench_n8.err; cut -c1-300 gpurun_out/bench_c5_n8.json; tail -n 2 gpurun_out/bench_c5_n8.err
