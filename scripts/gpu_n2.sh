#!/bin/bash
# two GPUs of one box: the C2 bench line (weak scaling) and C5 at a quarter of the named size (each rank holds the
# share it would hold on 8 GPUs at full size: 1.25 G residues), NCCL top-k merge in both
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $T bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?" >> gpurun_out/bench_n2.err
timeout 900 $T bench.py --gpus 2 --workload c5 --scale 0.25 --steps 1 --warmup 3 > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err; echo "rc=$?" >> gpurun_out/bench_c5_n2.err
cut -c1-300 gpurun_out/bench_n2.json; tail -n 2 gpurun_out/bench_n2.err; cut -c1-300 gpurun_out/bench_c5_n2.json; tail -n 2 gpurun_out/bench_c5_n2.err
