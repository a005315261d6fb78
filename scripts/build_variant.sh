#!/bin/bash
# Build a variant of libswgpu.so with extra -D flags for in-box A/B runs (scripts/ab_lib.py):
#   scripts/build_variant.sh <name> [-DK1S_...=...]...   ->  seq-align-gpu_b200/libswgpu_<name>.so
set -e
name=$1; shift
cd "$(dirname "$0")/../seq-align-gpu_b200"
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fopenmp,-Wall"
make -s csrc/db_pack.o >/dev/null 2>&1 || make -s >/dev/null
$NV "$@" -Xptxas -O1 -c csrc/sw_k1s_o1.cu -o /tmp/sw_k1s_o1_$name.o
$NV "$@" -shared -o libswgpu_$name.so csrc/swgpu.cu csrc/swgpu_topk.cu csrc/swgpu_traceback.cu csrc/swgpu_synth.cu /tmp/sw_k1s_o1_$name.o csrc/db_pack.o -lgomp
ls -la libswgpu_$name.so
