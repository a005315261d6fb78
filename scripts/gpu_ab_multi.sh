#!/bin/bash
# A/B of several builds of libswgpu.so (scripts/build_variant.sh) on one box: LIBS="v0 v1 ..." (first = baseline)
mkdir -p gpurun_out
L=""; for v in $LIBS; do L="$L,seq-align-gpu_b200/libswgpu_$v.so"; done
timeout 1500 python scripts/ab_lib.py --libs ${L#,} --rounds ${ROUNDS:-2} > gpurun_out/ab_multi.log 2>&1
cat gpurun_out/ab_multi.log
