#!/bin/bash
# A/B of libswgpu.so against every seq-align-gpu_b200/libswgpu_v*.so present (built by scripts/build_variant.sh), interleaved on one box
mkdir -p gpurun_out
L=seq-align-gpu_b200; LIBS=$L/libswgpu.so
for f in $L/libswgpu_v*.so; do [ -f "$f" ] && LIBS=$LIBS,$f; done
timeout 2400 python scripts/ab_lib.py --libs $LIBS --rounds ${ROUNDS:-2} > gpurun_out/r2_ab_variants.log 2>&1
echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2_ab_variants.log
