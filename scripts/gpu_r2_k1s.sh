#!/bin/bash
# round 2: the rewritten K1S (gap specialisations, unroll, multi-stripe hand-over rings): parity first, then A/B against the round-1 build
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_plans.py tests/test_gpu_parity.py tests/test_gpu_many.py -x -q --durations=5 > gpurun_out/r2_pytest_k1s.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_k1s.log
tail -n 12 gpurun_out/r2_pytest_k1s.log
L=seq-align-gpu_b200
LIBS=$L/libswgpu_old.so,$L/libswgpu.so
for f in $L/libswgpu_v*.so; do [ -f "$f" ] && LIBS=$LIBS,$f; done
timeout 1500 python scripts/ab_lib.py --libs $LIBS --rounds ${ROUNDS:-2} > gpurun_out/r2_ab3.log 2>&1
echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2_ab3.log
