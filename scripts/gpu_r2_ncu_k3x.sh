#!/bin/bash
# round 2: ncu --set full of k3x<false> in its two regimes (one pair; 64 targets), after the same command ran plain
mkdir -p gpurun_out
timeout 300 python scripts/prof_k3x_two.py > gpurun_out/r2l_k3x_two_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2l_k3x_two_plain.log; exit 1; }
cat gpurun_out/r2l_k3x_two_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k3x$" -c 2 -f -o gpurun_out/r2l_k3x python scripts/prof_k3x_two.py > gpurun_out/r2l_ncu_k3x.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/r2l_ncu_k3x.log; ls -la gpurun_out/r2l_k3x.ncu-rep
