#!/bin/bash
# round 2: ncu --set full of the K1S launch of C4 (every stream an X item, one CTA per SM), after the same command ran plain
mkdir -p gpurun_out
timeout 300 python scripts/prof_c4.py --reps 2 > gpurun_out/r2p_c4_plain.log 2>&1 || { echo "plain failed"; tail -5 gpurun_out/r2p_c4_plain.log; exit 1; }
tail -1 gpurun_out/r2p_c4_plain.log | cut -c1-300
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k1s$" -s 1 -c 1 -f -o gpurun_out/r2p_k1s_c4 python scripts/prof_c4.py --reps 2 > gpurun_out/r2p_ncu_c4.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/r2p_ncu_c4.log; ls -la gpurun_out/r2p_k1s_c4.ncu-rep
