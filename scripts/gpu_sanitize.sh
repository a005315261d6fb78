#!/bin/bash
# ONE compute-sanitizer tool per gpurun call (B200_PROFILING.md): scripts/gpu_sanitize.sh memcheck|racecheck|synccheck|initcheck
# The case first runs plain; the tool only runs if that exited 0.  Output: gpurun_out/r2_sanitize_<tool>.txt
tool=${1:-memcheck}
mkdir -p gpurun_out
timeout 300 python scripts/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -n 20 gpurun_out/sanitize_plain.log; exit 1; }
tail -n 1 gpurun_out/sanitize_plain.log
timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python scripts/sanitize_case.py > gpurun_out/r2_sanitize_$tool.txt 2>&1
echo "compute-sanitizer --tool $tool rc=$?" | tee -a gpurun_out/r2_sanitize_$tool.txt
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_case:|========= (Error|Warning|Invalid|Race|Barrier|Uninit)" gpurun_out/r2_sanitize_$tool.txt | head -n 30
