#!/bin/bash
# The GPU suite and the sanitizer case under the guard-zone debug mode (SWGPU_GUARD=1: every engine allocation carries a 512-byte
# tail that is checked after every test and at swgpu_db_destroy): stands in for compute-sanitizer memcheck, which this pool refuses.
mkdir -p gpurun_out
SWGPU_GUARD=1 timeout 200 python scripts/sanitize_case.py > gpurun_out/r2_guard_sanitize_case.txt 2>&1; echo "sanitize_case under guard rc=$?"
SWGPU_GUARD=1 timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r2_guard_pytest.txt 2>&1; echo "pytest under guard rc=$?"; tail -3 gpurun_out/r2_guard_pytest.txt
