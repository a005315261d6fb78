#!/bin/bash
# round 2, batch u: planner tables from the round-2 sweep + the planner guard test on two other database shapes; C2 A/B line
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_planner.py tests/test_gpu_plans.py -m gpu -x -q -s > gpurun_out/r2u_pytest_planner.txt 2>&1; echo "pytest rc=$?"; grep -E "planner on|passed|failed|Error|assert" gpurun_out/r2u_pytest_planner.txt | cut -c1-400 | tail -8
timeout 900 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c2', round(d['value'],1), round(d['ms_per_step'],2), 'frac', round(d['roofline']['frac'],4), 'k1', round(d['roofline']['k1_ms_per_step'],2), d['parity']['mismatch'], d['topk_crc'])"
timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'k1', round(d['roofline']['k1_ms_per_step'],2), d['parity']['mismatch'])"
