#!/bin/bash
# reversed coordinate pass: levels sized from the hits' scores: C4, C3, C2 top-k and the top-k tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_topk.py tests/test_gpu_traceback.py tests/test_gpu_upstream.py tests/test_gpu_many.py tests/test_gpu_c5.py -m gpu -x -q > gpurun_out/r2m_pytest.txt 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r2m_pytest.txt
for w in c4 c3 c2; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$w', round(d['value'],1), round(d['ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'], d['topk_crc'])"
done
