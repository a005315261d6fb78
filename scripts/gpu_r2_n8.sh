#!/bin/bash
# round 2: N = 8 sanity line with the final code (gpurun --gpus 8; short: one timed step)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 1 --warmup 3 --no-cpu > gpurun_out/r2n_bench_n8.json 2> gpurun_out/r2n_bench_n8.err; echo "n8 rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2n_bench_n8.json"))
print("n8", round(d["value"],1), "n_gpus", d["n_gpus"], "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],1), "merge", d["merge"]["ms_per_step"], "parity", d["parity"]["mismatch"], "clocks", d["clocks"].get("reasons"))
PY
