#!/bin/bash
# round 2, final N = 2 lines (gpurun --gpus 2): the multi-GPU tests, weak scaling on C2 with the reference arm beside it, ONE database dealt over two ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2n_pytest_multi.txt 2>&1; echo "multi rc=$?"; tail -1 gpurun_out/r2n_pytest_multi.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r2n_bench_n1.json 2> gpurun_out/r2n_bench_n1.err; echo "n1 rc=$?"
timeout 900 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2n_bench_n2.json 2> gpurun_out/r2n_bench_n2.err; echo "n2 rc=$?"
timeout 900 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/r2n_bench_n2_ref.json 2> gpurun_out/r2n_bench_n2_ref.err; echo "n2 ref rc=$?"
timeout 900 python bench.py --sharded --shards-total 2 --steps 2 --warmup 3 --no-cpu > gpurun_out/r2n_bench_sharded_n1.json 2> gpurun_out/r2n_bench_sharded_n1.err; echo "sharded n1 rc=$?"
timeout 900 $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --sharded --shards-total 2 --steps 2 --warmup 3 --no-cpu > gpurun_out/r2n_bench_sharded_n2.json 2> gpurun_out/r2n_bench_sharded_n2.err; echo "sharded n2 rc=$?"
python - <<'PY'
import json
for f in ("n1","n2","n2_ref","sharded_n1","sharded_n2"):
    try:
        d=json.load(open(f"gpurun_out/r2n_bench_{f}.json"))
        print(f, round(d["value"],1), d["unit"], "n_gpus", d["n_gpus"], "ms", round(d["ms_per_step"],2), "cores", d.get("cpu_baseline",{}).get("cores"), "crc", d.get("topk_crc"), "merge", d.get("merge",{}).get("ms_per_step"), "parity", d.get("parity",{}).get("mismatch"))
    except Exception as e: print(f, "ERR", e)
PY
