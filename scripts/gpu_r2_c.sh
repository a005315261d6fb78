#!/bin/bash
# round 2, batch c (2 GPUs): tests touched by the k3x tag / packer / counter changes, u8x4 microbenchmark, then the
# multi-GPU benches: weak scaling (one C2 shard per rank) and --sharded (one 2 x C2 database dealt over the ranks) at N = 1 and 2
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_topk.py tests/test_gpu_traceback.py tests/test_gpu_c2.py tests/test_gpu_plans.py -m gpu -x -q > gpurun_out/r2_pytest_c.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_c.log
tail -n 4 gpurun_out/r2_pytest_c.log
timeout 120 bench_micro/ubench_u8x4 > gpurun_out/r2_ubench_u8x4.jsonl 2>&1; cat gpurun_out/r2_ubench_u8x4.jsonl
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "n2 rc=$?"
timeout 900 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_n2_ref.json 2> gpurun_out/r2_bench_n2_ref.err; echo "n2 ref rc=$?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_n1_ref.json 2> gpurun_out/r2_bench_n1_ref.err; echo "n1 ref rc=$?"
timeout 900 python bench.py --sharded --shards-total 2 --steps 2 --warmup 3 > gpurun_out/r2_bench_sharded_n1.json 2> gpurun_out/r2_bench_sharded_n1.err; echo "sharded n1 rc=$?"
timeout 900 $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --sharded --shards-total 2 --steps 2 --warmup 3 > gpurun_out/r2_bench_sharded_n2.json 2> gpurun_out/r2_bench_sharded_n2.err; echo "sharded n2 rc=$?"
python - <<'PY'
import json
for f in ("n2","n2_ref","n1_ref","sharded_n1","sharded_n2"):
    try:
        d=json.load(open(f"gpurun_out/r2_bench_{f}.json"))
        print(f, round(d["value"],1), d.get("ms_per_step") and round(d["ms_per_step"],1), "cores", (d.get("cpu_baseline") or {}).get("cores"), "merge", (d.get("merge") or {}).get("ms_per_step"), "crc", d.get("topk_crc"), "parity", d.get("parity") and (d["parity"]["checked"], d["parity"]["mismatch"]))
    except Exception as e:
        print(f, "ERR", e)
PY
tail -n 3 gpurun_out/r2_bench_n2.err gpurun_out/r2_bench_sharded_n2.err
