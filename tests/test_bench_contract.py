"""CPU suite: the reference arm of bench.py (the reference's own AVX2/OpenMP kernel on the host cores, oracle/_ref, or the
C restatement when that library is absent) prints exactly ONE JSON line on stdout with the keys the driver reads.
The GPU arm cannot run here; its line is produced by the same code path and differs in `impl` and the measured values."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-seconds", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "impl", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == "GCUPS" and d["unit"] == "GCUPS" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert "C2" in d["config"]["workload"] and d["vs_baseline"] is None
    # the arm says what it ran (a bounded sample, four query lengths incl. 5478) and times a one-thread row
    assert "5478" in d["ran"]["sample"] and d["cpu_baseline"]["one_thread"]["cores"] == 1
    assert d["ms_per_step"] > 0


def test_reference_arm_ignores_torchrun_omp_num_threads():
    """torchrun exports OMP_NUM_THREADS=1: the arm must still use every core it may run on (round-1 defect: the N > 1
    reference values were single-threaded)."""
    avail = len(os.sched_getaffinity(0))
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-seconds", "1", "--workload", "c1"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["cpu_baseline"]["cores"] == avail and d["ran"]["omp_num_threads_env"] == "1"
    assert "C1" in d["config"]["workload"]


def test_reference_arm_other_ranks_stay_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_c5_reference_arm_says_unavailable():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c5"],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d
