"""GPU suite (-m gpu): the C product's own multi-GPU path -- SWGPU_GPUS=<n> in bin/smith_waterman
(host/alignment_cmdline.c: length-balanced deal of the database over n shards, one host thread and one engine handle per
shard, host-side merge of the per-shard scores / top-k lists).  The listing, the top-k list with coordinates and the
CIGAR lines must be byte-equal to the single-GPU run for every shard count (the tie-break rule makes the merge
independent of the sharding).  On a box with fewer GPUs than shards SWGPU_GPUS_SHARE_DEVICE=1 lets the shards share the
devices, so the shard / merge logic is exercised on one GPU too; with >= 2 GPUs the same test also runs on real devices."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, SCORING

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "bin", "smith_waterman")


def run(args, **env):
    e = {k: v for k, v in os.environ.items() if not k.startswith("SWGPU_")}
    e.update(env)
    r = subprocess.run([BIN, *args], capture_output=True, text=True, timeout=600, env=e)
    assert r.returncode == 0, r.stderr[-2000:]
    return re.sub(r"Total Time: [0-9.]+", "Total Time: T", r.stdout)


def test_swgpu_gpus_sharded_cli_equals_single_gpu(tmp_path):
    import swgpu
    from swgpu import synth
    ndev = swgpu.lib().swgpu_device_count()
    g = synth.rng(4242)
    n = 2003                                                   # not a multiple of the shard counts, nor of 16
    lens = synth.lognormal_lengths(g, n, lo=1, hi=5000)
    lens[7] = 5000; lens[100] = 0
    off = np.zeros(n + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, off[-1])
    qs = [synth.make_query(51, 300), synth.make_query(52, 1100)]
    for t in (3, 500, 1500, 1501):                             # related targets, two of them with EQUAL content: a score tie across shards
        L = min(int(lens[t]), 280)
        codes[off[t]:off[t] + L] = qs[0][:L]
    codes[off[1501]:off[1501 + 1]] = np.resize(codes[off[1500]:off[1500 + 1]], int(lens[1501]))
    synth.to_fasta(str(tmp_path / "db.fa"), codes, off)
    qo = np.array([0, 300, 1400], dtype=np.int64)
    synth.to_fasta(str(tmp_path / "q.fa"), np.concatenate(qs), qo, prefix="query")
    base = ["--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--gapopen", "-10", "--gapextend", "-1",
            "--files", str(tmp_path / "q.fa"), str(tmp_path / "db.fa")]
    modes = [dict(), dict(SWGPU_TOPK="25", SWGPU_CIGAR="1", SWGPU_ALL_QUERIES="1")]
    for m in modes:
        want = run(base, **m)
        assert ("Hit #24" in want) if m else ("Total Entries: 2003" in want)
        for shards in (2, 3, 5):
            got = run(base, SWGPU_GPUS=str(shards), SWGPU_GPUS_SHARE_DEVICE="1", **m)
            assert got == want, (m, shards)
        if ndev >= 2:                                          # real devices, no sharing
            assert run(base, SWGPU_GPUS=str(min(ndev, 8)), **m) == want
