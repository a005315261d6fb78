"""N > 1 host logic on CPU: world_size-2 gloo run of shard dealing + top-k gather/merge.
Scores here are a deterministic stand-in (a hash of the entry index) -- the point is the partition and the
merge rule, not the kernel."""
import os
import subprocess
import sys
import textwrap

import numpy as np

from conftest import ROOT


def fake_scores(entries):
    e = np.asarray(entries, dtype=np.int64)
    return ((e * 2654435761) % 97).astype(np.int64)          # many ties -> exercises the entry-ascending rule


def test_deal_shards_balances_residues_and_length_mix():
    from swgpu import synth
    from swgpu.shard import deal_shards, take_shard
    codes, off = synth.make_db(9, 5000, sort=False)
    for world in (2, 4, 8):
        shards = deal_shards(off, world)
        assert sorted(np.concatenate(shards).tolist()) == list(range(5000))
        lens = off[1:] - off[:-1]
        res = [int(lens[s].sum()) for s in shards]
        assert (max(res) - min(res)) / (sum(res) / world) < 0.02
        assert max(int(lens[s].max()) for s in shards) / min(int(lens[s].max()) for s in shards) < 2.0
    c, o = take_shard(codes, off, shards[1])
    e = shards[1][7]
    assert np.array_equal(c[o[7]:o[8]], codes[off[e]:off[e + 1]])


def test_merge_is_independent_of_world_size():
    from swgpu.shard import deal_shards, merge_topk
    off = np.arange(0, 3001 * 10, 10)
    n = 3000
    want = None
    for world in (1, 2, 3, 8):
        lists = []
        for s in deal_shards(off, world):
            sc = fake_scores(s)
            o = np.lexsort((s, -sc))[:50]
            lists.append(np.stack([sc[o], s[o]], axis=1))
        got = merge_topk(lists, 50)
        if want is None:
            allsc = fake_scores(np.arange(n))
            o = np.lexsort((np.arange(n), -allsc))[:50]
            want = np.stack([allsc[o], np.arange(n)[o]], axis=1)
        assert np.array_equal(got, want), world


WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np
    sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "seq-align-gpu_b200"))
    sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
    import torch.distributed as dist
    from swgpu.shard import deal_shards, gather_topk
    from test_multi_gloo import fake_scores
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    off = np.arange(0, 2001 * 7, 7)
    mine = deal_shards(off, world)[rank]
    sc = fake_scores(mine)
    o = np.lexsort((mine, -sc))[:25]
    merged = gather_topk(np.stack([sc[o], mine[o]], axis=1), 25, dist)
    allsc = fake_scores(np.arange(2000))
    oo = np.lexsort((np.arange(2000), -allsc))[:25]
    want = np.stack([allsc[oo], np.arange(2000)[oo]], axis=1)
    assert np.array_equal(merged, want), (rank, merged[:5], want[:5])
    # the batched form bench.py uses: ONE all_gather for the hit lists of every query of a step
    from swgpu.shard import gather_topk_batch
    nq = 5
    loc = np.full((nq, 25, 2), -1, dtype=np.int64)
    wants = []
    for qi in range(nq):
        scq = (fake_scores(mine) * (qi + 3)) % 89
        oq = np.lexsort((mine, -scq))[: 25 - qi]                 # ragged: fewer than k hits on some queries
        loc[qi, : len(oq), 0] = scq[oq]; loc[qi, : len(oq), 1] = mine[oq]
        alls = (fake_scores(np.arange(2000)) * (qi + 3)) % 89
        parts = []
        for r in range(world):
            e = deal_shards(off, world)[r]; s_ = alls[e]; o_ = np.lexsort((e, -s_))[: 25 - qi]
            parts.append(np.stack([s_[o_], e[o_]], axis=1))
        cat = np.concatenate(parts); oo = np.lexsort((cat[:, 1], -cat[:, 0]))[:25]
        wants.append(cat[oo])
    got, ex_ms, mg_ms = gather_topk_batch(loc, 25, dist)
    for qi in range(nq):
        g = got[qi][got[qi][:, 1] >= 0]
        assert np.array_equal(g, wants[qi]), (rank, qi, g[:4], wants[qi][:4])
    assert ex_ms >= 0 and mg_ms >= 0
    dist.barrier(); dist.destroy_process_group()
    open(os.path.join(sys.argv[2], f"ok_{rank}"), "w").write("ok")
""")


def test_gloo_world2_gather_and_merge(tmp_path):
    import socket
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(w), ROOT, str(tmp_path)],
                       capture_output=True, text=True, timeout=180, env={**os.environ, "OMP_NUM_THREADS": "1"})
    assert r.returncode == 0, r.stdout + r.stderr
    assert (tmp_path / "ok_0").exists() and (tmp_path / "ok_1").exists()
