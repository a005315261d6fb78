"""GPU suite (-m gpu): the GPU's coordinates and alignment paths against the upstream seq-align binary's recorded output
(tests/golden/upstream_hits.npz; gating rules in tests/test_upstream_pin.py): the only evidence for swgpu_topk /
swgpu_traceback that is not this repo's own oracle."""
import numpy as np
import pytest

from test_upstream_pin import aln_to_cigar, load_pairs

pytestmark = pytest.mark.gpu


def test_gpu_hits_and_paths_against_upstream_binary(oracle):
    import swgpu
    pairs = load_pairs()
    n_end = n_start = n_path = n_path_same = 0
    for setting in sorted({(id(p[2]), p[3]) for p in pairs}, key=lambda x: x[1]):
        grp = [p for p in pairs if (id(p[2]), p[3]) == setting]
        tab, (go, ge) = grp[0][2], grp[0][3]
        # every pair is its own one-target database (the query differs per pair)
        for q, t, _, _, up in grp:
            db = swgpu.Database(t, np.array([0, len(t)], dtype=np.int64))
            got, _ = db.search(q, tab, go, ge)
            assert int(got[0]) == up["score"]
            h = db.topk(1)[0]
            assert h["score"] == up["score"]
            if oracle.count_best(q, t, tab, go, ge) == 1:
                n_end += 1
                assert (h["q_end"], h["t_end"]) == (up["q_end"], up["t_end"]), (h, up)
                if oracle.count_best(q[:h["q_end"] + 1][::-1], t[:h["t_end"] + 1][::-1], tab, go, ge) == 1:
                    n_start += 1
                    assert (h["q_start"], h["t_start"]) == (up["q_start"], up["t_start"]), (h, up)
            cg = db.traceback(q, tab, go, ge, [h])[0]
            qs, qe, ts, te = h["q_start"], h["q_end"], h["t_start"], h["t_end"]
            assert oracle.cigar_score(cg, q[qs:qe + 1], t[ts:te + 1], tab, go, ge) == up["score"]
            if go == -10 and (qs, qe, ts, te) == (up["q_start"], up["q_end"], up["t_start"], up["t_end"]):
                n_path += 1
                n_path_same += cg == aln_to_cigar(up["q_aln"], up["t_aln"])
            db.close()
    assert n_end >= 120 and n_start >= 100 and n_path >= 90 and n_path_same >= 0.9 * n_path, (n_end, n_start, n_path, n_path_same)
