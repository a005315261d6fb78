"""GPU suite: the guard-zone debug mode (SWGPU_GUARD=1, include/swgpu.h) sees a one-byte overrun of an engine allocation, and a
search + top-k + traceback on a small database leaves every guard intact.  Runs in a child process: the mode is read from the
environment when the library makes its first allocation."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(%r, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth
L = swgpu.lib()
assert L.swgpu_guard_check() == 0                      # mode on, nothing allocated yet
assert L.swgpu_guard_selftest() == 1                   # a one-byte overrun is reported, and only while the buffer lives
tab = synth.load_matrix(os.path.join(%r, "scoring", "BLOSUM62.txt"))[0]
codes, off = synth.make_db(5, 3000, median=200.0, lo=1, hi=6000)
db = swgpu.Database(codes, off)
for lq in (60, 400, 1500):
    q = synth.make_query(6 + lq, lq)
    db.search(q, tab, -10, -1)
    db.topk(50)
    assert L.swgpu_guard_check() == 0, lq
db.search_many([synth.make_query(9, 300), synth.make_query(10, 2100)], tab, -10, -1, k=20)
assert L.swgpu_guard_check() == 0
db.close()
print("guard OK")
"""


def test_guard_mode_reports_an_overrun_and_a_clean_search_leaves_none():
    env = dict(os.environ, SWGPU_GUARD="1")
    r = subprocess.run([sys.executable, "-c", CHILD % (ROOT, ROOT)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "guard OK" in r.stdout, r.stdout + r.stderr
    assert "written past its end" in r.stderr           # the selftest's deliberate overrun was described


def test_guard_mode_is_off_by_default():
    import swgpu
    if not os.environ.get("SWGPU_GUARD"):
        assert swgpu.lib().swgpu_guard_check() == -1
