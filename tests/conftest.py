import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "seq-align-gpu_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")
SCORING = os.path.join(ROOT, "scoring")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.oracle()          # builds liboracle.so if missing
    return O


@pytest.fixture(scope="session")
def matrices():
    from swgpu import synth
    return {n: synth.load_matrix(os.path.join(SCORING, n + ".txt"))[0] for n in ("BLOSUM62", "BLOSUM45", "PAM250")}


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(autouse=True)
def _guard_zones(request):
    """Debug mode SWGPU_GUARD=1 (include/swgpu.h): after every GPU test, no live engine allocation may have been written past
    its end.  Off by default; scripts/gpu_guard.sh runs the GPU suite with it (this pool refuses compute-sanitizer)."""
    yield
    if os.environ.get("SWGPU_GUARD") and request.node.get_closest_marker("gpu"):
        import swgpu
        assert swgpu.lib().swgpu_guard_check() == 0, "a device buffer of the engine was written past its end (see stderr)"
