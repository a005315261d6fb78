"""CPU suite: the C-ABI library loads and exports every symbol include/*.h declares; compute entry
points fail loudly (no CPU fallback) when there is no CUDA device.  No compute is launched here."""
import ctypes as C
import glob
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def declared_functions():
    names = []
    for h in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        txt = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        txt = re.sub(r"//[^\n]*", "", txt)
        txt = re.sub(r"^\s*#.*$", "", txt, flags=re.M)
        for m in re.finditer(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{}()]*(?:\([^()]*\)[^;{}()]*)*\)\s*;", txt):
            n = m.group(1)
            if n not in ("defined", "alignas", "sizeof", "__attribute__"):
                names.append((os.path.basename(h), n))
    return names


def libs():
    import swgpu
    out = {"swgpu.h": swgpu.lib()}
    host = os.path.join(ROOT, "seq-align-gpu_b200", "libalign.so")
    if os.path.exists(host):
        out["host"] = C.CDLL(host)
    return out


def test_every_declared_symbol_is_exported():
    L = libs()
    decl = declared_functions()
    assert len([d for d in decl if d[0] == "swgpu.h"]) >= 12
    missing = []
    for header, name in decl:
        lib = L["swgpu.h"] if header == "swgpu.h" else L.get("host")
        if lib is None:
            missing.append((header, name, "host library not built"))
            continue
        if not hasattr(lib, name):
            missing.append((header, name))
    assert not missing, missing


def test_no_cpu_fallback_without_device():
    import swgpu
    L = swgpu.lib()
    if L.swgpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    codes = np.array([1, 2, 3, 4], dtype=np.int8)
    off = np.array([0, 4], dtype=np.int64)
    with pytest.raises(swgpu.SwgpuError) as e:
        swgpu.Database(codes, off)
    assert "swgpu error -1" in str(e.value) or "swgpu error -2" in str(e.value)     # ENODEV / ECUDA, never a result


def test_version_string():
    import swgpu
    assert b"sm_100a" in swgpu.lib().swgpu_version()


def test_oracle_is_not_linked_into_the_product():
    """The product libraries must not reference the checker (oracle/) in any way."""
    for so in glob.glob(os.path.join(ROOT, "seq-align-gpu_b200", "*.so")):
        blob = open(so, "rb").read()
        assert b"swo_" not in blob and b"liboracle" not in blob and b"refh_" not in blob, so
    for src in glob.glob(os.path.join(ROOT, "seq-align-gpu_b200", "**", "*"), recursive=True):
        if os.path.isfile(src) and src.endswith((".c", ".cu", ".cuh", ".h", ".py")):
            txt = open(src, errors="ignore").read()
            for pat in ("from oracle", "import oracle", "liboracle", "sw_oracle", "oracle/", "libalign_ref"):
                assert pat not in txt, (src, pat)
