"""CPU: the C-ABI library loads and exports every symbol include/orb200.h declares; without a
CUDA device it refuses to create a context (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT, has_gpu
from optix_renderer_b200 import api


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "orb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = ctypes.CDLL(api.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), "liborb200.so does not export " + n
    assert lib.orb_api_version() == 2


def test_struct_layouts_match_header():
    assert ctypes.sizeof(api.OrbRenderConfig) == 32
    assert ctypes.sizeof(api.OrbDenoiseConfig) == 16
    assert ctypes.sizeof(api.OrbCounters) == 144
    assert ctypes.sizeof(api.OrbBvhOptions) == 16
    assert api.OrbMeshDesc.index.offset == 32 and api.OrbMeshDesc.diffuse.offset == 44


def test_no_cpu_fallback():
    if has_gpu():
        pytest.skip("a GPU is present")
    with pytest.raises(api.OrbError) as e:
        api.Renderer()
    assert "no CUDA device" in str(e.value)


def test_product_does_not_touch_the_oracle():
    """Nothing under optix_renderer_b200/ or host/ may import, link or open oracle/."""
    bad = []
    for top in ("optix_renderer_b200", "host", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp", "Makefile")):
                    s = open(os.path.join(d, f), errors="ignore").read()
                    if re.search(r"liboracle|_oracle|oracle_api|libref_|trace_cpu\.h", s) and "oracle/" in s:
                        for ln in s.splitlines():
                            if re.search(r"#include|import|CDLL|dlopen|-l", ln) and re.search(r"oracle|libref_", ln):
                                bad.append((f, ln.strip()))
    assert not bad, bad
