"""The C-ABI library loads on a machine without a GPU and exports every symbol include/nori_gpu.h
declares; struct layouts of the ctypes mirror match the header; no compute without a device."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def header_functions():
    src = open(os.path.join(ROOT, "include", "nori_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nori_gpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from nori_b200 import gpu
    L = gpu.load_library()
    names = header_functions()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(gpu.EXPORTS) == names


def test_struct_sizes_match_header():
    from nori_b200 import gpu
    assert C.sizeof(gpu.MeshDesc) == 4 * 8 + 2 * 4 + 4 + 6 * 4 + 4 + 3 * 4 + 4      # pointers, counts, bsdf, params, emitter, radiance, tail padding
    assert C.sizeof(gpu.CameraDesc) == 32 * 4 + 4 * 4
    assert C.sizeof(gpu.RenderDesc) == 16 * 4
    assert C.sizeof(gpu.AccelStats) == 7 * 4


def test_no_cpu_fallback_without_device():
    """On a box without CUDA the product refuses to run instead of computing on the CPU."""
    import torch
    from nori_b200 import gpu
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(gpu.NoriGpuError, match="no CUDA device"):
        gpu.Engine(0)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under nori_b200/ may import, load or link it."""
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "nori_b200")):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle|#\s*include\s*[<\"][^>\"]*oracle|liboracle\.so|libnori_ref", txt, flags=re.M):
                    bad.append(f)
    assert not bad, bad
