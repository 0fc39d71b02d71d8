"""Builds nori_b200/libnori_b200.so (the C-ABI engine, include/nori_gpu.h) with nvcc for sm_100a and
nori_b200/bin/nori_b200 (the C++ host front end above the C ABI, nori_b200/host) with g++.

In-tree build: the binaries travel with the repo snapshot to the GPU box. nvcc cross-compiles
without a GPU. `python -m nori_b200.build [--force] [-v]`.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libnori_b200.so")
SOURCES = ["capi.cu", "render.cu", "operators.cu", "bvh.cu", "pathgraph.cu", "pathgraph_scatter.cu", "objdedup.cu"]
HEADERS = ["engine.h", "bvh.h", "check.cuh", "nmath.cuh", "scene.cuh", "traverse.cuh", "traverse_persistent.cuh", "pathgraph.h", "pathgraph_bsdf.cuh", os.path.join("..", "..", "include", "nori_gpu.h")]
# -fmad=false: the arithmetic order written in nmath.cuh is the order executed (bit parity with
# the x86-64 reference build, which has no FMA). Division and sqrt stay IEEE (nvcc defaults).
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-Xcudafe", "--diag_suppress=177",
              "-I", os.path.join(HERE, "..", "include")]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


LIB_CHECKED = os.path.join(HERE, "libnori_b200_checked.so")


def build_variant(name: str, defines, verbose: bool = False) -> str:
    """Experiment builds: the same sources with extra -D flags -> libnori_b200_<name>.so (select with NORI_GPU_LIB)."""
    objdir = os.path.join(HERE, "build_" + name)
    lib = os.path.join(HERE, "libnori_b200_%s.so" % name)
    os.makedirs(objdir, exist_ok=True)
    objs, procs = [], []
    for s in SOURCES:
        src, obj = os.path.join(CSRC, s), os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [_nvcc()] + NVCC_FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {s}\n{out}\n")
        if p.returncode != 0:
            raise RuntimeError("nvcc failed building " + lib)
    subprocess.check_call([_nvcc(), "-shared", "-o", lib] + objs + ["-lcudart_static", "-lpthread", "-ldl", "-lrt"])
    return lib


def build(force: bool = False, verbose: bool = False, checked: bool = False) -> str:
    """checked=True: the same sources with -DNORI_GPU_CHECKED (csrc/check.cuh: every pool / record / queue /
    frame index range-checked on the device) -> libnori_b200_checked.so, loaded when NORI_GPU_LIB names it."""
    objdir = os.path.join(HERE, "build_checked" if checked else "build")
    lib = LIB_CHECKED if checked else LIB
    flags = NVCC_FLAGS + (["-DNORI_GPU_CHECKED"] if checked else [])
    os.makedirs(objdir, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    objs, procs = [], []
    for s in SOURCES:
        src, obj = os.path.join(CSRC, s), os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = [_nvcc()] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {s}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed building libnori_b200.so")
    if force or procs or _stale(lib, objs):
        subprocess.check_call([_nvcc(), "-shared", "-o", lib] + objs + ["-lcudart_static", "-lpthread", "-ldl", "-lrt"])
    return lib


HOST = os.path.join(HERE, "host")
HOST_BIN = os.path.join(HERE, "bin", "nori_b200")
HOST_SOURCES = ["common.cpp", "object.cpp", "xml.cpp", "parser.cpp", "plugins.cpp", "mesh.cpp", "scene.cpp", "bitmap.cpp", "tests.cpp", "main.cpp"]
# No FMA contraction and no fast-math: matrices and filter tables must carry the values the
# reference's x86-64 build computes.
HOST_FLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-Wall", "-Wextra", "-Wno-unused-parameter",
              "-I", HOST, "-I", os.path.join(HERE, "..", "include")]


def build_host(force: bool = False, verbose: bool = False) -> str:
    """g++ build of the host front end; links against libnori_b200.so next to it ($ORIGIN rpath)."""
    objdir = os.path.join(HERE, "build", "host")
    os.makedirs(objdir, exist_ok=True)
    os.makedirs(os.path.dirname(HOST_BIN), exist_ok=True)
    hdrs = [os.path.join(HOST, "nori", h) for h in sorted(os.listdir(os.path.join(HOST, "nori")))] + [os.path.join(HOST, "src", "xml.h"),
            os.path.join(HERE, "..", "include", "nori_gpu.h")]
    cxx = os.environ.get("CXX", "g++")
    objs, procs = [], []
    for s in HOST_SOURCES:
        src, obj = os.path.join(HOST, "src", s), os.path.join(objdir, s.replace(".cpp", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            procs.append((s, subprocess.Popen([cxx] + HOST_FLAGS + ["-c", src, "-o", obj], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or (verbose and out.strip()):
            sys.stderr.write(f"--- g++ {s}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("g++ failed building the host front end")
    if force or procs or _stale(HOST_BIN, objs + [LIB]):
        subprocess.check_call([cxx, "-o", HOST_BIN] + objs + ["-L", HERE, "-lnori_b200", "-Wl,-rpath,$ORIGIN/..", "-lpthread", "-ldl"])
    return HOST_BIN


if __name__ == "__main__":
    force, verbose = "--force" in sys.argv, "-v" in sys.argv
    print(build(force=force, verbose=verbose))
    if "--checked" in sys.argv or os.environ.get("NORI_GPU_CHECKED"):
        print(build(force=force, verbose=verbose, checked=True))
    print(build_host(force=force, verbose=verbose))

