"""Reader for the descriptor dump of the C++ host front end (`nori_b200 --dump FILE scene.xml`,
nori_b200/host/src/main.cpp:dumpScene): returns a FlatScene, i.e. exactly what crosses the C ABI."""
from __future__ import annotations

import os
import subprocess

import numpy as np

from .scene import FlatMesh, FlatScene

HOST_BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bin", "nori_b200")


def read_dump(path: str) -> FlatScene:
    buf = open(path, "rb").read()
    assert buf[:8] == b"NB2DUMP1", "not a nori_b200 dump"
    off = 8

    def take(dtype, n=1):
        nonlocal off
        a = np.frombuffer(buf, dtype=dtype, count=n, offset=off)
        off += a.nbytes
        return a

    n_meshes, W, H = [int(x) for x in take("<i4", 3)]
    near, far, fov = [float(x) for x in take("<f4", 3)]
    s2c, c2w = take("<f4", 16).reshape(4, 4).copy(), take("<f4", 16).reshape(4, 4).copy()
    radius = float(take("<f4")[0]); table = take("<f4", 33).copy()
    integrator, rr_start, max_depth, spp = [int(x) for x in take("<i4", 4)]
    rr_max = float(take("<f4")[0]); light_pos, light_energy = take("<f4", 3).copy(), take("<f4", 3).copy()
    meshes = []
    for _ in range(n_meshes):
        nV, nF, hasN, hasUV = [int(x) for x in take("<u4", 4)]
        btype = int(take("<i4")[0]); params = take("<f4", 6).copy()
        emitter = int(take("<i4")[0]); radiance = take("<f4", 3).copy()
        V = take("<f4", 3 * nV).reshape(nV, 3).copy()
        N = take("<f4", 3 * nV).reshape(nV, 3).copy() if hasN else None
        UV = take("<f4", 2 * nV).reshape(nV, 2).copy() if hasUV else None
        F = take("<u4", 3 * nF).reshape(nF, 3).copy()
        meshes.append(FlatMesh(V=V, F=F, N=N, UV=UV, bsdf_type=btype, bsdf_params=params, is_emitter=bool(emitter), radiance=radiance))
    assert off == len(buf), "trailing bytes in dump"
    return FlatScene(meshes=meshes, sample_to_camera=s2c, camera_to_world=c2w, width=W, height=H, near_clip=near, far_clip=far, fov=fov,
                     filter_radius=radius, filter_table=table, integrator=integrator, rr_start=rr_start, rr_max=rr_max, max_depth=max_depth,
                     light_pos=light_pos, light_energy=light_energy, sample_count=spp)


def parse_scene(xml_path: str, dump_path: str) -> FlatScene:
    """Runs the C++ front end on a scene description (no GPU needed) and reads its descriptors back."""
    r = subprocess.run([HOST_BIN, "--dump", dump_path, xml_path], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nori_b200 --dump failed ({r.returncode}): {r.stderr.strip()}")
    return read_dump(dump_path)
