"""Flattened scene description: the host-side POD view of a Nori scene graph.

A Nori scene (Scene -> Mesh[] {BSDF, Emitter}, Camera {ReconstructionFilter}, Sampler,
Integrator; reference include/nori/scene.h:75-127) is flattened into plain arrays before
it crosses the C ABI (include/nori_gpu.h). The arrays keep the reference's own layouts:

* per mesh ``V`` (3 x nV, column-major == numpy (nV, 3) C-order), ``N``, ``UV``, ``F``
  (3 x nF uint32) exactly as ``Mesh::m_V/m_N/m_UV/m_F`` (include/nori/mesh.h:172-175);
* camera ``sampleToCamera`` / ``cameraToWorld`` 4x4 row-major as computed by
  ``PerspectiveCamera::activate`` (src/perspective.cpp:53-86);
* the 33-entry filter table of ``ImageBlock`` (src/block.cpp:31-39).

``FlatScene`` round-trips through ``.npz`` so scenes travel to machines where the
XML/OBJ sources are absent (tests/golden/scenes/*.npz).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

# BSDF / integrator / stream enums shared with include/nori_gpu.h
BSDF_DIFFUSE, BSDF_MIRROR, BSDF_DIELECTRIC, BSDF_MICROFACET = 0, 1, 2, 3
INTEGRATOR_NAMES = {
    "normals": 0, "path_mats": 1, "path_ems": 2, "path_mis": 3,
    "whitted": 4, "primary_shadow": 5, "simple": 6, "ao": 7,
}
STREAM_BLOCK, STREAM_PER_SAMPLE = 0, 1


@dataclass
class FlatMesh:
    V: np.ndarray                        # (nV, 3) float32
    F: np.ndarray                        # (nF, 3) uint32
    N: Optional[np.ndarray] = None       # (nV, 3) float32 or None
    UV: Optional[np.ndarray] = None      # (nV, 2) float32 or None
    bsdf_type: int = BSDF_DIFFUSE
    bsdf_params: np.ndarray = field(default_factory=lambda: np.array([.5, .5, .5, 0, 0, 0], np.float32))
    is_emitter: bool = False
    radiance: np.ndarray = field(default_factory=lambda: np.zeros(3, np.float32))

    def __post_init__(self):
        self.V = np.ascontiguousarray(self.V, np.float32).reshape(-1, 3)
        self.F = np.ascontiguousarray(self.F, np.uint32).reshape(-1, 3)
        if self.N is not None:
            self.N = np.ascontiguousarray(self.N, np.float32).reshape(-1, 3)
        if self.UV is not None:
            self.UV = np.ascontiguousarray(self.UV, np.float32).reshape(-1, 2)
        self.bsdf_params = np.ascontiguousarray(self.bsdf_params, np.float32).reshape(6)
        self.radiance = np.ascontiguousarray(self.radiance, np.float32).reshape(3)


@dataclass
class FlatScene:
    meshes: List[FlatMesh]
    sample_to_camera: np.ndarray         # (4, 4) float32 row-major
    camera_to_world: np.ndarray          # (4, 4) float32 row-major
    width: int
    height: int
    near_clip: float = 1e-4
    far_clip: float = 1e4
    fov: float = 30.0
    filter_radius: float = 2.0
    filter_table: np.ndarray = field(default_factory=lambda: np.zeros(33, np.float32))
    integrator: int = 3                  # INTEGRATOR_NAMES value
    rr_start: int = 3
    rr_max: float = 0.99
    max_depth: int = 0
    light_pos: np.ndarray = field(default_factory=lambda: np.zeros(3, np.float32))
    light_energy: np.ndarray = field(default_factory=lambda: np.zeros(3, np.float32))
    sample_count: int = 1

    @property
    def border(self) -> int:             # src/block.cpp:32
        return int(np.ceil(np.float32(self.filter_radius) - np.float32(0.5)))

    @property
    def n_triangles(self) -> int:
        return int(sum(m.F.shape[0] for m in self.meshes))

    def block_shape(self):
        """(rows, cols, 4) of the full-frame ImageBlock incl. border (src/block.cpp:48)."""
        b = self.border
        return (self.height + 2 * b, self.width + 2 * b, 4)

    # ---- npz round trip -------------------------------------------------------------
    def save(self, path: str) -> None:
        d = dict(
            n_meshes=np.int32(len(self.meshes)),
            s2c=np.asarray(self.sample_to_camera, np.float32).reshape(4, 4),
            c2w=np.asarray(self.camera_to_world, np.float32).reshape(4, 4),
            size=np.array([self.width, self.height], np.int32),
            clip_fov=np.array([self.near_clip, self.far_clip, self.fov], np.float32),
            filter_radius=np.float32(self.filter_radius),
            filter_table=np.asarray(self.filter_table, np.float32),
            integrator=np.array([self.integrator, self.rr_start, self.max_depth, self.sample_count], np.int32),
            integrator_f=np.concatenate([[np.float32(self.rr_max)], self.light_pos, self.light_energy]).astype(np.float32),
            bsdf_type=np.array([m.bsdf_type for m in self.meshes], np.int32),
            bsdf_params=np.stack([m.bsdf_params for m in self.meshes]).astype(np.float32),
            is_emitter=np.array([m.is_emitter for m in self.meshes], np.int32),
            radiance=np.stack([m.radiance for m in self.meshes]).astype(np.float32),
        )
        for i, m in enumerate(self.meshes):
            d[f"V{i}"] = m.V
            d[f"F{i}"] = m.F
            if m.N is not None:
                d[f"N{i}"] = m.N
            if m.UV is not None:
                d[f"UV{i}"] = m.UV
        np.savez_compressed(path, **d)

    @staticmethod
    def load(path: str) -> "FlatScene":
        z = np.load(path)
        meshes = []
        for i in range(int(z["n_meshes"])):
            meshes.append(FlatMesh(
                V=z[f"V{i}"], F=z[f"F{i}"],
                N=z[f"N{i}"] if f"N{i}" in z.files else None,
                UV=z[f"UV{i}"] if f"UV{i}" in z.files else None,
                bsdf_type=int(z["bsdf_type"][i]), bsdf_params=z["bsdf_params"][i],
                is_emitter=bool(z["is_emitter"][i]), radiance=z["radiance"][i]))
        it, itf = z["integrator"], z["integrator_f"]
        return FlatScene(
            meshes=meshes, sample_to_camera=z["s2c"], camera_to_world=z["c2w"],
            width=int(z["size"][0]), height=int(z["size"][1]),
            near_clip=float(z["clip_fov"][0]), far_clip=float(z["clip_fov"][1]), fov=float(z["clip_fov"][2]),
            filter_radius=float(z["filter_radius"]), filter_table=z["filter_table"],
            integrator=int(it[0]), rr_start=int(it[1]), max_depth=int(it[2]), sample_count=int(it[3]),
            rr_max=float(itf[0]), light_pos=itf[1:4].copy(), light_energy=itf[4:7].copy())
