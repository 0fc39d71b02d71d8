"""Dev-time check (build container only: reads /root/reference): the C++ host front end against the
compiled reference's own parser on the reference's scene files. Prints max ulp distance per field."""
import os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nori_b200.hostdump import parse_scene
from oracle.pyoracle import Ref

def ulps(a, b):
    a = np.ascontiguousarray(a, np.float32).ravel(); b = np.ascontiguousarray(b, np.float32).ravel()
    ia = a.view(np.int32).astype(np.int64); ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia); ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    d = np.abs(ia - ib)
    return int(d.max()) if d.size else 0, int((d != 0).sum())

ref = Ref(build=False)
scenes = sys.argv[1:] or ["pa5/cbox/cbox_mis.xml", "pa5/table/table_mis.xml", "pa5/veach_mi/veach_mis.xml", "pa1/bunny.xml", "pa4/cbox/cbox-whitted.xml"]
for rel in scenes:
    path = os.path.join("/root/reference/scenes", rel)
    want = ref.load_xml(path).flatten()
    with tempfile.TemporaryDirectory() as td:
        got = parse_scene(path, os.path.join(td, "d.bin"))
    print(rel, "meshes", len(got.meshes), len(want.meshes), "size", (got.width, got.height), "integrator", got.integrator, want.integrator, "spp", got.sample_count, want.sample_count)
    print("   s2c", ulps(got.sample_to_camera, want.sample_to_camera), "c2w", ulps(got.camera_to_world, want.camera_to_world),
          "table", ulps(got.filter_table, want.filter_table), "clip/fov", (got.near_clip, got.far_clip, got.fov) == (want.near_clip, want.far_clip, want.fov))
    for i, (g, w) in enumerate(zip(got.meshes, want.meshes)):
        same_f = g.F.shape == w.F.shape and np.array_equal(g.F, w.F)
        print(f"   mesh {i}: nV {g.V.shape[0]}/{w.V.shape[0]} F equal {same_f} V {ulps(g.V, w.V) if g.V.shape == w.V.shape else 'shape'}",
              "N", (ulps(g.N, w.N) if g.N is not None and w.N is not None and g.N.shape == w.N.shape else (g.N is None, w.N is None)),
              "UV", (ulps(g.UV, w.UV) if g.UV is not None and w.UV is not None else (g.UV is None, w.UV is None)),
              "bsdf", g.bsdf_type == w.bsdf_type, ulps(g.bsdf_params, w.bsdf_params), "emit", g.is_emitter == w.is_emitter, ulps(g.radiance, w.radiance))
