"""For the record: the intersection code the reference actually ships — its UNMODIFIED
single-mesh brute-force Accel (src/accel.cpp:35-111, compiled as it is into oracle/_ref) — on the bunny-class scene
(2 000 triangles), next to the device operator on the same rays. One host thread (the stock class is what each TBB
worker calls); the GPU leg is skipped without a device.   python scripts/bench_stock_accel.py [n_rays]"""
import json, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nori_b200.scene import FlatScene
from oracle.export_xml import write_scene_files            # checker / baseline leg only
from oracle.pyoracle import Ref

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
scene = FlatScene.load(os.path.join(ROOT, "tests", "golden", "scenes", "bunny_normals.npz"))
rs = Ref(build=False).load_xml(write_scene_files(scene, tempfile.mkdtemp(prefix="nori_stock_")))
rng = np.random.default_rng(9)
rays = rs.camera_rays((rng.random((n, 2)) * [scene.width, scene.height]).astype(np.float32))
rs.stock_intersect(0, rays[:500])
t0 = time.perf_counter(); out, valid = rs.stock_intersect(0, rays); dt = time.perf_counter() - t0
line = {"scene": "bunny-class, %d triangles" % scene.n_triangles, "rays": n, "hit_fraction": float(valid.mean()),
        "stock_brute_force_accel": {"Mrays_per_s_one_thread": round(n / dt / 1e6, 4), "seconds": round(dt, 3),
                                    "ray_triangle_tests_per_s": round(n * scene.n_triangles / dt / 1e6, 1), "unit_tests": "M/s"}}
try:
    import torch
    from nori_b200.gpu import Engine
    e = Engine(0); e.load_scene(rs.flatten())            # the geometry exactly as the reference's OBJ loader holds it
    got, gvalid = e.intersect(rays)
    same = bool((gvalid == valid).all() and (got[valid].view(np.uint32) == out[valid].view(np.uint32)).all())
    big = np.tile(rays, (max(1, (1 << 22) // n), 1))
    dr = torch.from_numpy(big).cuda(); dh = torch.empty((len(big), 4), dtype=torch.float32, device="cuda")
    ms = [e.trace_device(dr.data_ptr(), len(big), dh.data_ptr()) for _ in range(5)][2:]
    line["gpu"] = {"Intersection_records_bit_equal_to_the_stock_accel": same, "Mrays_per_s_kernel_only": round(len(big) / np.mean(ms) / 1e3, 1), "rays": len(big)}
except Exception as ex:                                  # no device here: the CPU number stands alone
    line["gpu"] = {"skipped": str(ex)[:120]}
print(json.dumps(line))
